#!/usr/bin/env python
"""bench.py -- frames/s of the GF-ORB-SLAM tracking front end (ORB extract + frame-to-frame match + good-feature
select) on synthetic EuRoC-shaped 752x480 frames, 1000 keypoints (BASELINE.json config C2 + C4).

  python bench.py --gpus N --steps K --warmup W          our arm (under torchrun for N > 1; frames sharded, no collective)
  python bench.py --impl reference ...                   the CPU path (oracle port) on the host cores, same metric

One step = the 2000-frame sequence of config C2 through the whole front end, sharded over the GPUs in contiguous chunks
(--batch frames per launch).  Prints ONE JSON line (rank 0)."""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np

METRIC = "frames/sec ORB extract+match+GF-select (752x480, 1000 kp)"
UNIT = "frames/s"
W_, H_, NFEAT, NGOOD = 752, 480, 1000, 100
WORKLOAD = ("C2+C4: EuRoC-shaped 752x480 synthetic sequence (2000-frame schedule), ORBextractor(1000,1.2,8,FAST 20/7) + "
            "SearchForInitialization(prev,cur,window 100,nnratio 0.9) + info blocks + lazier-greedy pick 100 of <=1000 (eps 0.1)")


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_facts():
    """per-stage facts taken from the committed ncu --set full capture (profiles/ncu_facts.json, written by
    tools/ncu_facts.py): DRAM bytes per frame of each stage's kernels and their busiest pipe"""
    p = os.path.join(ROOT, "profiles", "ncu_facts.json")
    try:
        return json.load(open(p))
    except Exception:
        return {}


class ClockSampler:
    """SM clock and clock-event (throttle) reasons sampled while the timed region runs: NVML every 10 ms (a timed region
    is only tens of milliseconds long), `nvidia-smi -lms 200` when pynvml is missing."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    BITS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []
        self.nv, self.samples, self.stop_flag, self.mx = None, [], False, None

    def _poll(self):
        nv, h = self.nv
        while not self.stop_flag:
            try:
                clk = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                try:
                    rs = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.samples.append((clk, rs))
            except Exception:
                pass
            time.sleep(0.01)

    def start(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            # NVML enumerates physical devices: honour CUDA_VISIBLE_DEVICES when it lists indices
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            ids = [int(x) for x in vis.split(",") if x.strip().isdigit()]
            phys = ids[self.index] if self.index < len(ids) else self.index
            h = nv.nvmlDeviceGetHandleByIndex(phys)
            self.mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            self.nv = (nv, h)
            self.th = threading.Thread(target=self._poll, daemon=True)
            self.th.start()
            return
        except Exception:
            self.nv = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if self.nv is not None:
            self.stop_flag = True
            self.th.join(timeout=1)
            sm = [c for c, _ in self.samples]
            reasons = set()
            for _, rs in self.samples:
                for bit, nm in self.BITS.items():
                    if rs & bit:
                        reasons.add(nm)
            return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": float(self.mx) if self.mx else None,
                    "reasons": sorted(reasons), "samples": len(sm), "source": "nvml, 10 ms"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvidia-smi -lms 200"}


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_leg(frames, threads, fe_consts):
    """time the oracle port of the same chain on `threads` host threads; returns (frames/s, seconds)"""
    import oracle
    world, score, xv, sigma2, k, scale = fe_consts
    t0 = time.perf_counter()
    oracle.front_end_batch(frames, threads, world, score, xv, sigma2, k, scale)
    dt = time.perf_counter() - t0
    return frames.shape[0] / dt, dt


def cv2_stage_leg(nframes, workers):
    """tools/cv2_stage_baseline.py (the extractor's OpenCV calls through cv2, one frame per worker
    process) run as a subprocess; returns its JSON dict, or {"unavailable": reason} -- never silently nothing."""
    tool = os.path.join(os.path.dirname(os.path.abspath(__file__)), "tools", "cv2_stage_baseline.py")
    try:
        r = subprocess.run([sys.executable, tool, str(nframes), str(workers)], capture_output=True, text=True, timeout=240)
    except subprocess.TimeoutExpired:
        return {"unavailable": "tools/cv2_stage_baseline.py timed out after 240 s"}
    except Exception as e:                                   # noqa: BLE001
        return {"unavailable": "could not start tools/cv2_stage_baseline.py: %r" % (e,)}
    lines = [ln for ln in r.stdout.strip().splitlines() if ln.startswith("{")]
    if r.returncode != 0 or not lines:
        return {"unavailable": "tools/cv2_stage_baseline.py rc=%d: %s" % (r.returncode, (r.stderr or r.stdout).strip()[-300:])}
    try:
        d = json.loads(lines[-1])
    except ValueError as e:
        return {"unavailable": "unparsable output of tools/cv2_stage_baseline.py: %r" % (e,)}
    return d if "value" in d else {"unavailable": "no 'value' in the tool's output: %s" % lines[-1][:200]}


def consts():
    from gf_orb_slam_b200 import synth
    from gf_orb_slam_b200.pipeline_consts import SIMU_XV, level_sigma2, sample_scale_for, simu_visible_from
    world, _, score, _ = synth.landmark_cloud(NFEAT, 7, visible_from=simu_visible_from(W_, H_))
    return world, score, SIMU_XV, level_sigma2(), NGOOD, sample_scale_for(NFEAT, NGOOD, 0.1)


def run_reference(args):
    """--impl reference: the CPU path (oracle port; the reference itself cannot be compiled here) on all
    host cores.  Each step = a bounded sample of args.ref_frames consecutive frames of the same workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from gf_orb_slam_b200 import synth
    cores = host_cores()
    seq = synth.Sequence(W_, H_)
    nfr = args.ref_frames if args.ref_frames > 0 else 8 * cores     # several frames per thread: no idle tail
    frames = seq.frames(0, nfr)
    c = consts()
    for _ in range(args.warmup):
        cpu_leg(frames[:max(cores, 8)], cores, c)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_leg(frames, cores, c)
    dt = time.perf_counter() - t0
    val = nfr * args.steps / dt
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD, "frames_per_step": args.seq_frames, "sample_frames_per_step": nfr},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{nfr} consecutive frames per step x {args.steps} steps, oracle port (g++ -O3 -march=native), {cores} threads"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# SURVEY 8(d) integer-op model per frame (P = sum of level pixels, P0 = level 0, N = keypoints): FAST 16 loads + ~50 int ops
# per pixel = 55 P; blur 14 IMAD per pixel; bilinear resize ~10 ops per output pixel; IC_Angle 2 x (mul + add) over 749
# pixels; steered BRIEF 512 taps x ~6 ops.  Denominator: the measured integer ALU-pipe rate (tools/ubench/pipes.cu on
# B200, profiles/r2_int_pipe_peak.jsonl): 62.5 lane-ops/clk/SM (SHF; IMAD 61.5, IDP 62.0, VIMNMX3 53.7) x 148 SMs x 1.965 GHz.
INT_PIPE_PEAK_TOPS = 18.16
INT_PIPE_PEAK_SRC = "measured: profiles/r2_int_pipe_peak.jsonl (shf 18.16 Top/s = 62.5 lane-ops/clk/SM; imad 17.90, dp4a 18.03, vimnmx3 15.6, popc 4.53)"


def int_ops_per_frame(P, P0, N):
    return {"pyramid": 10 * (P - P0), "detect": 55 * P, "blur": 14 * P, "orient": 4 * 749 * N, "describe": 6 * 512 * N, "select": 0}


def run_ours(args):
    import torch
    import torch.distributed as dist
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the B200 path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world_size > 1 and os.environ.get("GFORB_NO_CPU_AFFINITY") is None:
        try:
            import pynvml as nv
            nv.nvmlInit()
            vis = [int(x) for x in os.environ.get("CUDA_VISIBLE_DEVICES", "").split(",") if x.strip().isdigit()]
            nv.nvmlDeviceSetCpuAffinity(nv.nvmlDeviceGetHandleByIndex(vis[local] if local < len(vis) else local))
        except Exception:
            pass
    if world_size > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from gf_orb_slam_b200 import synth
    from gf_orb_slam_b200.pipeline import TrackingFrontEnd
    # config C2 as written: ONE step = the 2000-frame sequence, sharded over the ranks in contiguous chunks; a rank other
    # than 0 also re-extracts the last frame of the previous chunk (1-frame halo) so that every counted frame
    # is matched against its true predecessor without any transfer between GPUs
    SEQ = args.seq_frames
    c0, c1 = SEQ * rank // world_size, SEQ * (rank + 1) // world_size
    mine = c1 - c0                                                    # frames this rank is credited with per step
    lo = max(c0 - 1, 0)
    nloc = c1 - lo                                                    # frames it processes per step
    nb = max(1, -(-mine // args.batch))                               # batches per step
    B = -(-nloc // nb)                                                # frames per batch (the last one may be shorter)
    seq = synth.Sequence(W_, H_)
    host = torch.empty((nloc, H_, W_), dtype=torch.uint8).pin_memory()
    hv = host.numpy()
    for i in range(nloc):
        hv[i] = seq.frame(lo + i)
    dev_frames = host.cuda(non_blocking=False)
    fe = TrackingFrontEnd(W_, H_, NFEAT, B, n_good=NGOOD, device=local, overlap=not args.no_overlap)
    batches = [(b0, min(b0 + B, nloc)) for b0 in range(0, nloc, B)]
    L = _lib_handles(fe)

    def barrier():
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_resident():
        for b0, b1 in batches:
            fe.step_device(dev_frames[b0:b1], b1 - b0)

    NSETS = 4
    outs = _staging_sets(fe, NSETS)
    issued = [0]

    def step_e2e():
        # frames of a step are in pinned host memory; results land in four rotating sets of pinned buffers, a set is
        # handed back to the library only after the step that last used it has delivered (gforb_frontend_wait_results)
        for b0, b1 in batches:
            if issued[0] >= NSETS:
                L.check(L.lib.gforb_frontend_wait_results(fe._h, NSETS - 1), "gforb_frontend_wait_results")
            L.check(L.lib.gforb_frontend_step_host(fe._h, L.vp(host[b0:b1].data_ptr()), b1 - b0, L.byref(outs[issued[0] % NSETS][0])),
                    "gforb_frontend_step_host")
            issued[0] += 1

    def timed(fn, steps):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(steps):
            fn()
        fe.sync()            # joins every stream of the front end and raises if any kernel recorded an error
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world_size > 1:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    W = max(args.warmup, 3)
    for _ in range(W):
        step_resident()
    fe.sync()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    fe.set_profiling(True)
    ms = timed(step_resident, args.steps)
    stages, pipe_stages = fe.extract_stage_times(), fe.stage_times()
    fe.set_profiling(False)
    launches = fe.launches_per_step * len(batches) * args.steps * world_size        # whole job, like `value`
    for _ in range(2):
        step_e2e()
    fe.sync()
    ms_e2e = timed(step_e2e, args.steps)
    clocks = sampler.stop() if rank == 0 else None
    # where the end-to-end step spends its time: CUDA events around the upload, the bulk download and the three compute stages of
    # a few more (untimed) steps, every rank at once
    barrier()
    fe.set_profiling(True)
    for _ in range(3):
        step_e2e()
    fe.sync()
    e2e_stages = {n: round(t, 4) for n, t in fe.stage_times()}
    fe.set_profiling(False)
    barrier()
    # weak-scaling companion (secondary): every rank walks a whole SEQ-frame schedule of its own (its chunk repeated)
    ms_weak = None
    if world_size > 1 and not args.no_weak:
        reps = max(1, SEQ // max(mine, 1))

        def step_weak():
            for _ in range(reps):
                step_resident()
        step_weak(); fe.sync()
        ms_weak = timed(step_weak, max(2, args.steps // 4)) / max(2, args.steps // 4)
    # the same kernels on ONE stream with nothing co-running: per-stage durations for the roofline (same run, CUDA events)
    ser = TrackingFrontEnd(W_, H_, NFEAT, B, n_good=NGOOD, device=local, overlap=False)
    b0, b1 = batches[0]
    for _ in range(2):
        ser.step_device(dev_frames[b0:b1], b1 - b0)
    ser.sync()
    ser.set_profiling(True)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    nser = max(4, min(16, args.steps))
    e0.record()
    for _ in range(nser):
        ser.step_device(dev_frames[b0:b1], b1 - b0)
    ser.sync(); e1.record(); torch.cuda.synchronize()
    ms_serial_batch = e0.elapsed_time(e1) / nser
    stages_serial, pipe_serial = ser.extract_stage_times(), ser.stage_times()
    ser.close()
    del ser
    # host link figures under the end-to-end number: one rank's upload alone, and every rank uploading at once (no kernels)
    h2d_alone, h2d_all = _h2d_rates(torch, dist, host, dev_frames, world_size)
    total_frames = mine * args.steps
    if world_size > 1:
        t = torch.tensor([float(total_frames)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t)
        total_frames = float(t.item())
    value = total_frames / (ms * 1e-3)
    e2e = total_frames / (ms_e2e * 1e-3)
    c5 = None
    if not args.no_extra:
        fe.close()
        del dev_frames
        torch.cuda.empty_cache()
        c5 = _guard(lambda: c5_leg(torch, dist, world_size, rank, local))
    if rank != 0:
        if world_size > 1:
            dist.destroy_process_group()
        return
    peak, peak_src = peaks()
    alg = fe.algorithmic_bytes()
    P = alg["pixels"]
    P0 = W_ * H_
    ops = int_ops_per_frame(P, P0, NFEAT)
    nB = batches[0][1] - batches[0][0]                                 # frames per launch of the serial pass
    dom = max(stages_serial, key=lambda s: s[1]) if stages_serial else ("none", 0.0)     # dominant kernel by its own duration
    dom_ms = dom[1]
    dom_ops = ops.get(dom[0], 0) * nB
    dom_bytes = alg["per_stage"].get(dom[0], 0) * nB
    ach_ops = dom_ops / (dom_ms * 1e-3) / 1e12 if dom_ms > 0 else 0.0
    ach_gbs = dom_bytes / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else 0.0
    prof = ncu_facts()
    kfacts = prof.get("stages", {}).get(dom[0], {})
    traffic = kfacts.get("dram_bytes_per_frame")
    ext_ms = sum(t for _, t in stages_serial)
    per_stage = {}
    for n, t in stages_serial:
        if t > 0:
            per_stage[n] = {"ms": round(t, 4), "int_frac": round(ops.get(n, 0) * nB / (t * 1e-3) / 1e12 / INT_PIPE_PEAK_TOPS, 4),
                            "hbm_frac": round(alg["per_stage"].get(n, 0) * nB / (t * 1e-3) / 1e9 / peak, 4)}
    step_ops = sum(ops.values())
    roof = {"bound": "int_pipe", "kernel": dom[0], "achieved": ach_ops, "peak": INT_PIPE_PEAK_TOPS, "unit": "Tops/s",
            "frac": ach_ops / INT_PIPE_PEAK_TOPS, "peak_source": INT_PIPE_PEAK_SRC,
            "traffic": traffic * nB if traffic is not None else None,
            "algorithmic_int_ops_per_launch": dom_ops, "frames_per_launch": nB,
            "op_model": "SURVEY 8(d): FAST 55 P, blur 14 P, resize 10 (P - P0), IC_Angle 4 x 749 N, BRIEF 6 x 512 N int ops per frame "
                        "(P = %d level pixels, N = %d); ncu of the dominant kernel: %s" % (P, NFEAT, kfacts.get("ncu", "see profiles/")),
            "hbm": {"achieved": ach_gbs, "peak": peak, "unit": "GB/s", "frac": ach_gbs / peak, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": dom_bytes},
            "per_stage_serial": per_stage,
            "extract": {"ms_per_launch": round(ext_ms, 4),
                        "int_frac": (step_ops * nB / (ext_ms * 1e-3) / 1e12 / INT_PIPE_PEAK_TOPS) if ext_ms > 0 else None,
                        "hbm_frac": (alg["extract_frame"] * nB / (ext_ms * 1e-3) / 1e9 / peak) if ext_ms > 0 else None},
            "whole_step": {"int_frac": step_ops * total_frames / (ms * 1e-3) / 1e12 / (INT_PIPE_PEAK_TOPS * world_size),
                           "hbm_frac": alg["extract_frame"] * total_frames / (ms * 1e-3) / 1e9 / (peak * world_size)},
            "ncu_source": prof.get("source"),
            "stage_ms": {n: round(t, 4) for n, t in stages}, "pipeline_ms": {n: round(t, 4) for n, t in pipe_stages},
            "stage_ms_serial": {n: round(t, 4) for n, t in stages_serial},
            "pipeline_ms_serial": {n: round(t, 4) for n, t in pipe_serial}}
    h2d_step = host.numel() if world_size == 1 else None
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world_size, "steps": args.steps, "warmup": W,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": {"workload": WORKLOAD, "frames_per_step": SEQ, "frames_per_step_rank0": mine, "batches_per_step_rank0": len(batches),
                   "frames_per_batch": B,
                   "l2": "inputs larger than L2: per step each GPU reads %d MB of frames and touches %d MB of pyramid planes per batch" % (
                       nloc * W_ * H_ >> 20, fe.pyramid_bytes() >> 20),
                   "sharding": "the %d-frame sequence in contiguous chunks per GPU, 1-frame halo re-extracted, no collective" % SEQ,
                   "dependency": "information blocks + selection of a batch start after its matcher (frontend.cu)",
                   "host_code": "C++ (csrc/frontend.cu behind gforb_frontend_*); bench.py only issues the steps",
                   "ms_per_batch_one_stream_no_overlap": round(ms_serial_batch, 4)},
        "roofline": roof,
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": nloc * W_ * H_, "d2h_bytes_per_step": fe.d2h_bytes() // fe.B * nloc,
                "ms_per_step": ms_e2e / args.steps,
                "path": "gforb_frontend_step_host: plain host pointers (pinned), cudaMemcpyAsync up / down inside the library",
                "h2d_GBps_in_step": nloc * W_ * H_ / (ms_e2e / args.steps * 1e-3) / 1e9, "h2d_GBps_copy_alone": h2d_alone,
                "h2d_GBps_per_rank_all_ranks_copying": h2d_all, "stage_ms_per_batch_rank0": e2e_stages},
        "gpu_launches": launches,
        "clocks": clocks,
    }
    if ms_weak is not None:
        out["weak_scaling"] = {"frames_per_step_per_gpu": SEQ // max(mine, 1) * mine, "ms_per_step": ms_weak,
                               "value": SEQ // max(mine, 1) * mine * world_size / (ms_weak * 1e-3), "unit": UNIT}
    if world_size == 1 and not args.no_cpu:
        cores = host_cores()
        nfr = args.cpu_frames
        cf = seq.frames(0, nfr)
        v, dt = cpu_leg(cf, cores, consts())
        out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                               "sample": f"{nfr} consecutive frames of the same sequence, oracle port (g++ -O3 -march=native -ffp-contract=off), {cores} threads, {dt:.1f} s wall"}
        cvl = cv2_stage_leg(nfr, cores)
        if "value" in cvl:
            out["cpu_baseline"]["opencv_stages"] = {
                "value": cvl["value"], "unit": UNIT, "cores": cores,
                "what": "only the OpenCV-backed stages of the extractor (copyMakeBorder, resize chain, per-cell FAST 20/7, "
                        "GaussianBlur) through cv2 %s, one frame per worker process, %d frames; the reference's own scalar stages "
                        "(orientation, descriptors, matching, selection) are not in it" % (cvl["cv2"], cvl["frames"])}
        else:
            out["cpu_baseline"]["opencv_stages"] = cvl
    if not args.no_extra:
        out["extra"] = extras(torch, local, world_size, rank)
        out["extra"]["C5"] = c5
    print(json.dumps(out))
    if world_size > 1:
        dist.destroy_process_group()


def _guard(fn):
    """an extra leg either reports numbers or says why not -- never silently absent"""
    try:
        return fn()
    except Exception as e:                                    # noqa: BLE001
        return {"unavailable": "%s: %s" % (type(e).__name__, str(e)[:300])}


def c5_leg(torch, dist, world_size, rank, local):
    """config 5 (all ranks): 2000 queries per rank all-gathered, all-pairs Hamming top-2 against a 2^20-row database sharded by
    rows over the ranks, NCCL all-gather of the 16-byte keys, merge.  One GPU: tensor-core kernel and integer-pipe kernel."""
    from gf_orb_slam_b200.matcher import ORBmatcher
    from gf_orb_slam_b200.sharded import ShardedDescriptorDB, db_shard
    NQ, NDB = 2000, 1 << 20
    g = torch.Generator().manual_seed(42)
    full = torch.randint(0, 256, (NDB, 32), dtype=torch.uint8, generator=g)        # identical on every rank
    q = torch.randint(0, 256, (NQ, 32), dtype=torch.uint8, generator=torch.Generator().manual_seed(100 + rank)).cuda()
    row0, n = db_shard(NDB, world_size, rank)
    shard = full[row0:row0 + n].cuda()
    del full
    m = ORBmatcher(0.9, True, max_batch=1, max_keypoints=NQ, max_queries=NQ, device=local)
    sdb = ShardedDescriptorDB(shard, row0, matcher=m)
    res = {"queries_per_step": NQ * world_size, "db_rows": NDB, "shard_rows": n}
    kernels = ("tc", "popc") if world_size == 1 else ("tc",)
    keys_by = {}
    for kern in kernels:
        os.environ["GFORB_ALLPAIRS"] = kern

        def step():
            return sdb.top2(sdb.gather_queries(q))
        for _ in range(2):
            keys = step()
        torch.cuda.synchronize()
        if world_size > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        iters = 5
        e0.record()
        for _ in range(iters):
            keys = step()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64, device="cuda")
        if world_size > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        msq = float(t.item())
        keys_by[kern] = keys
        res[kern] = {"ms_per_step": msq, "pairs_per_s": NQ * world_size * NDB / (msq * 1e-3), "frames_per_s": world_size / (msq * 1e-3)}
    os.environ.pop("GFORB_ALLPAIRS", None)
    if len(keys_by) == 2:
        res["keys_identical_tc_vs_popc"] = bool(torch.equal(keys_by["tc"], keys_by["popc"]))
    m.close()
    return res


def extras(torch, local, world_size, rank):
    """driver-run numbers for the other BASELINE.json configs (C1, C3; C5 is measured by c5_leg on all ranks)"""
    import time as _t
    from gf_orb_slam_b200 import synth
    out = {}

    def c1():
        # config 1: ONE 640x480 frame through gforb_extract with HOST pointers (flags = 0: the call a C++ Tracking.cc makes)
        from gf_orb_slam_b200 import ORBextractor
        import oracle
        img = synth.textured_frame(640, 480, seed=3)
        ex = ORBextractor(1000, 1.2, 8, max_width=640, max_height=480, max_batch=1, device=local)
        for _ in range(5):
            kps, desc, counts = ex.extract_batch(img[None])
        n = 50
        t0 = _t.perf_counter()
        for _ in range(n):
            kps, desc, counts = ex.extract_batch(img[None])
        us = (_t.perf_counter() - t0) / n * 1e6
        oex = oracle.ORBextractor(1000, 1.2, 8)
        oex(img)
        t0 = _t.perf_counter()
        for _ in range(5):
            ok, od = oex(img)
        oms = (_t.perf_counter() - t0) / 5 * 1e3
        same = int(counts[0]) == len(ok) and bool((desc[0, :len(ok)] == od).all())
        return {"what": "ORBextractor::operator() on one 640x480 frame, batch 1, host pointers in and out (H2D + 22 launches + D2H + sync)",
                "latency_us": us, "frames_per_s": 1e6 / us, "keypoints": int(counts[0]), "oracle_ms_per_frame_1_thread": oms,
                "bit_exact_vs_oracle": same}

    def c3():
        # config 3: 256 batched SearchByProjection(F, MPs) calls, 1500 keypoints x 10 000 projected map points each
        from gf_orb_slam_b200 import Frames, MapPoints, ORBextractor, ORBmatcher
        Bc, M = 256, 10000
        img = synth.textured_frame(W_, H_, seed=1)
        ex = ORBextractor(1500, 1.2, 8, max_width=W_, max_height=H_, max_batch=1, device=local)
        kps, desc, counts = ex.extract_batch(img[None])
        n = int(counts[0])
        mp = [synth.map_points_c3(kps[0, :n], desc[0, :n], M, seed=50 + b, w=W_, h=H_) for b in range(8)]
        import numpy as _np
        stack = lambda k: torch.from_numpy(_np.ascontiguousarray(_np.stack([mp[b % 8][k] for b in range(Bc)]))).cuda()
        mps = MapPoints(stack("proj_x"), stack("proj_y"), stack("scale_level"), stack("view_cos"), stack("flags"), stack("descriptors"),
                        torch.full((Bc,), M, dtype=torch.int32).cuda())
        kview = torch.from_numpy(_np.repeat(kps.view(_np.float32).reshape(1, 1500, 7), Bc, 0)).cuda()
        F = Frames(kview, torch.from_numpy(_np.repeat(desc, Bc, 0)).cuda(), torch.from_numpy(_np.repeat(counts, Bc)).cuda(), W_, H_)
        m = ORBmatcher(0.8, True, max_batch=Bc, max_keypoints=1500, max_queries=M, cand_pool=64 * M, device=local)
        fmp = torch.full((Bc, 1500), -1, dtype=torch.int32).cuda(); fsc = torch.full((Bc, 1500), 999, dtype=torch.int32).cuda()
        nm = torch.zeros(Bc, dtype=torch.int32).cuda()
        st = torch.cuda.current_stream().cuda_stream
        for _ in range(2):
            fmp.fill_(-1); fsc.fill_(999)
            m.SearchByProjection(F, mps, 1.0, fmp, fsc, nmatches=nm, stream=st, sync=False)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        it = 10
        e0.record()
        for _ in range(it):
            fmp.fill_(-1); fsc.fill_(999)
            m.SearchByProjection(F, mps, 1.0, fmp, fsc, nmatches=nm, stream=st, sync=False)
        e1.record()
        torch.cuda.synchronize()
        msb = e0.elapsed_time(e1) / it
        return {"what": "256 batched ORBmatcher::SearchByProjection(F, vpMapPoints, th=1) calls, 1500 keypoints x 10 000 map points each, device-resident",
                "ms_per_256_calls": msb, "calls_per_s": Bc / (msb * 1e-3), "map_points_per_s": Bc * M / (msb * 1e-3),
                "matches_per_call": float(nm.float().mean().item()), "launches": m.last_launches()}

    def lc():
        # the loop-closing / mapping overloads: 256 batched ORBmatcher::Fuse matching loops, 1000 map points projected into
        # a 1000-keypoint keyframe each, device-resident; the oracle's Fuse on one thread beside it
        from gf_orb_slam_b200 import Frames, ORBextractor, ORBmatcher, WorldPoints
        import numpy as _np
        import oracle
        Bc, L = 256, 1000
        img = synth.textured_frame(W_, H_, seed=1)
        ex = ORBextractor(1000, 1.2, 8, max_width=W_, max_height=H_, max_batch=1, device=local)
        kps, desc, counts = ex.extract_batch(img[None])
        n = int(counts[0])
        fx = fy = 458.0; cx, cy = W_ / 2.0, H_ / 2.0
        u01 = synth.uniform01(77, 2 * L).reshape(L, 2)
        pick = _np.minimum((u01[:, 0] * n).astype(_np.int64), n - 1)
        z = (2.0 + 6.0 * u01[:, 1]).astype(_np.float32)
        world = _np.stack([(kps["x"][0, pick] + 1.0 - cx) / fx * z, (kps["y"][0, pick] - 1.0 - cy) / fy * z, z], 1).astype(_np.float32)
        d3 = _np.linalg.norm(world.astype(_np.float64), axis=1)
        scale = (_np.float32(1.2) ** _np.arange(8)).astype(_np.float32)
        mind = (d3 / (scale[_np.clip(kps["octave"][0, pick], 0, 7)] * 0.99)).astype(_np.float32); maxd = (mind * 5).astype(_np.float32)
        nrm = (world / d3[:, None]).astype(_np.float32)
        mpd = desc[0, pick].copy()
        rep = lambda a: torch.from_numpy(_np.ascontiguousarray(_np.repeat(a[None], Bc, 0))).cuda()
        pts = WorldPoints(rep(world), rep(nrm), rep(mind), rep(maxd), rep(mpd), torch.zeros((Bc, L), dtype=torch.uint8).cuda(),
                          torch.full((Bc,), L, dtype=torch.int32).cuda())
        kview = torch.from_numpy(_np.repeat(kps.view(_np.float32).reshape(1, 1000, 7), Bc, 0)).cuda()
        KF = Frames(kview, torch.from_numpy(_np.repeat(desc, Bc, 0)).cuda(), torch.from_numpy(_np.repeat(counts, Bc)).cuda(), W_, H_)
        T = torch.eye(4, dtype=torch.float32).reshape(1, 16).repeat(Bc, 1).cuda()
        m = ORBmatcher(0.6, True, max_batch=Bc, max_keypoints=1000, max_queries=L, device=local)
        bi = torch.empty((Bc, L), dtype=torch.int32).cuda(); bd = torch.empty((Bc, L), dtype=torch.int32).cuda()
        nf = torch.zeros(Bc, dtype=torch.int32).cuda()
        st = torch.cuda.current_stream().cuda_stream
        for _ in range(2):
            m.Fuse(KF, T, (fx, fy, cx, cy), pts, 3.0, bi, bd, nfused=nf, stream=st, sync=False)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        it = 10
        e0.record()
        for _ in range(it):
            m.Fuse(KF, T, (fx, fy, cx, cy), pts, 3.0, bi, bd, nfused=nf, stream=st, sync=False)
        e1.record()
        torch.cuda.synchronize()
        msb = e0.elapsed_time(e1) / it
        okf = oracle.Frame(kps[0, :n], desc[0, :n], W_, H_)
        okf.set_camera(fx, fy, cx, cy, _np.eye(4, dtype=_np.float32))
        slot = _np.full(n, -1, _np.int32); bad = _np.zeros(L, _np.uint8); ik = _np.full(L, -1, _np.int32)
        t0 = _t.perf_counter()
        for _ in range(3):
            en, ebi, ebd, _, _, _ = oracle.fuse(okf, None, _np.arange(L, dtype=_np.int32), world, nrm, mind, maxd, mpd, 3.0, slot, bad, ik)
        oms = (_t.perf_counter() - t0) / 3 * 1e3
        same = bool((bi[0].cpu().numpy() == ebi).all() and (bd[0].cpu().numpy() == ebd).all())
        return {"what": "256 batched matching loops of ORBmatcher::Fuse(pKF, vpMapPoints, th=3): 1000 map points into a 1000-keypoint keyframe each, device-resident",
                "ms_per_256_calls": msb, "calls_per_s": Bc / (msb * 1e-3), "fused_per_call": float(nf.float().mean().item()),
                "launches": m.last_launches(), "oracle_ms_per_call_1_thread": oms, "bit_exact_vs_oracle": same}

    out["C1"] = _guard(c1)
    out["C3"] = _guard(c3)
    out["LC_fuse"] = _guard(lc)
    return out


class _LibHandles:
    def __init__(self):
        import ctypes as C
        from gf_orb_slam_b200 import _lib
        self.lib, self.check, self.byref, self.vp = _lib.lib(), _lib.check, C.byref, C.c_void_p


def _lib_handles(fe):
    return _LibHandles()


def _staging_sets(fe, n):
    """n sets of pinned host result buffers (GforbFrontEndOutputs) for the end-to-end leg"""
    import torch
    from gf_orb_slam_b200._lib import FrontEndOutputs
    sets = []
    for _ in range(n):
        t = [torch.zeros((fe.B, fe.nf, 7), dtype=torch.float32).pin_memory(), torch.zeros((fe.B, fe.nf, 32), dtype=torch.uint8).pin_memory(),
             torch.zeros(fe.B, dtype=torch.int32).pin_memory(), torch.zeros((fe.B, fe.nf), dtype=torch.int32).pin_memory(),
             torch.zeros(fe.B, dtype=torch.int32).pin_memory(), torch.zeros((fe.B, fe.k), dtype=torch.int32).pin_memory(),
             torch.zeros((fe.B, fe.k), dtype=torch.float64).pin_memory(), torch.zeros(fe.B, dtype=torch.int32).pin_memory()]
        sets.append((FrontEndOutputs(*[x.data_ptr() for x in t]), t))
    return sets


def _h2d_rates(torch, dist, host, dev, world_size):
    """GB/s of one plain pinned host -> device copy of this rank's frames: rank 0 alone (the others idle), then every rank
    at the same time with no kernels running (the aggregate host-link ceiling of the box)"""
    rank = int(os.environ.get("RANK", "0"))
    nbytes = host.numel()

    def rate():
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dev.copy_(host, non_blocking=True)
        torch.cuda.synchronize()
        c0.record()
        for _ in range(3):
            dev.copy_(host, non_blocking=True)
        c1.record()
        torch.cuda.synchronize()
        return nbytes * 3 / (c0.elapsed_time(c1) * 1e-3) / 1e9
    alone = None
    if world_size > 1:
        dist.barrier()
    if rank == 0:
        alone = rate()
    if world_size == 1:
        return alone, None
    dist.barrier()
    r = rate()
    t = torch.tensor([r], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return alone, float(t.item())

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--seq-frames", type=int, default=2000, help="frames of the sequence one step walks through (config C2: 2000)")
    ap.add_argument("--batch", type=int, default=296, help="most frames per front-end batch (a step is ceil(chunk / batch) equal batches per GPU; 296 = 2 x 148 SMs, one wave of the selection kernel at two CTAs per SM: 7 batches of 286 frames at N = 1)")
    ap.add_argument("--cpu-frames", type=int, default=256, help="frames of the cpu_baseline sample")
    ap.add_argument("--ref-frames", type=int, default=0, help="frames per step of the --impl reference arm (0: 8 per host thread)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the C1 / C3 / C5 legs")
    ap.add_argument("--no-weak", action="store_true", help="skip the weak-scaling companion leg (N > 1)")
    ap.add_argument("--no-overlap", action="store_true", help="run matcher and selection on the main stream")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
