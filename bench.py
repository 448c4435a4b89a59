#!/usr/bin/env python
"""bench.py -- frames/s of the GF-ORB-SLAM tracking front end (ORB extract + frame-to-frame match + good-feature
select) on synthetic EuRoC-shaped 752x480 frames, 1000 keypoints (BASELINE.json config C2 + C4).

  python bench.py --gpus N --steps K --warmup W          our arm (under torchrun for N > 1; frames sharded, no collective)
  python bench.py --impl reference ...                   the CPU path (oracle port) on the host cores, same metric

One step = one batch of --batch frames per GPU through the whole front end.  Prints ONE JSON line (rank 0)."""
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
    process) run as a subprocess; returns its JSON dict or None."""
    import subprocess
    tool = os.path.join(os.path.dirname(os.path.abspath(__file__)), "tools", "cv2_stage_baseline.py")
    try:
        r = subprocess.run([sys.executable, tool, str(nframes), str(workers)], capture_output=True, text=True, timeout=240)
        d = json.loads(r.stdout.strip().splitlines()[-1])
        return d if "value" in d else None
    except Exception:
        return None


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
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD, "frames_per_step": nfr},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{nfr} consecutive frames per step x {args.steps} steps, oracle port (g++ -O2), {cores} threads"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


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
        # one process per GPU on a multi-socket host: run (and first-touch the pinned staging buffers) on the CPUs NVML
        # reports as local to this GPU, so that the H2D / D2H traffic of the end-to-end leg does not cross sockets
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
    from gf_orb_slam_b200.pipeline import LanedFrontEnd
    B = args.batch
    fe = LanedFrontEnd(W_, H_, NFEAT, B, lanes=args.lanes, n_good=NGOOD, device=local, overlap=not args.no_overlap)
    # frames sharded across ranks: rank r takes a contiguous chunk of the 2000-frame schedule (1-frame halo inside)
    seq = synth.Sequence(W_, H_)
    chunk = 2000 // max(world_size, 1)
    t0f = rank * chunk
    ndist = min(args.distinct, B)
    frames = seq.frames(t0f, ndist)
    reps = (B + ndist - 1) // ndist
    host = np.concatenate([frames] * reps)[:B]
    fe.load_host_frames(host)
    torch.cuda.synchronize()

    def barrier():
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(steps):
            fn()
        fe.sync()            # the main stream joins the selection and copy streams before the closing event
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world_size > 1:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    for _ in range(max(args.warmup, 3)):
        fe.step_device()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    fe.set_extract_profiling(True)
    fe.profile = True
    ms = timed(fe.step_device, args.steps)
    fe.profile = False
    stages = fe.extract_stage_times()
    pipe_stages = fe.stage_times()
    launches = fe.launches_per_step * args.steps * world_size        # whole job, like `value`
    # the same stage events with the selection on the main stream (no co-running kernel): per-kernel durations that are
    # not inflated by the selection kernel of the previous step sharing the SMs
    ser = LanedFrontEnd(W_, H_, NFEAT, B, lanes=1, n_good=NGOOD, device=local, overlap=False)
    ser.load_host_frames(host)
    for _ in range(2):
        ser.step_device()
    ser.sync(); torch.cuda.synchronize()
    ser.set_extract_profiling(True)
    ser.profile = True
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        ser.step_device()
    ser.sync(); e1.record(); torch.cuda.synchronize()
    ms_serial = e0.elapsed_time(e1) / args.steps
    ser.profile = False
    stages_serial = ser.extract_stage_times()
    pipe_serial = ser.stage_times()
    del ser
    fe.set_extract_profiling(False)
    # end to end: host (pinned) frames in, every result out, copies inside the timed region
    for _ in range(2):
        fe.step_host()
    ms_e2e = timed(fe.step_host, args.steps)
    clocks = sampler.stop() if rank == 0 else None
    # the host link under the end-to-end number: one pinned host -> device copy of a step's input alone on the GPU
    h2d_alone = None
    if rank == 0:
        hsrc = torch.empty(fe.h2d_bytes(), dtype=torch.uint8).pin_memory()
        hdst = torch.empty(fe.h2d_bytes(), dtype=torch.uint8, device="cuda")
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        hdst.copy_(hsrc, non_blocking=True)
        torch.cuda.synchronize()
        c0.record()
        for _ in range(5):
            hdst.copy_(hsrc, non_blocking=True)
        c1.record()
        torch.cuda.synchronize()
        h2d_alone = fe.h2d_bytes() * 5 / (c0.elapsed_time(c1) * 1e-3) / 1e9
        del hsrc, hdst
    total_frames = B * args.steps * world_size
    value = total_frames / (ms * 1e-3)
    e2e = total_frames / (ms_e2e * 1e-3)
    if rank != 0:
        if world_size > 1:
            dist.destroy_process_group()
        return
    peak, peak_src = peaks()
    # dominant kernel of the step (DESIGN.md "Kernels"): algorithmic bytes per launch / its average launch time
    alg = fe.algorithmic_bytes()
    in_step = dict(stages)
    dom = max(stages_serial, key=lambda s: s[1]) if stages_serial else ("none", 0.0)     # dominant kernel by its own duration
    dom_ms = dom[1]        # its duration on one stream with nothing co-running (same run, CUDA events); stage_ms holds the
                           # in-step figures, which include time shared with the other lane's and the selection's kernels
    dom_bytes = alg["per_stage"].get(dom[0], 0) * B
    ach = dom_bytes / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else 0.0
    prof = ncu_facts()
    kfacts = prof.get("stages", {}).get(dom[0], {})
    traffic = kfacts.get("dram_bytes_per_frame")
    roof = {"bound": "hbm", "kernel": dom[0], "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
            "traffic": traffic * B if traffic is not None else None, "peak_source": peak_src,
            "algorithmic_bytes_per_launch": dom_bytes,
            "note": "the dominant kernel is bound by the integer ALU pipe, not HBM (ncu: %s); frac is its algorithmic bytes "
                    "over the measured copy bandwidth" % kfacts.get("ncu", "see profiles/"),
            "ncu_source": prof.get("source"),
            "stage_ms": {n: round(t, 4) for n, t in stages}, "pipeline_ms": {n: round(t, 4) for n, t in pipe_stages},
            "stage_ms_serial": {n: round(t, 4) for n, t in stages_serial},
            "pipeline_ms_serial": {n: round(t, 4) for n, t in pipe_serial},
            "extract_frame_bytes": alg["extract_frame"],
            "extract_frac_of_hbm": (alg["extract_frame"] * B / (sum(t for _, t in stages_serial) * 1e-3) / 1e9 / peak) if stages_serial else None}
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world_size, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": {"workload": WORKLOAD, "frames_per_step_per_gpu": B, "distinct_frames": ndist,
                   "l2": "inputs larger than L2: per step each GPU touches %d MB of frames + %d MB of pyramid planes" % (
                       B * W_ * H_ >> 20, fe.pyramid_bytes() >> 20),
                   "sharding": "contiguous frame chunks per GPU, 1-frame halo, no collective",
                   "lanes_per_gpu": args.lanes, "ms_per_step_one_stream_no_overlap": round(ms_serial, 4)},
        "roofline": roof,
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": fe.h2d_bytes(), "d2h_bytes_per_step": fe.d2h_bytes(),
                "ms_per_step": ms_e2e / args.steps,
                "h2d_GBps_in_step": fe.h2d_bytes() / (ms_e2e / args.steps * 1e-3) / 1e9, "h2d_GBps_copy_alone": h2d_alone,
                "note": "h2d_GBps_in_step against h2d_GBps_copy_alone shows how much of the host link the step uses: the "
                        "copies run on their own stream under the kernels, so the end-to-end rate is set by the kernels"},
        "gpu_launches": launches,
        "clocks": clocks,
    }
    if world_size == 1 and not args.no_cpu:
        cores = host_cores()
        nfr = args.cpu_frames
        cf = seq.frames(0, nfr)
        v, dt = cpu_leg(cf, cores, consts())
        out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                               "sample": f"{nfr} consecutive frames of the same sequence, oracle port (g++ -O2), {cores} threads, {dt:.1f} s wall"}
        cvl = cv2_stage_leg(nfr, cores)
        if cvl:
            out["cpu_baseline"]["opencv_stages"] = {
                "value": cvl["value"], "unit": UNIT, "cores": cores,
                "what": "only the OpenCV-backed stages of the extractor (copyMakeBorder, resize chain, per-cell FAST 20/7, "
                        "GaussianBlur) through cv2 %s, one frame per worker process, %d frames; the reference's own scalar stages "
                        "(orientation, descriptors, matching, selection) are not in it" % (cvl["cv2"], cvl["frames"])}
    print(json.dumps(out))
    if world_size > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=256, help="frames per step per GPU")
    ap.add_argument("--distinct", type=int, default=256, help="distinct synthetic frames generated per GPU (tiled to --batch)")
    ap.add_argument("--cpu-frames", type=int, default=128, help="frames of the cpu_baseline sample")
    ap.add_argument("--ref-frames", type=int, default=0, help="frames per step of the --impl reference arm (0: 8 per host thread)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-overlap", action="store_true", help="run the good-feature selection on the main stream")
    ap.add_argument("--lanes", type=int, default=1, help="independent half-batch pipelines (CUDA streams) per GPU")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
