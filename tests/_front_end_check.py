"""Checker side of the front-end parity tests (TEST INFRASTRUCTURE: drives the CPU oracle; nothing under
gf_orb_slam_b200/ imports this).  cpu_front_end = the per-frame chain of gf_orb_slam_b200.pipeline on the oracle;
oracle_front_end_batch = the same for many frames with the extraction in C++ threads; smoke_check = one small batch
through the CUDA front end compared bit for bit (used by __graft_entry__.smoke())."""
from __future__ import annotations

from concurrent.futures import ThreadPoolExecutor

import numpy as np

from gf_orb_slam_b200 import synth
from gf_orb_slam_b200.pipeline_consts import EUROC_CAM


def cpu_front_end(oracle, frames: np.ndarray, prev_kd, world, score, xv, sigma2, k, sample_scale, window=100, nnratio=0.9,
                  cam=None):
    """The same per-frame chain on the CPU oracle (checker / cpu_baseline only).  prev_kd = (kps, desc) of the frame
    before frames[0] or None.  Returns per frame (kps, desc, matches12 or None, nmatches, sel_idx, sel_logdet)."""
    h, w = frames.shape[1:]
    ex = oracle.ORBextractor(1000 if world is None else world.shape[0], 1.2, 8)
    cam = dict(EUROC_CAM) if cam is None else cam
    ob = oracle.Observability(**cam)
    out = []
    for t in range(frames.shape[0]):
        kps, desc = ex(frames[t])
        m12, nm = None, 0
        if prev_kd is not None:
            f1 = oracle.Frame(prev_kd[0], prev_kd[1], w, h); f2 = oracle.Frame(kps, desc, w, h)
            pm = np.stack([prev_kd[0]["x"], prev_kd[0]["y"]], -1).astype(np.float32)
            nm, m12, _ = oracle.search_for_initialization(f1, f2, pm, window, nnratio, True)
        n = len(kps)
        ob.set_xv(xv); ob.predict(0.1, 1)
        blocks, obs, _ = ob.build_frame(world[:n], kps["octave"], score[:n], np.ones(n, np.uint8), sigma2)
        si, sl = oracle.select(blocks, k, 2, sample_scale, 1, scores=obs) if n > 0 else (np.zeros(0, np.int32), np.zeros(0))
        out.append((kps, desc, m12, nm, si, sl))
        prev_kd = (kps, desc)
    return out


def smoke_check(oracle, device=0, batch=3):
    """One small batch through the CUDA front end, checked bit-exactly against the CPU oracle."""
    seq = synth.Sequence()
    frames = seq.frames(0, 2 * batch)
    import torch
    from gf_orb_slam_b200.pipeline import TrackingFrontEnd
    fe = TrackingFrontEnd(batch=batch, device=device)
    res = []
    for s in range(2):
        fe.load_host_frames(frames[s * batch:(s + 1) * batch])
        fe.step_device()
        fe.sync()
        res.append([t.cpu().numpy().copy() for t in (fe.kps, fe.desc, fe.counts, fe.m12, fe.nmatch, fe.sel_idx, fe.sel_ld, fe.sel_n)])
    exp = cpu_front_end(oracle, frames, None, fe.world_h, fe.score_h, fe.xv_h, fe.sigma2_h, fe.k, fe.sample_scale)
    from gf_orb_slam_b200._lib import KP_DTYPE
    nmatched = 0
    for s in range(2):
        kps, desc, counts, m12, nmatch, sidx, sld, sn = res[s]
        for b in range(batch):
            t = s * batch + b
            ek, ed, em12, enm, esi, esl = exp[t]
            n = counts[b]
            assert n == len(ek), ("keypoint count", t, n, len(ek))
            gk = kps[b, :n].copy().view(KP_DTYPE).reshape(-1)
            assert gk.tobytes() == ek.tobytes(), ("keypoints", t)
            assert np.array_equal(desc[b, :n], ed), ("descriptors", t)
            if t > 0:
                assert nmatch[b] == enm and np.array_equal(m12[b, :len(em12)], em12), ("matches", t)
                nmatched += int(enm)
            assert sn[b] == len(esi) and np.array_equal(sidx[b, :sn[b]], esi), ("good-feature subset", t)
            assert np.array_equal(sld[b, :sn[b]], esl), ("log-det", t)
    return dict(frames=2 * batch, launches_per_step=fe.launches_per_step, matches=nmatched, selected=int(res[1][7].sum()))


def oracle_front_end_batch(oracle, frames: np.ndarray, world, score, xv, sigma2, k, sample_scale, window=100, nnratio=0.9,
                           threads=8, cam=None):
    """cpu_front_end for a whole batch: extraction through the oracle's multi-threaded run_batch, then matching and
    selection per frame on a thread pool (ctypes releases the GIL).  Frame 0 has no predecessor (m12 None)."""
    nf, h, w = frames.shape
    ex = oracle.ORBextractor(world.shape[0], 1.2, 8)
    kps_all, desc_all, counts = ex.run_batch(frames, threads)
    cam = dict(EUROC_CAM) if cam is None else cam

    def one(t):
        n = int(counts[t])
        kps, desc = kps_all[t, :n].copy(), desc_all[t, :n].copy()
        m12, nm = None, 0
        if t > 0:
            pn = int(counts[t - 1])
            pk, pd = kps_all[t - 1, :pn].copy(), desc_all[t - 1, :pn].copy()
            f1 = oracle.Frame(pk, pd, w, h); f2 = oracle.Frame(kps, desc, w, h)
            pm = np.stack([pk["x"], pk["y"]], -1).astype(np.float32)
            nm, m12, _ = oracle.search_for_initialization(f1, f2, pm, window, nnratio, True)
        ob = oracle.Observability(**cam)
        ob.set_xv(xv); ob.predict(0.1, 1)
        blocks, obs, _ = ob.build_frame(world[:n], kps["octave"], score[:n], np.ones(n, np.uint8), sigma2)
        si, sl = oracle.select(blocks, k, 2, sample_scale, 1, scores=obs) if n > 0 else (np.zeros(0, np.int32), np.zeros(0))
        return kps, desc, m12, nm, si, sl
    with ThreadPoolExecutor(threads) as pool:
        return list(pool.map(one, range(nf)))
