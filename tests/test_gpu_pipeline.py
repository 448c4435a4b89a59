"""GPU tests of the whole front end at the bench's full size (256 frames of 752x480, 1000 keypoints, pick 100) through
size-independent properties, plus a bit-exact comparison of all 256 frames against the CPU oracle, and agreement of the selection
kernel's launch variants."""
import hashlib
import os

import numpy as np
import pytest

from gf_orb_slam_b200 import synth

pytestmark = pytest.mark.gpu

W, H, NF, K = 752, 480, 1000, 100


def _run(frames, batch, host=False):
    """all frames through the front end in steps of `batch`; returns per-frame numpy results.  host=True: through
    gforb_frontend_step_host (pinned host frames in, results copied out by the library) instead of device-resident frames"""
    import torch
    from gf_orb_slam_b200.pipeline import TrackingFrontEnd
    fe = TrackingFrontEnd(W, H, NF, batch, n_good=K)
    out = []
    for s in range(0, frames.shape[0], batch):
        if host:
            fe.h_images.copy_(torch.from_numpy(frames[s:s + batch]))
            fe.step_host()
            fe.sync()
            res = [t.numpy().copy() for t in (fe.h_kps, fe.h_desc, fe.h_counts, fe.h_m12, fe.h_nmatch, fe.h_sel_idx, fe.h_sel_ld,
                                              fe.h_sel_n)]
        else:
            fe.load_host_frames(frames[s:s + batch])
            fe.step_device()
            fe.sync()
            res = [t.cpu().numpy().copy() for t in (fe.kps, fe.desc, fe.counts, fe.m12, fe.nmatch, fe.sel_idx, fe.sel_ld, fe.sel_n)]
        for b in range(batch):
            out.append(tuple(r[b] for r in res))
    fe.close()
    return out


def _sha(res, skip_first_match=True):
    h = hashlib.sha1()
    for t, (kps, desc, cnt, m12, nm, sidx, sld, sn) in enumerate(res):
        c = int(cnt)
        h.update(kps[:c].tobytes()); h.update(desc[:c].tobytes())
        if t > 0:
            h.update(m12.tobytes()); h.update(np.int32(nm).tobytes())
        h.update(sidx[:int(sn)].tobytes()); h.update(sld[:int(sn)].tobytes())
    return h.hexdigest()


@pytest.fixture(scope="module")
def full():
    frames = synth.Sequence(W, H).frames(0, 256)
    return frames, _run(frames, 256)


def test_full_size_invariants(full):
    frames, res = full
    assert len(res) == 256
    for t, (kps, desc, cnt, m12, nm, sidx, sld, sn) in enumerate(res):
        c = int(cnt)
        assert 0 < c <= NF
        x, y, octv = kps[:c, 0], kps[:c, 1], kps[:c].view(np.int32)[:, 5]
        assert (x >= 0).all() and (x < W).all() and (y >= 0).all() and (y < H).all()
        assert (octv >= 0).all() and (octv < 8).all() and (np.diff(octv) >= 0).all()      # level-major order
        ang = kps[:c, 3]
        assert (ang >= 0).all() and (ang < 360).all()
        if t > 0:
            m = m12[m12 >= 0]
            assert len(m) == int(nm) and len(np.unique(m)) == len(m) and (m < c).all()     # vnMatches12 is injective
            assert int(nm) > 50                                                           # consecutive frames do match
        s = sidx[:int(sn)]
        assert int(sn) == min(K, c) and len(np.unique(s)) == len(s) and (s >= 0).all() and (s < c).all()
        if c > K:
            assert (np.diff(sld[:int(sn)]) >= 0).all()                                    # log det grows with every pick


def test_batch_size_and_host_path_do_not_change_results(full):
    """the same 256 frames in 4 steps of 64 (1-frame halo between steps) and through the host-buffer entry point
    (gforb_frontend_step_host: H2D / D2H inside the library, steps pipelined) give the same bytes"""
    frames, res = full
    ref = _sha(res)
    assert _sha(_run(frames, 64)) == ref
    assert _sha(_run(frames, 64, host=True)) == ref
    assert _sha(_run(frames, 256)) == ref                                                 # and the run is deterministic


def test_all_frames_match_oracle(orc, full):
    """every one of the 256 bench-size frames against the CPU oracle: keypoints, descriptors, matches, subset and the
    running log-dets, all bit for bit"""
    from gf_orb_slam_b200._lib import KP_DTYPE
    from gf_orb_slam_b200.pipeline import TrackingFrontEnd
    from tests._front_end_check import oracle_front_end_batch
    frames, res = full
    fe = TrackingFrontEnd(W, H, NF, 1, n_good=K)            # only for the constants of the synthetic map points
    exp = oracle_front_end_batch(orc, frames, fe.world_h, fe.score_h, fe.xv_h, fe.sigma2_h, K, fe.sample_scale)
    assert len(exp) == len(res) == 256
    for t in range(256):
        ek, ed, em12, enm, esi, esl = exp[t]
        kps, desc, cnt, m12, nm, sidx, sld, sn = res[t]
        c = int(cnt)
        assert c == len(ek) and kps[:c].copy().view(KP_DTYPE).reshape(-1).tobytes() == ek.tobytes(), t
        assert np.array_equal(desc[:c], ed), t
        if t > 0:
            assert int(nm) == enm and np.array_equal(m12[:len(em12)], em12), t
        assert int(sn) == len(esi) and np.array_equal(sidx[:int(sn)], esi), t
        assert np.array_equal(sld[:int(sn)], esl), t           # north_star asks 1e-5 relative; they are bit-equal


def _select_once(B, n, k, mtd):
    from gf_orb_slam_b200.observability import Observability
    from gf_orb_slam_b200.pipeline_consts import EUROC_CAM, SIMU_XV, level_sigma2, simu_visible_from
    ob = Observability(**EUROC_CAM, max_batch=B, max_points=n)
    ob.set_state(SIMU_XV); ob.predictPWLSVec(0.1, 1)
    Xv = np.repeat(ob.kinematic(0)["Xv"][None], B, 0)
    cl = [synth.landmark_cloud(n, 300 + b, visible_from=simu_visible_from(W, H)) for b in range(B)]
    world = np.stack([c[0] for c in cl]); octave = np.stack([c[1] for c in cl]).astype(np.int32)
    score = np.stack([c[2] for c in cl]).astype(np.int32)
    valid = np.ones((B, n), np.uint8)
    valid[B - 1, ::13] = 0                   # the last problem has invalid pool entries -> in-order kernel
    counts = np.array([n, n - 5, n, n // 2, n, n], np.int32)
    blocks, osc = ob.build_info_frame(Xv, world, octave, score, valid, counts, level_sigma2())
    idx, ld, cnt = ob.select(blocks, osc, counts, k, mtd, 9.0, np.arange(1, B + 1, dtype=np.uint32))
    return idx.copy(), ld.copy(), cnt.copy()


@pytest.mark.parametrize("mtd,k", [(2, 100), (5, 700), (3, 120)])
def test_selection_variants_agree(mtd, k):
    """threads per problem (64 / 128 / 256), CTAs per SM and the sampler-ahead vs in-order kernels are launch choices only"""
    B, n = 6, 900
    outs = []
    keys = ("GFORB_SELECT_THREADS", "GFORB_SELECT_NO_AHEAD", "GFORB_SELECT_MINB")
    for env in ({"GFORB_SELECT_THREADS": "256"}, {"GFORB_SELECT_THREADS": "128", "GFORB_SELECT_MINB": "2"},
                {"GFORB_SELECT_THREADS": "128", "GFORB_SELECT_MINB": "3"}, {"GFORB_SELECT_THREADS": "64"},
                {"GFORB_SELECT_NO_AHEAD": "1"}):
        old = {key: os.environ.get(key) for key in keys}
        for key in keys:
            os.environ.pop(key, None)
        os.environ.update(env)
        try:
            outs.append(_select_once(B, n, k, mtd))
        finally:
            for key, v in old.items():
                os.environ.pop(key, None)
                if v is not None:
                    os.environ[key] = v
    for o in outs[1:]:
        assert np.array_equal(o[2], outs[0][2])
        for b in range(B):
            c = int(o[2][b])
            assert np.array_equal(o[0][b, :c], outs[0][0][b, :c]) and np.array_equal(o[1][b, :c], outs[0][1][b, :c])


def test_cpp_caller_through_the_c_abi(tmp_path):
    """tests/cpp/frontend_cli.cpp -- a plain C++ program over include/gforb.h (cudaHostAlloc'ed frames in, host results
    out, two staging sets, steps enqueued back to back) -- delivers the same bytes as the Python-driven front end."""
    import json
    import subprocess
    import torch
    from gf_orb_slam_b200._lib import KP_DTYPE
    from gf_orb_slam_b200.pipeline import TrackingFrontEnd
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "tests", "cpp", "frontend_cli.bin")
    assert os.path.exists(exe), "run __graft_entry__.build()"
    B, steps = 32, 3
    frames = synth.Sequence(W, H).frames(0, 2 * B)
    fe = TrackingFrontEnd(W, H, NF, B, n_good=K)
    (tmp_path / "frames.bin").write_bytes(frames.tobytes())
    (tmp_path / "map.bin").write_bytes(fe.world_h.tobytes() + fe.score_h.tobytes() + np.ascontiguousarray(fe.xv_h, np.float64).tobytes()
                                       + np.ascontiguousarray(fe.sigma2_h, np.float32).tobytes())
    r = subprocess.run([exe, str(tmp_path / "frames.bin"), str(tmp_path / "map.bin"), str(W), str(H), str(NF), str(B), str(K),
                        repr(fe.sample_scale), str(steps), str(tmp_path / "out.bin")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    info = json.loads(r.stdout.strip().splitlines()[-1])
    assert info["frames_per_s"] > 0 and info["keypoints_last_step"] > 0
    for s in range(steps + 1):                                   # the same step sequence (warm-up step 0 + `steps`)
        idx = [(s * B + b) % (2 * B) for b in range(B)]
        fe.h_images.copy_(torch.from_numpy(frames[idx]))
        fe.step_host()
        fe.sync()
    raw = (tmp_path / "out.bin").read_bytes()
    off = 0

    def take(dtype, shape):
        nonlocal off
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        a = np.frombuffer(raw, dtype, int(np.prod(shape)), off).reshape(shape)
        off += n
        return a
    kps = take(KP_DTYPE, (B, NF)); desc = take(np.uint8, (B, NF, 32)); cnt = take(np.int32, (B,)); m12 = take(np.int32, (B, NF))
    nm = take(np.int32, (B,)); gi = take(np.int32, (B, K)); gl = take(np.float64, (B, K)); gn = take(np.int32, (B,))
    assert off == len(raw)
    assert np.array_equal(cnt, fe.h_counts.numpy()) and np.array_equal(nm, fe.h_nmatch.numpy()) and np.array_equal(gn, fe.h_sel_n.numpy())
    hk = fe.h_kps.numpy().view(np.int32)
    for b in range(B):
        c = int(cnt[b])
        assert kps[b, :c].tobytes() == hk[b, :c].tobytes() and np.array_equal(desc[b, :c], fe.h_desc.numpy()[b, :c])
        assert np.array_equal(m12[b], fe.h_m12.numpy()[b])
        assert np.array_equal(gi[b, :gn[b]], fe.h_sel_idx.numpy()[b, :gn[b]]) and np.array_equal(gl[b, :gn[b]], fe.h_sel_ld.numpy()[b, :gn[b]])
    fe.close()
