"""CPU checks of the oracle's keyframe-side matcher restatements (oracle/orb_matcher.cpp): SearchByBoW against an
independent pure-Python walk of the same rule on a small case, and the CV_32F transposed gemm behind `Ow = -Rcw.t()*tcw`
against cv2 4.13 (fixture tests/golden/cv413_gemm_t.npz, generator tests/golden/make_gemm_golden.py)."""
import os

import numpy as np
import pytest

from gf_orb_slam_b200 import synth

HERE = os.path.dirname(os.path.abspath(__file__))
KP = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])


@pytest.fixture(scope="module")
def orc():
    import oracle
    oracle.build()
    return oracle


def _frame(orc, n, seed, base=None):
    u = synth.uniform01(seed, n * 4).reshape(n, 4)
    k = np.zeros(n, KP)
    k["x"] = 20 + 600 * u[:, 0]; k["y"] = 20 + 400 * u[:, 1]; k["angle"] = 360 * u[:, 2]; k["octave"] = (u[:, 3] * 8).astype(np.int32)
    bits = synth.splitmix64(seed + 1, n * 4).view(np.uint8).reshape(n, 32).copy()
    if base is not None:                       # a noisy copy of another frame's descriptors and angles
        m = min(n, base[1].shape[0])
        bits[:m] = base[1][:m]
        flip = synth.uniform01(seed + 2, m * 32).reshape(m, 32) < 0.05
        bits[:m][flip] ^= 0x04
        k["angle"][:m] = (base[0]["angle"][:m] + 3.0) % 360
    return k, bits, orc.Frame(k, bits, 752, 480)


def _py_bow(kf_k, kf_d, kf_mp, kn, f_k, f_d, fn, ratio, ori):
    pop = np.array([bin(i).count("1") for i in range(256)], np.int32)
    out = np.full(len(f_k), -1, np.int32)
    hist = [[] for _ in range(30)]
    nm = 0
    order = sorted([i for i in range(len(kf_k)) if kn[i] >= 0], key=lambda i: (kn[i], i))
    for i in order:
        if kf_mp[i] < 0:
            continue
        cands = [j for j in range(len(f_k)) if fn[j] == kn[i]]
        b1 = b2 = 2**31 - 1; bj = -1
        for j in cands:
            if out[j] >= 0:
                continue
            d = int(pop[kf_d[i] ^ f_d[j]].sum())
            if d < b1:
                b2 = b1; b1 = d; bj = j
            elif d < b2:
                b2 = d
        if b1 <= 50 and np.float32(b1) < np.float32(ratio) * np.float32(b2):
            out[bj] = kf_mp[i]; nm += 1
            if ori:
                rot = np.float32(kf_k["angle"][i]) - np.float32(f_k["angle"][bj])
                if rot < 0:
                    rot = np.float32(rot + np.float32(360.0))
                b = int(np.round(np.float32(rot * np.float32(1.0 / 30))))      # halves do not occur with these angles
                hist[0 if b == 30 else b].append(bj)
    if ori:
        cnt = [len(h) for h in hist]
        m1 = m2 = m3 = 0; i1 = i2 = i3 = -1
        for i, s in enumerate(cnt):
            if s > m1:
                m3, m2, m1 = m2, m1, s; i3, i2, i1 = i2, i1, i
            elif s > m2:
                m3, m2 = m2, s; i3, i2 = i2, i
            elif s > m3:
                m3 = s; i3 = i
        if m2 < 0.1 * np.float32(m1):
            i2 = i3 = -1
        elif m3 < 0.1 * np.float32(m1):
            i3 = -1
        for i in range(30):
            if i not in (i1, i2, i3):
                for j in hist[i]:
                    out[j] = -1; nm -= 1
    return nm, out


@pytest.mark.parametrize("ratio,ori", [(0.7, True), (0.9, False)])
def test_search_by_bow_matches_python_walk(orc, ratio, ori):
    n1, n2 = 160, 150
    k1, d1, F1 = _frame(orc, n1, 11)
    k2, d2, F2 = _frame(orc, n2, 21, base=(k1, d1))
    mp = np.where(synth.uniform01(31, n1) < 0.8, 400 + np.arange(n1), -1).astype(np.int32)
    kn = (50 + 3 * (d1[:, 0] & 7)).astype(np.int32); kn[::17] = -1
    fn = (50 + 3 * (d2[:, 0] & 7)).astype(np.int32); fn[5::19] = -1; fn[d2[:, 0] & 7 == 6] = 999       # a node only the frame has
    F1.set_mappoints(mp)
    n, m = orc.search_by_bow(F1, kn, F2, fn, ratio, ori)
    en, em = _py_bow(k1, d1, mp, kn, k2, d2, fn, ratio, ori)
    assert n == en and np.array_equal(m, em)
    assert n > 20


def test_small_gemm_restatements_match_cv2_golden(orc):
    """`Rcw*x3Dw+tcw` and `-Rcw.t()*tcw` as cv::MatExpr evaluates them (one cv::gemm each, small-matrix path: float accumulation,
    alpha / beta applied in double) -- the forms the oracle (oracle.hpp: cvGemmRxPlusT, cvNegRtTimesT) and the kernels use,
    bit-for-bit against cv2 4.13 (tests/golden/cv413_gemm_small.npz, generator make_gemm_golden.py)"""
    g = np.load(os.path.join(HERE, "golden", "cv413_gemm_small.npz"))
    R, x, t, y, Ow = g["R"], g["x"], g["t"], g["y"], g["Ow"]
    for i in range(R.shape[0]):
        assert np.array_equal(orc.cv_rx_plus_t(R[i], x[i], t[i]), y[i]), i
        assert np.array_equal(orc.cv_neg_rt_t(R[i], t[i]), Ow[i]), i


def test_general_gemm_path_differs_from_the_small_path(orc):
    """cv::gemm with a transpose flag takes the general path (double accumulators): its results (cv413_gemm_t.npz) are not
    what `-Rcw.t()*tcw` evaluates to -- the MatExpr materialises the transpose first.  Guards against mixing the two up."""
    g = np.load(os.path.join(HERE, "golden", "cv413_gemm_t.npz"))
    R, t, general = g["R"], g["t"], g["Ow"]
    diff = 0
    for i in range(R.shape[0]):
        got = np.zeros(3, np.float32)
        for c in range(3):
            s = 0.0
            for k in range(3):
                s += float(R[i, k, c]) * float(t[i, k])
            got[c] = np.float32(s * -1.0)
        assert np.array_equal(got, general[i])
        diff += not np.array_equal(orc.cv_neg_rt_t(R[i], t[i]), general[i])
    assert diff > 0


# ---------------------------------------------------------------- mapping / loop-closing overloads (oracle/orb_matcher_kf.cpp)
POP = np.array([bin(i).count("1") for i in range(256)], np.int32)


def _py_triangulation(k1, d1, mp1, n1, k2, d2, mp2, n2, F, sigma2):
    """independent walk of src/ORBmatcher.cc:1426-1547 without the rotation histogram: (node, index) order, unmatched
    candidates without a map point at distance <= 50, sorted by (distance, index), first within twice the best that passes the
    epipolar test"""
    out = np.full(len(k1), -1, np.int32)
    taken = np.zeros(len(k2), bool)
    F = F.astype(np.float32).reshape(3, 3)
    order = sorted([i for i in range(len(k1)) if n1[i] >= 0], key=lambda i: (n1[i], i))
    for i in order:
        if mp1[i] >= 0:
            continue
        cand = []
        for j in range(len(k2)):
            if n2[j] != n1[i] or taken[j] or mp2[j] >= 0:
                continue
            d = int(POP[d1[i] ^ d2[j]].sum())
            if d <= 50:
                cand.append((d, j))
        if not cand:
            continue
        cand.sort()
        x1, y1 = np.float32(k1["x"][i]), np.float32(k1["y"][i])
        a = np.float32(np.float32(x1 * F[0, 0]) + np.float32(y1 * F[1, 0])) + F[2, 0]
        b = np.float32(np.float32(x1 * F[0, 1]) + np.float32(y1 * F[1, 1])) + F[2, 1]
        c = np.float32(np.float32(x1 * F[0, 2]) + np.float32(y1 * F[1, 2])) + F[2, 2]
        for d, j in cand:
            if d > 2 * cand[0][0]:
                break
            x2, y2 = np.float32(k2["x"][j]), np.float32(k2["y"][j])
            num = np.float32(np.float32(a * x2) + np.float32(b * y2)) + c
            den = np.float32(a * a) + np.float32(b * b)
            if den == 0:
                continue
            dsqr = np.float32(np.float32(num * num) / den)
            if float(dsqr) < 3.84 * float(sigma2[k2["octave"][j]]):
                taken[j] = True; out[i] = j
                break
    return int((out >= 0).sum()), out


def test_search_for_triangulation_matches_python_walk(orc):
    n1, n2 = 170, 160
    k1, d1, F1 = _frame(orc, n1, 111)
    k2, d2, F2 = _frame(orc, n2, 121, base=(k1, d1))
    k2["x"][:n2] = k1["x"][:n2] + 2.0; k2["y"][:n2] = k1["y"][:n2] + np.where(synth.uniform01(5, n2) < 0.3, 4.0, 0.5)
    F2 = orc.Frame(k2, d2, 752, 480)
    mp1 = np.where(synth.uniform01(131, n1) < 0.3, 10 + np.arange(n1), -1).astype(np.int32)
    mp2 = np.where(synth.uniform01(141, n2) < 0.3, 900 + np.arange(n2), -1).astype(np.int32)
    nd1 = (50 + 3 * (d1[:, 0] & 3)).astype(np.int32); nd1[::17] = -1
    nd2 = (50 + 3 * (d2[:, 0] & 3)).astype(np.int32); nd2[5::19] = -1
    F1.set_mappoints(mp1); F2.set_mappoints(mp2)
    F = np.array([0, 2e-5, 0, -1e-5, 0, -1, 1e-5, 1, 0.1], np.float32)
    n, m = orc.search_for_triangulation(F1, nd1, F2, nd2, F, checkOri=False)
    sigma2 = (np.float32(1.2) ** np.arange(8)).astype(np.float32) ** 2
    s = np.ones(8, np.float32)
    for i in range(1, 8):
        s[i] = np.float32(s[i - 1] * np.float32(1.2))
    en, em = _py_triangulation(k1, d1, mp1, nd1, k2, d2, mp2, nd2, F, (s * s).astype(np.float32))
    assert n == en and np.array_equal(m, em)
    assert 20 < n < (mp1 < 0).sum()
    # with the rotation histogram the result is a subset of it
    n_o, m_o = orc.search_for_triangulation(F1, nd1, F2, nd2, F, checkOri=True)
    assert n_o <= n and np.all((m_o == m) | (m_o == -1))


def test_search_by_bow_keyframes_properties(orc):
    """SearchByBoW(KF1, KF2): every match joins features of one node, both with a map point, distance < 50, no KF2 point twice"""
    n1, n2 = 200, 190
    k1, d1, F1 = _frame(orc, n1, 211)
    k2, d2, F2 = _frame(orc, n2, 221, base=(k1, d1))
    mp1 = np.where(synth.uniform01(231, n1) < 0.8, 10 + np.arange(n1), -1).astype(np.int32)
    mp2 = np.where(synth.uniform01(241, n2) < 0.8, 900 + np.arange(n2), -1).astype(np.int32)
    nd1 = (50 + 3 * (d1[:, 0] & 7)).astype(np.int32); nd2 = (50 + 3 * (d2[:, 0] & 7)).astype(np.int32)
    F1.set_mappoints(mp1); F2.set_mappoints(mp2)
    n, m = orc.search_by_bow_keyframes(F1, nd1, F2, nd2, 0.9, False)
    assert n == (m >= 0).sum() and n > 40
    used = m[m >= 0]
    assert len(set(used.tolist())) == len(used)
    for i in np.nonzero(m >= 0)[0]:
        j = int(m[i]) - 900
        assert mp1[i] >= 0 and mp2[j] >= 0 and nd1[i] == nd2[j] and int(POP[d1[i] ^ d2[j]].sum()) < 50


def test_sim3_decomposition_and_projection_properties(orc):
    """decomposeSim3 recovers (R, t, Ow) of Scw = s [R | t] to float accuracy; Fuse through Scw = s Tcw finds what Fuse with Tcw finds
    when the scale is a power of two (every product then scales exactly)"""
    a = 0.3
    R = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]], np.float32)
    t = np.array([0.3, -0.2, 0.5], np.float32)
    S = np.eye(4, dtype=np.float32); S[:3, :3] = 1.75 * R; S[:3, 3] = 1.75 * t
    Rd, td, Ow = orc.sim3_decompose(S)
    assert np.allclose(Rd, R, atol=2e-7) and np.allclose(td, t, atol=2e-7)
    assert np.allclose(Ow, -(R.T.astype(np.float64) @ t), atol=1e-6)
    n = 300
    k, d, F = _frame(orc, n, 311)
    fx = fy = 458.0; cx, cy = 376.0, 240.0
    z = (2 + 4 * synth.uniform01(312, n)).astype(np.float32)
    Xc = np.stack([(k["x"] + 1.0 - cx) / fx * z, (k["y"] - 1.0 - cy) / fy * z, z], 1).astype(np.float32)
    Xw = ((Xc - t) @ R).astype(np.float32)
    T = np.eye(4, dtype=np.float32); T[:3, :3] = R; T[:3, 3] = t
    Owd = -(R.T.astype(np.float64) @ t)
    d3 = np.linalg.norm(Xw - Owd, axis=1)
    scale = (np.float32(1.2) ** np.arange(8))
    mind = (d3 / (scale[k["octave"]] * 0.99)).astype(np.float32); maxd = (mind * 5).astype(np.float32)
    nrm = ((Xw - Owd) / d3[:, None]).astype(np.float32)
    F.set_camera(fx, fy, cx, cy, T)
    pts = np.arange(n, dtype=np.int32)
    slot = np.full(n, -1, np.int32); bad = np.zeros(n, np.uint8); ik = np.full(n, -1, np.int32)
    n1, bi1, bd1, s1, b1, i1 = orc.fuse(F, None, pts, Xw, nrm, mind, maxd, d, 3.0, slot, bad, ik)
    S2 = T.copy(); S2[:3, :] *= 2.0
    n2, bi2, bd2, s2, b2, i2 = orc.fuse(F, S2, pts, Xw, nrm, mind, maxd, d, 3.0, slot, bad, ik)
    assert n1 == n2 and n1 > 200
    assert np.array_equal(bi1, bi2) and np.array_equal(bd1, bd2) and np.array_equal(s1, s2)
    # every fused point took the keypoint it was generated from (descriptor distance 0) and now sits in that slot
    fused = bd1 <= 50
    assert np.array_equal(bi1[fused], pts[fused]) and (bd1[fused] == 0).all() and np.array_equal(s1[fused], pts[fused])


@pytest.mark.parametrize("dy,octave,expect", [(0.0, 0, 1), (1.9, 0, 1), (2.0, 0, 0), (3.0, 0, 0), (3.0, 3, 1), (3.4, 3, 0), (-3.0, 3, 1)])
def test_epipolar_gate_known_answers(orc, dy, octave, expect):
    """CheckDistEpipolarLine (src/ORBmatcher.cc:705-722) by hand: with F12 = [t]x for a translation along x the line of (x1, y1) is
    y = y1, the squared distance of (x2, y1 + dy) is dy^2, and the gate is dy^2 < 3.84 * sigma2(octave of kp2) with
    sigma2 = 1.2^(2 octave): 3.84 at octave 0 (|dy| < 1.96), 11.47 at octave 3 (|dy| < 3.39)"""
    k1 = np.zeros(1, KP); k2 = np.zeros(1, KP)
    k1["x"] = 100.0; k1["y"] = 50.0; k2["x"] = 80.0; k2["y"] = 50.0 + dy; k2["octave"] = octave
    d = np.arange(32, dtype=np.uint8).reshape(1, 32)
    F1 = orc.Frame(k1, d, 752, 480); F2 = orc.Frame(k2, d, 752, 480)
    F1.set_mappoints(np.array([-1], np.int32)); F2.set_mappoints(np.array([-1], np.int32))
    F = np.array([0, 0, 0, 0, 0, -1, 0, 1, 0], np.float32)
    n, m = orc.search_for_triangulation(F1, np.array([7], np.int32), F2, np.array([7], np.int32), F, checkOri=False)
    assert n == expect and m[0] == (0 if expect else -1)
