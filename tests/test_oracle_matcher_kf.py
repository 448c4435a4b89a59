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
