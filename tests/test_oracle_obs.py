"""Pins the oracle's Observability restatement against the reference's own golden vectors
(test/test_Jacobian.cpp, test/test_Kine_1.cpp, test/test_Kine_2.cpp) with the reference's own tolerances."""
import numpy as np
import pytest

from tests.golden import obs_reference_vectors as G


def infnorm(a):   # arma::norm(M, "inf") = max row sum of |.|
    return np.abs(np.asarray(a)).sum(axis=1).max()


def test_jacobian_golden(orc):
    obs = orc.Observability(**G.JAC_CAMERA)
    obs.set_xv(G.JAC_XV)
    obs.predict(0.1, 1)
    xv = obs.kine(0)["Xv"]
    for i, (Y, Z, H13g, H47g) in enumerate(G.JAC_CASES):
        Yf = np.asarray(Y, np.float32).astype(np.float64)     # reference stores the landmark in a CV_32F Mat
        Zf = np.asarray(Z, np.float32).astype(np.float64)
        H13, H47 = obs.H_complete(xv, Yf, Zf)
        assert infnorm(H13 - np.asarray(H13g)) < G.JAC_TOL, (i, H13)
        tol47 = G.JAC_TOL_CASE3_H47 if i == 3 else G.JAC_TOL
        assert infnorm(H47 - np.asarray(H47g)) < tol47, (i, H47)
        # k1=k2=0 -> the "simplied" variant used by the batch builders must agree with "complete"
        H13s, H47s, uv = obs.H_simplified(xv, Yf)
        np.testing.assert_allclose(H13s, H13, rtol=1e-12, atol=1e-12)
        np.testing.assert_allclose(H47s, H47, rtol=1e-12, atol=1e-12)
        assert abs(uv[0] - Z[0]) < 10 and abs(uv[1] - Z[1]) < 10      # test_Jacobian.cpp:258-259


def test_kine1_golden(orc):
    obs = orc.Observability(**G.KINE1_CAMERA)
    obs.set_xv(G.KINE1_XV)
    obs.predict(0.1, 1)
    k = obs.kine(0)
    for name, tol in G.KINE_TOL.items():
        assert infnorm(k[name] - np.asarray(G.KINE1[name])) < tol, name


def test_kine2_golden(orc):
    obs = orc.Observability(**G.JAC_CAMERA)
    obs.set_xv(G.JAC_XV)
    obs.predict(0.1, 3)
    for seg in range(3):
        k = obs.kine(seg)
        for name, tol in G.KINE_TOL.items():
            assert infnorm(k[name] - np.asarray(G.KINE2[seg][name])) < tol, (seg, name)
        F = k["F"]
        assert np.allclose(F[3:7, 3:7], k["F_Q"]) and np.allclose(F[3:7, 10:13], k["F_Omg"])
        assert F[0, 7] == pytest.approx(0.1, rel=1e-6)


def test_logdet_matches_numpy(orc):
    rng = np.random.default_rng(0)
    for n in (7, 13):
        for _ in range(20):
            A = rng.standard_normal((n, 2 * n))
            M = A @ A.T + 1e-5 * np.eye(n)
            s, ld = np.linalg.slogdet(M)
            assert orc.logdet(M) == pytest.approx(ld, rel=1e-10)
    # non-PD -> LU fallback returns log|det|
    M = np.diag([1.0, -2.0, 3.0, 1, 1, 1, 1])
    assert orc.logdet(M) == pytest.approx(np.log(6.0), rel=1e-12)


def test_info_block_is_HtH_over_sigma2(orc):
    obs = orc.Observability(**G.JAC_CAMERA)
    obs.set_xv(G.JAC_XV)
    obs.predict(0.1, 1)
    world = np.asarray([c[0] for c in G.JAC_CASES], np.float32)
    octave = np.array([0, 1, 2, 3, 7], np.int32)
    sc = np.array([10, 20, 30, 40, 50], np.int32)
    valid = np.array([1, 1, 0, 1, 1], np.uint8)
    s = np.ones(8, np.float32)
    for i in range(1, 8):
        s[i] = s[i - 1] * np.float32(1.2)
    sig2 = (s * s).astype(np.float32)
    blocks, obsScore, H = obs.build_frame(world, octave, sc, valid, sig2)
    for i in range(5):
        if not valid[i]:
            assert obsScore[i] == -1
            continue
        np.testing.assert_allclose(blocks[i], H[i].T @ H[i] / float(sig2[octave[i]]), rtol=1e-12)
        assert obsScore[i] == sc[i]
