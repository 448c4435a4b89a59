"""Optimizer::PoseOptimization restatement (oracle/pose_optimizer.cpp, src/Optimizer.cc:279-413): the reference holds no
vectors for it (its g2o/Eigen arithmetic is pinned by nothing), so the restatement is checked through the properties the
reference relies on: it recovers a perturbed pose from noisy projections to sub-pixel reprojection error, flags exactly the
gross outliers (chi2 > 5.991 at sigma^2 = level scale^2), leaves unmatched keypoints' flags alone, and returns
nInitialCorrespondences - nBad."""
import numpy as np

from gf_orb_slam_b200 import synth


def make_scene(n, seed, n_out=40, noise=0.5):
    u = synth.uniform01(seed, 8 * n).reshape(n, 8)
    fx, fy, cx, cy = 458.0, 457.0, 376.0, 240.0
    a = 0.1
    c, s = np.cos(a), np.sin(a)
    R = np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]]) @ np.array([[1, 0, 0], [0, np.cos(0.05), -np.sin(0.05)], [0, np.sin(0.05), np.cos(0.05)]])
    t = np.array([0.2, -0.1, 0.3])
    uv = np.stack([u[:, 0] * 752, u[:, 1] * 480], 1)
    z = 2.0 + 6.0 * u[:, 2]
    Xc = np.stack([(uv[:, 0] - cx) / fx * z, (uv[:, 1] - cy) / fy * z, z], 1)
    Xw = (Xc - t) @ R                      # Xc = R Xw + t
    octave = np.minimum((u[:, 3] * 8).astype(np.int32), 7)
    sig = 1.2 ** octave
    obs = uv + (u[:, 4:6] * 2 - 1) * noise * sig[:, None]
    out_idx = np.arange(n)[u[:, 6] < n_out / n]
    obs[out_idx] += np.where(u[out_idx, 7:8] < 0.5, 1.0, -1.0) * (25.0 + 40.0 * u[out_idx, 4:6])
    mp = np.where(u[:, 7] < 0.9, np.arange(n), -1).astype(np.int32)
    T = np.eye(4, dtype=np.float32); T[:3, :3] = R; T[:3, 3] = t
    # perturbed start: small rotation about z and a translation offset
    b = 0.02
    Rz = np.array([[np.cos(b), -np.sin(b), 0], [np.sin(b), np.cos(b), 0], [0, 0, 1]])
    T0 = np.eye(4, dtype=np.float32); T0[:3, :3] = Rz @ R; T0[:3, 3] = t + np.array([0.05, -0.03, 0.04])
    s2 = np.ones(8, np.float32)
    for i in range(1, 8):
        s2[i] = s2[i - 1] * np.float32(1.2)
    inv_sigma2 = (np.float32(1.0) / (s2 * s2)).astype(np.float32)
    return dict(T_true=T, T0=T0, xy=obs.astype(np.float32), octave=octave, mp=mp, world=Xw.astype(np.float32), inv_sigma2=inv_sigma2,
                K=(fx, fy, cx, cy), gross=out_idx)


def test_pose_optimization_recovers_pose_and_flags_outliers(orc):
    for seed, n in ((1, 1000), (2, 300), (3, 60)):
        S = make_scene(n, seed, n_out=max(4, n // 25))
        pre = np.full(n, 7, np.uint8)
        ninl, T, out = orc.pose_optimization(S["T0"], S["xy"], S["octave"], S["mp"], S["world"], S["inv_sigma2"], S["K"])
        matched = S["mp"] >= 0
        assert np.abs(T[:3, :3] - S["T_true"][:3, :3]).max() < 2e-3 and np.abs(T[:3, 3] - S["T_true"][:3, 3]).max() < 2e-2
        gross = np.zeros(n, bool); gross[S["gross"]] = True
        assert out[matched & gross].all()                               # every gross outlier among the matches is flagged
        assert out[matched & ~gross].mean() < 0.05                      # (a 2-sigma noise tail may exceed chi2 5.991)
        assert ninl == matched.sum() - out[matched].sum()
        assert (out[~matched] == 0).all()


def test_pose_optimization_degenerate(orc):
    S = make_scene(50, 5)
    S["mp"][:] = -1
    ninl, T, out = orc.pose_optimization(S["T0"], S["xy"], S["octave"], S["mp"], S["world"], S["inv_sigma2"], S["K"])
    assert ninl == 0 and np.array_equal(T, S["T0"].reshape(4, 4))
    S = make_scene(50, 6)
    S["mp"][6:] = -1                     # fewer than 10 edges: one optimisation round only (src/Optimizer.cc:400-401)
    ninl, T, out = orc.pose_optimization(S["T0"], S["xy"], S["octave"], S["mp"], S["world"], S["inv_sigma2"], S["K"])
    assert 0 <= ninl <= 6
