"""Torch-free constants of the synthetic front-end workload (shared by pipeline.py and bench.py's CPU legs)."""
from __future__ import annotations

import math

import numpy as np

EUROC_CAM = dict(f=5.1369248, nRows=480, nCols=752, Cx=367.215, Cy=248.375, k1=0.0, k2=0.0, dx=0.01123325985,
                 dy=0.01123325985)                                    # tools/simu_greedy.cc:33-41
SIMU_XV = np.array([0.3, -0.1, 1.0, 0.9992, 0.0131, 0.0314, -0.0209, 3.0, -1.0, 10.0, 0.2618, 0.6283, -0.4189])   # :67


def level_sigma2(nlevels=8, scale_factor=1.2) -> np.ndarray:
    """Frame::mvLevelSigma2 (src/Frame.cc:104-112)"""
    s = np.ones(nlevels, np.float32)
    for i in range(1, nlevels):
        s[i] = s[i - 1] * np.float32(scale_factor)
    return (s * s).astype(np.float32)


def sample_scale_for(N_frame: int, k: int, error_bound: float) -> float:
    """random_sample_scale of setSelction_Number (src/Observability.cc:826)"""
    return float(int(float(np.float32(N_frame) / np.float32(k)) * math.log(1.0 / error_bound)))


def level_sizes(w: int, h: int, nlevels=8, scale_factor=1.2):
    """ComputePyramid level sizes (src/ORBextractor.cc:926-927) with the ctor's float tables (:469-478)"""
    inv = np.float32(1.0)
    f = np.float32(1.0) / np.float32(scale_factor)
    out = []
    for _ in range(nlevels):
        out.append((int(np.rint(np.float32(w) * inv)), int(np.rint(np.float32(h) * inv))))
        inv = np.float32(inv * f)
    return out


def simu_visible_from(width=752, height=480):
    """(Xv, fu, fv, Cx, Cy, nCols, nRows) of the tools/simu_greedy.cc fixture, for synth.landmark_cloud(visible_from=)"""
    c = EUROC_CAM
    return (SIMU_XV, c["f"] / c["dx"], c["f"] / c["dy"], c["Cx"], c["Cy"], width, height)
