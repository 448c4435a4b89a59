"""Deterministic synthetic workloads.

Library-independent generators so every arm (CUDA path, CPU oracle, cv2 cross-checks) sees identical bytes.
RNG = counter-based splitmix64: output i (0-based) of stream `seed` is mix(seed + (i+1)*GAMMA).
"""
from __future__ import annotations

import numpy as np

_GAMMA = np.uint64(0x9E3779B97F4A7C15)
_M1 = np.uint64(0xBF58476D1CE4E5B9)
_M2 = np.uint64(0x94D049BB133111EB)


def splitmix64(seed: int, n: int, offset: int = 0) -> np.ndarray:
    """n outputs of the splitmix64 stream `seed`, starting at output index `offset`."""
    with np.errstate(over="ignore"):
        i = np.arange(offset + 1, offset + n + 1, dtype=np.uint64)
        z = np.uint64(seed & 0xFFFFFFFFFFFFFFFF) + i * _GAMMA
        z = (z ^ (z >> np.uint64(30))) * _M1
        z = (z ^ (z >> np.uint64(27))) * _M2
        return z ^ (z >> np.uint64(31))


def _binomial_axis(a: np.ndarray, axis: int) -> np.ndarray:
    """[1,4,6,4,1] pass along `axis`, reflect-101 border, (sum + 8) >> 4."""
    a = a.astype(np.int32)
    pad = [(0, 0), (0, 0)]
    pad[axis] = (2, 2)
    p = np.pad(a, pad, mode="reflect")
    n = a.shape[axis]
    sl = lambda k: tuple(slice(k, k + n) if ax == axis else slice(None) for ax in range(2))
    s = p[sl(0)] + 4 * p[sl(1)] + 6 * p[sl(2)] + 4 * p[sl(3)] + p[sl(4)]
    return (s + 8) >> 4


def textured_frame(w: int, h: int, seed: int, band: bool = True) -> np.ndarray:
    """8-bit textured frame: 4x nearest-upsampled random blocks, binomial smoothing, +-6 integer noise, plus a
    low-contrast band (middle third of the rows: weak checkerboard on the left half -> FAST-7 fallback cells, flat
    on the right half -> empty cells)."""
    bw, bh = -(-w // 4), -(-h // 4)
    base = (splitmix64(seed, bw * bh) >> np.uint64(56)).astype(np.uint8).reshape(bh, bw)
    img = np.repeat(np.repeat(base, 4, axis=0), 4, axis=1)[:h, :w]
    img = _binomial_axis(_binomial_axis(img, 1), 0)
    noise = ((splitmix64(seed, w * h, offset=bw * bh) >> np.uint64(56)).astype(np.int32) % 13 - 6).reshape(h, w)
    img = np.clip(img + noise, 0, 255).astype(np.uint8)
    if band:
        y0, y1 = h // 2 - h // 6, h // 2 + h // 6
        yy, xx = np.mgrid[y0:y1, 0:w]
        weak = (96 + 12 * (((xx >> 4) + (yy >> 4)) & 1)).astype(np.uint8)
        img[y0:y1, :] = np.where(xx < w // 2, weak, np.uint8(128))
    return np.ascontiguousarray(img)


class Sequence:
    """EuRoC-shaped sequence (config C2): integer-shifted crops of one master texture plus fresh +-3 noise."""

    def __init__(self, w: int = 752, h: int = 480, master_seed: int = 0):
        self.w, self.h = w, h
        self.mw, self.mh = w + 528, h + 288      # 1280 x 768 for 752 x 480
        self.master = textured_frame(self.mw, self.mh, master_seed)

    def offset(self, t: int) -> tuple[int, int]:
        return t % 512, (t // 4) % 256

    def frame(self, t: int) -> np.ndarray:
        x0, y0 = self.offset(t)
        crop = self.master[y0:y0 + self.h, x0:x0 + self.w].astype(np.int32)
        noise = ((splitmix64(1000 + t, self.w * self.h) >> np.uint64(56)).astype(np.int32) % 7 - 3)
        return np.clip(crop + noise.reshape(self.h, self.w), 0, 255).astype(np.uint8)

    def frames(self, t0: int, n: int) -> np.ndarray:
        return np.stack([self.frame(t) for t in range(t0, t0 + n)])


def landmark_cloud(n: int, seed: int, visible_from=None):
    """tools/simu_greedy.cc:30-145 recipe with splitmix64 instead of the unseeded rand(): landmarks x,y in [-4,4],
    z in [0,8]; octave = round(7 U); match score = round(100 U).  Returns (world[n,3] f32, octave[n] i32,
    score[n] i32, U[n,2] noise in [-1,1))."""
    if visible_from is not None:
        # tools/simu_greedy.cc:108-116 keeps a landmark only if project_Point_To_Frame (include/Observability.h:
        # 299-333) puts it in front of the camera and inside the image; draw until n such landmarks exist
        xv, fu, fv, cx, cy, ncols, nrows = visible_from
        q0, qx, qy, qz = [float(v) for v in xv[3:7]]
        R = np.array([[q0 * q0 + qx * qx - qy * qy - qz * qz, 2 * (qx * qy - q0 * qz), 2 * (qz * qx + q0 * qy)],
                      [2 * (qx * qy + q0 * qz), q0 * q0 - qx * qx + qy * qy - qz * qz, 2 * (qy * qz - q0 * qx)],
                      [2 * (qz * qx - q0 * qy), 2 * (qy * qz + q0 * qx), q0 * q0 - qx * qx - qy * qy + qz * qz]])
        Rinv = np.linalg.inv(R)
        m = 8 * n + 64
        uu = ((splitmix64(seed, 7 * m) >> np.uint64(11)).astype(np.float64) / float(1 << 53)).reshape(m, 7)
        w = np.stack([uu[:, 0] * 8 - 4, uu[:, 1] * 8 - 4, uu[:, 2] * 8], axis=1).astype(np.float32).astype(np.float64)
        h = (w - np.asarray(xv[:3], np.float64)) @ Rinv.T
        with np.errstate(divide="ignore", invalid="ignore"):
            pu = fu * h[:, 0] / h[:, 2] + cx
            pv = fv * h[:, 1] / h[:, 2] + cy
        ok = (h[:, 2] > 0) & (pu >= 0) & (pu <= ncols) & (pv >= 0) & (pv <= nrows)
        u = uu[ok][:n]
        assert u.shape[0] == n, "not enough visible landmarks"
    else:
        u = (splitmix64(seed, 7 * n) >> np.uint64(11)).astype(np.float64) / float(1 << 53)
        u = u.reshape(n, 7)
    world = np.stack([u[:, 0] * 8 - 4, u[:, 1] * 8 - 4, u[:, 2] * 8], axis=1).astype(np.float32)
    octave = np.round(u[:, 3] * 7).astype(np.int32)
    score = np.round(u[:, 4] * 100).astype(np.int32)
    pix_noise = (u[:, 5:7] * 2 - 1).astype(np.float32)
    return world, octave, score, pix_noise


def uniform01(seed: int, n: int, offset: int = 0) -> np.ndarray:
    return (splitmix64(seed, n, offset) >> np.uint64(11)).astype(np.float64) / float(1 << 53)


def map_points_c3(kps: np.ndarray, desc: np.ndarray, M: int, seed: int, w: int, h: int, nlevels: int = 8):
    """Config C3: M projected map points against one frame's keypoints: uniform (u, v) for 40 % of the
    points, the rest placed within a few pixels of a random keypoint; level in U{0..7}; viewCos in {0.9, 0.999};
    descriptor = a random frame descriptor with k~U{0..60} flipped bits (70 %) or fully random (30 %); 3 % of the
    points are not in view and 2 % are bad.  Returns dict of SoA arrays."""
    n = len(kps)
    u01 = uniform01(seed, 8 * M).reshape(M, 8)
    pick = np.minimum((u01[:, 0] * n).astype(np.int64), n - 1)
    near = u01[:, 1] < 0.6
    jit = (u01[:, 2:4] * 2 - 1) * 6.0
    px = np.where(near, kps["x"][pick] + jit[:, 0], u01[:, 2] * w).astype(np.float32)
    py = np.where(near, kps["y"][pick] + jit[:, 1], u01[:, 3] * h).astype(np.float32)
    level = np.where(near & (u01[:, 4] < 0.8), kps["octave"][pick] + (u01[:, 4] < 0.3), np.floor(u01[:, 4] * nlevels))
    level = np.clip(level, 0, nlevels - 1).astype(np.int32)
    view_cos = np.where(u01[:, 5] < 0.5, 0.9, 0.999).astype(np.float32)
    flags = np.ones(M, np.uint8)
    flags[u01[:, 6] < 0.03] = 0
    flags[(u01[:, 6] >= 0.03) & (u01[:, 6] < 0.05)] = 3
    d = desc[pick].copy()
    bits = splitmix64(seed + 1, M * 4).view(np.uint8).reshape(M, 32)
    rnd = u01[:, 7] < 0.3
    d[rnd] = bits[rnd]
    nflip = np.floor(uniform01(seed + 2, M) * 61).astype(np.int64)
    pos = (uniform01(seed + 3, M * 60).reshape(M, 60) * 256).astype(np.int64)
    for k in range(60):
        sel = (~rnd) & (nflip > k)
        b = pos[sel, k]
        d[sel, b >> 3] ^= (1 << (b & 7)).astype(np.uint8)
    return dict(proj_x=px, proj_y=py, scale_level=level, view_cos=view_cos, flags=flags, descriptors=np.ascontiguousarray(d))
