"""Host-side mirror of ORB_SLAM::Optimizer::PoseOptimization (reference src/Optimizer.cc:279-413) over the C ABI
(gforb_pose_optimization): batched motion-only bundle adjustment, one CTA per frame (csrc/poseopt.cu)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import check, is_device, ptr


class Optimizer:
    def __init__(self, *, max_batch=1, max_points=2048, device=0):
        h = C.c_void_p()
        check(_lib.lib().gforb_poseopt_create(int(max_batch), int(max_points), int(device), C.byref(h)), "gforb_poseopt_create")
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            _lib.lib().gforb_poseopt_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def PoseOptimization(self, Tcw, keys_un, map_points, world, counts, inv_level_sigma2, K, outlier=None, *, stream=None, sync=True):
        """Tcw (B, 16) float32 in/out; keys_un (B, S, 7) float32 view of cv::KeyPoint (or the structured KP_DTYPE array);
        map_points (B, S) int32 ids (-1 = no map point); world (B, S, 3) float32; counts (B,) int32; inv_level_sigma2 host
        float32 array.  Returns (n_inliers (B,), outlier (B, S) uint8); Tcw is updated in place."""
        B, S = int(map_points.shape[0]), int(map_points.shape[1])
        dev = is_device(map_points)
        if outlier is None:
            if dev:
                import torch
                outlier = torch.zeros((B, S), dtype=torch.uint8, device=map_points.device)
            else:
                outlier = np.zeros((B, S), np.uint8)
        if dev:
            import torch
            ninl = torch.zeros(B, dtype=torch.int32, device=map_points.device)
        else:
            ninl = np.zeros(B, np.int32)
        s2 = np.ascontiguousarray(inv_level_sigma2, np.float32)
        fx, fy, cx, cy = [float(v) for v in K]
        flags = (_lib.GFORB_IN_DEVICE | _lib.GFORB_OUT_DEVICE if dev else 0) | (0 if sync else _lib.GFORB_ASYNC)
        check(_lib.lib().gforb_pose_optimization(self._h, ptr(Tcw), ptr(keys_un), ptr(map_points), ptr(world), ptr(counts), S, B, ptr(s2),
                                                 int(s2.shape[0]), fx, fy, cx, cy, ptr(outlier), ptr(ninl), flags,
                                                 C.c_void_p(stream or 0)), "gforb_pose_optimization")
        return ninl, outlier
