"""The per-frame tracking front end as one batched, pipelined device path (BASELINE.json metric: extract + match +
GF-select) -- a thin ctypes caller of gforb_frontend_* (include/gforb.h; the scheduling lives in C++, csrc/frontend.cu).

For every frame t of a step:  ORBextractor::operator() -> ORBmatcher::SearchForInitialization(frame t-1, frame t,
window 100, nnratio 0.9, checkOri)  (the map-free frame-to-frame matcher) -> after the matcher,
Observability::runMatrixBuilding(FRAME_INFO_MATRIX) + maxVolSelection_LazierGreedy (config C4: <=1000 matched points,
pick 100, eps 0.1).  Frame t-1 of a step's first frame is the last frame of the previous step (1-frame halo).
torch only owns the frame buffers handed in and wraps the result buffers the library owns."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib, synth
from ._lib import FrontEndConfig, FrontEndOutputs, check
from .observability import Observability
from .pipeline_consts import (EUROC_CAM, SIMU_XV, level_sigma2, level_sizes, sample_scale_for,  # noqa: F401
                              simu_visible_from)


class _DevArray:
    """a device buffer owned by libgforb.so, exposed through __cuda_array_interface__ so torch can view it"""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


class TrackingFrontEnd:
    def __init__(self, width=752, height=480, nfeatures=1000, batch=64, *, n_good=100, error_bound=0.1, window=100,
                 nnratio=0.9, device=0, cloud_seed=7, overlap=True):
        self.w, self.h, self.nf, self.B, self.k = width, height, nfeatures, batch, n_good
        self.dev = torch.device("cuda", device)
        torch.cuda.set_device(self.dev)
        cam = dict(EUROC_CAM); cam["nRows"], cam["nCols"] = height, width
        ob = Observability(**cam, max_batch=1, max_points=16, device=device)        # host-side kinematics only
        ob.set_state(SIMU_XV)
        ob.predictPWLSVec(0.1, 1)
        self.xv_h = ob.kinematic(0)["Xv"]
        ob.close()
        self.sample_scale = sample_scale_for(nfeatures, n_good, error_bound)
        self.cfg = FrontEndConfig(width, height, nfeatures, 8, 1.2, 20, 7, batch, window, nnratio, 1, n_good, self.sample_scale,
                                  cam["f"] / cam["dx"], cam["f"] / cam["dy"], cam["Cx"], cam["Cy"], cam["k1"], cam["k2"],
                                  cam["dx"], cam["dy"], 1 if overlap else 0, device)
        h = C.c_void_p()
        check(_lib.lib().gforb_frontend_create(C.byref(self.cfg), C.byref(h)), "gforb_frontend_create")
        self._h = h
        # synthetic matched map points of every keypoint slot (tools/simu_greedy.cc recipe, fixed per slot)
        world, _, score, _ = synth.landmark_cloud(nfeatures, cloud_seed, visible_from=simu_visible_from(width, height))
        self.world_h, self.score_h = np.ascontiguousarray(world, np.float32), np.ascontiguousarray(score, np.int32)
        self.sigma2_h = level_sigma2()
        xv = np.ascontiguousarray(self.xv_h, np.float64)
        check(_lib.lib().gforb_frontend_set_map(self._h, _lib.ptr(self.world_h), _lib.ptr(self.score_h), _lib.ptr(xv),
                                                _lib.ptr(self.sigma2_h)), "gforb_frontend_set_map")
        self.images = torch.zeros((batch, height, width), dtype=torch.uint8, device=self.dev)
        # pinned host staging of the end-to-end path: frames in, every result a Tracking caller consumes out
        pin = lambda *s, dt=torch.float32: torch.zeros(s, dtype=dt).pin_memory()
        self.h_images = pin(batch, height, width, dt=torch.uint8)
        self.h_kps = pin(batch, nfeatures, 7)
        self.h_desc = pin(batch, nfeatures, 32, dt=torch.uint8)
        self.h_counts = pin(batch, dt=torch.int32)
        self.h_m12 = pin(batch, nfeatures, dt=torch.int32)
        self.h_nmatch = pin(batch, dt=torch.int32)
        self.h_sel_idx = pin(batch, n_good, dt=torch.int32)
        self.h_sel_ld = pin(batch, n_good, dt=torch.float64)
        self.h_sel_n = pin(batch, dt=torch.int32)
        self.h_out = FrontEndOutputs(self.h_kps.data_ptr(), self.h_desc.data_ptr(), self.h_counts.data_ptr(), self.h_m12.data_ptr(),
                                     self.h_nmatch.data_ptr(), self.h_sel_idx.data_ptr(), self.h_sel_ld.data_ptr(),
                                     self.h_sel_n.data_ptr())
        self.launches_per_step = 0

    def close(self):
        if getattr(self, "_h", None):
            _lib.lib().gforb_frontend_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- steps (asynchronous; sync() joins and raises on any error a kernel recorded)
    def load_host_frames(self, host: np.ndarray):
        n = host.shape[0]
        self.h_images[:n].copy_(torch.from_numpy(host))
        self.images[:n].copy_(self.h_images[:n])
        torch.cuda.synchronize()

    def step_device(self, images: torch.Tensor | None = None, batch: int | None = None):
        """one step over frames resident in HBM"""
        img = self.images if images is None else images
        b = int(batch if batch is not None else img.shape[0])
        check(_lib.lib().gforb_frontend_step_device(self._h, C.c_void_p(img.data_ptr()), b), "gforb_frontend_step_device")
        self.launches_per_step = _lib.lib().gforb_frontend_launches_per_step(self._h)
        return self.launches_per_step

    def step_host(self, batch: int | None = None):
        """the same step through host buffers: H2D of self.h_images, D2H of every result into the pinned h_* arrays"""
        b = int(batch if batch is not None else self.B)
        check(_lib.lib().gforb_frontend_step_host(self._h, C.c_void_p(self.h_images.data_ptr()), b, C.byref(self.h_out)),
              "gforb_frontend_step_host")
        self.launches_per_step = _lib.lib().gforb_frontend_launches_per_step(self._h)

    def sync(self):
        check(_lib.lib().gforb_frontend_sync(self._h), "gforb_frontend_sync")

    # ---- device results of the most recently issued step (valid after sync())
    def _results(self):
        o = FrontEndOutputs()
        check(_lib.lib().gforb_frontend_results(self._h, C.byref(o)), "gforb_frontend_results")
        return o

    def _view(self, ptr, shape, typestr):
        return torch.as_tensor(_DevArray(ptr, shape, typestr), device=self.dev)

    @property
    def kps(self): return self._view(self._results().keypoints, (self.B, self.nf, 7), "<f4")
    @property
    def desc(self): return self._view(self._results().descriptors, (self.B, self.nf, 32), "|u1")
    @property
    def counts(self): return self._view(self._results().counts, (self.B,), "<i4")
    @property
    def m12(self): return self._view(self._results().matches12, (self.B, self.nf), "<i4")
    @property
    def nmatch(self): return self._view(self._results().nmatches, (self.B,), "<i4")
    @property
    def sel_idx(self): return self._view(self._results().good_idx, (self.B, self.k), "<i4")
    @property
    def sel_ld(self): return self._view(self._results().good_logdet, (self.B, self.k), "<f8")
    @property
    def sel_n(self): return self._view(self._results().good_n, (self.B,), "<i4")

    # ---- profiling (CUDA events inside the library, on each stage's own stream)
    def set_profiling(self, on: bool):
        check(_lib.lib().gforb_frontend_set_profiling(self._h, 1 if on else 0), "gforb_frontend_set_profiling")

    def _stages(self):
        L = _lib.lib()
        n = L.gforb_frontend_stage_time(self._h, -1, None, None)
        out = []
        for i in range(max(n, 0)):
            name = C.c_char_p(); ms = C.c_float()
            L.gforb_frontend_stage_time(self._h, i, C.byref(name), C.byref(ms))
            out.append((name.value.decode() if name.value else "?", float(ms.value)))
        return out

    def _split(self):
        st = self._stages()
        own = next((i for i, (n, _) in enumerate(st) if n == "pyramid"), len(st))
        return st[:own], st[own:]

    def stage_times(self):
        """[(stage, ms)]: extract / match / info+select of the pipeline (+ h2d and d2h after a profiled step_host)"""
        return [x for x in self._split()[0] if x[0]]

    def extract_stage_times(self):
        """[(stage, ms)]: the extractor's own stages (pyramid, detect, select, orient, blur, describe)"""
        return self._split()[1]

    # ---- sizes
    def pyramid_bytes(self) -> int:
        return 2 * self.B * sum((w + 32) * (h + 32) for w, h in level_sizes(self.w, self.h))

    def algorithmic_bytes(self):
        """SURVEY 8(d) materialised-stage model per frame: B = 5P - P0 - P7 + 60 N, split per extractor stage."""
        px = [w * h for w, h in level_sizes(self.w, self.h)]
        P = sum(px)
        per = {"pyramid": (P - px[-1]) + (P - px[0]), "detect": P, "select": 0, "orient": 0, "blur": 2 * P,
               "describe": 60 * self.nf}
        return {"per_stage": per, "extract_frame": sum(per.values()), "pixels": P}

    def h2d_bytes(self) -> int:
        return self.h_images.numel()

    def d2h_bytes(self) -> int:
        return sum(t.numel() * t.element_size() for t in (self.h_kps, self.h_desc, self.h_counts, self.h_m12, self.h_nmatch,
                                                           self.h_sel_idx, self.h_sel_ld, self.h_sel_n))
