"""Host-side mirror of ORB_SLAM::ORBextractor (reference include/ORBextractor.h:51-104) over the C ABI."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import KP_DTYPE, ExtractorConfig, check, ptr

HARRIS_SCORE, FAST_SCORE = 0, 1


class ORBextractor:
    """ORBextractor(nfeatures=1000, scaleFactor=1.2f, nlevels=8, scoreType=FAST_SCORE, fastTh=20).

    `ex(image)` == `operator()(image, mask, keypoints, descriptors)` for one frame; `ex.extract_batch(images)` runs a
    batch of equally sized frames.  iniThFAST/minThFAST are the north-star aliases of fastTh / the hard-coded 7."""

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, scoreType=FAST_SCORE, fastTh=20, *, iniThFAST=None,
                 minThFAST=7, blur_variant=0, retain_variant=0, max_width=752, max_height=480, max_batch=1, device=0):
        cfg = ExtractorConfig(nfeatures, scaleFactor, nlevels, scoreType, fastTh if iniThFAST is None else iniThFAST,
                              minThFAST, blur_variant, retain_variant, max_width, max_height, max_batch, device)
        self.cfg = cfg
        self.nfeatures, self.nlevels, self.max_batch = nfeatures, nlevels, max_batch
        h = C.c_void_p()
        check(_lib.lib().gforb_extractor_create(C.byref(cfg), C.byref(h)), "gforb_extractor_create")
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            _lib.lib().gforb_extractor_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:      # interpreter shutdown
            pass

    def GetLevels(self) -> int:
        return _lib.lib().gforb_extractor_levels(self._h)

    def GetScaleFactor(self) -> float:
        return _lib.lib().gforb_extractor_scale_factor(self._h)

    def tables(self):
        sf = np.zeros(self.nlevels, np.float32); fpl = np.zeros(self.nlevels, np.int32)
        check(_lib.lib().gforb_extractor_tables(self._h, ptr(sf), ptr(fpl)), "gforb_extractor_tables")
        return sf, fpl

    # -- single frame, host in / host out (the reference call shape)
    def __call__(self, image: np.ndarray, mask=None):
        if image is None or image.size == 0:
            return np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)       # silent return on empty image
        assert image.dtype == np.uint8 and image.ndim == 2, "image.type() == CV_8UC1"
        k, d, n = self.extract_batch(image[None])
        return k[0, :n[0]].copy(), d[0, :n[0]].copy()

    def extract_batch(self, images: np.ndarray):
        images = np.ascontiguousarray(images, np.uint8)
        B, h, w = images.shape
        kps = np.zeros((B, self.nfeatures), KP_DTYPE)
        desc = np.zeros((B, self.nfeatures, 32), np.uint8)
        counts = np.zeros(B, np.int32)
        check(_lib.lib().gforb_extract(self._h, ptr(images), w, h, w, w * h, B, ptr(kps), ptr(desc), ptr(counts), 0, None),
              "gforb_extract")
        return kps, desc, counts

    # -- device in / device out (torch tensors own the HBM buffers; no copies, optional async on `stream`)
    def extract_device(self, images, kps, desc, counts, stream=None, sync=True):
        B, h, w = images.shape
        flags = _lib.GFORB_IN_DEVICE | _lib.GFORB_OUT_DEVICE | (0 if sync else _lib.GFORB_ASYNC)
        check(_lib.lib().gforb_extract(self._h, ptr(images), w, h, images.stride(1), images.stride(0), B, ptr(kps),
                                       ptr(desc), ptr(counts), flags, C.c_void_p(stream or 0)),
              "gforb_extract")

    def debug_level(self, b: int, level: int, which: int = 0) -> np.ndarray:
        w = C.c_int(); h = C.c_int(); p = C.c_int()
        L = _lib.lib()
        check(L.gforb_extractor_debug_level(self._h, b, level, which, None, 0, C.byref(w), C.byref(h), C.byref(p)), "debug_level")
        buf = np.zeros((h.value + 32, p.value), np.uint8)
        check(L.gforb_extractor_debug_level(self._h, b, level, which, ptr(buf), buf.size, C.byref(w), C.byref(h), C.byref(p)), "debug_level")
        return buf[:, :w.value + 32]

    def last_launches(self) -> int:
        return _lib.lib().gforb_extractor_last_launches(self._h)

    def set_distribution(self, mode: int):
        """0: the reference's cell-quota keypoint distribution (default); 1: ORB-SLAM2's quadtree (non-default mode)"""
        check(_lib.lib().gforb_extractor_set_distribution(self._h, int(mode)), "gforb_extractor_set_distribution")

    def set_profiling(self, on=True):
        _lib.lib().gforb_extractor_set_profiling(self._h, int(on))

    def stage_times(self):
        L = _lib.lib()
        n = L.gforb_extractor_stage_time(self._h, -1, None, None)
        out = []
        for i in range(n):
            name = C.c_char_p(); ms = C.c_float()
            L.gforb_extractor_stage_time(self._h, i, C.byref(name), C.byref(ms))
            out.append((name.value.decode(), ms.value))
        return out
