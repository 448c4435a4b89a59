"""ctypes binding of libgforb.so (C ABI: include/gforb.h).  No fallback: if the CUDA library is missing or no
device is usable, the operators raise."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgforb.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])

GFORB_IN_DEVICE, GFORB_OUT_DEVICE, GFORB_ASYNC = 1, 2, 4
ERRORS = {-1: "GFORB_ERR_CUDA", -2: "GFORB_ERR_ARG", -3: "GFORB_ERR_CAPACITY", -4: "GFORB_ERR_GEOMETRY",
          -5: "GFORB_ERR_OVERFLOW"}


class GforbError(RuntimeError):
    def __init__(self, code: int, what: str):
        super().__init__(f"{what}: {ERRORS.get(code, code)}")
        self.code = code


class ExtractorConfig(C.Structure):
    _fields_ = [("nfeatures", C.c_int32), ("scaleFactor", C.c_float), ("nlevels", C.c_int32), ("scoreType", C.c_int32),
                ("iniThFAST", C.c_int32), ("minThFAST", C.c_int32), ("blur_variant", C.c_int32),
                ("retain_variant", C.c_int32), ("max_width", C.c_int32), ("max_height", C.c_int32),
                ("max_batch", C.c_int32), ("device", C.c_int32)]


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(there is no CPU fallback for the B200 path)")
        L = C.CDLL(LIB_PATH)
        L.gforb_version.restype = C.c_char_p
        L.gforb_extractor_scale_factor.restype = C.c_float
        vp, i32, sz = C.c_void_p, C.c_int, C.c_size_t
        L.gforb_extractor_create.argtypes = [C.POINTER(ExtractorConfig), C.POINTER(vp)]
        L.gforb_extractor_destroy.argtypes = [vp]
        L.gforb_extract.argtypes = [vp, vp, i32, i32, i32, sz, i32, vp, vp, vp, i32, vp]
        L.gforb_extractor_levels.argtypes = [vp]
        L.gforb_extractor_scale_factor.argtypes = [vp]
        L.gforb_extractor_tables.argtypes = [vp, vp, vp]
        L.gforb_extractor_debug_level.argtypes = [vp, i32, i32, i32, vp, sz, vp, vp, vp]
        L.gforb_extractor_last_launches.argtypes = [vp]
        L.gforb_extractor_set_profiling.argtypes = [vp, i32]
        L.gforb_extractor_stage_time.argtypes = [vp, i32, C.POINTER(C.c_char_p), C.POINTER(C.c_float)]
        _lib = L
    return _lib


def check(rc: int, what: str):
    if rc != 0:
        raise GforbError(rc, what)


def ptr(a):
    """device or host pointer of a numpy array / torch tensor / None"""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(C.c_void_p)
    return C.c_void_p(a.data_ptr())          # torch tensor


def is_device(a) -> bool:
    return a is not None and not isinstance(a, np.ndarray) and getattr(a, "is_cuda", False)
