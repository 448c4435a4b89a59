"""ctypes binding of libgforb.so (C ABI: include/gforb.h).  No fallback: if the CUDA library is missing or no
device is usable, the operators raise."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GFORB_LIB") or os.path.join(_HERE, "libgforb.so")     # GFORB_LIB: A/B builds of the same library

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])

GFORB_IN_DEVICE, GFORB_OUT_DEVICE, GFORB_ASYNC = 1, 2, 4
ERRORS = {-1: "GFORB_ERR_CUDA", -2: "GFORB_ERR_ARG", -3: "GFORB_ERR_CAPACITY", -4: "GFORB_ERR_GEOMETRY",
          -5: "GFORB_ERR_OVERFLOW"}


class GforbError(RuntimeError):
    def __init__(self, code: int, what: str):
        super().__init__(f"{what}: {ERRORS.get(code, code)}")
        self.code = code


class FrameSet(C.Structure):
    _fields_ = [("keypoints", C.c_void_p), ("descriptors", C.c_void_p), ("counts", C.c_void_p), ("kp_stride", C.c_int32),
                ("min_x", C.c_int32), ("max_x", C.c_int32), ("min_y", C.c_int32), ("max_y", C.c_int32)]


class MapPointSet(C.Structure):
    _fields_ = [("proj_x", C.c_void_p), ("proj_y", C.c_void_p), ("scale_level", C.c_void_p), ("view_cos", C.c_void_p),
                ("flags", C.c_void_p), ("descriptors", C.c_void_p), ("counts", C.c_void_p), ("mp_stride", C.c_int32)]


class WorldPointSet(C.Structure):
    _fields_ = [("world", C.c_void_p), ("normal", C.c_void_p), ("min_distance", C.c_void_p), ("max_distance", C.c_void_p),
                ("descriptors", C.c_void_p), ("skip", C.c_void_p), ("counts", C.c_void_p), ("mp_stride", C.c_int32)]


class MatcherConfig(C.Structure):
    _fields_ = [("max_batch", C.c_int32), ("max_keypoints", C.c_int32), ("max_queries", C.c_int32),
                ("cand_pool", C.c_int32), ("nlevels", C.c_int32), ("scaleFactor", C.c_float), ("device", C.c_int32)]


class ObsConfig(C.Structure):
    _fields_ = [("fu", C.c_double), ("fv", C.c_double), ("nRows", C.c_int32), ("nCols", C.c_int32), ("Cx", C.c_double),
                ("Cy", C.c_double), ("k1", C.c_double), ("k2", C.c_double), ("dx", C.c_double), ("dy", C.c_double),
                ("max_batch", C.c_int32), ("max_points", C.c_int32), ("device", C.c_int32)]


class ExtractorConfig(C.Structure):
    _fields_ = [("nfeatures", C.c_int32), ("scaleFactor", C.c_float), ("nlevels", C.c_int32), ("scoreType", C.c_int32),
                ("iniThFAST", C.c_int32), ("minThFAST", C.c_int32), ("blur_variant", C.c_int32),
                ("retain_variant", C.c_int32), ("max_width", C.c_int32), ("max_height", C.c_int32),
                ("max_batch", C.c_int32), ("device", C.c_int32)]


class FrontEndConfig(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("nfeatures", C.c_int32), ("nlevels", C.c_int32),
                ("scaleFactor", C.c_float), ("iniThFAST", C.c_int32), ("minThFAST", C.c_int32), ("max_batch", C.c_int32),
                ("window_size", C.c_int32), ("nnratio", C.c_float), ("check_orientation", C.c_int32), ("n_good", C.c_int32),
                ("sample_scale", C.c_double), ("fu", C.c_double), ("fv", C.c_double), ("Cx", C.c_double), ("Cy", C.c_double),
                ("k1", C.c_double), ("k2", C.c_double), ("dx", C.c_double), ("dy", C.c_double), ("overlap", C.c_int32),
                ("device", C.c_int32)]


class FrontEndOutputs(C.Structure):
    _fields_ = [("keypoints", C.c_void_p), ("descriptors", C.c_void_p), ("counts", C.c_void_p), ("matches12", C.c_void_p),
                ("nmatches", C.c_void_p), ("good_idx", C.c_void_p), ("good_logdet", C.c_void_p), ("good_n", C.c_void_p)]


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
        L.gforb_extractor_set_concurrency.argtypes = [vp, i32]
        L.gforb_extractor_set_distribution.argtypes = [vp, i32]
        L.gforb_extractor_stage_time.argtypes = [vp, i32, C.POINTER(C.c_char_p), C.POINTER(C.c_float)]
        f32, f64, i64, u32 = C.c_float, C.c_double, C.c_int64, C.c_uint32
        FS, MS = C.POINTER(FrameSet), C.POINTER(MapPointSet)
        L.gforb_matcher_create.argtypes = [C.POINTER(MatcherConfig), C.POINTER(vp)]
        L.gforb_matcher_destroy.argtypes = [vp]
        L.gforb_matcher_last_launches.argtypes = [vp]
        L.gforb_descriptor_distance.argtypes = [vp, vp, vp, i32, vp, i32, vp]
        L.gforb_features_in_area.argtypes = [vp, FS, i32, i32, vp, vp, vp, vp, vp, i32, vp, i32, vp, i32, vp]
        L.gforb_search_for_initialization.argtypes = [vp, FS, FS, i32, vp, vp, vp, i32, f32, i32, i32, vp]
        L.gforb_search_by_projection.argtypes = [vp, FS, MS, i32, f32, f32, vp, vp, vp, i32, vp]
        L.gforb_search_by_projection_frames.argtypes = [vp, FS, FS, i32, vp, vp, vp, vp, f32, f32, f32, f32, f32, f32,
                                                        i32, vp, vp, vp, i32, vp]
        L.gforb_search_by_projection_good_feature.argtypes = [vp, FS, MS, i32, f32, f32, f32, vp, vp, i32, vp]
        L.gforb_search_by_projection_frames_good_feature.argtypes = [vp, FS, FS, i32, vp, vp, vp, vp, vp, f32, f32, f32, f32, f32,
                                                                     f32, f32, i32, vp, vp, i32, vp]
        L.gforb_mappoint_distinctive_descriptors.argtypes = [vp, i32, vp, vp, vp, vp, i32, vp]
        L.gforb_frame_is_in_frustum.argtypes = [vp, i32, vp, vp, vp, f32, f32, f32, f32, i32, i32, i32, i32, vp, vp, vp, vp, vp, i32,
                                                f32, vp, vp, vp, vp, vp, i32, vp]
        L.gforb_frame_undistort_keypoints.argtypes = [vp, i32, vp, vp, i32, f32, f32, f32, f32, vp, i32, vp, i32, vp]
        L.gforb_window_search.argtypes = [vp, FS, FS, i32, vp, i32, i32, i32, f32, i32, vp, vp, i32, vp]
        L.gforb_search_by_projection_window.argtypes = [vp, FS, FS, i32, vp, vp, vp, f32, f32, f32, f32, i32, f32, vp, vp, vp, i32, vp]
        L.gforb_search_by_bow.argtypes = [vp, FS, vp, vp, FS, vp, i32, f32, i32, vp, vp, i32, vp]
        L.gforb_search_by_projection_keyframe.argtypes = [vp, FS, FS, i32, vp, vp, vp, vp, vp, vp, f32, f32, f32, f32, f32, i32, i32, vp, vp, vp, i32, vp]
        WS = C.POINTER(WorldPointSet)
        L.gforb_search_by_bow_keyframes.argtypes = [vp, FS, vp, vp, FS, vp, vp, i32, f32, i32, vp, vp, i32, vp]
        L.gforb_search_for_triangulation.argtypes = [vp, FS, vp, vp, FS, vp, vp, i32, vp, i32, vp, vp, i32, vp]
        L.gforb_fuse.argtypes = [vp, FS, i32, vp, i32, f32, f32, f32, f32, WS, f32, vp, vp, vp, i32, vp]
        L.gforb_search_by_projection_sim3.argtypes = [vp, FS, i32, vp, f32, f32, f32, f32, WS, vp, f32, vp, vp, i32, vp]
        L.gforb_search_by_sim3.argtypes = [vp, FS, FS, i32, WS, WS, vp, vp, vp, vp, vp, vp, f32, f32, f32, f32, f32, vp, vp, i32, vp]
        L.gforb_allpairs_top2.argtypes = [vp, vp, i32, vp, i64, i64, vp, i32, vp]
        L.gforb_top2_merge.argtypes = [vp, vp, i32, i32, vp, i32, vp]
        L.gforb_allpairs_top2_static.argtypes = [i32, vp, i32, vp, i64, i64, vp, i32, vp]
        L.gforb_matcher_poll_error.argtypes = [vp, vp]
        L.gforb_poseopt_create.argtypes = [i32, i32, i32, C.POINTER(vp)]
        L.gforb_poseopt_destroy.argtypes = [vp]
        L.gforb_pose_optimization.argtypes = [vp, vp, vp, vp, vp, vp, i32, i32, vp, i32, f32, f32, f32, f32, vp, vp, i32, vp]
        L.gforb_extractor_poll_error.argtypes = [vp, vp]
        L.gforb_obs_poll_error.argtypes = [vp, vp]
        FO = C.POINTER(FrontEndOutputs)
        L.gforb_frontend_create.argtypes = [C.POINTER(FrontEndConfig), C.POINTER(vp)]
        L.gforb_frontend_destroy.argtypes = [vp]
        L.gforb_frontend_set_map.argtypes = [vp, vp, vp, vp, vp]
        L.gforb_frontend_step_device.argtypes = [vp, vp, i32]
        L.gforb_frontend_step_host.argtypes = [vp, vp, i32, FO]
        L.gforb_frontend_sync.argtypes = [vp]
        L.gforb_frontend_wait_results.argtypes = [vp, i32]
        L.gforb_frontend_results.argtypes = [vp, FO]
        L.gforb_frontend_launches_per_step.argtypes = [vp]
        L.gforb_frontend_set_profiling.argtypes = [vp, i32]
        L.gforb_frontend_stage_time.argtypes = [vp, i32, C.POINTER(C.c_char_p), C.POINTER(C.c_float)]
        L.gforb_obs_create.argtypes = [C.POINTER(ObsConfig), C.POINTER(vp)]
        L.gforb_obs_destroy.argtypes = [vp]
        L.gforb_obs_last_launches.argtypes = [vp]
        L.gforb_obs_set_select_residency.argtypes = [vp, i32]
        L.gforb_obs_set_state.argtypes = [vp, vp]
        L.gforb_obs_get_state.argtypes = [vp, vp]
        L.gforb_obs_update_pwls.argtypes = [vp, f64, vp, f64, vp]
        L.gforb_obs_predict_pwls.argtypes = [vp, f64, i32]
        L.gforb_obs_num_segments.argtypes = [vp]
        L.gforb_obs_get_segment.argtypes = [vp, i32, vp, vp, vp, vp, vp, vp, vp]
        L.gforb_obs_build_info_frame.argtypes = [vp, vp, vp, vp, vp, vp, vp, i32, i32, vp, i32, vp, vp, vp, i32, vp]
        L.gforb_obs_build_info_map.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp, vp, vp, vp, i32, vp]
        L.gforb_obs_logdet.argtypes = [vp, vp, i32, vp, i32, vp]
        L.gforb_obs_rand_stream.argtypes = [vp, u32, i32, vp]
        L.gforb_obs_select.argtypes = [vp, vp, vp, vp, i32, i32, i32, i32, f64, vp, vp, vp, vp, i32, i32, vp]
        L.gforb_obs_active_match.argtypes = [vp, vp, FS, MS, i32, vp, vp, vp, vp, vp, i32, i32, f32, f32, i32, vp, vp, vp, vp,
                                             vp, vp, vp, i32, vp]
        _lib = L
    return _lib

# every symbol include/gforb.h declares (checked by tests/test_abi.py against the header and the built library)
EXPORTS = """gforb_version gforb_device_count gforb_extractor_create gforb_extractor_destroy gforb_extract
gforb_extractor_levels gforb_extractor_scale_factor gforb_extractor_tables gforb_extractor_debug_level
gforb_extractor_last_launches gforb_extractor_set_profiling gforb_extractor_set_concurrency gforb_extractor_set_distribution gforb_extractor_stage_time gforb_matcher_create
gforb_matcher_destroy gforb_matcher_last_launches gforb_descriptor_distance gforb_features_in_area
gforb_search_for_initialization gforb_search_by_projection gforb_search_by_projection_frames gforb_window_search gforb_search_by_projection_window gforb_search_by_bow gforb_search_by_projection_keyframe
gforb_search_by_bow_keyframes gforb_search_for_triangulation gforb_fuse gforb_search_by_projection_sim3 gforb_search_by_sim3
gforb_allpairs_top2 gforb_top2_merge gforb_allpairs_top2_static gforb_obs_create gforb_obs_destroy gforb_obs_last_launches gforb_obs_set_select_residency gforb_obs_set_state
gforb_obs_get_state gforb_obs_update_pwls gforb_obs_predict_pwls gforb_obs_num_segments gforb_obs_get_segment
gforb_obs_build_info_frame gforb_obs_build_info_map gforb_obs_logdet gforb_obs_rand_stream gforb_obs_select
gforb_obs_active_match gforb_frame_is_in_frustum gforb_frame_undistort_keypoints
gforb_search_by_projection_good_feature gforb_search_by_projection_frames_good_feature
gforb_mappoint_distinctive_descriptors gforb_matcher_poll_error gforb_extractor_poll_error gforb_obs_poll_error
gforb_poseopt_create gforb_poseopt_destroy gforb_pose_optimization
gforb_frontend_create gforb_frontend_destroy gforb_frontend_set_map gforb_frontend_step_device gforb_frontend_step_host
gforb_frontend_sync gforb_frontend_wait_results gforb_frontend_results gforb_frontend_launches_per_step gforb_frontend_set_profiling
gforb_frontend_stage_time""".split()


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
