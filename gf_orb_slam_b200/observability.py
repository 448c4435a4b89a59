"""Host-side mirror of ORB_SLAM::Observability (reference include/Observability.h:113-770) over the C ABI:
kinematics on the host, batched information-block build and log-det subset selection on the GPU."""
from __future__ import annotations

import ctypes as C
import math

import numpy as np

from . import _lib
from ._lib import ObsConfig, check, is_device, ptr

BASELINE_GREEDY, LAZIER_GREEDY, AUTOMATIC_LAZIER, DELETION_LAZIER = 1, 2, 3, 5
MAP_INFO_MATRIX, FRAME_INFO_MATRIX, MAP_HYBRID_MATRIX, FRAME_HYBRID_MATRIX = 0, 1, 2, 3   # include/Observability.h:94-102


def _like(ref, shape, dtype_np, dtype_t):
    if is_device(ref):
        import torch
        return torch.empty(shape, dtype=getattr(torch, dtype_t), device=ref.device)
    return np.empty(shape, dtype_np)


class Observability:
    """Observability(fu, fv, nRows, nCols, Cx, Cy, k1, k2) (Tracking's ctor) or, with f/dx/dy given,
    Observability(f, nRows, nCols, Cx, Cy, k1, k2, dx, dy) (tests/tools: fu = f/dx, fv = f/dy)."""

    def __init__(self, *, fu=None, fv=None, f=None, nRows=480, nCols=752, Cx=0.0, Cy=0.0, k1=0.0, k2=0.0, dx=None, dy=None,
                 max_batch=1, max_points=2048, device=0):
        if f is not None:
            fu, fv = f / dx, f / dy                       # include/Util.hpp:134-145
        else:
            dx = 0.0112; dy = fu * dx / fv                # include/Util.hpp:148-158
        self.cfg = ObsConfig(fu, fv, nRows, nCols, Cx, Cy, k1, k2, dx, dy, max_batch, max_points, device)
        h = C.c_void_p()
        check(_lib.lib().gforb_obs_create(C.byref(self.cfg), C.byref(h)), "gforb_obs_create")
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            _lib.lib().gforb_obs_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def last_launches(self) -> int:
        return _lib.lib().gforb_obs_last_launches(self._h)

    def set_select_residency(self, ctas_per_sm: int):
        """register budget of the lazier-greedy kernel: 2 or 3 resident CTAs per SM, 0 = from the batch size (scheduling only)"""
        check(_lib.lib().gforb_obs_set_select_residency(self._h, int(ctas_per_sm)), "gforb_obs_set_select_residency")

    # ---- kinematics (host)
    def set_state(self, Xv):
        Xv = np.ascontiguousarray(Xv, np.float64)
        check(_lib.lib().gforb_obs_set_state(self._h, ptr(Xv)), "gforb_obs_set_state")

    def get_state(self):
        Xv = np.empty(13, np.float64)
        check(_lib.lib().gforb_obs_get_state(self._h, ptr(Xv)), "gforb_obs_get_state")
        return Xv

    def updatePWLSVec(self, t_prev, Tcw_prev, t_cur, Twc_cur):
        a = np.ascontiguousarray(Tcw_prev, np.float32); b = np.ascontiguousarray(Twc_cur, np.float32)
        check(_lib.lib().gforb_obs_update_pwls(self._h, float(t_prev), ptr(a), float(t_cur), ptr(b)), "gforb_obs_update_pwls")

    def predictPWLSVec(self, dt, num_seg_pred):
        check(_lib.lib().gforb_obs_predict_pwls(self._h, float(dt), int(num_seg_pred)), "gforb_obs_predict_pwls")

    def kinematic(self, seg=0):
        Xv = np.empty(13); F = np.empty((13, 13)); FQ = np.empty((4, 4)); FO = np.empty((4, 3))
        FQi = np.empty((4, 4)); FOi = np.empty((4, 3)); T = np.empty((4, 4), np.float32)
        check(_lib.lib().gforb_obs_get_segment(self._h, seg, ptr(Xv), ptr(F), ptr(FQ), ptr(FO), ptr(FQi), ptr(FOi), ptr(T)),
              "gforb_obs_get_segment")
        return dict(Xv=Xv, F=F, F_Q=FQ, F_Omg=FO, F_Q_inSeg=FQi, F_Omg_inSeg=FOi, Tcw=T)

    # ---- runMatrixBuilding(FRAME_INFO_MATRIX): batch x stride points
    def build_info_frame(self, Xv, world, octave, match_score, valid, counts, level_sigma2, *, want_H=False,
                         blocks=None, obs_score=None, stream=None, sync=True):
        B, S = int(world.shape[0]), int(world.shape[1])
        dev = is_device(world)
        if blocks is None:
            blocks = _like(world, (B, S, 49), np.float64, "float64")
            obs_score = _like(world, (B, S), np.float64, "float64")
        H = _like(world, (B, S, 14), np.float64, "float64") if want_H else None
        flags = (_lib.GFORB_IN_DEVICE | _lib.GFORB_OUT_DEVICE if dev else 0) | (0 if sync else _lib.GFORB_ASYNC)
        nl = int(level_sigma2.shape[0])
        check(_lib.lib().gforb_obs_build_info_frame(self._h, ptr(Xv), ptr(world), ptr(octave), ptr(match_score), ptr(valid),
                                                    ptr(counts), S, B, ptr(level_sigma2), nl, ptr(blocks), ptr(obs_score),
                                                    ptr(H), flags, C.c_void_p(stream or 0)), "gforb_obs_build_info_frame")
        return (blocks, obs_score, H) if want_H else (blocks, obs_score)

    # ---- runMatrixBuilding(MAP_INFO_MATRIX)
    def build_info_map(self, Xv, world, track_in_view, counts, check_viz, bounds, *, want_H=False, stream=None, sync=True):
        B, S = int(world.shape[0]), int(world.shape[1])
        dev = is_device(world)
        blocks = _like(world, (B, S, 49), np.float64, "float64")
        obs_score = _like(world, (B, S), np.float64, "float64")
        uv = _like(world, (B, S, 2), np.float32, "float32")
        H = _like(world, (B, S, 14), np.float64, "float64") if want_H else None
        flags = (_lib.GFORB_IN_DEVICE | _lib.GFORB_OUT_DEVICE if dev else 0) | (0 if sync else _lib.GFORB_ASYNC)
        check(_lib.lib().gforb_obs_build_info_map(self._h, ptr(Xv), ptr(world), ptr(track_in_view), ptr(counts), S, B,
                                                  int(check_viz), *[int(v) for v in bounds], ptr(blocks), ptr(obs_score),
                                                  ptr(uv), ptr(H), flags, C.c_void_p(stream or 0)), "gforb_obs_build_info_map")
        return (blocks, obs_score, uv, H) if want_H else (blocks, obs_score, uv)

    def logDet(self, mats):
        dev = is_device(mats)
        n = int(mats.shape[0])
        out = _like(mats, (n,), np.float64, "float64")
        flags = _lib.GFORB_IN_DEVICE | _lib.GFORB_OUT_DEVICE if dev else 0
        check(_lib.lib().gforb_obs_logdet(self._h, ptr(mats), n, ptr(out), flags, None), "gforb_obs_logdet")
        return out

    def rand_stream(self, seed: int, n: int) -> np.ndarray:
        out = np.empty(n, np.int32)
        check(_lib.lib().gforb_obs_rand_stream(self._h, seed, n, ptr(out)), "gforb_obs_rand_stream")
        return out

    # ---- setSelction_Number's selection stage / maxVol*_LazierGreedy over batch x stride pools
    def select(self, blocks, obs_score, counts, k, mtd=LAZIER_GREEDY, sample_scale=6.0, seeds=None, *, out_idx=None,
               out_logdet=None, out_n=None, stream=None, sync=True):
        B, S = int(blocks.shape[0]), int(blocks.shape[1])
        dev = is_device(blocks)
        out_stride = S if mtd in (AUTOMATIC_LAZIER, DELETION_LAZIER) else max(int(k), 1)
        if out_idx is None:
            out_idx = _like(blocks, (B, out_stride), np.int32, "int32")
            out_logdet = _like(blocks, (B, out_stride), np.float64, "float64")
            out_n = _like(blocks, (B,), np.int32, "int32")
        else:
            out_stride = int(out_idx.shape[1])
        flags = (_lib.GFORB_IN_DEVICE | _lib.GFORB_OUT_DEVICE if dev else 0) | (0 if sync else _lib.GFORB_ASYNC)
        check(_lib.lib().gforb_obs_select(self._h, ptr(blocks), ptr(obs_score), ptr(counts), S, B, int(k), int(mtd),
                                          float(sample_scale), ptr(seeds), ptr(out_idx), ptr(out_logdet), ptr(out_n),
                                          out_stride, flags, C.c_void_p(stream or 0)), "gforb_obs_select")
        return out_idx, out_logdet, out_n

    def select_map_mtd3(self, blocks, obs_score, k, seed=1, num_threads=8):
        """Selection stage of setSelction_Number(k, 3, T, vector<MapPoint*>*, out) (src/Observability.cc:1085-1163) over ONE
        pool (n x 49 numpy blocks whose entries all have ObsScore >= 0): sample scale 6.0; pools of >= 2000 points that
        exceed 1.2 k by more than 10 are cut into min(round(n / 1000), num_threads) ranges selected in one batched call
        (range i draws from rand() stream seed + 1 + i), their picks are merged and the final k picked with stream `seed`."""
        n = int(blocks.shape[0])
        blocks = np.ascontiguousarray(blocks, np.float64).reshape(n, 49)
        sc = np.ones(n) if obs_score is None else np.ascontiguousarray(obs_score, np.float64)
        thread_needed = 1
        if not (n - 1.2 * float(np.float32(k)) <= 10 or n < 2000):        # (size_t - double * float) as written at :1086
            thread_needed = min(int(np.round(np.float32(n) / np.float32(1000))), int(num_threads))
        if thread_needed <= 1:
            idx, _, cnt = self.select(blocks[None], sc[None], np.array([n], np.int32), k, AUTOMATIC_LAZIER, 6.0, np.array([seed], np.uint32))
            return idx[0, :cnt[0]].copy()
        sub = int(np.ceil(float(np.float32(k) / np.float32(thread_needed)) * 1.2))       # float division, double product (:1110)
        per = int(np.ceil(float(np.float32(n) / np.float32(thread_needed))))
        pb = np.zeros((thread_needed, per, 49)); ps = np.full((thread_needed, per), -1.0); cnts = np.zeros(thread_needed, np.int32)
        for i in range(thread_needed):
            st, ed = i * per, min((i + 1) * per, n)
            pb[i, :ed - st] = blocks[st:ed]; ps[i, :ed - st] = sc[st:ed]; cnts[i] = max(ed - st, 0)
        seeds = (seed + 1 + np.arange(thread_needed)).astype(np.uint32)
        idx, _, cnt = self.select(pb, ps, cnts, sub, AUTOMATIC_LAZIER, 6.0, seeds)
        merged = np.concatenate([i * per + idx[i, :cnt[i]] for i in range(thread_needed)]).astype(np.int64)
        m = len(merged)
        idx2, _, cnt2 = self.select(blocks[merged][None], sc[merged][None], np.array([m], np.int32), k, AUTOMATIC_LAZIER, 6.0,
                                    np.array([seed], np.uint32))
        return merged[idx2[0, :cnt2[0]]].astype(np.int32)

    # ---- int runActiveMapMatching(Frame*, mat_type, mBaseInfoMat, th, ORBmatcher&, num_to_match, time_for_match)
    def runActiveMapMatching(self, F, mps, mat_type, base_info, th, matcher, num_to_match, *, obs_score, obs_mat, H_meas,
                             level_sigma2, frame_mappoints, match_score, seeds=None, stream=None, sync=True):
        """src/Observability.cc:1249-1524 for a batch of frames.  F: matcher.Frames, mps: matcher.MapPoints (mMapPoints
        flattened), obs_score / obs_mat / H_meas: outputs of build_info_map(want_H=True), base_info: (B, 49) or (B, 7, 7),
        frame_mappoints / match_score: (B, kp_stride) int32, updated in place.  Returns (nMatched[B], left_idx[B, stride],
        left_n[B], info[B, 49])."""
        B, S = mps.batch, mps.stride
        dev = is_device(obs_mat)
        nmatched = _like(obs_mat, (B,), np.int32, "int32")
        left = _like(obs_mat, (B, S), np.int32, "int32")
        left_n = _like(obs_mat, (B,), np.int32, "int32")
        info = _like(obs_mat, (B, 49), np.float64, "float64")
        fs, ms = F.cstruct(), mps.cstruct()
        flags = (_lib.GFORB_IN_DEVICE | _lib.GFORB_OUT_DEVICE if dev else 0) | (0 if sync else _lib.GFORB_ASYNC)
        check(_lib.lib().gforb_obs_active_match(self._h, matcher._h, C.byref(fs), C.byref(ms), B, ptr(obs_score), ptr(obs_mat),
                                                ptr(H_meas), ptr(base_info), ptr(level_sigma2), int(level_sigma2.shape[0]),
                                                int(mat_type), float(th), float(matcher.mfNNratio), int(num_to_match),
                                                ptr(seeds), ptr(frame_mappoints), ptr(match_score), ptr(nmatched), ptr(left),
                                                ptr(left_n), ptr(info), flags, C.c_void_p(stream or 0)),
              "gforb_obs_active_match")
        return nmatched, left, left_n, info

    @staticmethod
    def sample_scale_for(N_frame: int, k: int, error_bound: float) -> float:
        """random_sample_scale of setSelction_Number (src/Observability.cc:826):
        static_cast<size_t>(float(N)/float(k) * log(1.0/error_bound))."""
        return float(int(float(np.float32(N_frame) / np.float32(k)) * math.log(1.0 / error_bound)))
