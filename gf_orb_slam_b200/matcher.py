"""Host-side mirror of ORB_SLAM::ORBmatcher and of the Frame fields it reads (reference include/ORBmatcher.h:39-294,
include/Frame.h) over the C ABI.  Arrays may be numpy (host; the library stages them) or torch CUDA tensors (device;
zero copies).  All calls are batch-first: a `Frames` object holds B frames."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import KP_DTYPE, FrameSet, MapPointSet, MatcherConfig, WorldPointSet, check, is_device, ptr

TH_LOW, TH_HIGH, HISTO_LENGTH = 50, 100, 30      # src/ORBmatcher.cc:40-42


def _raw(a):
    p = ptr(a)
    return p.value if p is not None else None


class Frames:
    """B frames' mvKeysUn / mDescriptors / N plus the static image bounds (Frame::mnMinX.. , src/Frame.cc:489-493).
    keypoints: (B, stride) KP_DTYPE (numpy) or (B, stride, 7) float32 torch tensor with the same 28-byte layout."""

    def __init__(self, keypoints, descriptors, counts, width: int, height: int, min_x=0, min_y=0):
        self.keypoints, self.descriptors, self.counts = keypoints, descriptors, counts
        self.batch = int(keypoints.shape[0])
        self.stride = int(keypoints.shape[1])
        self.bounds = (int(min_x), int(width), int(min_y), int(height))
        self.device = is_device(keypoints)
        assert is_device(descriptors) == self.device and is_device(counts) == self.device

    @classmethod
    def single(cls, kps: np.ndarray, desc: np.ndarray, width: int, height: int):
        kps = np.ascontiguousarray(kps, KP_DTYPE)
        n = max(kps.size, 1)
        k = np.zeros((1, n), KP_DTYPE); k[0, :kps.size] = kps
        d = np.zeros((1, n, 32), np.uint8); d[0, :kps.size] = desc
        return cls(k, d, np.array([kps.size], np.int32), width, height)

    def cstruct(self) -> FrameSet:
        b = self.bounds
        return FrameSet(_raw(self.keypoints), _raw(self.descriptors), _raw(self.counts), self.stride, b[0], b[1], b[2], b[3])


class MapPoints:
    """SoA of the MapPoint fields SearchByProjection reads; arrays (B, stride); flags bit0 mbTrackInView, bit1 isBad."""

    def __init__(self, proj_x, proj_y, scale_level, view_cos, flags, descriptors, counts):
        self.a = (proj_x, proj_y, scale_level, view_cos, flags, descriptors, counts)
        self.batch, self.stride = int(proj_x.shape[0]), int(proj_x.shape[1])
        self.device = is_device(proj_x)

    def cstruct(self) -> MapPointSet:
        return MapPointSet(*[_raw(x) for x in self.a], self.stride)


class WorldPoints:
    """SoA of the MapPoint fields the mapping / loop-closing overloads read (GforbWorldPointSet); arrays (B, stride[, 3 | 32]).
    skip non-zero = the reference `continue`s before projecting (NULL, isBad(), already found / in the keyframe)."""

    def __init__(self, world, normal, min_distance, max_distance, descriptors, skip, counts):
        self.a = (world, normal, min_distance, max_distance, descriptors, skip, counts)
        self.batch, self.stride = int(world.shape[0]), int(world.shape[1])
        self.device = is_device(world)

    def cstruct(self) -> WorldPointSet:
        return WorldPointSet(*[_raw(x) for x in self.a], self.stride)


class ORBmatcher:
    """ORBmatcher(nnratio=0.6, checkOri=true).  One instance owns the device scratch for up to max_batch problems."""

    def __init__(self, nnratio=0.6, checkOri=True, *, max_batch=1, max_keypoints=2048, max_queries=2048, cand_pool=0,
                 nlevels=8, scaleFactor=1.2, device=0):
        self.mfNNratio, self.mbCheckOrientation = float(nnratio), bool(checkOri)
        cfg = MatcherConfig(max_batch, max_keypoints, max_queries, cand_pool, nlevels, scaleFactor, device)
        self.cfg = cfg
        h = C.c_void_p()
        check(_lib.lib().gforb_matcher_create(C.byref(cfg), C.byref(h)), "gforb_matcher_create")
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            _lib.lib().gforb_matcher_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def last_launches(self) -> int:
        return _lib.lib().gforb_matcher_last_launches(self._h)

    @staticmethod
    def _flags(in_dev: bool, out_dev: bool, sync=True) -> int:
        return (_lib.GFORB_IN_DEVICE if in_dev else 0) | (_lib.GFORB_OUT_DEVICE if out_dev else 0) | (0 if sync else _lib.GFORB_ASYNC)

    # -- static int DescriptorDistance(a, b): here for n pairs
    def DescriptorDistance(self, a, b):
        dev = is_device(a)
        if dev:
            import torch
            n = a.numel() // 32
            out = torch.empty(n, dtype=torch.int32, device=a.device)
        else:
            a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
            n = a.size // 32
            out = np.empty(n, np.int32)
        check(_lib.lib().gforb_descriptor_distance(self._h, ptr(a), ptr(b), n, ptr(out), self._flags(dev, dev), None),
              "gforb_descriptor_distance")
        return out if (dev or n != 1) else int(out[0])

    # -- Frame::GetFeaturesInArea (parity tap)
    def GetFeaturesInArea(self, F: Frames, frame: int, x, y, r, minLevel, maxLevel, cap=512):
        x = np.ascontiguousarray(np.atleast_1d(x), np.float32); y = np.ascontiguousarray(np.atleast_1d(y), np.float32)
        r = np.ascontiguousarray(np.atleast_1d(r), np.float32)
        mn = np.ascontiguousarray(np.atleast_1d(minLevel), np.int32); mx = np.ascontiguousarray(np.atleast_1d(maxLevel), np.int32)
        n = x.size
        idx = np.full((n, cap), -1, np.int32); cnt = np.zeros(n, np.int32)
        fs = F.cstruct()
        check(_lib.lib().gforb_features_in_area(self._h, C.byref(fs), F.batch, frame, ptr(x), ptr(y), ptr(r), ptr(mn), ptr(mx),
                                                n, ptr(idx), cap, ptr(cnt), self._flags(F.device, False), None),
              "gforb_features_in_area")
        return [idx[i, :min(cnt[i], cap)].copy() for i in range(n)], cnt

    # -- bool Frame::isInFrustum(MapPoint*, viewingCosLimit), batched over every map point of B frames
    def isInFrustum(self, Rcw, tcw, Ow, fx, fy, cx, cy, bounds, world, normal, min_distance, max_distance, counts,
                    viewingCosLimit, mps: "MapPoints", *, stream=None, sync=True):
        """src/Frame.cc:166-227.  Rcw (B, 9), tcw / Ow (B, 3), world / normal (B, stride, 3), min/max_distance (B, stride);
        bounds = (mnMinX, mnMaxX, mnMinY, mnMaxY).  Fills mps.proj_x / proj_y / scale_level / view_cos and bit0 of
        mps.flags in place."""
        px, py, lvl, vc, fl = mps.a[0], mps.a[1], mps.a[2], mps.a[3], mps.a[4]
        dev = is_device(world)
        check(_lib.lib().gforb_frame_is_in_frustum(self._h, mps.batch, ptr(Rcw), ptr(tcw), ptr(Ow), float(fx), float(fy),
                                                   float(cx), float(cy), *[int(b) for b in bounds], ptr(world), ptr(normal),
                                                   ptr(min_distance), ptr(max_distance), ptr(counts), mps.stride,
                                                   float(viewingCosLimit), ptr(px), ptr(py), ptr(lvl), ptr(vc), ptr(fl),
                                                   self._flags(dev, is_device(px), sync), C.c_void_p(stream or 0)),
              "gforb_frame_is_in_frustum")
        return fl

    # -- void MapPoint::ComputeDistinctiveDescriptors(), batched over map points
    def ComputeDistinctiveDescriptors(self, descriptors, offsets, out_desc=None, *, stream=None, sync=True):
        """src/MapPoint.cc:196-262.  descriptors (total, 32) uint8 = the observation descriptors of all points concatenated,
        offsets (n_points + 1) int32.  Returns (best row per point or -1, the representative descriptors (n_points, 32))."""
        dev = is_device(descriptors)
        n = int(offsets.shape[0]) - 1
        if dev:
            import torch
            best = torch.empty(n, dtype=torch.int32, device=descriptors.device)
            out = torch.zeros((n, 32), dtype=torch.uint8, device=descriptors.device) if out_desc is None else out_desc
        else:
            best = np.empty(n, np.int32)
            out = np.zeros((n, 32), np.uint8) if out_desc is None else out_desc
        check(_lib.lib().gforb_mappoint_distinctive_descriptors(self._h, n, ptr(descriptors), ptr(offsets), ptr(best), ptr(out),
                                                                self._flags(dev, dev, sync), C.c_void_p(stream or 0)),
              "gforb_mappoint_distinctive_descriptors")
        return best, out

    # -- void Frame::UndistortKeyPoints(), batched over the keypoints of B frames
    def UndistortKeyPoints(self, keys, counts, fx, fy, cx, cy, dist_coef, keys_un=None, *, stream=None, sync=True):
        """src/Frame.cc:389-423.  keys (B, stride) KP_DTYPE records (numpy) or (B, stride, 7) float32 (torch, device);
        dist_coef = Frame::mDistCoef (k1, k2, p1, p2[, k3]) on the host.  Returns mvKeysUn (same shape; entries past
        counts[b] are left as allocated)."""
        dev = is_device(keys)
        B, stride = int(keys.shape[0]), int(keys.shape[1])
        if keys_un is None:
            if dev:
                import torch
                keys_un = torch.zeros_like(keys)
            else:
                keys_un = np.zeros_like(keys)
        d = np.ascontiguousarray(dist_coef, np.float32).reshape(-1)
        check(_lib.lib().gforb_frame_undistort_keypoints(self._h, B, ptr(keys), ptr(counts), stride, float(fx), float(fy),
                                                         float(cx), float(cy), ptr(d), int(d.size), ptr(keys_un),
                                                         self._flags(dev, is_device(keys_un), sync), C.c_void_p(stream or 0)),
              "gforb_frame_undistort_keypoints")
        return keys_un

    # -- int SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize=10)
    def SearchForInitialization(self, F1: Frames, F2: Frames, vbPrevMatched, windowSize=10, *, matches12=None,
                                nmatches=None, stream=None, sync=True):
        """vbPrevMatched (B, stride1, 2) float32 is updated in place; returns (nmatches[B], vnMatches12[B, stride1])."""
        out_dev = is_device(vbPrevMatched)
        if matches12 is None:
            if out_dev:
                import torch
                matches12 = torch.empty((F1.batch, F1.stride), dtype=torch.int32, device=vbPrevMatched.device)
                nmatches = torch.empty(F1.batch, dtype=torch.int32, device=vbPrevMatched.device)
            else:
                matches12 = np.empty((F1.batch, F1.stride), np.int32); nmatches = np.empty(F1.batch, np.int32)
        f1, f2 = F1.cstruct(), F2.cstruct()
        check(_lib.lib().gforb_search_for_initialization(
            self._h, C.byref(f1), C.byref(f2), F1.batch, ptr(vbPrevMatched), ptr(matches12), ptr(nmatches), int(windowSize),
            self.mfNNratio, int(self.mbCheckOrientation), self._flags(F1.device, out_dev, sync), C.c_void_p(stream or 0)),
            "gforb_search_for_initialization")
        return nmatches, matches12

    # -- int SearchByProjection(Frame &F, const vector<MapPoint*> &vpMapPoints, th=3)
    def SearchByProjection(self, F: Frames, mps: MapPoints, th, mvpMapPoints, mvpMatchScore, *, nmatches=None,
                           stream=None, sync=True):
        """mvpMapPoints / mvpMatchScore (B, stride) int32 are the frame's in/out arrays (ids, -1 = NULL; 999 default)."""
        out_dev = is_device(mvpMapPoints)
        if nmatches is None:
            if out_dev:
                import torch
                nmatches = torch.empty(F.batch, dtype=torch.int32, device=mvpMapPoints.device)
            else:
                nmatches = np.empty(F.batch, np.int32)
        fs, ms = F.cstruct(), mps.cstruct()
        check(_lib.lib().gforb_search_by_projection(self._h, C.byref(fs), C.byref(ms), F.batch, float(th), self.mfNNratio,
                                                    ptr(mvpMapPoints), ptr(mvpMatchScore), ptr(nmatches),
                                                    self._flags(F.device, out_dev, sync), C.c_void_p(stream or 0)),
              "gforb_search_by_projection")
        return nmatches

    # -- int SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, float th)
    def SearchByProjectionFrames(self, Cur: Frames, Last: Frames, last_mappoints, last_outlier, last_world, Tcw, K, th,
                                 cur_mappoints, cur_score, *, nmatches=None, stream=None, sync=True):
        out_dev = is_device(cur_mappoints)
        if nmatches is None:
            if out_dev:
                import torch
                nmatches = torch.empty(Cur.batch, dtype=torch.int32, device=cur_mappoints.device)
            else:
                nmatches = np.empty(Cur.batch, np.int32)
        fx, fy, cx, cy = [float(v) for v in K]
        c, l = Cur.cstruct(), Last.cstruct()
        check(_lib.lib().gforb_search_by_projection_frames(
            self._h, C.byref(c), C.byref(l), Cur.batch, ptr(last_mappoints), ptr(last_outlier), ptr(last_world), ptr(Tcw),
            fx, fy, cx, cy, float(th), self.mfNNratio, int(self.mbCheckOrientation), ptr(cur_mappoints), ptr(cur_score),
            ptr(nmatches), self._flags(Cur.device, out_dev, sync), C.c_void_p(stream or 0)),
            "gforb_search_by_projection_frames")
        return nmatches

    # -- int SearchByProjection_GoodFeature(Frame &F, const vector<MapPoint*>&, th_good, th_rest)
    def SearchByProjection_GoodFeature(self, F: Frames, mps: MapPoints, th_good, th_rest, mvpMapPoints, *, nmatches=None,
                                       stream=None, sync=True):
        """src/ORBmatcher.cc:51-140.  Bit2 of mps.flags marks the good features (goodAtFrameId == F.mnId); mvpMatchScore is
        not written by this overload."""
        out_dev = is_device(mvpMapPoints)
        if nmatches is None:
            if out_dev:
                import torch
                nmatches = torch.empty(F.batch, dtype=torch.int32, device=mvpMapPoints.device)
            else:
                nmatches = np.empty(F.batch, np.int32)
        fs, ms = F.cstruct(), mps.cstruct()
        check(_lib.lib().gforb_search_by_projection_good_feature(
            self._h, C.byref(fs), C.byref(ms), F.batch, float(th_good), float(th_rest), self.mfNNratio, ptr(mvpMapPoints),
            ptr(nmatches), self._flags(F.device, out_dev, sync), C.c_void_p(stream or 0)),
            "gforb_search_by_projection_good_feature")
        return nmatches

    # -- int SearchByProjection_Budget(Frame &F, const vector<MapPoint*>&, th, time_constr)
    def SearchByProjection_Budget(self, F: Frames, mps: MapPoints, th, time_constr, mvpMapPoints, mvpMatchScore, **kw):
        """src/ORBmatcher.cc:276-379: nothing for a non-positive budget, else the plain search (budgets are unlimited here)."""
        if time_constr <= 0:
            return np.zeros(F.batch, np.int32)
        return self.SearchByProjection(F, mps, th, mvpMapPoints, mvpMatchScore, **kw)

    # -- int SearchByProjection_GoodFeature(Frame &CurrentFrame, const Frame &LastFrame, th_good, th_rest)
    def SearchByProjectionFrames_GoodFeature(self, Cur: Frames, Last: Frames, last_mappoints, last_outlier, last_good, last_world,
                                             Tcw, K, th_good, th_rest, cur_mappoints, *, nmatches=None, stream=None, sync=True):
        """src/ORBmatcher.cc:143-274.  last_good (B, stride) uint8: goodAtFrameId == CurrentFrame.mnId per last-frame map point."""
        out_dev = is_device(cur_mappoints)
        if nmatches is None:
            if out_dev:
                import torch
                nmatches = torch.empty(Cur.batch, dtype=torch.int32, device=cur_mappoints.device)
            else:
                nmatches = np.empty(Cur.batch, np.int32)
        fx, fy, cx, cy = [float(v) for v in K]
        c, l = Cur.cstruct(), Last.cstruct()
        check(_lib.lib().gforb_search_by_projection_frames_good_feature(
            self._h, C.byref(c), C.byref(l), Cur.batch, ptr(last_mappoints), ptr(last_outlier), ptr(last_good), ptr(last_world),
            ptr(Tcw), fx, fy, cx, cy, float(th_good), float(th_rest), self.mfNNratio, int(self.mbCheckOrientation),
            ptr(cur_mappoints), ptr(nmatches), self._flags(Cur.device, out_dev, sync), C.c_void_p(stream or 0)),
            "gforb_search_by_projection_frames_good_feature")
        return nmatches

    # -- int WindowSearch(F1, F2, windowSize, vpMapPointMatches2, minScaleLevel=-1, maxScaleLevel=INT_MAX)
    def WindowSearch(self, F1: Frames, F2: Frames, f1_mappoints, windowSize, minScaleLevel=-1, maxScaleLevel=0x7FFFFFFF,
                     *, stream=None, sync=True):
        dev = F1.device
        if dev:
            import torch
            m2 = torch.empty((F2.batch, F2.stride), dtype=torch.int32, device=F1.keypoints.device)
            nm = torch.empty(F2.batch, dtype=torch.int32, device=F1.keypoints.device)
        else:
            m2 = np.empty((F2.batch, F2.stride), np.int32); nm = np.empty(F2.batch, np.int32)
        a, b = F1.cstruct(), F2.cstruct()
        check(_lib.lib().gforb_window_search(self._h, C.byref(a), C.byref(b), F1.batch, ptr(f1_mappoints), int(windowSize),
                                             int(minScaleLevel), int(maxScaleLevel), self.mfNNratio,
                                             int(self.mbCheckOrientation), ptr(m2), ptr(nm), self._flags(dev, dev, sync),
                                             C.c_void_p(stream or 0)), "gforb_window_search")
        return nm, m2

    # -- int SearchByProjection(Frame &F1, Frame &F2, int windowSize, vector<MapPoint*> &vpMapPointMatches2)
    def SearchByProjectionWindow(self, F1: Frames, F2: Frames, f1_mappoints, f1_world, Tcw2, K, windowSize, matches2, f2_score, *,
                                 nmatches=None, stream=None, sync=True):
        """src/ORBmatcher.cc:1089-1168.  matches2 (B, F2.stride) int32: in = F2.mvpMapPoints, out = vpMapPointMatches2;
        f2_score: F2.mvpMatchScore, updated in place."""
        out_dev = is_device(matches2)
        if nmatches is None:
            if out_dev:
                import torch
                nmatches = torch.empty(F2.batch, dtype=torch.int32, device=matches2.device)
            else:
                nmatches = np.empty(F2.batch, np.int32)
        fx, fy, cx, cy = [float(v) for v in K]
        a, b = F1.cstruct(), F2.cstruct()
        check(_lib.lib().gforb_search_by_projection_window(
            self._h, C.byref(a), C.byref(b), F1.batch, ptr(f1_mappoints), ptr(f1_world), ptr(Tcw2), fx, fy, cx, cy, int(windowSize),
            self.mfNNratio, ptr(matches2), ptr(f2_score), ptr(nmatches), self._flags(F1.device, out_dev, sync),
            C.c_void_p(stream or 0)), "gforb_search_by_projection_window")
        return nmatches

    def _nm(self, n, like, out_dev):
        if out_dev:
            import torch
            return torch.empty(n, dtype=torch.int32, device=like.device)
        return np.empty(n, np.int32)

    def SearchByBoW(self, KF: Frames, kf_mappoints, kf_nodes, F: Frames, f_nodes, matches, *, nmatches=None, stream=None, sync=True):
        """ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches) (src/ORBmatcher.cc:724-853).  *_nodes: DBoW2
        FeatureVector node per feature (-1 = none); matches (B, F.stride) int32 out."""
        out_dev = is_device(matches)
        if nmatches is None:
            nmatches = self._nm(F.batch, matches, out_dev)
        a, b = KF.cstruct(), F.cstruct()
        check(_lib.lib().gforb_search_by_bow(self._h, C.byref(a), ptr(kf_mappoints), ptr(kf_nodes), C.byref(b), ptr(f_nodes), F.batch,
                                             self.mfNNratio, int(self.mbCheckOrientation), ptr(matches), ptr(nmatches),
                                             self._flags(F.device, out_dev, sync), C.c_void_p(stream or 0)), "gforb_search_by_bow")
        return nmatches

    def SearchByProjectionKeyFrame(self, Cur: Frames, KF: Frames, kf_mappoints, kf_skip, kf_world, kf_min_distance, kf_mp_desc, Tcw, K,
                                   th, ORBdist, cur_mappoints, cur_score, *, nmatches=None, stream=None, sync=True):
        """ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (src/ORBmatcher.cc:2204-2336);
        kf_skip non-zero = isBad() or in sAlreadyFound; cur_mappoints / cur_score are updated in place."""
        out_dev = is_device(cur_mappoints)
        if nmatches is None:
            nmatches = self._nm(Cur.batch, cur_mappoints, out_dev)
        fx, fy, cx, cy = [float(v) for v in K]
        a, b = Cur.cstruct(), KF.cstruct()
        check(_lib.lib().gforb_search_by_projection_keyframe(
            self._h, C.byref(a), C.byref(b), Cur.batch, ptr(kf_mappoints), ptr(kf_skip), ptr(kf_world), ptr(kf_min_distance),
            ptr(kf_mp_desc), ptr(Tcw), fx, fy, cx, cy, float(th), int(ORBdist), int(self.mbCheckOrientation), ptr(cur_mappoints),
            ptr(cur_score), ptr(nmatches), self._flags(Cur.device, out_dev, sync), C.c_void_p(stream or 0)),
            "gforb_search_by_projection_keyframe")
        return nmatches

    # -- the overloads of the LocalMapping / LoopClosing threads
    def SearchByBoWKeyFrames(self, KF1: Frames, kf1_mappoints, kf1_nodes, KF2: Frames, kf2_mappoints, kf2_nodes, matches12, *,
                             nmatches=None, stream=None, sync=True):
        """ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12) (src/ORBmatcher.cc:1289-1424); matches12 (B, KF1.stride)
        int32 out = id of the matched KF2 map point or -1."""
        out_dev = is_device(matches12)
        if nmatches is None:
            nmatches = self._nm(KF1.batch, matches12, out_dev)
        a, b = KF1.cstruct(), KF2.cstruct()
        check(_lib.lib().gforb_search_by_bow_keyframes(
            self._h, C.byref(a), ptr(kf1_mappoints), ptr(kf1_nodes), C.byref(b), ptr(kf2_mappoints), ptr(kf2_nodes), KF1.batch,
            self.mfNNratio, int(self.mbCheckOrientation), ptr(matches12), ptr(nmatches), self._flags(KF1.device, out_dev, sync),
            C.c_void_p(stream or 0)), "gforb_search_by_bow_keyframes")
        return nmatches

    def SearchForTriangulation(self, KF1: Frames, kf1_mappoints, kf1_nodes, KF2: Frames, kf2_mappoints, kf2_nodes, F12, matches12, *,
                               nmatches=None, stream=None, sync=True):
        """ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:1426-1588); F12 (B, 9); matches12 (B, KF1.stride) out = index
        into KF2 or -1 (vMatchedPairs = the pairs (i, matches12[i]) in ascending i)."""
        out_dev = is_device(matches12)
        if nmatches is None:
            nmatches = self._nm(KF1.batch, matches12, out_dev)
        a, b = KF1.cstruct(), KF2.cstruct()
        check(_lib.lib().gforb_search_for_triangulation(
            self._h, C.byref(a), ptr(kf1_mappoints), ptr(kf1_nodes), C.byref(b), ptr(kf2_mappoints), ptr(kf2_nodes), KF1.batch,
            ptr(F12), int(self.mbCheckOrientation), ptr(matches12), ptr(nmatches), self._flags(KF1.device, out_dev, sync),
            C.c_void_p(stream or 0)), "gforb_search_for_triangulation")
        return nmatches

    def Fuse(self, KF: Frames, pose, K, points: WorldPoints, th, best_idx, best_dist, *, sim3=False, nfused=None, stream=None, sync=True):
        """The matching loop of ORBmatcher::Fuse(pKF, vpMapPoints, th) (pose = Tcw, src/ORBmatcher.cc:1590-1708) or, with
        sim3=True, Fuse(pKF, Scw, vpPoints, th) (:1710-1839): best keypoint and distance per list entry; the caller applies
        Replace / AddObservation in list order."""
        out_dev = is_device(best_idx)
        if nfused is None:
            nfused = self._nm(KF.batch, best_idx, out_dev)
        fx, fy, cx, cy = [float(v) for v in K]
        a, w = KF.cstruct(), points.cstruct()
        check(_lib.lib().gforb_fuse(self._h, C.byref(a), KF.batch, ptr(pose), int(sim3), fx, fy, cx, cy, C.byref(w), float(th),
                                    ptr(best_idx), ptr(best_dist), ptr(nfused), self._flags(KF.device, out_dev, sync),
                                    C.c_void_p(stream or 0)), "gforb_fuse")
        return nfused

    def SearchByProjectionSim3(self, KF: Frames, Scw, K, points: WorldPoints, point_ids, th, matched, *, nmatches=None, stream=None,
                               sync=True):
        """ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (src/ORBmatcher.cc:855-977); matched (B, KF.stride)
        in/out ids."""
        out_dev = is_device(matched)
        if nmatches is None:
            nmatches = self._nm(KF.batch, matched, out_dev)
        fx, fy, cx, cy = [float(v) for v in K]
        a, w = KF.cstruct(), points.cstruct()
        check(_lib.lib().gforb_search_by_projection_sim3(self._h, C.byref(a), KF.batch, ptr(Scw), fx, fy, cx, cy, C.byref(w),
                                                         ptr(point_ids), float(th), ptr(matched), ptr(nmatches),
                                                         self._flags(KF.device, out_dev, sync), C.c_void_p(stream or 0)),
              "gforb_search_by_projection_sim3")
        return nmatches

    def SearchBySim3(self, KF1: Frames, KF2: Frames, points1: WorldPoints, points2: WorldPoints, kf2_mappoints, T1w, T2w, s12, R12, t12,
                     K, th, matches12, *, nfound=None, stream=None, sync=True):
        """ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1841-2079); matches12 (B, KF1.stride) in/out ids."""
        out_dev = is_device(matches12)
        if nfound is None:
            nfound = self._nm(KF1.batch, matches12, out_dev)
        fx, fy, cx, cy = [float(v) for v in K]
        a, b, w1, w2 = KF1.cstruct(), KF2.cstruct(), points1.cstruct(), points2.cstruct()
        check(_lib.lib().gforb_search_by_sim3(self._h, C.byref(a), C.byref(b), KF1.batch, C.byref(w1), C.byref(w2), ptr(kf2_mappoints),
                                              ptr(T1w), ptr(T2w), ptr(s12), ptr(R12), ptr(t12), fx, fy, cx, cy, float(th),
                                              ptr(matches12), ptr(nfound), self._flags(KF1.device, out_dev, sync),
                                              C.c_void_p(stream or 0)), "gforb_search_by_sim3")
        return nfound

    # -- all-pairs top-2 against a database shard resident in HBM (torch CUDA uint8 tensors)
    def AllPairsTop2(self, queries, db, idx_base=0, *, out=None, stream=None, sync=True):
        import torch
        nq, ndb = queries.shape[0], db.shape[0]
        if out is None:
            out = torch.empty((nq, 2), dtype=torch.int64, device=queries.device)
        check(_lib.lib().gforb_allpairs_top2(self._h, ptr(queries), nq, ptr(db), ndb, int(idx_base), ptr(out),
                                             self._flags(True, True, sync), C.c_void_p(stream or 0)), "gforb_allpairs_top2")
        return out

    def Top2Merge(self, gathered, *, out=None, stream=None, sync=True):
        import torch
        ns, nq = gathered.shape[0], gathered.shape[1]
        if out is None:
            out = torch.empty((nq, 2), dtype=torch.int64, device=gathered.device)
        check(_lib.lib().gforb_top2_merge(self._h, ptr(gathered), ns, nq, ptr(out), self._flags(True, True, sync),
                                          C.c_void_p(stream or 0)), "gforb_top2_merge")
        return out
