"""TEST INFRASTRUCTURE ONLY: ctypes access to the CPU oracle (oracle/liboracle.so).

Importable only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
The product package (gf_orb_slam_b200) never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])


def host_signature() -> str:
    """the CPU feature flags of this host: liboracle.so is built -O3 -march=native (the reference's own flags,
    CMakeLists.txt:13-31), so a copy built on another machine must be rebuilt here before it is loaded"""
    import hashlib
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("flags"):
                return hashlib.sha1(" ".join(sorted(ln.split(":", 1)[1].split())).encode()).hexdigest()
    except OSError:
        pass
    return "unknown"


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "liboracle.so")
    tag = so + ".host"
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".cpp", ".hpp")) or f == "Makefile"]
    sig = host_signature()
    stale = not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs)
    foreign = not os.path.exists(tag) or open(tag).read().strip() != sig
    if force or stale or foreign:
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"], stdout=subprocess.DEVNULL)
        with open(tag, "w") as f:
            f.write(sig + "\n")
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        L = _LIB
        L.orc_obs_create_f.restype = C.c_void_p
        L.orc_obs_create_fufv.restype = C.c_void_p
        L.orc_extractor_create.restype = C.c_void_p
        L.orc_frame_create.restype = C.c_void_p
        L.orc_logdet.restype = C.c_double
        L.orc_ic_angle_at.restype = C.c_float
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


# ---------------------------------------------------------------- primitives
def resize(src: np.ndarray, dw: int, dh: int) -> np.ndarray:
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.empty((dh, dw), np.uint8)
    lib().orc_resize(_p(src), src.shape[1], src.shape[0], src.shape[1], _p(dst), dw, dh, dw)
    return dst


def border101(src: np.ndarray, b: int) -> np.ndarray:
    src = np.ascontiguousarray(src, np.uint8)
    h, w = src.shape
    dst = np.empty((h + 2 * b, w + 2 * b), np.uint8)
    lib().orc_border101(_p(src), w, h, w, _p(dst), w + 2 * b, b)
    return dst


def blur(src: np.ndarray, variant: int = 0) -> np.ndarray:
    src = np.ascontiguousarray(src, np.uint8)
    h, w = src.shape
    dst = np.empty_like(src)
    lib().orc_blur(_p(src), w, h, w, _p(dst), w, variant)
    return dst


def fast(img: np.ndarray, th: int, nonmax: bool = True) -> np.ndarray:
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    cap = max(16, w * h // 4)
    out = np.zeros(cap, KP_DTYPE)
    n = lib().orc_fast(_p(img), w, h, w, th, int(nonmax), _p(out), cap)
    return out[:n]


def fast_atan2(y: np.ndarray, x: np.ndarray) -> np.ndarray:
    y = np.ascontiguousarray(y, np.float32); x = np.ascontiguousarray(x, np.float32)
    out = np.empty_like(y)
    lib().orc_fastatan2(_p(y), _p(x), _p(out), y.size)
    return out


def retain_best(resp: np.ndarray, n_keep: int, variant: int = 0) -> np.ndarray:
    resp = np.ascontiguousarray(resp, np.float32)
    out = np.empty(resp.size, np.int32)
    n = lib().orc_retain_best(_p(resp), resp.size, n_keep, variant, _p(out))
    return out[:n]


def sincosf(x: np.ndarray):
    x = np.ascontiguousarray(x, np.float32)
    s = np.empty_like(x); c = np.empty_like(x)
    lib().orc_sincosf(_p(x), _p(s), _p(c), x.size)
    return s, c


def rand_stream(seed: int, n: int) -> np.ndarray:
    out = np.empty(n, np.int32)
    lib().orc_rand_stream(C.c_uint(seed), n, _p(out))
    return out


def ic_angle_at(img: np.ndarray, x: float, y: float) -> float:
    """IC_Angle (src/ORBextractor.cc:131-158) of the keypoint at (x, y) of `img` (needs 15 pixels around it)"""
    img = np.ascontiguousarray(img, np.uint8)
    return float(lib().orc_ic_angle_at(_p(img), img.shape[1], C.c_float(x), C.c_float(y)))


def describe_at(img: np.ndarray, x: float, y: float, angle_deg: float) -> np.ndarray:
    """computeOrbDescriptor (src/ORBextractor.cc:161-201) of the keypoint at (x, y) with the given angle on `img`"""
    img = np.ascontiguousarray(img, np.uint8)
    d = np.zeros(32, np.uint8)
    lib().orc_describe_at(_p(img), img.shape[1], C.c_float(x), C.c_float(y), C.c_float(angle_deg), _p(d))
    return d


def descriptor_distance(a: np.ndarray, b: np.ndarray) -> int:
    a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
    return lib().orc_descriptor_distance(_p(a), _p(b))


# ---------------------------------------------------------------- extractor
class ORBextractor:
    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, scoreType=1, fastTh=20, minFastTh=7,
                 blurVariant=0, retainVariant=0):
        self.args = (nfeatures, scaleFactor, nlevels, scoreType, fastTh, minFastTh, blurVariant, retainVariant)
        self.nfeatures, self.nlevels = nfeatures, nlevels
        self.h = C.c_void_p(lib().orc_extractor_create(nfeatures, C.c_float(scaleFactor), nlevels, scoreType, fastTh,
                                                        minFastTh, blurVariant, retainVariant))

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_extractor_destroy(self.h)
            self.h = None

    def __call__(self, img: np.ndarray):
        img = np.ascontiguousarray(img, np.uint8)
        cap = self.nfeatures + 64
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = lib().orc_extractor_run(self.h, _p(img), img.shape[1], img.shape[0], img.shape[1], _p(kps), _p(desc), cap)
        if n < 0:
            raise RuntimeError(f"oracle extractor error {n}")
        return kps[:n].copy(), desc[:n].copy()

    def run_batch(self, imgs: np.ndarray, threads: int = 1):
        imgs = np.ascontiguousarray(imgs, np.uint8)
        nf, h, w = imgs.shape
        cap = self.nfeatures
        kps = np.zeros((nf, cap), KP_DTYPE)
        desc = np.zeros((nf, cap, 32), np.uint8)
        counts = np.zeros(nf, np.int32)
        a = self.args
        lib().orc_extractor_run_batch(a[0], C.c_float(a[1]), a[2], a[4], a[5], a[6], a[7], _p(imgs), w, h, nf, threads,
                                      _p(kps), _p(desc), _p(counts), cap)
        return kps, desc, counts

    def set_distribution(self, mode: int):
        """0: the reference's cell-quota distribution, 1: ORB-SLAM2's quadtree (oracle/orb_quadtree.cpp)"""
        lib().orc_extractor_set_distribution(self.h, int(mode))

    def pyramid(self, img: np.ndarray):
        img = np.ascontiguousarray(img, np.uint8)
        lib().orc_extractor_pyramid(self.h, _p(img), img.shape[1], img.shape[0], img.shape[1])
        return [self.level(l) for l in range(self.nlevels)]

    def level(self, l: int) -> np.ndarray:
        """bordered plane (h+32, w+32) of level l as currently held (blurred in place after a full run)"""
        w = C.c_int(); h = C.c_int(); s = C.c_int(); nf = C.c_int()
        lib().orc_extractor_level_info(self.h, l, C.byref(w), C.byref(h), C.byref(s), C.byref(nf))
        buf = np.empty((h.value + 32, s.value), np.uint8)
        lib().orc_extractor_level_copy(self.h, l, _p(buf))
        return buf

    def features_per_level(self):
        out = []
        for l in range(self.nlevels):
            w = C.c_int(); h = C.c_int(); s = C.c_int(); nf = C.c_int()
            lib().orc_extractor_level_info(self.h, l, C.byref(w), C.byref(h), C.byref(s), C.byref(nf))
            out.append(nf.value)
        return out

    def tables(self):
        sc = np.zeros(self.nlevels, np.float32); inv = np.zeros(self.nlevels, np.float32); um = np.zeros(16, np.int32)
        lib().orc_extractor_tables(self.h, _p(sc), _p(inv), _p(um))
        return sc, inv, um

    def level_keypoints(self, l: int) -> np.ndarray:
        cap = 4 * self.nfeatures + 64
        out = np.zeros(cap, KP_DTYPE)
        n = lib().orc_extractor_level_keypoints(self.h, l, _p(out), cap)
        return out[:n].copy()


# ---------------------------------------------------------------- frame + matcher
class Frame:
    def __init__(self, kps: np.ndarray, desc: np.ndarray, w: int, h: int, nlevels: int = 8, scaleFactor: float = 1.2):
        self.kps = np.ascontiguousarray(kps, KP_DTYPE)
        self.desc = np.ascontiguousarray(desc, np.uint8)
        self.N = self.kps.size
        self.h = C.c_void_p(lib().orc_frame_create(_p(self.kps), _p(self.desc), self.N, w, h, nlevels, C.c_float(scaleFactor)))

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_frame_destroy(self.h)
            self.h = None

    def set_camera(self, fx, fy, cx, cy, Tcw=None):
        T = None if Tcw is None else np.ascontiguousarray(Tcw, np.float32)
        lib().orc_frame_set_camera(self.h, C.c_float(fx), C.c_float(fy), C.c_float(cx), C.c_float(cy), _p(T))

    def set_mappoints(self, mp: np.ndarray, outlier: np.ndarray | None = None):
        mp = np.ascontiguousarray(mp, np.int32)
        o = None if outlier is None else np.ascontiguousarray(outlier, np.uint8)
        lib().orc_frame_set_mappoints(self.h, _p(mp), _p(o))

    def matches(self):
        mp = np.empty(self.N, np.int32); sc = np.empty(self.N, np.int32)
        lib().orc_frame_get_matches(self.h, _p(mp), _p(sc))
        return mp, sc

    def features_in_area(self, x, y, r, minL=-1, maxL=-1) -> np.ndarray:
        out = np.empty(max(self.N, 1), np.int32)
        n = lib().orc_frame_features_in_area(self.h, C.c_float(x), C.c_float(y), C.c_float(r), minL, maxL, _p(out), out.size)
        return out[:n].copy()


def search_by_projection(F: Frame, u, v, level, viewCos, flags, desc, th: float, nnratio: float) -> int:
    u = np.ascontiguousarray(u, np.float32); v = np.ascontiguousarray(v, np.float32)
    level = np.ascontiguousarray(level, np.int32); viewCos = np.ascontiguousarray(viewCos, np.float32)
    flags = np.ascontiguousarray(flags, np.uint8); desc = np.ascontiguousarray(desc, np.uint8)
    return lib().orc_search_by_projection(F.h, u.size, _p(u), _p(v), _p(level), _p(viewCos), _p(flags), _p(desc),
                                          C.c_float(th), C.c_float(nnratio))


def search_by_projection_frames(cur: Frame, last: Frame, last_world: np.ndarray, th: float, nnratio=0.9, checkOri=True) -> int:
    lw = np.ascontiguousarray(last_world, np.float32)
    return lib().orc_search_by_projection_frames(cur.h, last.h, _p(lw), C.c_float(th), C.c_float(nnratio), int(checkOri))


def search_by_projection_good_feature(F: Frame, u, v, level, viewCos, flags, desc, th_good: float, th_rest: float, nnratio: float) -> int:
    """ORBmatcher::SearchByProjection_GoodFeature(F, MPs, th_good, th_rest) (src/ORBmatcher.cc:51-140); flags bit2 = good"""
    u = np.ascontiguousarray(u, np.float32); v = np.ascontiguousarray(v, np.float32)
    level = np.ascontiguousarray(level, np.int32); viewCos = np.ascontiguousarray(viewCos, np.float32)
    flags = np.ascontiguousarray(flags, np.uint8); desc = np.ascontiguousarray(desc, np.uint8)
    return lib().orc_search_by_projection_good(F.h, u.size, _p(u), _p(v), _p(level), _p(viewCos), _p(flags), _p(desc),
                                               C.c_float(th_good), C.c_float(th_rest), C.c_float(nnratio))


def search_by_projection_frames_good_feature(cur: Frame, last: Frame, last_world, good, th_good: float, th_rest: float,
                                             nnratio=0.9, checkOri=True) -> int:
    """ORBmatcher::SearchByProjection_GoodFeature(Cur, Last, th_good, th_rest) (src/ORBmatcher.cc:143-274)"""
    lw = np.ascontiguousarray(last_world, np.float32); g = np.ascontiguousarray(good, np.uint8)
    return lib().orc_search_by_projection_frames_good(cur.h, last.h, _p(lw), _p(g), C.c_float(th_good), C.c_float(th_rest),
                                                      C.c_float(nnratio), int(checkOri))


def search_for_initialization(F1: Frame, F2: Frame, prev_matched: np.ndarray, window: int, nnratio=0.9, checkOri=True):
    pm = np.ascontiguousarray(prev_matched, np.float32).copy()
    m12 = np.empty(F1.N, np.int32)
    n = lib().orc_search_for_initialization(F1.h, F2.h, _p(pm), _p(m12), window, C.c_float(nnratio), int(checkOri))
    return n, m12, pm


def search_by_projection_window(F1: Frame, F2: Frame, f1_world: np.ndarray, window: int, nnratio=0.9):
    """ORBmatcher::SearchByProjection(F1, F2, windowSize, vpMapPointMatches2) (src/ORBmatcher.cc:1089-1168); F1 / F2 carry
    their map-point ids (set_mappoints), F2 its camera and pose (set_camera).  Returns (nmatches, matches2); F2's match
    scores are updated in place (F2.matches())."""
    w = np.ascontiguousarray(f1_world, np.float32)
    m2 = np.full(F2.N, -1, np.int32)
    n = lib().orc_search_by_projection_window(F1.h, F2.h, _p(w), int(window), _p(m2), C.c_float(nnratio))
    return n, m2


def search_by_bow(KF: Frame, kf_nodes, F: Frame, f_nodes, nnratio=0.7, checkOri=True):
    """ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches) (src/ORBmatcher.cc:724-853); KF carries the keyframe's
    map-point ids (set_mappoints, bad points as -1); *_nodes = DBoW2 FeatureVector node id per feature (-1 = none).
    Returns (nmatches, matches per F keypoint)."""
    kn = np.ascontiguousarray(kf_nodes, np.int32); fn = np.ascontiguousarray(f_nodes, np.int32)
    m = np.full(F.N, -1, np.int32)
    n = lib().orc_search_by_bow(KF.h, _p(kn), F.h, _p(fn), _p(m), C.c_float(nnratio), int(checkOri))
    return n, m


def search_by_projection_keyframe(Cur: Frame, KF: Frame, kf_skip, kf_world, kf_min_dist, kf_mp_desc, th: float, orb_dist: int,
                                  checkOri=True):
    """ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (src/ORBmatcher.cc:2204-2336); Cur carries
    camera + pose (set_camera) and its map points; updated in place (Cur.matches()).  Returns nmatches."""
    sk = np.ascontiguousarray(kf_skip, np.uint8); w = np.ascontiguousarray(kf_world, np.float32)
    md = np.ascontiguousarray(kf_min_dist, np.float32); d = np.ascontiguousarray(kf_mp_desc, np.uint8)
    return lib().orc_search_by_projection_keyframe(Cur.h, KF.h, _p(sk), _p(w), _p(md), _p(d), C.c_float(th), int(orb_dist),
                                                   C.c_float(0.9), int(checkOri))


def cv_rx_plus_t(R, x, t):
    """cv::Mat R*x + t on CV_32F (3x3, 3x1, 3x1) as the matchers evaluate it (oracle.hpp: cvGemmRxPlusT)"""
    R = np.ascontiguousarray(R, np.float32); x = np.ascontiguousarray(x, np.float32); t = np.ascontiguousarray(t, np.float32)
    y = np.empty(3, np.float32)
    lib().orc_cv_rx_plus_t(_p(R), _p(x), _p(t), _p(y))
    return y


def cv_neg_rt_t(R, t):
    """cv::Mat -R.t()*t on CV_32F (oracle.hpp: cvNegRtTimesT)"""
    R = np.ascontiguousarray(R, np.float32); t = np.ascontiguousarray(t, np.float32)
    y = np.empty(3, np.float32)
    lib().orc_cv_neg_rt_t(_p(R), _p(t), _p(y))
    return y


def sim3_decompose(Scw):
    """Scw -> (Rcw, tcw, Ow) as src/ORBmatcher.cc:866-871 computes them"""
    S = np.ascontiguousarray(Scw, np.float32)
    R = np.empty((3, 3), np.float32); t = np.empty(3, np.float32); Ow = np.empty(3, np.float32)
    lib().orc_sim3_decompose(_p(S), _p(R), _p(t), _p(Ow))
    return R, t, Ow


def search_by_bow_keyframes(KF1: Frame, nodes1, KF2: Frame, nodes2, nnratio=0.8, checkOri=True):
    """ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12) (src/ORBmatcher.cc:1289-1424); both keyframes carry their
    map-point ids (set_mappoints; NULL / bad = -1).  Returns (nmatches, vpMatches12 as ids of KF2's map points)."""
    n1 = np.ascontiguousarray(nodes1, np.int32); n2 = np.ascontiguousarray(nodes2, np.int32)
    m = np.full(KF1.N, -1, np.int32)
    n = lib().orc_search_by_bow_kf(KF1.h, _p(n1), KF2.h, _p(n2), _p(m), C.c_float(nnratio), int(checkOri))
    return n, m


def search_for_triangulation(KF1: Frame, nodes1, KF2: Frame, nodes2, F12, checkOri=True):
    """ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:1426-1588).  Returns (nmatches, vMatches12 = index into KF2 or -1)."""
    n1 = np.ascontiguousarray(nodes1, np.int32); n2 = np.ascontiguousarray(nodes2, np.int32)
    F = np.ascontiguousarray(F12, np.float32)
    m = np.full(KF1.N, -1, np.int32)
    n = lib().orc_search_for_triangulation(KF1.h, _p(n1), KF2.h, _p(n2), _p(F), _p(m), int(checkOri))
    return n, m


def _geom(world, normal, min_dist, max_dist, desc):
    w = np.ascontiguousarray(world, np.float32)
    nr = None if normal is None else np.ascontiguousarray(normal, np.float32)
    return (w, nr, np.ascontiguousarray(min_dist, np.float32), np.ascontiguousarray(max_dist, np.float32),
            np.ascontiguousarray(desc, np.uint8))


def fuse(KF: Frame, Scw, points, world, normal, min_dist, max_dist, desc, th, kf_slot, bad, idx_in_kf):
    """ORBmatcher::Fuse(pKF, vpMapPoints, th) (Scw None; src/ORBmatcher.cc:1590-1708) / Fuse(pKF, Scw, vpPoints, th) (:1710-1839).
    points: ids (-1 = NULL); geometry arrays indexed by id; kf_slot (KF.N), bad / idx_in_kf (per id) are the map state, returned
    updated.  Returns (nFused, bestIdx, bestDist per list position, kf_slot, bad, idx_in_kf)."""
    pts = np.ascontiguousarray(points, np.int32)
    w, nr, mn, mx, d = _geom(world, normal, min_dist, max_dist, desc)
    slot = np.ascontiguousarray(kf_slot, np.int32).copy(); b = np.ascontiguousarray(bad, np.uint8).copy()
    ik = np.ascontiguousarray(idx_in_kf, np.int32).copy()
    bi = np.empty(pts.size, np.int32); bd = np.empty(pts.size, np.int32)
    S = None if Scw is None else np.ascontiguousarray(Scw, np.float32)
    n = lib().orc_fuse(KF.h, _p(S), _p(pts), pts.size, mn.size, _p(w), _p(nr), _p(mn), _p(mx), _p(d), C.c_float(th), _p(slot), _p(b),
                       _p(ik), _p(bi), _p(bd))
    return n, bi, bd, slot, b, ik


def search_by_projection_sim3(KF: Frame, Scw, points, world, normal, min_dist, max_dist, desc, bad, matched, th: int):
    """ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (src/ORBmatcher.cc:855-977).  Returns (nmatches, vpMatched)."""
    pts = np.ascontiguousarray(points, np.int32)
    w, nr, mn, mx, d = _geom(world, normal, min_dist, max_dist, desc)
    b = np.ascontiguousarray(bad, np.uint8); mt = np.ascontiguousarray(matched, np.int32).copy()
    S = np.ascontiguousarray(Scw, np.float32)
    n = lib().orc_search_by_projection_sim3(KF.h, _p(S), _p(pts), pts.size, mn.size, _p(w), _p(nr), _p(mn), _p(mx), _p(d), _p(b), _p(mt), int(th))
    return n, mt


def search_by_sim3(KF1: Frame, KF2: Frame, matches12, idx2_in_kf2, world, min_dist, max_dist, desc, bad, s12, R12, t12, th):
    """ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1841-2079); KF1 / KF2 carry map-point ids, cameras and poses.  Returns
    (nFound, vpMatches12 as ids)."""
    w, _, mn, mx, d = _geom(world, None, min_dist, max_dist, desc)
    b = np.ascontiguousarray(bad, np.uint8); m = np.ascontiguousarray(matches12, np.int32).copy()
    i2 = np.ascontiguousarray(idx2_in_kf2, np.int32)
    R = np.ascontiguousarray(R12, np.float32); t = np.ascontiguousarray(t12, np.float32)
    n = lib().orc_search_by_sim3(KF1.h, KF2.h, _p(m), _p(i2), mn.size, _p(w), _p(mn), _p(mx), _p(d), _p(b), C.c_float(s12), _p(R), _p(t),
                                 C.c_float(th))
    return n, m


def window_search(F1: Frame, F2: Frame, window: int, minL=-1, maxL=0x7FFFFFFF, nnratio=0.9, checkOri=True):
    m2 = np.empty(F2.N, np.int32)
    n = lib().orc_window_search(F1.h, F2.h, window, _p(m2), minL, maxL, C.c_float(nnratio), int(checkOri))
    return n, m2


def active_match(F: Frame, u, v, level, viewCos, flags, desc, in_pool, obs_mat, H_meas, base_info, th: float, nnratio: float,
                 num_to_match: int, seed: int = 1):
    """Observability::runActiveMapMatching (src/Observability.cc:1249-1524): returns (nMatched, left map-point indices,
    accumulated information matrix 7x7); F.matches() holds the updated mvpMapPoints / mvpMatchScore."""
    u = np.ascontiguousarray(u, np.float32); v = np.ascontiguousarray(v, np.float32)
    level = np.ascontiguousarray(level, np.int32); viewCos = np.ascontiguousarray(viewCos, np.float32)
    flags = np.ascontiguousarray(flags, np.uint8); desc = np.ascontiguousarray(desc, np.uint8)
    in_pool = np.ascontiguousarray(in_pool, np.uint8)
    obs_mat = np.ascontiguousarray(obs_mat, np.float64); H_meas = np.ascontiguousarray(H_meas, np.float64)
    cur = np.ascontiguousarray(base_info, np.float64).reshape(49).copy()
    left = np.empty(max(u.size, 1), np.int32); nleft = C.c_int(0)
    n = lib().orc_active_match(F.h, u.size, _p(u), _p(v), _p(level), _p(viewCos), _p(flags), _p(desc), _p(in_pool), _p(obs_mat),
                               _p(H_meas), _p(cur), C.c_float(th), C.c_float(nnratio), int(num_to_match), C.c_long(seed),
                               _p(left), C.byref(nleft))
    return n, left[:nleft.value].copy(), cur.reshape(7, 7)


def is_in_frustum(Rcw, tcw, Ow, fx, fy, cx, cy, bounds, world, normal, min_dist, max_dist, cos_limit, nlevels=8, scaleFactor=1.2,
                  flags=None):
    """Frame::isInFrustum (src/Frame.cc:166-227) for M map points -> (u, v, level, viewCos, flags)"""
    f32 = lambda a: np.ascontiguousarray(a, np.float32)
    Rcw, tcw, Ow, world, normal, min_dist, max_dist = map(f32, (Rcw, tcw, Ow, world, normal, min_dist, max_dist))
    M = world.shape[0]
    u = np.zeros(M, np.float32); v = np.zeros(M, np.float32); lvl = np.zeros(M, np.int32); vc = np.zeros(M, np.float32)
    fl = np.zeros(M, np.uint8) if flags is None else np.ascontiguousarray(flags, np.uint8).copy()
    cf = C.c_float
    lib().orc_is_in_frustum(_p(Rcw), _p(tcw), _p(Ow), cf(fx), cf(fy), cf(cx), cf(cy), *[int(b) for b in bounds], nlevels,
                            cf(scaleFactor), M, _p(world), _p(normal), _p(min_dist), _p(max_dist), cf(cos_limit), _p(u), _p(v),
                            _p(lvl), _p(vc), _p(fl))
    return u, v, lvl, vc, fl


def distinctive_descriptors(descriptors, offsets):
    """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:196-262) per map point -> best row index (or -1)"""
    d = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 32)
    off = np.ascontiguousarray(offsets, np.int32)
    out = np.empty(off.size - 1, np.int32)
    lib().orc_distinctive_descriptors(off.size - 1, _p(d), _p(off), _p(out))
    return out


def undistort_keypoints(xy, fx, fy, cx, cy, dist):
    """Frame::UndistortKeyPoints (src/Frame.cc:389-423) on (N, 2) float32 keypoint positions -> (N, 2) float32"""
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    dist = np.ascontiguousarray(dist, np.float32).reshape(-1)
    out = np.empty_like(xy)
    cf = C.c_float
    lib().orc_undistort_keypoints(_p(xy), xy.shape[0], cf(fx), cf(fy), cf(cx), cf(cy), _p(dist), dist.size, _p(out))
    return out


# ---------------------------------------------------------------- observability
class Observability:
    def __init__(self, *, f=None, fu=None, fv=None, nRows=480, nCols=752, Cx=0.0, Cy=0.0, k1=0.0, k2=0.0, dx=None, dy=None):
        d = C.c_double
        if f is not None:
            self.h = C.c_void_p(lib().orc_obs_create_f(d(f), nRows, nCols, d(Cx), d(Cy), d(k1), d(k2), d(dx), d(dy)))
        else:
            self.h = C.c_void_p(lib().orc_obs_create_fufv(d(fu), d(fv), nRows, nCols, d(Cx), d(Cy), d(k1), d(k2)))

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_obs_destroy(self.h)
            self.h = None

    def set_xv(self, xv):
        xv = np.ascontiguousarray(xv, np.float64)
        lib().orc_obs_set_xv(self.h, _p(xv))

    def get_xv(self):
        xv = np.empty(13, np.float64)
        lib().orc_obs_get_xv(self.h, _p(xv))
        return xv

    def update(self, t0, Tcw_prev, t1, Twc_cur):
        a = np.ascontiguousarray(Tcw_prev, np.float32); b = np.ascontiguousarray(Twc_cur, np.float32)
        lib().orc_obs_update(self.h, C.c_double(t0), _p(a), C.c_double(t1), _p(b))

    def predict(self, dt, nseg):
        lib().orc_obs_predict(self.h, C.c_double(dt), nseg)

    def kine(self, seg=0):
        Xv = np.empty(13); FQ = np.empty((4, 4)); FO = np.empty((4, 3)); F = np.empty((13, 13))
        FQi = np.empty((4, 4)); FOi = np.empty((4, 3)); T = np.empty((4, 4), np.float32)
        lib().orc_obs_get_kine(self.h, seg, _p(Xv), _p(FQ), _p(FO), _p(F), _p(FQi), _p(FOi), _p(T))
        return dict(Xv=Xv, F_Q=FQ, F_Omg=FO, F=F, F_Q_inSeg=FQi, F_Omg_inSeg=FOi, Tcw=T)

    def H_complete(self, Xv, yi, zi):
        Xv = np.ascontiguousarray(Xv, np.float64); yi = np.ascontiguousarray(yi, np.float64); zi = np.ascontiguousarray(zi, np.float64)
        H13 = np.empty((2, 3)); H47 = np.empty((2, 4))
        lib().orc_obs_H_complete(self.h, _p(Xv), _p(yi), _p(zi), _p(H13), _p(H47))
        return H13, H47

    def H_simplified(self, Xv, yi):
        Xv = np.ascontiguousarray(Xv, np.float64); yi = np.ascontiguousarray(yi, np.float64)
        H13 = np.empty((2, 3)); H47 = np.empty((2, 4)); uv = np.empty(2, np.float32)
        lib().orc_obs_H_simplified(self.h, _p(Xv), _p(yi), _p(H13), _p(H47), _p(uv))
        return H13, H47, uv

    def build_frame(self, world, octave, score, valid, level_sigma2, seg=0):
        world = np.ascontiguousarray(world, np.float32); octave = np.ascontiguousarray(octave, np.int32)
        score = np.ascontiguousarray(score, np.int32); valid = np.ascontiguousarray(valid, np.uint8)
        ls = np.ascontiguousarray(level_sigma2, np.float32)
        n = world.shape[0]
        blocks = np.zeros((n, 7, 7)); obs = np.zeros(n); H = np.zeros((n, 2, 7))
        lib().orc_obs_build_frame(self.h, seg, _p(world), _p(octave), _p(score), _p(valid), n, _p(ls), _p(blocks), _p(obs), _p(H))
        return blocks, obs, H

    def build_map(self, world, track_in_view, check_viz, bounds, seg=0, want_H=False):
        world = np.ascontiguousarray(world, np.float32); tv = np.ascontiguousarray(track_in_view, np.uint8)
        n = world.shape[0]
        blocks = np.zeros((n, 7, 7)); obs = np.zeros(n); uv = np.zeros((n, 2), np.float32)
        H = np.zeros((n, 2, 7)) if want_H else None
        lib().orc_obs_build_map(self.h, seg, _p(world), _p(tv), n, int(check_viz), *[int(b) for b in bounds], _p(blocks), _p(obs), _p(uv),
                                _p(H))
        return (blocks, obs, uv, H) if want_H else (blocks, obs, uv)


def logdet(M: np.ndarray) -> float:
    M = np.ascontiguousarray(M, np.float64)
    return lib().orc_logdet(_p(M), M.shape[0])


def select(blocks: np.ndarray, k: int, mtd: int = 2, sample_scale: float = 23.0, seed: int = 1, scores=None):
    blocks = np.ascontiguousarray(blocks, np.float64)
    n = blocks.shape[0]
    sc = None if scores is None else np.ascontiguousarray(scores, np.float64)
    idx = np.empty(n, np.int32); ld = np.zeros(n, np.float64)
    m = lib().orc_obs_select(_p(blocks), _p(sc), n, k, mtd, C.c_double(sample_scale), C.c_long(seed), _p(idx), _p(ld))
    return idx[:m].copy(), ld[:m].copy()


def select_map_mtd3(blocks: np.ndarray, k: int, seed: int = 1, num_threads: int = 8, scores=None):
    """selection stage of the local-map setSelction_Number, greedy_mtd 3 incl. the partitioned pool (src/Observability.cc:1085-1163)"""
    blocks = np.ascontiguousarray(blocks, np.float64)
    n = blocks.shape[0]
    sc = None if scores is None else np.ascontiguousarray(scores, np.float64)
    idx = np.empty(n, np.int32)
    m = lib().orc_obs_select_map_mtd3(_p(blocks), _p(sc), n, k, C.c_long(seed), num_threads, _p(idx))
    return idx[:m].copy()


def pose_optimization(Tcw, kp_xy, octave, map_point, world, inv_level_sigma2, K):
    """Optimizer::PoseOptimization (src/Optimizer.cc:279-413).  Returns (n_inliers, Tcw_out (4x4 f32), outlier (N u8))."""
    T = np.ascontiguousarray(Tcw, np.float32).reshape(4, 4).copy()
    xy = np.ascontiguousarray(kp_xy, np.float32); oc = np.ascontiguousarray(octave, np.int32)
    mp = np.ascontiguousarray(map_point, np.int32); w = np.ascontiguousarray(world, np.float32)
    s2 = np.ascontiguousarray(inv_level_sigma2, np.float32)
    out = np.zeros(len(mp), np.uint8)
    fx, fy, cx, cy = [C.c_float(v) for v in K]
    n = lib().orc_pose_optimization(_p(T), len(mp), _p(xy), _p(oc), _p(mp), _p(w), _p(s2), fx, fy, cx, cy, _p(out))
    return n, T, out


def front_end_batch(imgs: np.ndarray, threads: int, world, score, xv, sigma2, k: int, sample_scale: float, window=100,
                    nnratio=0.9, cam=None):
    """extract + SearchForInitialization(prev, cur) + info blocks + lazier-greedy selection for every frame, frames
    split over `threads` C++ threads (cpu_baseline / --impl reference leg of bench.py)."""
    imgs = np.ascontiguousarray(imgs, np.uint8)
    nf, h, w = imgs.shape
    world = np.ascontiguousarray(world, np.float32); score = np.ascontiguousarray(score, np.int32)
    xv = np.ascontiguousarray(xv, np.float64); sigma2 = np.ascontiguousarray(sigma2, np.float32)
    c = cam or dict(f=5.1369248, nRows=h, nCols=w, Cx=367.215, Cy=248.375, k1=0.0, k2=0.0, dx=0.01123325985, dy=0.01123325985)
    cam9 = np.array([c["f"], c["nRows"], c["nCols"], c["Cx"], c["Cy"], c["k1"], c["k2"], c["dx"], c["dy"]], np.float64)
    counts = np.zeros(nf, np.int32); nmatch = np.zeros(nf, np.int32)
    sel = np.full((nf, k), -1, np.int32); seln = np.zeros(nf, np.int32)
    lib().orc_front_end_batch(_p(imgs), w, h, nf, threads, world.shape[0], _p(world), _p(score), _p(xv), _p(sigma2), k,
                              C.c_double(sample_scale), window, C.c_float(nnratio), _p(cam9), _p(counts), _p(nmatch),
                              _p(sel), _p(seln))
    return counts, nmatch, sel, seln
