"""GPU parity of the ORBmatcher overloads of the LocalMapping / LoopClosing threads through the C ABI against the
CPU oracle (oracle/orb_matcher_kf.cpp): SearchByBoW(KF, KF), SearchForTriangulation, both Fuse overloads,
SearchByProjection(KF, Scw, ..) and SearchBySim3 -- indices, ids, distances and counts bit-exact."""
import numpy as np
import pytest

from gf_orb_slam_b200 import synth
from tests.test_gpu_matcher import W, H, _bow_nodes, _oframe, _pose_and_world, seq_frames  # noqa: F401 (fixture)

pytestmark = pytest.mark.gpu
S = 1000
INT_MAX = 2**31 - 1
SCALE = (np.float32(1.2) ** np.arange(8)).astype(np.float32)


def _pair_inputs(kps, desc, counts, B, p_mp, seed):
    mp1 = np.full((B, S), -1, np.int32); mp2 = np.full((B, S), -1, np.int32)
    n1_ = np.full((B, S), -1, np.int32); n2_ = np.full((B, S), -1, np.int32)
    for b in range(B):
        n1, n2 = int(counts[b]), int(counts[b + 1])
        mp1[b, :n1] = np.where(synth.uniform01(seed + b, n1) < p_mp, 9000 + np.arange(n1), -1)
        mp2[b, :n2] = np.where(synth.uniform01(seed + 50 + b, n2) < p_mp, 20000 + np.arange(n2), -1)
        n1_[b, :n1] = _bow_nodes(desc[b], n1, seed + 100 + b); n2_[b, :n2] = _bow_nodes(desc[b + 1], n2, seed + 150 + b)
    return mp1, mp2, n1_, n2_


@pytest.mark.parametrize("nnratio,check_ori", [(0.8, True), (0.95, False)])
def test_search_by_bow_keyframes(orc, seq_frames, nnratio, check_ori):
    """ORBmatcher::SearchByBoW(pKF1, pKF2, vpMatches12) (src/ORBmatcher.cc:1289-1424): only KF2 features that hold a map point,
    best < TH_LOW (strict), result per KF1 feature = the KF2 map point, histogram pruning by KF1 index"""
    from gf_orb_slam_b200 import ORBmatcher, Frames
    kps, desc, counts = seq_frames
    B = 3
    KF1 = Frames(kps[:B], desc[:B], counts[:B], W, H); KF2 = Frames(kps[1:B + 1], desc[1:B + 1], counts[1:B + 1], W, H)
    mp1, mp2, nd1, nd2 = _pair_inputs(kps, desc, counts, B, 0.8, 2100)
    out = np.full((B, S), -7, np.int32)
    m = ORBmatcher(nnratio, check_ori, max_batch=B, max_keypoints=S, max_queries=S)
    nm = m.SearchByBoWKeyFrames(KF1, mp1, nd1, KF2, mp2, nd2, out)
    tot = 0
    for b in range(B):
        n1, n2 = int(counts[b]), int(counts[b + 1])
        k1 = _oframe(orc, kps[b], desc[b], n1); k2 = _oframe(orc, kps[b + 1], desc[b + 1], n2)
        k1.set_mappoints(mp1[b, :n1]); k2.set_mappoints(mp2[b, :n2])
        en, em = orc.search_by_bow_keyframes(k1, nd1[b, :n1], k2, nd2[b, :n2], nnratio, check_ori)
        assert nm[b] == en, (b, nm[b], en)
        assert np.array_equal(out[b, :n1], em) and (out[b, n1:] == -1).all()
        tot += en
    assert tot > 100


@pytest.mark.parametrize("check_ori", [True, False])
def test_search_for_triangulation(orc, seq_frames, check_ori):
    """ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:1426-1588) + CheckDistEpipolarLine (:705-722): features without a
    map point, candidates in (distance, index) order up to twice the best distance, epipolar gate, claims, histogram"""
    from gf_orb_slam_b200 import ORBmatcher, Frames
    kps, desc, counts = seq_frames
    B = 3
    KF1 = Frames(kps[:B], desc[:B], counts[:B], W, H); KF2 = Frames(kps[1:B + 1], desc[1:B + 1], counts[1:B + 1], W, H)
    mp1, mp2, nd1, nd2 = _pair_inputs(kps, desc, counts, B, 0.35, 2300)
    # consecutive sequence frames differ by an image shift along x (and at most one row): the fundamental matrix of a pure
    # x translation puts the epipolar line of (x1, y1) at y = y1; a small perturbation exercises every term of the test
    F12 = np.zeros((B, 9), np.float32)
    for b in range(B):
        F = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float64) + 2e-4 * (synth.uniform01(2400 + b, 9).reshape(3, 3) - 0.5)
        F12[b] = F.astype(np.float32).reshape(9)
    out = np.full((B, S), -7, np.int32)
    m = ORBmatcher(0.6, check_ori, max_batch=B, max_keypoints=S, max_queries=S)
    nm = m.SearchForTriangulation(KF1, mp1, nd1, KF2, mp2, nd2, F12, out)
    tot = 0
    for b in range(B):
        n1, n2 = int(counts[b]), int(counts[b + 1])
        k1 = _oframe(orc, kps[b], desc[b], n1); k2 = _oframe(orc, kps[b + 1], desc[b + 1], n2)
        k1.set_mappoints(mp1[b, :n1]); k2.set_mappoints(mp2[b, :n2])
        en, em = orc.search_for_triangulation(k1, nd1[b, :n1], k2, nd2[b, :n2], F12[b], check_ori)
        assert nm[b] == en, (b, nm[b], en)
        assert np.array_equal(out[b, :n1], em) and (out[b, n1:] == -1).all()
        tot += en
    assert tot > 60
    # the gate does reject: with a matrix whose lines miss, nothing matches
    Fbad = np.tile(np.array([0, 0, 0, 0, 0, -1, 0, 1, 300], np.float32), (B, 1))
    nm0 = m.SearchForTriangulation(KF1, mp1, nd1, KF2, mp2, nd2, Fbad, out)
    assert int(np.sum(nm0)) == 0


def _geometry(kps_b, n, seed, shift):
    """map points that project near the keypoints of a keyframe: world positions, normals (towards the camera, a few turned
    away), distance-invariance bounds that put the predicted level at the keypoint's octave or next to it, descriptors"""
    K, T, w = _pose_and_world(kps_b, n, seed=seed, shift=shift)
    Ow = -(T[:3, :3].T.astype(np.float64) @ T[:3, 3].astype(np.float64))
    PO = w.astype(np.float64) - Ow
    d3 = np.linalg.norm(PO, axis=1)
    jitter = 0.93 + 0.14 * synth.uniform01(seed + 1, n)
    mind = (d3 / (SCALE[np.clip(kps_b["octave"][:n], 0, 7)] * jitter)).astype(np.float32)
    maxd = (mind * np.float32(1.2) ** 7 * np.where(synth.uniform01(seed + 2, n) < 0.05, 0.2, 1.3)).astype(np.float32)   # a few out of range
    nrm = (PO / d3[:, None]).astype(np.float32)
    away = synth.uniform01(seed + 3, n) < 0.07
    nrm[away] *= -1.0
    return K, T, w, nrm, mind, maxd


def _noisy_desc(d, seed, p=0.04):
    d = d.copy()
    flip = synth.uniform01(seed, d.size).reshape(d.shape) < p
    d[flip] ^= 0x10
    return d


def _apply_fuse(points, gi, gd, slot, bad, in_kf, sim3):
    """what the caller does with gforb_fuse's result: the map mutations of src/ORBmatcher.cc:1689-1704 / :1818-1833 in list
    order, modelled on this keyframe (MapPoint::Replace, AddObservation, KeyFrame::AddMapPoint)"""
    slot, bad, in_kf = slot.copy(), bad.copy(), in_kf.copy()
    found = {int(i) for i in slot if i >= 0 and not bad[i]}
    n = 0
    for i, pid in enumerate(points):
        if pid < 0:
            continue
        if bad[pid] or (pid in found if sim3 else in_kf[pid] >= 0):
            continue
        bi, bd = int(gi[i]), int(gd[i])
        if bi < 0 or bd > 50:
            continue
        other = int(slot[bi])
        if other >= 0:
            if not bad[other]:
                if not sim3:
                    bad[pid] = 1                                   # pMP->Replace(pMPinKF)
                elif other != pid:                                 # pMPinKF->Replace(pMP)
                    bad[other] = 1
                    at = int(in_kf[other]); in_kf[other] = -1
                    if at >= 0:
                        if in_kf[pid] < 0:
                            slot[at] = pid; in_kf[pid] = at
                        else:
                            slot[at] = -1
        else:
            in_kf[pid] = bi; slot[bi] = pid
        n += 1
    return n, slot, bad, in_kf


@pytest.mark.parametrize("sim3,th", [(False, 3.0), (True, 4.0), (False, 2.5)])
def test_fuse(orc, seq_frames, sim3, th):
    """ORBmatcher::Fuse(pKF, vpMapPoints, th) (src/ORBmatcher.cc:1590-1708) and Fuse(pKF, Scw, vpPoints, th) (:1710-1839): the
    matching loop on the GPU (best keypoint / distance per list entry), the map mutations applied in list order by the caller;
    lists with repeated and NULL entries, points already in the keyframe, bad points, points out of range / seen from behind"""
    from gf_orb_slam_b200 import ORBmatcher, Frames, WorldPoints
    kps, desc, counts = seq_frames
    B = 3
    KF = Frames(kps[:B], desc[:B], counts[:B], W, H)
    NID = 1100                                                     # ids 0 .. NID-1; the list draws from them with repeats
    L = 1000
    world = np.zeros((B, L, 3), np.float32); normal = np.zeros((B, L, 3), np.float32)
    mind = np.ones((B, L), np.float32); maxd = np.ones((B, L), np.float32); dsc = np.zeros((B, L, 32), np.uint8)
    skip = np.ones((B, L), np.uint8); cnt = np.zeros(B, np.int32); pose = np.zeros((B, 16), np.float32)
    state = []
    K = None
    for b in range(B):
        n = int(counts[b])
        K, T, w, nrm, mn, mx = _geometry(kps[b], n, 3100 + 10 * b, shift=(0.8, -0.6))
        gw = np.zeros((NID, 3), np.float32); gn = np.zeros((NID, 3), np.float32); gmn = np.ones(NID, np.float32); gmx = np.ones(NID, np.float32)
        gd = synth.splitmix64(3200 + b, NID * 4).view(np.uint8).reshape(NID, 32).copy()
        m_ = min(n, NID)
        gw[:m_] = w[:m_]; gn[:m_] = nrm[:m_]; gmn[:m_] = mn[:m_]; gmx[:m_] = mx[:m_]; gd[:m_] = _noisy_desc(desc[b, :m_], 3300 + b)
        if sim3:
            s = np.float32(1.7 + 0.1 * b)
            P_ = T.copy(); P_[:3, :] *= s
        else:
            P_ = T
        pose[b] = P_.reshape(16)
        # map state of the keyframe: 30 % of its keypoints hold a point (ids from the top of the range, some of them bad)
        slot = np.full(n, -1, np.int32)
        u = synth.uniform01(3400 + b, n)
        holder = np.nonzero(u < 0.3)[0]
        slot[holder] = (NID - 1 - np.arange(holder.size)) % NID
        bad = (synth.uniform01(3500 + b, NID) < 0.05).astype(np.uint8)
        in_kf = np.full(NID, -1, np.int32)
        in_kf[slot[holder]] = holder
        pts = np.minimum((synth.uniform01(3600 + b, L) * NID).astype(np.int32), NID - 1)
        pts[::17] = pts[1::17][: pts[::17].size]                   # repeated entries
        if not sim3:
            pts[5::23] = -1                                        # NULL entries (Fuse(pKF, vpMapPoints, th) tolerates them)
        found = {int(i) for i in slot if i >= 0 and not bad[i]}
        for i, pid in enumerate(pts):
            if pid < 0:
                continue
            world[b, i] = gw[pid]; normal[b, i] = gn[pid]; mind[b, i] = gmn[pid]; maxd[b, i] = gmx[pid]; dsc[b, i] = gd[pid]
            skip[b, i] = bad[pid] or ((pid in found) if sim3 else in_kf[pid] >= 0)
        cnt[b] = L
        state.append((pts, gw, gn, gmn, gmx, gd, slot, bad, in_kf, T))
    bi = np.full((B, L), -9, np.int32); bd = np.full((B, L), -9, np.int32)
    m = ORBmatcher(0.6, True, max_batch=B, max_keypoints=S, max_queries=L)
    nf = m.Fuse(KF, pose, K, WorldPoints(world, normal, mind, maxd, dsc, skip, cnt), th, bi, bd, sim3=sim3)
    tot = 0
    for b in range(B):
        n = int(counts[b])
        pts, gw, gn, gmn, gmx, gd, slot, bad, in_kf, T = state[b]
        kf = _oframe(orc, kps[b], desc[b], n)
        kf.set_camera(*K, T)
        en, ebi, ebd, eslot, ebad, ein = orc.fuse(kf, pose[b] if sim3 else None, pts, gw, gn, gmn, gmx, gd, th, slot, bad, in_kf)
        # stateless matching result: equal wherever the oracle's walk reached the matching loop
        reached = ebi >= 0
        assert reached.sum() > 100
        assert np.array_equal(bi[b][reached], ebi[reached]) and np.array_equal(bd[b][reached], ebd[reached])
        assert nf[b] == int(np.sum(bd[b] <= 50))
        # entries the walk skipped at entry are skipped on the GPU too; entries it skipped because an earlier entry changed
        # the map (repeats) are dropped by the in-order application
        gn_, gslot, gbad, gin = _apply_fuse(pts, bi[b], bd[b], slot, bad, in_kf, sim3)
        assert gn_ == en, (b, gn_, en)
        assert np.array_equal(gslot, eslot) and np.array_equal(gbad, ebad) and np.array_equal(gin, ein)
        tot += en
    assert tot > 150


def test_search_by_projection_sim3(orc, seq_frames):
    """ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (src/ORBmatcher.cc:855-977): Sim3 decomposition,
    gates, claimed keypoints skipped, repeated list entries claim a second keypoint"""
    from gf_orb_slam_b200 import ORBmatcher, Frames, WorldPoints
    kps, desc, counts = seq_frames
    B = 3
    KF = Frames(kps[:B], desc[:B], counts[:B], W, H)
    L = 1200
    world = np.zeros((B, L, 3), np.float32); normal = np.zeros((B, L, 3), np.float32)
    mind = np.ones((B, L), np.float32); maxd = np.ones((B, L), np.float32); dsc = np.zeros((B, L, 32), np.uint8)
    skip = np.ones((B, L), np.uint8); cnt = np.zeros(B, np.int32); Scw = np.zeros((B, 16), np.float32)
    ids = np.zeros((B, L), np.int32); matched = np.full((B, S), -1, np.int32)
    state = []
    K = None
    th = 10
    for b in range(B):
        n = int(counts[b])
        K, T, w, nrm, mn, mx = _geometry(kps[b], n, 4100 + 10 * b, shift=(1.5, 1.0))
        gd = _noisy_desc(desc[b, :n], 4200 + b)
        P_ = T.copy(); P_[:3, :] *= np.float32(0.6 + 0.2 * b)
        Scw[b] = P_.reshape(16)
        bad = (synth.uniform01(4300 + b, n) < 0.05).astype(np.uint8)
        m0 = np.where(synth.uniform01(4400 + b, n) < 0.15, np.minimum((synth.uniform01(4500 + b, n) * n).astype(np.int32), n - 1), -1).astype(np.int32)
        matched[b, :n] = m0
        pts = np.minimum((synth.uniform01(4600 + b, L) * n).astype(np.int32), n - 1)
        already = set(int(i) for i in m0 if i >= 0)
        world[b] = w[pts]; normal[b] = nrm[pts]; mind[b] = mn[pts]; maxd[b] = mx[pts]; dsc[b] = gd[pts]; ids[b] = pts
        skip[b] = [bad[p] or (int(p) in already) for p in pts]
        cnt[b] = L
        state.append((pts, w, nrm, mn, mx, gd, bad, m0))
    m = ORBmatcher(0.6, True, max_batch=B, max_keypoints=S, max_queries=L)
    nm = m.SearchByProjectionSim3(KF, Scw, K, WorldPoints(world, normal, mind, maxd, dsc, skip, cnt), ids, th, matched)
    tot = 0
    for b in range(B):
        n = int(counts[b])
        pts, w, nrm, mn, mx, gd, bad, m0 = state[b]
        kf = _oframe(orc, kps[b], desc[b], n)
        kf.set_camera(*K, np.eye(4, dtype=np.float32))
        en, em = orc.search_by_projection_sim3(kf, Scw[b], pts, w, nrm, mn, mx, gd, bad, m0, th)
        assert nm[b] == en, (b, nm[b], en)
        assert np.array_equal(matched[b, :n], em)
        tot += en
    assert tot > 300


@pytest.mark.parametrize("th", [7.5, 10.0])
def test_search_by_sim3(orc, seq_frames, th):
    """ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1841-2079): both projection directions through the similarity, mutual
    agreement, pairs matched before the call left alone"""
    from gf_orb_slam_b200 import ORBmatcher, Frames, WorldPoints
    kps, desc, counts = seq_frames
    B = 3
    KF1 = Frames(kps[:B], desc[:B], counts[:B], W, H); KF2 = Frames(kps[1:B + 1], desc[1:B + 1], counts[1:B + 1], W, H)
    z3 = lambda: np.zeros((B, S, 3), np.float32)   # noqa: E731
    w1, w2 = z3(), z3()
    mn1 = np.ones((B, S), np.float32); mx1 = np.ones((B, S), np.float32); mn2 = np.ones((B, S), np.float32); mx2 = np.ones((B, S), np.float32)
    d1 = np.zeros((B, S, 32), np.uint8); d2 = np.zeros((B, S, 32), np.uint8)
    sk1 = np.ones((B, S), np.uint8); sk2 = np.ones((B, S), np.uint8)
    mp1 = np.full((B, S), -1, np.int32); mp2 = np.full((B, S), -1, np.int32); m12 = np.full((B, S), -1, np.int32)
    T1 = np.zeros((B, 16), np.float32); T2 = np.zeros((B, 16), np.float32)
    s12 = np.zeros(B, np.float32); R12 = np.zeros((B, 9), np.float32); t12 = np.zeros((B, 3), np.float32)
    state = []
    K = None
    for b in range(B):
        n1, n2 = int(counts[b]), int(counts[b + 1])
        # both keyframes share one pose; the similarity is close to the identity, so KF1's points (placed on KF1's keypoints
        # shifted by the image motion) land near their KF2 matches and the other way round
        K, T, a1 = _pose_and_world(kps[b], n1, seed=5100 + b, shift=(-1.0, 0.0))
        _, _, a2 = _pose_and_world(kps[b + 1], n2, seed=5100 + b, shift=(1.0, 0.0))
        ang = 0.002 * (b + 1)
        R = np.array([[np.cos(ang), -np.sin(ang), 0], [np.sin(ang), np.cos(ang), 0], [0, 0, 1]], np.float32)
        s12[b] = 1.0 + 0.01 * b; R12[b] = R.reshape(9); t12[b] = np.array([0.004, -0.003, 0.01 * b], np.float32)
        T1[b] = T.reshape(16); T2[b] = T.reshape(16)
        c1 = a1.astype(np.float64) @ T[:3, :3].T.astype(np.float64) + T[:3, 3]
        c2 = a2.astype(np.float64) @ T[:3, :3].T.astype(np.float64) + T[:3, 3]
        j1 = 0.93 + 0.14 * synth.uniform01(5200 + b, n1); j2 = 0.93 + 0.14 * synth.uniform01(5300 + b, n2)
        w1[b, :n1] = a1; w2[b, :n2] = a2
        mn1[b, :n1] = np.linalg.norm(c1, axis=1) / (SCALE[np.clip(kps["octave"][b, :n1], 0, 7)] * j1)
        mn2[b, :n2] = np.linalg.norm(c2, axis=1) / (SCALE[np.clip(kps["octave"][b + 1, :n2], 0, 7)] * j2)
        mx1[b] = mn1[b] * np.float32(1.2) ** 7 * 1.3; mx2[b] = mn2[b] * np.float32(1.2) ** 7 * 1.3
        d1[b, :n1] = _noisy_desc(desc[b, :n1], 5400 + b); d2[b, :n2] = _noisy_desc(desc[b + 1, :n2], 5500 + b)
        mp1[b, :n1] = np.where(synth.uniform01(5600 + b, n1) < 0.9, np.arange(n1), -1)             # ids: KF1's points 0 .. n1-1
        mp2[b, :n2] = np.where(synth.uniform01(5700 + b, n2) < 0.9, 2000 + np.arange(n2), -1)      # KF2's points 2000 ..
        bad = (synth.uniform01(5800 + b, 4000) < 0.04).astype(np.uint8)
        # pairs matched before the call: vpMatches12[i1] = some KF2 point, whose index in KF2 marks vbAlreadyMatched2
        pre = np.nonzero((synth.uniform01(5900 + b, n1) < 0.1) & (mp1[b, :n1] >= 0))[0]
        idx2 = np.full(n1, -1, np.int32)
        for i1 in pre:
            j = int(synth.uniform01(6000 + b, 1, offset=int(i1))[0] * n2)
            if mp2[b, j] >= 0:
                m12[b, i1] = mp2[b, j]; idx2[i1] = j
        already2 = np.zeros(n2, bool); already2[idx2[idx2 >= 0]] = True
        sk1[b, :n1] = (mp1[b, :n1] < 0) | (m12[b, :n1] >= 0) | (bad[np.clip(mp1[b, :n1], 0, None)] > 0)
        sk2[b, :n2] = (mp2[b, :n2] < 0) | already2 | (bad[np.clip(mp2[b, :n2], 0, None)] > 0)
        state.append((idx2, bad, m12[b, :n1].copy()))
    cnt1 = counts[:B].astype(np.int32).copy(); cnt2 = counts[1:B + 1].astype(np.int32).copy()
    m = ORBmatcher(0.6, True, max_batch=B, max_keypoints=S, max_queries=S)
    nf = m.SearchBySim3(KF1, KF2, WorldPoints(w1, None, mn1, mx1, d1, sk1, cnt1), WorldPoints(w2, None, mn2, mx2, d2, sk2, cnt2), mp2,
                        T1, T2, s12, R12, t12, K, th, m12)
    tot = 0
    for b in range(B):
        n1, n2 = int(counts[b]), int(counts[b + 1])
        idx2, bad, m12_0 = state[b]
        k1 = _oframe(orc, kps[b], desc[b], n1); k2 = _oframe(orc, kps[b + 1], desc[b + 1], n2)
        k1.set_mappoints(mp1[b, :n1]); k2.set_mappoints(mp2[b, :n2])
        k1.set_camera(*K, T1[b].reshape(4, 4)); k2.set_camera(*K, T2[b].reshape(4, 4))
        # geometry by id: KF1's points at 0 .., KF2's at 2000 ..
        gw = np.zeros((4000, 3), np.float32); gmn = np.ones(4000, np.float32); gmx = np.ones(4000, np.float32); gd = np.zeros((4000, 32), np.uint8)
        gw[:n1] = w1[b, :n1]; gmn[:n1] = mn1[b, :n1]; gmx[:n1] = mx1[b, :n1]; gd[:n1] = d1[b, :n1]
        gw[2000:2000 + n2] = w2[b, :n2]; gmn[2000:2000 + n2] = mn2[b, :n2]; gmx[2000:2000 + n2] = mx2[b, :n2]; gd[2000:2000 + n2] = d2[b, :n2]
        en, em = orc.search_by_sim3(k1, k2, m12_0, idx2, gw, gmn, gmx, gd, bad, float(s12[b]), R12[b], t12[b], th)
        assert nf[b] == en, (b, nf[b], en)
        assert np.array_equal(m12[b, :n1], em)
        tot += en
    assert tot > 300


def test_edge_cases(orc, seq_frames):
    """empty keyframes, lists of skipped points only, features without vocabulary nodes, a batch of zero problems: no match,
    every output written, no error"""
    from gf_orb_slam_b200 import ORBmatcher, Frames, WorldPoints
    kps, desc, counts = seq_frames
    B = 2
    c0 = counts[:B].astype(np.int32).copy(); c0[1] = 0                      # the second keyframe is empty
    KF1 = Frames(kps[:B], desc[:B], c0, W, H); KF2 = Frames(kps[1:B + 1], desc[1:B + 1], counts[1:B + 1].astype(np.int32).copy(), W, H)
    m = ORBmatcher(0.8, True, max_batch=B, max_keypoints=S, max_queries=S)
    mp = np.arange(B * S, dtype=np.int32).reshape(B, S); none = np.full((B, S), -1, np.int32)
    out = np.full((B, S), -7, np.int32)
    nm = m.SearchByBoWKeyFrames(KF1, mp, none, KF2, mp, none, out)          # no feature sits in a node
    assert (np.asarray(nm) == 0).all() and (out == -1).all()
    out[:] = -7
    nodes = np.full((B, S), 5, np.int32)
    nm = m.SearchForTriangulation(KF1, mp, nodes, KF2, mp, nodes, np.zeros((B, 9), np.float32), out)   # every feature holds a map point
    assert (np.asarray(nm) == 0).all() and (out == -1).all()
    L = 64
    z = np.zeros((B, L, 3), np.float32)
    pts = WorldPoints(z, z, np.ones((B, L), np.float32), np.ones((B, L), np.float32), np.zeros((B, L, 32), np.uint8),
                      np.ones((B, L), np.uint8), np.full(B, L, np.int32))   # every entry skipped
    bi = np.full((B, L), -9, np.int32); bd = np.full((B, L), -9, np.int32)
    T = np.tile(np.eye(4, dtype=np.float32).reshape(16), (B, 1))
    nf = m.Fuse(KF1, T, (458.0, 458.0, 376.0, 240.0), pts, 3.0, bi, bd)
    assert (np.asarray(nf) == 0).all() and (bi == -1).all() and (bd == INT_MAX).all()
    # points that are not skipped but lie behind the camera / project outside the image / at z = 0
    w = np.zeros((B, L, 3), np.float32); w[:, :, 2] = -1.0; w[:, 1::3, 2] = 0.0; w[:, 2::3] = (1e6, 1e6, 1.0)
    pts = WorldPoints(w, z, np.ones((B, L), np.float32), np.full((B, L), 1e9, np.float32), np.zeros((B, L, 32), np.uint8),
                      np.zeros((B, L), np.uint8), np.full(B, L, np.int32))
    nf = m.Fuse(KF1, T, (458.0, 458.0, 376.0, 240.0), pts, 3.0, bi, bd, sim3=True)
    assert (np.asarray(nf) == 0).all() and (bi == -1).all()
    matched = np.full((B, S), -1, np.int32)
    nm = m.SearchByProjectionSim3(KF1, T, (458.0, 458.0, 376.0, 240.0), pts, np.zeros((B, L), np.int32), 10, matched)
    assert (np.asarray(nm) == 0).all() and (matched == -1).all()
    m0 = ORBmatcher(0.8, True, max_batch=1, max_keypoints=S, max_queries=S)
    E = Frames(kps[:0], desc[:0], counts[:0].astype(np.int32), W, H)
    from gf_orb_slam_b200 import _lib
    import ctypes as C
    a = E.cstruct()
    rc = _lib.lib().gforb_search_by_bow_keyframes(m0._h, C.byref(a), None, None, C.byref(a), None, None, 0, 0.8, 1, None, None, 0, None)
    assert rc in (0, -2)                                                     # zero problems: nothing to do (or rejected arguments)
