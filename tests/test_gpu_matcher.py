"""GPU parity: CUDA ORBmatcher / Frame grid (through the C ABI) vs the CPU oracle -- bit-exact indices, scores, counts."""
import numpy as np
import pytest

from gf_orb_slam_b200 import synth

pytestmark = pytest.mark.gpu

W, H = 752, 480


@pytest.fixture(scope="module")
def seq_frames():
    """6 consecutive C2 sequence frames extracted on the GPU (parity of the extractor itself: test_gpu_extractor)."""
    from gf_orb_slam_b200 import ORBextractor
    seq = synth.Sequence()
    imgs = seq.frames(0, 6)
    ex = ORBextractor(1000, 1.2, 8, max_batch=6)
    kps, desc, counts = ex.extract_batch(imgs)
    return kps, desc, counts


def _oframe(orc, kps, desc, n):
    return orc.Frame(kps[:n], desc[:n], W, H)


def test_descriptor_distance(orc):
    from gf_orb_slam_b200 import ORBmatcher
    m = ORBmatcher(max_batch=1)
    z = np.zeros(32, np.uint8); f = np.full(32, 255, np.uint8)
    one = z.copy(); one[17] = 0x10
    assert m.DescriptorDistance(z, z) == 0 and m.DescriptorDistance(z, f) == 256 and m.DescriptorDistance(z, one) == 1
    rng = np.random.default_rng(3)
    a = rng.integers(0, 256, (500, 32), dtype=np.uint8); b = rng.integers(0, 256, (500, 32), dtype=np.uint8)
    got = m.DescriptorDistance(a, b)
    exp = np.array([orc.descriptor_distance(a[i], b[i]) for i in range(500)])
    assert np.array_equal(got, exp)


def test_features_in_area(orc, seq_frames):
    from gf_orb_slam_b200 import ORBmatcher, Frames
    kps, desc, counts = seq_frames
    F = Frames(kps, desc, counts, W, H)
    m = ORBmatcher(max_batch=6, max_keypoints=1000, max_queries=1000)
    fo = _oframe(orc, kps[2], desc[2], counts[2])
    rng = np.random.default_rng(1)
    nq = 300
    x = rng.uniform(-30, W + 30, nq).astype(np.float32); y = rng.uniform(-30, H + 30, nq).astype(np.float32)
    r = rng.choice([2.5, 7.0, 15.0, 40.0, 100.0, 900.0], nq).astype(np.float32)
    lv = rng.integers(0, 8, nq)
    kind = rng.integers(0, 4, nq)
    mn = np.where(kind == 0, -1, np.where(kind == 1, lv, lv - 1)).astype(np.int32)
    mx = np.where(kind == 0, -1, np.where(kind == 1, lv, np.where(kind == 2, lv, lv + 1))).astype(np.int32)
    got, cnt = m.GetFeaturesInArea(F, 2, x, y, r, mn, mx, cap=1000)
    nonempty = 0
    for i in range(nq):
        exp = fo.features_in_area(x[i], y[i], r[i], int(mn[i]), int(mx[i]))
        assert cnt[i] == len(exp)
        assert np.array_equal(got[i], exp), i
        nonempty += len(exp) > 0
    assert nonempty > nq // 2


@pytest.mark.parametrize("window,nnratio,check_ori", [(100, 0.9, True), (30, 0.7, False), (10, 0.9, True)])
def test_search_for_initialization(orc, seq_frames, window, nnratio, check_ori):
    from gf_orb_slam_b200 import ORBmatcher, Frames
    kps, desc, counts = seq_frames
    B = 5
    F1 = Frames(kps[:B], desc[:B], counts[:B], W, H)
    F2 = Frames(kps[1:B + 1], desc[1:B + 1], counts[1:B + 1], W, H)
    prev = np.stack([kps["x"][:B], kps["y"][:B]], axis=-1).astype(np.float32).copy()
    prev0 = prev.copy()
    m = ORBmatcher(nnratio, check_ori, max_batch=B, max_keypoints=1000, max_queries=1000)
    nm, m12 = m.SearchForInitialization(F1, F2, prev, window)
    total = 0
    for b in range(B):
        f1 = _oframe(orc, kps[b], desc[b], counts[b]); f2 = _oframe(orc, kps[b + 1], desc[b + 1], counts[b + 1])
        n, e12, epm = orc.search_for_initialization(f1, f2, prev0[b, :counts[b]], window, nnratio, check_ori)
        assert nm[b] == n, (b, nm[b], n)
        assert np.array_equal(m12[b, :counts[b]], e12)
        assert np.array_equal(prev[b, :counts[b]], epm.reshape(-1, 2))
        total += n
    if window == 100:
        assert total > 50          # the workload really matches something
    # second call with the updated vbPrevMatched (the reference calls it repeatedly during initialisation)
    nm2, m12b = m.SearchForInitialization(F1, F2, prev, window)
    f1 = _oframe(orc, kps[0], desc[0], counts[0]); f2 = _oframe(orc, kps[1], desc[1], counts[1])
    _, _, pm1 = orc.search_for_initialization(f1, f2, prev0[0, :counts[0]], window, nnratio, check_ori)
    n2, e12b, _ = orc.search_for_initialization(f1, f2, pm1.reshape(-1, 2), window, nnratio, check_ori)
    assert nm2[0] == n2 and np.array_equal(m12b[0, :counts[0]], e12b)
    assert m.last_launches() == 3


@pytest.mark.parametrize("M,th,nnratio", [(2000, 1.0, 0.8), (10000, 1.0, 0.8), (3000, 5.0, 0.8), (1500, 3.0, 0.6)])
def test_search_by_projection_map(orc, M, th, nnratio):
    """config C3: 1500 frame keypoints against 10k projected map points (plus smaller / wider-window variants)."""
    from gf_orb_slam_b200 import ORBextractor, ORBmatcher, Frames, MapPoints
    img = synth.textured_frame(W, H, seed=1)
    ex = ORBextractor(1500, 1.2, 8, max_width=W, max_height=H, max_batch=1)
    kps, desc, counts = ex.extract_batch(img[None])
    n = int(counts[0])
    B = 2
    mp = [synth.map_points_c3(kps[0, :n], desc[0, :n], M, seed=50 + b, w=W, h=H) for b in range(B)]
    stack = lambda k: np.ascontiguousarray(np.stack([mp[b][k] for b in range(B)]))
    mps = MapPoints(stack("proj_x"), stack("proj_y"), stack("scale_level"), stack("view_cos"), stack("flags"),
                    stack("descriptors"), np.array([M, M - 7], np.int32))
    kB = np.repeat(kps, B, 0); dB = np.repeat(desc, B, 0); cB = np.repeat(counts, B)
    F = Frames(kB, dB, cB, W, H)
    fmp = np.full((B, 1500), -1, np.int32); fsc = np.full((B, 1500), 999, np.int32)
    fmp[1, 5:40:3] = 123456          # frame 1 arrives with some keypoints already claimed
    fmp0 = fmp.copy()
    m = ORBmatcher(nnratio, True, max_batch=B, max_keypoints=1500, max_queries=M, cand_pool=64 * M)
    nm = m.SearchByProjection(F, mps, th, fmp, fsc)
    for b in range(B):
        fo = _oframe(orc, kps[0], desc[0], n)
        fo.set_mappoints(fmp0[b, :n])
        Mb = M if b == 0 else M - 7
        en = orc.search_by_projection(fo, mp[b]["proj_x"][:Mb], mp[b]["proj_y"][:Mb], mp[b]["scale_level"][:Mb],
                                      mp[b]["view_cos"][:Mb], mp[b]["flags"][:Mb], mp[b]["descriptors"][:Mb], th, nnratio)
        emp, esc = fo.matches()
        assert nm[b] == en, (b, nm[b], en)
        assert np.array_equal(fmp[b, :n], emp) and np.array_equal(fsc[b, :n], esc)
    assert nm[0] > 100


def _pose_and_world(kps, n, seed, shift=(3.0, -2.0)):
    """world points that project (with Tcw) near the last frame's keypoints shifted by a few pixels"""
    fx = fy = 458.0; cx = W / 2.0; cy = H / 2.0
    u01 = synth.uniform01(seed, 2 * n).reshape(n, 2)
    z = (2.0 + 6.0 * u01[:, 0]).astype(np.float32)
    x = ((kps["x"][:n] + shift[0] - cx) / fx * z).astype(np.float32)
    y = ((kps["y"][:n] + shift[1] - cy) / fy * z).astype(np.float32)
    Xc = np.stack([x, y, z], 1)
    a = 0.02
    R = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]], np.float32)
    t = np.array([0.1, -0.05, 0.2], np.float32)
    Xw = ((Xc - t) @ R).astype(np.float32)               # Xc = R Xw + t
    Tcw = np.eye(4, dtype=np.float32); Tcw[:3, :3] = R; Tcw[:3, 3] = t
    return (fx, fy, cx, cy), Tcw, Xw


@pytest.mark.parametrize("th,check_ori", [(15.0, True), (50.0, False)])
def test_search_by_projection_frames(orc, seq_frames, th, check_ori):
    from gf_orb_slam_b200 import ORBmatcher, Frames
    kps, desc, counts = seq_frames
    B = 3
    Last = Frames(kps[:B], desc[:B], counts[:B], W, H)
    Cur = Frames(kps[1:B + 1], desc[1:B + 1], counts[1:B + 1], W, H)
    lmp = np.full((B, 1000), -1, np.int32); lout = np.zeros((B, 1000), np.uint8); lworld = np.zeros((B, 1000, 3), np.float32)
    Tcw = np.zeros((B, 4, 4), np.float32)
    K = None
    for b in range(B):
        n = counts[b]
        # sequence offset between frame b and b+1 is (-1, 0) px or (-1,-1): add it so projections land on the match
        K, Tcw[b], lworld[b, :n] = _pose_and_world(kps[b], n, seed=70 + b, shift=(-1.0, 0.0))
        u = synth.uniform01(90 + b, n)
        lmp[b, :n] = np.where(u < 0.8, 1000 + np.arange(n), -1)
        lout[b, :n] = (u > 0.75) & (u < 0.8)
    cmp_ = np.full((B, 1000), -1, np.int32); csc = np.full((B, 1000), 999, np.int32)
    cmp_[2, 0:30:4] = 777
    cmp0 = cmp_.copy()
    m = ORBmatcher(0.9, check_ori, max_batch=B, max_keypoints=1000, max_queries=1000)
    nm = m.SearchByProjectionFrames(Cur, Last, lmp, lout, lworld, Tcw, K, th, cmp_, csc)
    for b in range(B):
        fl = _oframe(orc, kps[b], desc[b], counts[b]); fc = _oframe(orc, kps[b + 1], desc[b + 1], counts[b + 1])
        fl.set_mappoints(lmp[b, :counts[b]], lout[b, :counts[b]])
        fc.set_mappoints(cmp0[b, :counts[b + 1]])
        fc.set_camera(*K, Tcw[b])
        en = orc.search_by_projection_frames(fc, fl, lworld[b, :counts[b]], th, 0.9, check_ori)
        emp, esc = fc.matches()
        assert nm[b] == en, (b, nm[b], en)
        assert np.array_equal(cmp_[b, :counts[b + 1]], emp) and np.array_equal(csc[b, :counts[b + 1]], esc)
    assert nm[0] > 50


@pytest.mark.parametrize("window,minL,maxL", [(100, -1, 0x7FFFFFFF), (200, 1, 4)])
def test_window_search(orc, seq_frames, window, minL, maxL):
    from gf_orb_slam_b200 import ORBmatcher, Frames
    kps, desc, counts = seq_frames
    B = 3
    F1 = Frames(kps[:B], desc[:B], counts[:B], W, H)
    F2 = Frames(kps[1:B + 1], desc[1:B + 1], counts[1:B + 1], W, H)
    f1mp = np.full((B, 1000), -1, np.int32)
    for b in range(B):
        u = synth.uniform01(33 + b, counts[b])
        f1mp[b, :counts[b]] = np.where(u < 0.85, 5000 + np.arange(counts[b]), -1)
    m = ORBmatcher(0.9, True, max_batch=B, max_keypoints=1000, max_queries=1000)
    nm, m2 = m.WindowSearch(F1, F2, f1mp, window, minL, maxL)
    for b in range(B):
        f1 = _oframe(orc, kps[b], desc[b], counts[b]); f2 = _oframe(orc, kps[b + 1], desc[b + 1], counts[b + 1])
        f1.set_mappoints(f1mp[b, :counts[b]])
        en, e2 = orc.window_search(f1, f2, window, minL, maxL, 0.9, True)
        assert nm[b] == en, (b, nm[b], en)
        assert np.array_equal(m2[b, :counts[b + 1]], e2)
    assert nm[0] > 50


@pytest.mark.parametrize("window,nnratio", [(15, 0.9), (50, 0.7)])
def test_search_by_projection_window(orc, seq_frames, window, nnratio):
    """ORBmatcher::SearchByProjection(F1, F2, windowSize, vpMapPointMatches2) (src/ORBmatcher.cc:1089-1168, TrackPreviousFrame):
    F2 arrives with map points of its own (claimed keypoints, ids F1 must skip), matches and F2's match scores bit-exact"""
    from gf_orb_slam_b200 import ORBmatcher, Frames
    kps, desc, counts = seq_frames
    B = 3
    F1 = Frames(kps[:B], desc[:B], counts[:B], W, H)
    F2 = Frames(kps[1:B + 1], desc[1:B + 1], counts[1:B + 1], W, H)
    K = None
    f1mp = np.full((B, 1000), -1, np.int32); world = np.zeros((B, 1000, 3), np.float32); Tcw = np.zeros((B, 16), np.float32)
    m2 = np.full((B, 1000), -1, np.int32); sc = np.full((B, 1000), 999, np.int32)
    for b in range(B):
        n1, n2 = int(counts[b]), int(counts[b + 1])
        u = synth.uniform01(133 + b, n1)
        f1mp[b, :n1] = np.where(u < 0.85, 7000 + np.arange(n1), -1)
        K, T, w = _pose_and_world(kps[b], n1, seed=70 + b, shift=(float(b) + 1.0, -2.0))
        world[b, :n1] = w; Tcw[b] = T.reshape(16)
        pre = synth.uniform01(233 + b, n2)
        m2[b, :n2] = np.where(pre < 0.10, 7000 + 3 * np.arange(n2), -1)      # F2 already holds some points, some shared with F1
    m20 = m2.copy()
    m = ORBmatcher(nnratio, True, max_batch=B, max_keypoints=1000, max_queries=1000)
    nm = m.SearchByProjectionWindow(F1, F2, f1mp, world, Tcw, K, window, m2, sc)
    tot = 0
    for b in range(B):
        n1, n2 = int(counts[b]), int(counts[b + 1])
        f1 = _oframe(orc, kps[b], desc[b], n1); f2 = _oframe(orc, kps[b + 1], desc[b + 1], n2)
        f1.set_mappoints(f1mp[b, :n1]); f2.set_mappoints(m20[b, :n2])
        f2.set_camera(*K, Tcw[b].reshape(4, 4))
        en, e2 = orc.search_by_projection_window(f1, f2, world[b, :n1], window, nnratio)
        _, esc = f2.matches()
        assert nm[b] == en, (b, nm[b], en)
        assert np.array_equal(m2[b, :n2], e2) and np.array_equal(sc[b, :n2], esc)
        tot += en
    assert tot > 100


def test_empty_and_ragged(orc, seq_frames):
    """frames with zero keypoints and different counts in one batch; zero map points"""
    from gf_orb_slam_b200 import ORBmatcher, Frames, MapPoints
    kps, desc, counts = seq_frames
    c1 = counts[:3].copy(); c1[1] = 0
    c2 = counts[1:4].copy(); c2[2] = 0; c2[0] = 17
    F1 = Frames(kps[:3], desc[:3], c1, W, H); F2 = Frames(kps[1:4], desc[1:4], c2, W, H)
    prev = np.stack([kps["x"][:3], kps["y"][:3]], axis=-1).astype(np.float32).copy()
    prev0 = prev.copy()
    m = ORBmatcher(0.9, True, max_batch=3, max_keypoints=1000, max_queries=1000)
    nm, m12 = m.SearchForInitialization(F1, F2, prev, 100)
    for b in range(3):
        f1 = _oframe(orc, kps[b], desc[b], c1[b]); f2 = _oframe(orc, kps[b + 1], desc[b + 1], c2[b])
        n, e12, _ = orc.search_for_initialization(f1, f2, prev0[b, :c1[b]], 100, 0.9, True)
        assert nm[b] == n and np.array_equal(m12[b, :c1[b]], e12)
    assert nm[1] == 0 and nm[2] == 0
    z = lambda dt, *s: np.zeros(s, dt)
    mps = MapPoints(z(np.float32, 3, 4), z(np.float32, 3, 4), z(np.int32, 3, 4), z(np.float32, 3, 4), z(np.uint8, 3, 4),
                    z(np.uint8, 3, 4, 32), np.zeros(3, np.int32))
    fmp = np.full((3, 1000), -1, np.int32); fsc = np.full((3, 1000), 999, np.int32)
    assert np.array_equal(m.SearchByProjection(F2, mps, 1.0, fmp, fsc), [0, 0, 0]) and (fmp == -1).all()


def _allpairs_ref(q, db, base):
    lut = np.array([bin(i).count("1") for i in range(256)], np.int32)
    out = np.full((len(q), 2), -1, np.int64)                    # ~0ull = "no such row"
    for s in range(0, len(q), 64):
        d = lut[q[s:s + 64, None, :] ^ db[None, :, :]].sum(-1)
        key = d.astype(np.int64) * (1 << 32) + (np.arange(len(db), dtype=np.int64) + base)[None, :]
        part = np.sort(key, axis=1)[:, :2]
        out[s:s + 64, :part.shape[1]] = part
    return out


@pytest.mark.parametrize("kernel,nq,ndb,splits", [("tc", 777, 20011, None), ("popc", 777, 20011, None), ("tc", 777, 20011, 1),
                                                  ("tc", 1, 1, None), ("tc", 3, 2, None), ("tc", 130, 255, None),
                                                  ("tc", 128, 256, None), ("tc", 129, 257, 2), ("tc", 2000, 70013, None),
                                                  ("tc", 300, 9000, 3)])
def test_allpairs_top2(kernel, nq, ndb, splits, monkeypatch):
    """config 5 brute force: tensor-core kernel (tcgen05 kind::i8, default) and integer-pipe kernel (GFORB_ALLPAIRS=popc)
    against numpy, packed (dist << 32 | row) keys bit for bit incl. ties (lower row wins), ragged tails, tiny inputs,
    one CTA walking many tiles (splits = 1: every pipeline stage wraps) and the shard merge"""
    import torch
    from gf_orb_slam_b200 import ORBmatcher
    monkeypatch.setenv("GFORB_ALLPAIRS", kernel)
    if splits:
        monkeypatch.setenv("GFORB_ALLPAIRS_SPLITS", str(splits))
    rng = np.random.default_rng(9 + nq)
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    db = rng.integers(0, 256, (ndb, 32), dtype=np.uint8)
    if ndb > 9000:
        db[100] = q[5]; db[9000] = q[5]; db[ndb - 1] = q[nq - 1]        # exact hits + a tie (lower index must win)
        db[300] = ~q[7]                                                  # distance 256
    if ndb > 600:
        db[512:520] = db[40]                                             # eight equal rows across a tile boundary region
    m = ORBmatcher(max_batch=1)
    qd, dd = torch.from_numpy(q).cuda(), torch.from_numpy(db).cuda()
    keys = m.AllPairsTop2(qd, dd, idx_base=1000).cpu().numpy().astype(np.int64)
    exp = _allpairs_ref(q, db, 1000)
    assert np.array_equal(keys, exp)
    if ndb > 9000:
        assert keys[5, 0] == 1100 and keys[5, 1] == 10000 and keys[nq - 1, 0] == 1000 + ndb - 1
    if ndb >= 4:
        h = ndb // 2                                                     # shard merge: two half-shards == the full result
        k0 = m.AllPairsTop2(qd, dd[:h], idx_base=1000).clone(); k1 = m.AllPairsTop2(qd, dd[h:], idx_base=1000 + h).clone()
        merged = m.Top2Merge(torch.stack([k0, k1])).cpu().numpy().astype(np.int64)
        assert np.array_equal(merged, exp)


def test_is_in_frustum(orc):
    """Frame::isInFrustum (src/Frame.cc:166-227) batched: flags, projections, predicted level and viewing cosine are
    bit-equal to the oracle for points inside, outside, behind the camera, too near / far and seen from behind."""
    from gf_orb_slam_b200 import ORBmatcher, MapPoints
    B, M = 3, 5000
    fx, fy, cx, cy = 457.3, 457.6, 367.2, 248.4
    bounds = (0, W, 0, H)
    rng = np.random.default_rng(5)
    Rs, ts, Os = [], [], []
    for b in range(B):
        a = 0.1 * (b + 1)
        c, s_ = np.cos(a), np.sin(a)
        R = (np.array([[c, 0, s_], [0, 1, 0], [-s_, 0, c]]) @ np.array([[1, 0, 0], [0, np.cos(0.05), -np.sin(0.05)], [0, np.sin(0.05), np.cos(0.05)]])).astype(np.float32)
        t = np.array([0.2 * b, -0.1, 0.3], np.float32)
        Rs.append(R.reshape(9)); ts.append(t); Os.append((-R.T.astype(np.float32) @ t).astype(np.float32))
    Rcw = np.stack(Rs); tcw = np.stack(ts); Ow = np.stack(Os)
    world = (rng.uniform(-6, 6, size=(B, M, 3)) * np.array([1, 1, 1.5])).astype(np.float32)
    normal = rng.normal(size=(B, M, 3)).astype(np.float32)
    normal[:, ::2] = world[:, ::2] - Ow[:, None, :]            # half of the normals look along the viewing ray
    normal /= np.linalg.norm(normal, axis=2, keepdims=True).astype(np.float32)
    dist = np.linalg.norm(world - Ow[:, None, :], axis=2).astype(np.float32)
    min_d = (dist * rng.uniform(0.2, 1.1, size=(B, M))).astype(np.float32)
    max_d = (dist * rng.uniform(0.9, 3.0, size=(B, M))).astype(np.float32)
    counts = np.array([M, M - 11, M], np.int32)
    z = lambda dt: np.zeros((B, M), dt)
    flags = (rng.integers(0, 4, size=(B, M)) & 2).astype(np.uint8)       # bit1 (bad) must survive
    flags0 = flags.copy()
    mps = MapPoints(z(np.float32), z(np.float32), z(np.int32), z(np.float32), flags, np.zeros((B, M, 32), np.uint8), counts)
    m = ORBmatcher(0.8, True, max_batch=B, max_keypoints=64, max_queries=M)
    m.isInFrustum(Rcw, tcw, Ow, fx, fy, cx, cy, bounds, world, normal, min_d, max_d, counts, 0.5, mps)
    nin = 0
    for b in range(B):
        n = int(counts[b])
        eu, ev, el, ec, ef = orc.is_in_frustum(Rcw[b], tcw[b], Ow[b], fx, fy, cx, cy, bounds, world[b, :n], normal[b, :n],
                                               min_d[b, :n], max_d[b, :n], 0.5, flags=flags0[b, :n])
        assert np.array_equal(mps.a[4][b, :n], ef), b
        assert np.array_equal(mps.a[4][b, n:], flags0[b, n:])
        k = (ef & 1) == 1
        nin += int(k.sum())
        for got, exp in ((mps.a[0], eu), (mps.a[1], ev), (mps.a[2], el), (mps.a[3], ec)):
            assert np.array_equal(got[b, :n][k], exp[k]), b
    assert 200 < nin < B * M // 2


def test_undistort_keypoints(orc):
    """Frame::UndistortKeyPoints (src/Frame.cc:389-423) batched: positions bit-equal to the oracle and to the committed
    cv2 4.13 vectors, every other KeyPoint field copied, entries past counts[b] untouched, k1 == 0 copies."""
    import os
    from gf_orb_slam_b200 import ORBmatcher
    from gf_orb_slam_b200._lib import KP_DTYPE
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "cv413_undistort.npz"))
    m = ORBmatcher(0.8, True, max_batch=3, max_keypoints=3000, max_queries=64)
    rng = np.random.default_rng(3)
    for name in ("euroc", "tum1", "k1only"):
        K, D, xy, und = g[name + "_K"], g[name + "_D"], g[name + "_xy"], g[name + "_und"]
        B, n = 3, xy.shape[0]
        keys = np.zeros((B, n), KP_DTYPE)
        for b in range(B):
            p = np.roll(xy, 7 * b, axis=0)
            keys[b]["x"], keys[b]["y"] = p[:, 0], p[:, 1]
        keys["size"] = 31.0; keys["angle"] = rng.uniform(0, 360, (B, n)).astype(np.float32)
        keys["response"] = rng.integers(7, 120, (B, n)); keys["octave"] = rng.integers(0, 8, (B, n)); keys["class_id"] = -1
        counts = np.array([n, n - 100, 17], np.int32)
        out = np.zeros((B, n), KP_DTYPE); out["class_id"] = 12345
        m.UndistortKeyPoints(keys, counts, K[0, 0], K[1, 1], K[0, 2], K[1, 2], D, out)
        for b in range(B):
            c = int(counts[b])
            exp = orc.undistort_keypoints(np.stack([keys[b]["x"][:c], keys[b]["y"][:c]], 1), K[0, 0], K[1, 1], K[0, 2], K[1, 2], D)
            assert np.array_equal(out[b]["x"][:c], exp[:, 0]) and np.array_equal(out[b]["y"][:c], exp[:, 1]), (name, b)
            if b == 0:
                assert np.array_equal(exp, und)                                  # == cv2.undistortPoints
            for f in ("size", "angle", "response", "octave", "class_id"):
                assert np.array_equal(out[b][f][:c], keys[b][f][:c])
            assert (out[b]["class_id"][c:] == 12345).all()
        same = m.UndistortKeyPoints(keys, counts, K[0, 0], K[1, 1], K[0, 2], K[1, 2], [0.0, 0.2, 0.0, 0.0])
        assert same[0].tobytes() == keys[0].tobytes()


@pytest.mark.parametrize("M,th_good,th_rest", [(10000, 1.0, 3.0), (3000, 2.0, 1.0)])
def test_search_by_projection_good_feature_map(orc, M, th_good, th_rest):
    """SearchByProjection_GoodFeature(F, MPs, th_good, th_rest) (src/ORBmatcher.cc:51-140): per-point window factor from
    the good-feature flag (bit2), match scores untouched; and SearchByProjection_Budget (:276-379)."""
    from gf_orb_slam_b200 import ORBextractor, ORBmatcher, Frames, MapPoints
    img = synth.textured_frame(W, H, seed=1)
    ex = ORBextractor(1500, 1.2, 8, max_width=W, max_height=H, max_batch=1)
    kps, desc, counts = ex.extract_batch(img[None])
    n = int(counts[0])
    B = 2
    mp = [synth.map_points_c3(kps[0, :n], desc[0, :n], M, seed=60 + b, w=W, h=H) for b in range(B)]
    for b in range(B):
        good = (synth.uniform01(300 + b, M) < 0.3).astype(np.uint8)
        mp[b]["flags"] = (mp[b]["flags"] | (good << 2)).astype(np.uint8)
    stack = lambda k: np.ascontiguousarray(np.stack([mp[b][k] for b in range(B)]))
    mps = MapPoints(stack("proj_x"), stack("proj_y"), stack("scale_level"), stack("view_cos"), stack("flags"),
                    stack("descriptors"), np.array([M, M - 7], np.int32))
    kB = np.repeat(kps, B, 0); dB = np.repeat(desc, B, 0); cB = np.repeat(counts, B)
    F = Frames(kB, dB, cB, W, H)
    fmp = np.full((B, 1500), -1, np.int32)
    fmp[1, 5:40:3] = 123456
    fmp0 = fmp.copy()
    m = ORBmatcher(0.8, True, max_batch=B, max_keypoints=1500, max_queries=M, cand_pool=64 * M)
    nm = m.SearchByProjection_GoodFeature(F, mps, th_good, th_rest, fmp)
    plain = fmp0.copy(); psc = np.full((B, 1500), 999, np.int32)
    nplain = m.SearchByProjection(F, mps, th_rest, plain, psc)
    for b in range(B):
        fo = _oframe(orc, kps[0], desc[0], n)
        fo.set_mappoints(fmp0[b, :n])
        Mb = M if b == 0 else M - 7
        en = orc.search_by_projection_good_feature(fo, mp[b]["proj_x"][:Mb], mp[b]["proj_y"][:Mb], mp[b]["scale_level"][:Mb],
                                                   mp[b]["view_cos"][:Mb], mp[b]["flags"][:Mb], mp[b]["descriptors"][:Mb],
                                                   th_good, th_rest, 0.8)
        emp, esc = fo.matches()
        assert nm[b] == en, (b, nm[b], en)
        assert np.array_equal(fmp[b, :n], emp)
        assert (esc == 999).all()                                  # the overload leaves mvpMatchScore alone
    assert nm[0] > 100 and not np.array_equal(fmp, plain)          # the good-feature windows change the associations
    # _Budget: a spent budget matches nothing, an open one is the plain search
    b0 = fmp0.copy(); s0 = np.full((B, 1500), 999, np.int32)
    assert (m.SearchByProjection_Budget(F, mps, th_rest, 0.0, b0, s0) == 0).all() and np.array_equal(b0, fmp0)
    nb = m.SearchByProjection_Budget(F, mps, th_rest, 5.0, b0, s0)
    assert np.array_equal(nb, nplain) and np.array_equal(b0, plain) and np.array_equal(s0, psc)


@pytest.mark.parametrize("th_good,th_rest,check_ori", [(7.0, 15.0, True), (30.0, 10.0, False)])
def test_search_by_projection_frames_good_feature(orc, seq_frames, th_good, th_rest, check_ori):
    """SearchByProjection_GoodFeature(Cur, Last, th_good, th_rest) (src/ORBmatcher.cc:143-274)"""
    from gf_orb_slam_b200 import ORBmatcher, Frames
    kps, desc, counts = seq_frames
    B = 3
    Last = Frames(kps[:B], desc[:B], counts[:B], W, H)
    Cur = Frames(kps[1:B + 1], desc[1:B + 1], counts[1:B + 1], W, H)
    lmp = np.full((B, 1000), -1, np.int32); lout = np.zeros((B, 1000), np.uint8); lworld = np.zeros((B, 1000, 3), np.float32)
    lgood = np.zeros((B, 1000), np.uint8)
    Tcw = np.zeros((B, 4, 4), np.float32)
    K = None
    for b in range(B):
        n = counts[b]
        K, Tcw[b], lworld[b, :n] = _pose_and_world(kps[b], n, seed=170 + b, shift=(-1.0, 0.0))
        u = synth.uniform01(190 + b, n)
        lmp[b, :n] = np.where(u < 0.8, 1000 + np.arange(n), -1)
        lout[b, :n] = (u > 0.75) & (u < 0.8)
        lgood[b, :n] = synth.uniform01(210 + b, n) < 0.4
    cmp_ = np.full((B, 1000), -1, np.int32)
    cmp_[2, 0:30:4] = 777
    cmp0 = cmp_.copy()
    m = ORBmatcher(0.9, check_ori, max_batch=B, max_keypoints=1000, max_queries=1000)
    nm = m.SearchByProjectionFrames_GoodFeature(Cur, Last, lmp, lout, lgood, lworld, Tcw, K, th_good, th_rest, cmp_)
    for b in range(B):
        fl = _oframe(orc, kps[b], desc[b], counts[b]); fc = _oframe(orc, kps[b + 1], desc[b + 1], counts[b + 1])
        fl.set_mappoints(lmp[b, :counts[b]], lout[b, :counts[b]])
        fc.set_mappoints(cmp0[b, :counts[b + 1]])
        fc.set_camera(*K, Tcw[b])
        en = orc.search_by_projection_frames_good_feature(fc, fl, lworld[b, :counts[b]], lgood[b, :counts[b]], th_good, th_rest,
                                                          0.9, check_ori)
        emp, esc = fc.matches()
        assert nm[b] == en, (b, nm[b], en)
        assert np.array_equal(cmp_[b, :counts[b + 1]], emp) and (esc == 999).all()
    assert nm[0] > 50


def test_compute_distinctive_descriptors(orc):
    """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:196-262) batched: least-median row per map point, ties to
    the first row, empty points -> -1 and an untouched descriptor; host and device inputs."""
    import torch
    from gf_orb_slam_b200 import ORBmatcher
    rng = np.random.default_rng(11)
    sizes = [0, 1, 2, 3, 4, 7, 33, 50, 64, 129, 300, 0, 5] + list(rng.integers(1, 40, 200))
    descs = []
    for n in sizes:
        base = rng.integers(0, 256, 32, dtype=np.uint8)
        d = np.repeat(base[None], n, 0)
        for r in range(n):                                        # observations = a base descriptor with a few flipped bits
            k = int(rng.integers(0, 60)) if r % 5 != 2 else 0      # every fifth is an exact copy -> ties between rows
            bits = rng.choice(256, k, replace=False)
            for b in bits:
                d[r, b >> 3] ^= np.uint8(1 << (b & 7))
        descs.append(d)
    allD = np.concatenate(descs).astype(np.uint8)
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    exp = orc.distinctive_descriptors(allD, off)
    m = ORBmatcher(0.8, True, max_batch=1, max_keypoints=4096, max_queries=4096)
    out0 = np.full((len(sizes), 32), 0xAB, np.uint8)
    best, out = m.ComputeDistinctiveDescriptors(allD, off, out0)
    assert np.array_equal(best, exp)
    for i, n in enumerate(sizes):
        if n == 0:
            assert exp[i] == -1 and (out[i] == 0xAB).all()
        else:
            assert np.array_equal(out[i], descs[i][exp[i]])
    assert len(set(exp[np.array(sizes) > 4].tolist())) > 3          # the winner is not a fixed row
    bd, od = m.ComputeDistinctiveDescriptors(torch.from_numpy(allD).cuda(), torch.from_numpy(off).cuda())
    assert np.array_equal(bd.cpu().numpy(), exp)
    nz = np.array(sizes) > 0
    assert np.array_equal(od.cpu().numpy()[nz], out[nz])


def _bow_nodes(desc, n, seed):
    """a stand-in for the DBoW2 transform: features with the same leading descriptor bits share a node; a few have none"""
    node = 300 + 7 * (desc[:n, 0].astype(np.int32) & 0x3F)
    node[synth.uniform01(seed, n) < 0.03] = -1
    return node


@pytest.mark.parametrize("nnratio,check_ori", [(0.7, True), (0.9, False)])
def test_search_by_bow(orc, seq_frames, nnratio, check_ori):
    """ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches) (src/ORBmatcher.cc:724-853): walk in (node, index) order,
    claims, ratio test, rotation histogram -- matches per frame keypoint bit-exact"""
    from gf_orb_slam_b200 import ORBmatcher, Frames
    kps, desc, counts = seq_frames
    B = 3
    KF = Frames(kps[:B], desc[:B], counts[:B], W, H)
    F = Frames(kps[1:B + 1], desc[1:B + 1], counts[1:B + 1], W, H)
    kfmp = np.full((B, 1000), -1, np.int32); kn = np.full((B, 1000), -1, np.int32); fn = np.full((B, 1000), -1, np.int32)
    for b in range(B):
        n1, n2 = int(counts[b]), int(counts[b + 1])
        kfmp[b, :n1] = np.where(synth.uniform01(433 + b, n1) < 0.85, 9000 + np.arange(n1), -1)
        kn[b, :n1] = _bow_nodes(desc[b], n1, 533 + b); fn[b, :n2] = _bow_nodes(desc[b + 1], n2, 633 + b)
    out = np.full((B, 1000), -7, np.int32)
    m = ORBmatcher(nnratio, check_ori, max_batch=B, max_keypoints=1000, max_queries=1000)
    nm = m.SearchByBoW(KF, kfmp, kn, F, fn, out)
    tot = 0
    for b in range(B):
        n1, n2 = int(counts[b]), int(counts[b + 1])
        kf = _oframe(orc, kps[b], desc[b], n1); f = _oframe(orc, kps[b + 1], desc[b + 1], n2)
        kf.set_mappoints(kfmp[b, :n1])
        en, em = orc.search_by_bow(kf, kn[b, :n1], f, fn[b, :n2], nnratio, check_ori)
        assert nm[b] == en, (b, nm[b], en)
        assert np.array_equal(out[b, :n2], em) and (out[b, n2:] == -1).all()
        tot += en
    assert tot > 100


@pytest.mark.parametrize("th,orb_dist,check_ori", [(10.0, 100, True), (3.0, 64, False)])
def test_search_by_projection_keyframe(orc, seq_frames, th, orb_dist, check_ori):
    """ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (src/ORBmatcher.cc:2204-2336): level
    predicted from the distance to the camera centre, map-point descriptors, skip flags, pre-claimed keypoints"""
    from gf_orb_slam_b200 import ORBmatcher, Frames
    kps, desc, counts = seq_frames
    B = 3
    KF = Frames(kps[:B], desc[:B], counts[:B], W, H)
    Cur = Frames(kps[1:B + 1], desc[1:B + 1], counts[1:B + 1], W, H)
    S = 1000
    kfmp = np.full((B, S), -1, np.int32); skip = np.zeros((B, S), np.uint8); world = np.zeros((B, S, 3), np.float32)
    mind = np.ones((B, S), np.float32); mpd = np.zeros((B, S, 32), np.uint8); Tcw = np.zeros((B, 16), np.float32)
    cmp_ = np.full((B, S), -1, np.int32); csc = np.full((B, S), 999, np.int32)
    K = None
    scale = (np.float32(1.2) ** np.arange(8)).astype(np.float32)
    for b in range(B):
        n1, n2 = int(counts[b]), int(counts[b + 1])
        kfmp[b, :n1] = np.where(synth.uniform01(733 + b, n1) < 0.9, 5000 + np.arange(n1), -1)
        skip[b, :n1] = synth.uniform01(833 + b, n1) < 0.1
        K, T, w = _pose_and_world(kps[b], n1, seed=170 + b, shift=(1.0 + b, 1.5))
        world[b, :n1] = w; Tcw[b] = T.reshape(16)
        Ow = -(T[:3, :3].T.astype(np.float64) @ T[:3, 3].astype(np.float64))
        d3 = np.linalg.norm(w.astype(np.float64) - Ow, axis=1)
        jitter = 0.93 + 0.14 * synth.uniform01(933 + b, n1)               # puts the predicted level at the octave or next to it
        mind[b, :n1] = (d3 / (scale[np.clip(kps["octave"][b, :n1], 0, 7)] * jitter)).astype(np.float32)
        d = desc[b, :n1].copy()
        flip = synth.uniform01(1033 + b, n1 * 32).reshape(n1, 32) < 0.04   # the map point's descriptor differs by a few bits
        d[flip] ^= 0x10
        mpd[b, :n1] = d
        cmp_[b, :n2] = np.where(synth.uniform01(1133 + b, n2) < 0.1, 100 + np.arange(n2), -1)
    cmp0 = cmp_.copy()
    m = ORBmatcher(0.9, check_ori, max_batch=B, max_keypoints=S, max_queries=S)
    nm = m.SearchByProjectionKeyFrame(Cur, KF, kfmp, skip, world, mind, mpd, Tcw, K, th, orb_dist, cmp_, csc)
    tot = 0
    for b in range(B):
        n1, n2 = int(counts[b]), int(counts[b + 1])
        kf = _oframe(orc, kps[b], desc[b], n1); cur = _oframe(orc, kps[b + 1], desc[b + 1], n2)
        kf.set_mappoints(kfmp[b, :n1]); cur.set_mappoints(cmp0[b, :n2])
        cur.set_camera(*K, Tcw[b].reshape(4, 4))
        en = orc.search_by_projection_keyframe(cur, kf, skip[b, :n1], world[b, :n1], mind[b, :n1], mpd[b, :n1], th, orb_dist, check_ori)
        emp, esc = cur.matches()
        assert nm[b] == en, (b, nm[b], en)
        assert np.array_equal(cmp_[b, :n2], emp) and np.array_equal(csc[b, :n2], esc)
        tot += en
    assert tot > 100


def test_init_resolve_variants_agree(monkeypatch, seq_frames):
    """SearchForInitialization resolved query by query with the candidates over the lanes (default) == the speculative
    32-queries-per-round resolver of the other modes (GFORB_RESOLVE=spec), incl. a repeated call on the updated vbPrevMatched"""
    from gf_orb_slam_b200 import ORBmatcher, Frames
    kps, desc, counts = seq_frames
    B = 4
    F1 = Frames(kps[:B], desc[:B], counts[:B], W, H); F2 = Frames(kps[1:B + 1], desc[1:B + 1], counts[1:B + 1], W, H)
    res = []
    for v in ("spec", "lanes"):
        monkeypatch.setenv("GFORB_RESOLVE", v)
        m = ORBmatcher(0.9, True, max_batch=B, max_keypoints=1000, max_queries=1000)
        prev = np.stack([kps["x"][:B], kps["y"][:B]], axis=-1).astype(np.float32).copy()
        na, m12a = m.SearchForInitialization(F1, F2, prev, 100)
        m12a = np.array(m12a).copy(); na = np.array(na).copy(); p1 = prev.copy()
        nb, m12b = m.SearchForInitialization(F1, F2, prev, 30)
        res.append((m12a, na, p1, np.array(m12b).copy(), np.array(nb).copy(), prev.copy()))
    for x, y in zip(*res):
        assert np.array_equal(x, y)
    assert res[0][1].sum() > 300
