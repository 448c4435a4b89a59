"""GPU parity: Observability::runActiveMapMatching (src/Observability.cc:1249-1524) through the C ABI vs the CPU oracle --
bit-exact matches, scores, leftover list and match count; accumulated information matrix within 1e-12 relative."""
import numpy as np
import pytest

from gf_orb_slam_b200 import synth

pytestmark = pytest.mark.gpu

W, H = 752, 480
EUROC = dict(f=5.1369248, nRows=480, nCols=752, Cx=367.215, Cy=248.375, k1=0.0, k2=0.0, dx=0.01123325985, dy=0.01123325985)
XV = np.array([0.3, -0.1, 1.0, 0.9992, 0.0131, 0.0314, -0.0209, 3.0, -1.0, 10.0, 0.2618, 0.6283, -0.4189])


def _sigma2():
    s = np.ones(8, np.float32)
    for i in range(1, 8):
        s[i] = s[i - 1] * np.float32(1.2)
    return (s * s).astype(np.float32)


@pytest.fixture(scope="module")
def frame():
    from gf_orb_slam_b200 import ORBextractor
    img = synth.textured_frame(W, H, seed=1)
    ex = ORBextractor(1500, 1.2, 8, max_width=W, max_height=H, max_batch=1)
    kps, desc, counts = ex.extract_batch(img[None])
    return kps, desc, counts


def _problem(frame, M, seed, random_desc=False):
    kps, desc, counts = frame
    n = int(counts[0])
    mp = synth.map_points_c3(kps[0, :n], desc[0, :n], M, seed=seed, w=W, h=H)
    if random_desc:      # most one-point matches fail: exercises the pop / re-draw path
        mp["descriptors"] = synth.splitmix64(seed + 9, M * 4).view(np.uint8).reshape(M, 32).copy()
    world, _, _, _ = synth.landmark_cloud(M, seed + 1)
    return mp, world


@pytest.mark.parametrize("M,num,th,nnratio,random_desc", [
    (2000, 100, 1.0, 0.8, False),      # the default-compiled GF mode at local-map scale
    (3000, 60, 3.0, 0.8, False),       # larger random subsets (50 per round), wider windows
    (1500, 120, 1.0, 0.6, True),       # mostly failing matches
    (60, 20, 1.0, 0.8, False),         # tiny pool: "went through all map points"
    (50, 100, 1.0, 0.8, False),        # szLazierSubset == 0: "early termination" in round 0
    (300, 0, 1.0, 0.8, False),         # num_to_match <= 0: every visible point is left over
])
def test_run_active_map_matching(orc, frame, M, num, th, nnratio, random_desc):
    from gf_orb_slam_b200 import Frames, MapPoints, Observability, ORBmatcher
    from gf_orb_slam_b200.observability import MAP_INFO_MATRIX
    kps, desc, counts = frame
    n = int(counts[0])
    B = 3
    probs = [_problem(frame, M, 70 + 11 * b, random_desc) for b in range(B)]
    stack = lambda k: np.ascontiguousarray(np.stack([probs[b][0][k] for b in range(B)]))
    mcounts = np.array([M, M - 7, M], np.int32)
    mps = MapPoints(stack("proj_x"), stack("proj_y"), stack("scale_level"), stack("view_cos"), stack("flags"),
                    stack("descriptors"), mcounts)
    world = np.ascontiguousarray(np.stack([p[1] for p in probs]))
    kB = np.repeat(kps, B, 0); dB = np.repeat(desc, B, 0); cB = np.repeat(counts, B)
    F = Frames(kB, dB, cB, W, H)
    ob = Observability(**EUROC, max_batch=B, max_points=M)
    ob.set_state(XV); ob.predictPWLSVec(0.1, 1)
    Xv = np.repeat(ob.kinematic(0)["Xv"][None], B, 0)
    tv = np.ascontiguousarray(stack("flags") & 1)
    blocks, osc, uv, Hm = ob.build_info_map(Xv, world, tv, mcounts, 0, (0, W, 0, H), want_H=True)
    osc[2, ::5] = -1.0                  # frame 2: a fifth of the points were not refreshed at this frame
    fmp = np.full((B, 1500), -1, np.int32); fsc = np.full((B, 1500), 999, np.int32)
    fmp[1, 3:600:4] = 777777            # frame 1 arrives with keypoints already matched (TrackWithMotionModel)
    fmp0 = fmp.copy()
    base = np.zeros((B, 49)); base[:, ::8] = 1e-5
    base[2] = (np.eye(7) * 0.5 + 0.01).reshape(49)
    seeds = np.array([1, 1, 12345], np.uint32)
    m = ORBmatcher(nnratio, True, max_batch=B, max_keypoints=1500, max_queries=M, cand_pool=64 * M)
    nm, left, left_n, info = ob.runActiveMapMatching(F, mps, MAP_INFO_MATRIX, base, th, m, num, obs_score=osc, obs_mat=blocks,
                                                     H_meas=Hm, level_sigma2=_sigma2(), frame_mappoints=fmp, match_score=fsc,
                                                     seeds=seeds)
    total = 0
    for b in range(B):
        Mb = int(mcounts[b])
        fo = orc.Frame(kps[0, :n], desc[0, :n], W, H)
        fo.set_mappoints(fmp0[b, :n])
        p = probs[b][0]
        en, eleft, einfo = orc.active_match(fo, p["proj_x"][:Mb], p["proj_y"][:Mb], p["scale_level"][:Mb], p["view_cos"][:Mb],
                                            p["flags"][:Mb], p["descriptors"][:Mb], (osc[b, :Mb] > 0).astype(np.uint8),
                                            blocks[b, :Mb], Hm[b, :Mb], base[b], th, nnratio, num, seed=int(seeds[b]))
        emp, esc = fo.matches()
        assert nm[b] == en, (b, nm[b], en)
        assert np.array_equal(fmp[b, :n], emp) and np.array_equal(fsc[b, :n], esc), b
        assert left_n[b] == len(eleft) and np.array_equal(left[b, :left_n[b]], eleft), b
        assert np.allclose(info[b].reshape(7, 7), einfo, rtol=1e-12, atol=0), b
        total += en
    if num > 0 and M >= 1000 and not random_desc:
        assert total > 0.5 * B * num          # the workload really matches


def test_active_match_rejects_hybrid(frame):
    from gf_orb_slam_b200 import Frames, MapPoints, Observability, ORBmatcher
    from gf_orb_slam_b200._lib import GforbError
    from gf_orb_slam_b200.observability import FRAME_HYBRID_MATRIX
    kps, desc, counts = frame
    M = 64
    mp, world = _problem(frame, M, 5)
    mps = MapPoints(*[mp[k][None] for k in ("proj_x", "proj_y", "scale_level", "view_cos", "flags", "descriptors")],
                    np.array([M], np.int32))
    F = Frames(kps, desc, counts, W, H)
    ob = Observability(**EUROC, max_batch=1, max_points=M)
    m = ORBmatcher(0.8, True, max_batch=1, max_keypoints=1500, max_queries=M)
    z = np.zeros
    with pytest.raises(GforbError):
        ob.runActiveMapMatching(F, mps, FRAME_HYBRID_MATRIX, z((1, 49)), 1.0, m, 10, obs_score=z((1, M)), obs_mat=z((1, M, 49)),
                                H_meas=z((1, M, 14)), level_sigma2=_sigma2(), frame_mappoints=np.full((1, 1500), -1, np.int32),
                                match_score=np.full((1, 1500), 999, np.int32))
