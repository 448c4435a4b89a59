"""The reference's own acceptance test for good-feature selection, test/test_Greedy.cpp:209-294 (TEST_F(TestLazierGreedy,
Selection)), restated: 320 random visible landmarks (fixture :74-206), k = 60, 70 ... 140; the baseline greedy
(greedy_mtd 1) gives the base subset, then 100 repeats of the lazier greedy (greedy_mtd 3, error_bound 0.1 ->
random_sample_scale = (size_t)(float(N)/k * ln 10), src/Observability.cc:826) on a continuing rand() stream must each
differ from it by at most ceil(0.2 k) landmarks (set_difference of the sorted index lists, EXPECT_NEAR(diff, 0, ...)).
The reference's fixture draws from an unseeded rand(); here the cloud comes from splitmix64 (synth.landmark_cloud) and
the 100 repeats use rand() seeds 1..100.  Run on the CPU oracle (pins the restatement with the reference's own
property) and on the GPU (same property + subsets equal to the oracle's)."""
import math

import numpy as np
import pytest

from gf_orb_slam_b200 import synth
from gf_orb_slam_b200.pipeline_consts import EUROC_CAM, SIMU_XV, level_sigma2, sample_scale_for, simu_visible_from

N_LMK, REPEATS = 320, 100
KS = list(range(60, 141, 10))


def _fixture(orc, seed=11):
    oo = orc.Observability(**EUROC_CAM)
    oo.set_xv(SIMU_XV); oo.predict(0.1, 1)
    world, octave, score, _ = synth.landmark_cloud(N_LMK, seed, visible_from=simu_visible_from())
    blocks, obs, _ = oo.build_frame(world, octave, score, np.ones(N_LMK, np.uint8), level_sigma2())
    return blocks.reshape(N_LMK, 49), obs


def _diff(base, cur):
    return len(set(base.tolist()) - set(cur.tolist()))


def test_oracle_lazier_vs_baseline_greedy(orc):
    blocks, obs = _fixture(orc)
    for k in KS:
        base, _ = orc.select(blocks.reshape(-1, 7, 7), k, 1, 0.0, 1, scores=obs)
        assert len(base) == k
        scale = sample_scale_for(N_LMK, k, 0.1)
        worst = 0
        for rep in range(1, REPEATS + 1):
            cur, _ = orc.select(blocks.reshape(-1, 7, 7), k, 3, scale, rep, scores=obs)
            assert len(cur) == k
            worst = max(worst, _diff(base, cur))
        assert worst <= math.ceil(0.2 * k), (k, worst)


@pytest.mark.gpu
def test_gpu_lazier_vs_baseline_greedy(orc):
    from gf_orb_slam_b200 import Observability
    blocks, obs = _fixture(orc)
    ob = Observability(**EUROC_CAM, max_batch=REPEATS, max_points=N_LMK)
    B = REPEATS
    bb = np.ascontiguousarray(np.broadcast_to(blocks, (B, N_LMK, 49)))
    oo = np.ascontiguousarray(np.broadcast_to(obs, (B, N_LMK)))
    counts = np.full(B, N_LMK, np.int32)
    seeds = np.arange(1, B + 1, dtype=np.uint32)
    for k in KS:
        bidx, _, bn = ob.select(bb[:1], oo[:1], counts[:1], k, 1, 0.0, seeds[:1])
        assert bn[0] == k
        base = bidx[0, :k]
        ebase, _ = orc.select(blocks.reshape(-1, 7, 7), k, 1, 0.0, 1, scores=obs)
        assert set(base.tolist()) == set(ebase.tolist())
        scale = sample_scale_for(N_LMK, k, 0.1)
        idx, _, cnt = ob.select(bb, oo, counts, k, 3, scale, seeds)
        worst = 0
        for rep in range(B):
            assert cnt[rep] == k
            worst = max(worst, _diff(base, idx[rep, :k]))
        assert worst <= math.ceil(0.2 * k), (k, worst)
        for rep in (0, 17, 99):                                   # and the GPU subsets are the oracle's
            cur, _ = orc.select(blocks.reshape(-1, 7, 7), k, 3, scale, rep + 1, scores=obs)
            assert np.array_equal(np.sort(idx[rep, :k]), np.sort(cur)), (k, rep)
