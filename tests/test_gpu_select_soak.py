"""Selection soak: 2048 C4-shaped lazier-greedy problems (1000 -> 100, eps 0.1 => sampleScale 23; 64 landmark clouds x
32 rand() seeds, seeds 1..2048 all distinct): the GPU subset must equal the CPU oracle's pick for pick, and the running
log-dets bit for bit (the arg-max compares glibc-log values with strict `>`, src/Observability.cc:2950-2960; the device
restates that log, csrc/glibc_log.h)."""
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest

from gf_orb_slam_b200 import synth
from gf_orb_slam_b200.pipeline_consts import EUROC_CAM, SIMU_XV, level_sigma2, simu_visible_from

pytestmark = pytest.mark.gpu

N, K, SCALE, CLOUDS, SEEDS_PER = 1000, 100, 23.0, 64, 32


def test_selection_soak_2048_problems(orc):
    from gf_orb_slam_b200 import Observability
    oo = orc.Observability(**EUROC_CAM)
    oo.set_xv(SIMU_XV); oo.predict(0.1, 1)
    pools = []
    for c in range(CLOUDS):
        world, octave, score, _ = synth.landmark_cloud(N, 5000 + c, visible_from=simu_visible_from())
        blocks, obs, _ = oo.build_frame(world, octave, score, np.ones(N, np.uint8), level_sigma2())
        pools.append((blocks.reshape(N, 49), obs))
    B = 256
    ob = Observability(**EUROC_CAM, max_batch=B, max_points=N)
    total = 0
    for chunk in range(CLOUDS * SEEDS_PER // B):
        ids = np.arange(chunk * B, (chunk + 1) * B)
        cloud = ids // SEEDS_PER
        seeds = (ids + 1).astype(np.uint32)
        bb = np.stack([pools[c][0] for c in cloud]); sc = np.stack([pools[c][1] for c in cloud])
        idx, ld, cnt = ob.select(bb, sc, np.full(B, N, np.int32), K, 2, SCALE, seeds)

        def ref(j):
            return orc.select(pools[cloud[j]][0].reshape(N, 7, 7), K, 2, SCALE, int(seeds[j]), scores=pools[cloud[j]][1])
        with ThreadPoolExecutor(8) as ex:
            exp = list(ex.map(ref, range(B)))
        for j in range(B):
            ei, el = exp[j]
            assert cnt[j] == K == len(ei)
            assert np.array_equal(idx[j, :K], ei), ("subset", int(ids[j]))
            assert np.array_equal(ld[j, :K], el), ("log-det", int(ids[j]))
        total += B
    assert total == 2048
