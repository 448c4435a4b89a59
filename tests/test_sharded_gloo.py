"""Host-side logic of the multi-GPU layer (gf_orb_slam_b200/sharded.py) on CPU: world_size-2 gloo process group, the
CUDA operators replaced by numpy restatements (checker only; the product defaults are the CUDA kernels)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gf_orb_slam_b200.sharded import ShardedDescriptorDB, db_shard, frame_shard, unpack_keys


def np_top2(q: np.ndarray, db: np.ndarray, base: int) -> np.ndarray:
    """brute-force packed top-2 keys (dist << 32 | global row), ties -> lower row"""
    lut = np.array([bin(i).count("1") for i in range(256)], np.int64)
    d = lut[(q[:, None, :] ^ db[None, :, :])].sum(-1)                       # n_q x n_db
    keys = (d << 32) | (np.arange(db.shape[0], dtype=np.int64)[None, :] + base)
    keys = np.sort(keys, axis=1)[:, :2]
    if keys.shape[1] < 2:
        keys = np.concatenate([keys, np.full((q.shape[0], 2 - keys.shape[1]), -1, np.int64)], 1)
    return keys


def np_merge(g: np.ndarray) -> np.ndarray:
    allk = g.transpose(1, 0, 2).reshape(g.shape[1], -1).astype(np.uint64)
    return np.sort(allk, axis=1)[:, :2].astype(np.int64)


def _worker(rank, world, port, q, db, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        row0, n = db_shard(db.shape[0], world, rank)
        shard = torch.from_numpy(db[row0:row0 + n].copy())
        sdb = ShardedDescriptorDB(shard, row0,
                                  local_top2=lambda qq, dd, base: torch.from_numpy(np_top2(qq.numpy(), dd.numpy(), base)),
                                  merge=lambda g: torch.from_numpy(np_merge(g.numpy())))
        # every rank "extracted" half of the queries: replicate, then search
        nq = q.shape[0] // world
        mine = torch.from_numpy(q[rank * nq:(rank + 1) * nq].copy())
        allq = sdb.gather_queries(mine)
        keys = sdb.top2(allq)
        ret[rank] = (allq.numpy().copy(), keys.numpy().copy())
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_sharded_top2_world2_gloo():
    rng = np.random.default_rng(5)
    q = rng.integers(0, 256, (64, 32), dtype=np.uint8)
    db = rng.integers(0, 256, (1001, 32), dtype=np.uint8)      # odd row count: shards of 501 and 500
    db[700] = q[3]                                              # exact hit in shard 1
    db[10] = q[3]; db[10, 0] ^= 1                               # distance-1 runner-up in shard 0
    db[20] = db[900]                                            # cross-shard tie -> lower global row wins
    q[5] = db[900]
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(2, _free_port(), q, db, ret), nprocs=2, join=True)
    want = np_top2(q, db, 0)
    for r in range(2):
        allq, keys = ret[r]
        assert np.array_equal(allq, q)
        assert np.array_equal(keys, want), r
    d, row = unpack_keys(want)
    assert d[3, 0] == 0 and row[3, 0] == 700 and d[3, 1] == 1 and row[3, 1] == 10
    assert d[5, 0] == 0 and row[5, 0] == 20 and row[5, 1] == 900


@pytest.mark.parametrize("n,world", [(2000, 1), (2000, 2), (2000, 8), (7, 8), (0, 4), (1001, 3)])
def test_frame_shard_covers_sequence(n, world):
    seen = []
    for r in range(world):
        first, own, halo = frame_shard(n, world, r)
        assert halo == (1 if (first > 0 and own > 0) else 0)
        seen += list(range(first, first + own))
    assert seen == list(range(n))
    sizes = [frame_shard(n, world, r)[1] for r in range(world)]
    assert max(sizes) - min(sizes) <= 1
