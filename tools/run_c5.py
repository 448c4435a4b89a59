#!/usr/bin/env python
"""BASELINE.json config 5 on N GPUs (torchrun): KITTI-shaped 1241x376 frames, 2000 keypoints per frame, all-pairs Hamming
top-2 of the frame descriptors against a 1 048 576-row descriptor database sharded by rows across the ranks; the only
collectives are the NCCL all-gathers of the query block and of the 16-byte top-2 keys.
Checks: (a) sharded result == single-shard result computed by rank 0 over the whole database (bit-exact keys),
(b) 16 sampled queries against a numpy brute force.  Prints one JSON line with timings (CUDA events, max over ranks)."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gf_orb_slam_b200 import synth  # noqa: E402
from gf_orb_slam_b200.extractor import ORBextractor  # noqa: E402
from gf_orb_slam_b200.matcher import ORBmatcher  # noqa: E402
from gf_orb_slam_b200.sharded import ShardedDescriptorDB, db_shard  # noqa: E402


def main():
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    W, H, NF, NDB = 1241, 376, 2000, 1 << 20
    steps = int(os.environ.get("C5_STEPS", "5"))
    ex = ORBextractor(NF, 1.2, 8, max_width=W, max_height=H, max_batch=1, device=local)
    m = ORBmatcher(0.9, True, max_batch=1, max_keypoints=NF, max_queries=NF, device=local)
    img = synth.textured_frame(W, H, seed=100 + rank)
    dimg = torch.from_numpy(img[None]).cuda()
    kps = torch.zeros((1, NF, 7), device="cuda"); desc = torch.zeros((1, NF, 32), dtype=torch.uint8, device="cuda")
    cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
    g = torch.Generator().manual_seed(42)
    full = torch.randint(0, 256, (NDB, 32), dtype=torch.uint8, generator=g)        # identical on every rank
    row0, n = db_shard(NDB, world, rank)
    shard = full[row0:row0 + n].cuda()
    sdb = ShardedDescriptorDB(shard, row0, matcher=m)
    st = torch.cuda.current_stream().cuda_stream

    def step():
        ex.extract_device(dimg, kps, desc, cnt, stream=st, sync=False)
        q = sdb.gather_queries(desc[0])                 # world * 2000 queries (rows past counts are zero descriptors)
        return q, sdb.top2(q)

    for _ in range(3):
        q, keys = step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        q, keys = step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / steps], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ok_full = ok_np = None
    if rank == 0:
        ref = m.AllPairsTop2(q, full.cuda(), idx_base=0)
        ref = m.Top2Merge(ref.reshape(1, *ref.shape))
        ok_full = bool(torch.equal(ref, keys))
        qs = q[:: max(1, q.shape[0] // 16)][:16].cpu().numpy(); sel = list(range(0, q.shape[0], max(1, q.shape[0] // 16)))[:16]
        lut = np.array([bin(i).count("1") for i in range(256)], np.int64)
        fn = full.numpy()
        ok_np = True
        for t, qi in zip(sel, qs):
            d = lut[fn ^ qi[None, :]].sum(1)
            k = np.sort((d << 32) | np.arange(NDB, dtype=np.int64))[:2]
            ok_np = ok_np and np.array_equal(k, keys[t].cpu().numpy())
        print(json.dumps({"config": "C5", "n_gpus": world, "frames_per_step": world, "queries_per_step": int(q.shape[0]),
                          "db_rows": NDB, "ms_per_step": float(ms.item()), "frames_per_s": world / (float(ms.item()) * 1e-3),
                          "pairs_per_s": q.shape[0] * NDB / (float(ms.item()) * 1e-3), "keypoints_rank0": int(cnt.item()),
                          "sharded_equals_single_shard": ok_full, "sampled_equals_numpy": bool(ok_np)}))
    if world > 1:
        dist.destroy_process_group()
    if rank == 0 and not (ok_full and ok_np):
        sys.exit(1)


if __name__ == "__main__":
    main()
