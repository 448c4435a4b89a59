#!/usr/bin/env python
"""Config 5 on one GPU, A/B of the two all-pairs Hamming top-2 kernels: tensor cores (tcgen05 kind::i8, default) against
the integer pipe (GFORB_ALLPAIRS=popc).  2000 queries x 2^20 database rows (BASELINE.json config 5 shape; --nq / --ndb).
Keys must be bit-identical; prints one JSON line with CUDA-event times."""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gf_orb_slam_b200.matcher import ORBmatcher  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nq", type=int, default=2000)
    ap.add_argument("--ndb", type=int, default=1 << 20)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--only", default="")
    a = ap.parse_args()
    g = torch.Generator().manual_seed(42)
    q = torch.randint(0, 256, (a.nq, 32), dtype=torch.uint8, generator=g).cuda()
    db = torch.randint(0, 256, (a.ndb, 32), dtype=torch.uint8, generator=g).cuda()
    m = ORBmatcher(0.9, True, max_batch=1, max_keypoints=max(a.nq, 16), max_queries=max(a.nq, 16))
    st = torch.cuda.current_stream().cuda_stream
    out, res = {}, {}
    for kern in ("tc", "popc"):
        if a.only and kern != a.only:
            continue
        os.environ["GFORB_ALLPAIRS"] = kern
        keys = m.AllPairsTop2(q, db)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.iters):
            keys = m.AllPairsTop2(q, db, stream=st, sync=False)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.iters
        res[kern] = keys.clone()
        out[kern] = {"ms": ms, "pairs_per_s": a.nq * a.ndb / (ms * 1e-3), "int8_tera_ops": 2 * 256 * a.nq * a.ndb / (ms * 1e-3) / 1e12}
    if len(res) == 2:
        out["keys_identical"] = bool(torch.equal(res["tc"], res["popc"]))
        out["speedup_tc_over_popc"] = out["popc"]["ms"] / out["tc"]["ms"]
    out["nq"], out["ndb"] = a.nq, a.ndb
    print(json.dumps(out))
    if len(res) == 2 and not out["keys_identical"]:
        sys.exit(1)


if __name__ == "__main__":
    main()
