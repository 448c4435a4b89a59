"""Times the good-feature selection kernel alone (config C4: 1000 matched points -> lazier-greedy pick of 100, eps 0.1)
for a batch of problems, and prints a checksum of the picked subsets (regression guard while tuning).
usage: python tools/bench_select.py [batch] [reps]"""
import hashlib
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from gf_orb_slam_b200 import synth
from gf_orb_slam_b200.observability import LAZIER_GREEDY, Observability
from gf_orb_slam_b200.pipeline_consts import EUROC_CAM, SIMU_XV, level_sigma2, sample_scale_for, simu_visible_from


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    nf, k = 1000, 100
    dev = torch.device("cuda", 0)
    cam = dict(EUROC_CAM)
    obs = Observability(**cam, max_batch=B, max_points=nf, device=0)
    obs.set_state(SIMU_XV); obs.predictPWLSVec(0.1, 1)
    xv = obs.kinematic(0)["Xv"]
    worlds, scores, octs = [], [], []
    for b in range(B):
        world, octave, score, _ = synth.landmark_cloud(nf, 7 + (b % 16), visible_from=simu_visible_from(752, 480))
        worlds.append(world); scores.append(score); octs.append(octave)
    world = torch.from_numpy(np.stack(worlds)).to(dev)
    score = torch.from_numpy(np.stack(scores)).to(dev)
    octave = torch.from_numpy(np.stack(octs).astype(np.int32)).to(dev)
    valid = torch.ones((B, nf), dtype=torch.uint8, device=dev)
    counts = torch.full((B,), nf, dtype=torch.int32, device=dev)
    Xv = torch.from_numpy(np.broadcast_to(xv, (B, 13)).copy()).to(dev)
    sig = torch.from_numpy(level_sigma2()).to(dev)
    blocks, osc = obs.build_info_frame(Xv, world, octave, score, valid, counts, sig)
    seeds = torch.ones(B, dtype=torch.int32, device=dev)
    ss = sample_scale_for(nf, k, 0.1)
    out = None
    for _ in range(2):
        out = obs.select(blocks, osc, counts, k, LAZIER_GREEDY, ss, seeds)
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = obs.select(blocks, osc, counts, k, LAZIER_GREEDY, ss, seeds, sync=False)
    e1.record(); torch.cuda.synchronize()
    idx = out[0].cpu().numpy(); ld = out[1].cpu().numpy()
    print("select batch=%d: %.3f ms per launch; subset sha=%s logdet[-1]=%.12f" % (
        B, e0.elapsed_time(e1) / reps, hashlib.sha1(idx.tobytes()).hexdigest()[:12], ld[0, -1]))


if __name__ == "__main__":
    main()
