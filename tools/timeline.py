"""Poor man's timeline of the front-end streams (no nsys in the image): CUDA events on the main, matcher and selection
streams of a few steady-state steps, printed relative to the first extraction start, plus the extractor's own stage
durations laid end to end from each extract_begin.  usage: python tools/timeline.py [batch] [steps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from gf_orb_slam_b200 import synth
from gf_orb_slam_b200.pipeline import LanedFrontEnd


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    fe = LanedFrontEnd(752, 480, 1000, B, lanes=1, n_good=100)
    fe.load_host_frames(synth.Sequence(752, 480).frames(0, min(B, 64)).repeat((B + 63) // 64, 0)[:B])
    for _ in range(5):
        fe.step_device()
    fe.sync(); torch.cuda.synchronize()
    ln = fe.lane[0]
    ln.ex.set_profiling(True)
    ln.timeline = []
    for _ in range(steps):
        fe.step_device()
    fe.sync(); torch.cuda.synchronize()
    stages = ln.ex.stage_times()
    tl = ln.timeline
    t0 = [e for n, t, e in tl if n == "extract_begin"][0]
    print("extractor stages (ms, average):", ", ".join("%s %.3f" % (n, ms) for n, ms in stages))
    for n, t, e in sorted(tl, key=lambda r: t0.elapsed_time(r[2])):
        print("%8.3f ms  step %d  %s" % (t0.elapsed_time(e), t, n))


if __name__ == "__main__":
    main()
