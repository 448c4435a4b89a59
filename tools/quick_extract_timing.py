"""Developer timing probe: per-stage device times of the batched extractor (not the contract bench)."""
import sys, time
import numpy as np
import torch
sys.path.insert(0, '.')
from gf_orb_slam_b200 import ORBextractor, synth, KP_DTYPE

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
seq = synth.Sequence()
imgs = seq.frames(0, min(B, 16))
imgs = np.concatenate([imgs] * ((B + len(imgs) - 1) // len(imgs)))[:B]
ex = ORBextractor(1000, 1.2, 8, max_batch=B)
d_img = torch.from_numpy(imgs).cuda()
d_kps = torch.zeros((B, 1000, 7), dtype=torch.float32, device='cuda')
d_desc = torch.zeros((B, 1000, 32), dtype=torch.uint8, device='cuda')
d_cnt = torch.zeros(B, dtype=torch.int32, device='cuda')
st = torch.cuda.current_stream().cuda_stream
for _ in range(3):
    ex.extract_device(d_img, d_kps, d_desc, d_cnt, stream=st)
ex.set_profiling(True)
ex.extract_device(d_img, d_kps, d_desc, d_cnt, stream=st)
ex.extract_device(d_img, d_kps, d_desc, d_cnt, stream=st)
tot = 0
for name, ms in ex.stage_times():
    print(f"{name:10s} {ms:8.3f} ms  ({ms / B * 1e3:7.1f} us/frame)")
    tot += ms
print(f"total {tot:.3f} ms -> {B / tot * 1e3:.0f} frames/s; counts {d_cnt[:4].tolist()}; launches {ex.last_launches()}")
ex.set_profiling(False)
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    ex.extract_device(d_img, d_kps, d_desc, d_cnt, stream=st, sync=False)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(f"event-timed {ms:.3f} ms/batch -> {B / ms * 1e3:.0f} frames/s")
