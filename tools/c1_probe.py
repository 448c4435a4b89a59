import sys, time
sys.path.insert(0, '.')
import numpy as np, torch
from gf_orb_slam_b200 import ORBextractor, synth
img = synth.textured_frame(640, 480, seed=3)
ex = ORBextractor(1000, 1.2, 8, max_width=640, max_height=480, max_batch=1, device=0)
for _ in range(5): ex.extract_batch(img[None])
n=200
t0=time.perf_counter()
for _ in range(n): ex.extract_batch(img[None])
print("host pointers: %.1f us" % ((time.perf_counter()-t0)/n*1e6))
d_img = torch.from_numpy(img[None]).cuda()
d_kps = torch.zeros((1, 1000, 7), dtype=torch.float32, device='cuda'); d_desc = torch.zeros((1, 1000, 32), dtype=torch.uint8, device='cuda'); d_cnt = torch.zeros(1, dtype=torch.int32, device='cuda')
st = torch.cuda.current_stream().cuda_stream
for _ in range(5): ex.extract_device(d_img, d_kps, d_desc, d_cnt, stream=st)
torch.cuda.synchronize()
t0=time.perf_counter()
for _ in range(n):
    ex.extract_device(d_img, d_kps, d_desc, d_cnt, stream=st); torch.cuda.synchronize()
print("device pointers + sync each: %.1f us" % ((time.perf_counter()-t0)/n*1e6))
e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(n): ex.extract_device(d_img, d_kps, d_desc, d_cnt, stream=st)
e1.record(); torch.cuda.synchronize()
print("device pointers back-to-back (GPU time per frame): %.1f us" % (e0.elapsed_time(e1)/n*1e3))
ex.set_concurrency(False) if hasattr(ex,'set_concurrency') else None
t0=time.perf_counter()
for _ in range(n):
    ex.extract_device(d_img, d_kps, d_desc, d_cnt, stream=st); torch.cuda.synchronize()
print("no blur fork, sync each: %.1f us" % ((time.perf_counter()-t0)/n*1e6))
