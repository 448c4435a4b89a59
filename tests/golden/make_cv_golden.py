"""Generates tests/golden/cv413_primitives.npz with OpenCV 4.13 (python cv2, the only OpenCV in the image) outputs
for the primitives ORBextractor relies on, on a small deterministic frame.  Run here (cv2 present); the fixture is
committed so the oracle stays pinned even where cv2 is unavailable."""
import os, sys
import numpy as np
import cv2

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from gf_orb_slam_b200 import synth

cv2.setNumThreads(1)
img = synth.textured_frame(200, 152, seed=11)
out = {"img": img, "cv_version": np.array(cv2.__version__)}
out["resize_167x127"] = cv2.resize(img, (167, 127), interpolation=cv2.INTER_LINEAR)
out["resize_139x106"] = cv2.resize(out["resize_167x127"], (139, 106), interpolation=cv2.INTER_LINEAR)
out["border16"] = cv2.copyMakeBorder(img, 16, 16, 16, 16, cv2.BORDER_REFLECT_101)
out["blur7"] = cv2.GaussianBlur(img, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
for t in (20, 7):
    fd = cv2.FastFeatureDetector_create(t, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    k = fd.detect(img)
    out[f"fast{t}"] = np.array([(p.pt[0], p.pt[1], p.response) for p in k], np.float32).reshape(-1, 3)
rng = np.random.default_rng(5)
y = rng.integers(-300000, 300000, 4096).astype(np.float32)
x = rng.integers(-300000, 300000, 4096).astype(np.float32)
y[:64] = 0; x[32:96] = 0
out["atan_y"], out["atan_x"] = y, x
out["atan"] = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(y, x)], np.float32)
np.savez_compressed(os.path.join(os.path.dirname(__file__), "cv413_primitives.npz"), **out)
print("written", {k: getattr(v, "shape", None) for k, v in out.items()})
