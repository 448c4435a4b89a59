"""Generates tests/golden/cv413_undistort.npz: cv2.undistortPoints (OpenCV 4.13, the only OpenCV in the image) on keypoint
positions with the call shape of Frame::UndistortKeyPoints (src/Frame.cc:389-423: R empty, P = K, CV_32F K and
coefficients).  Run here (cv2 present); the fixture is committed so the oracle stays pinned where cv2 is absent."""
import os
import numpy as np
import cv2

rng = np.random.default_rng(17)
out = {"cv_version": np.array(cv2.__version__)}
cams = {
    "euroc": (458.654, 457.296, 367.215, 248.375, [-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05]),
    "tum1": (517.306408, 516.469215, 318.643040, 255.313989, [0.262383, -0.953104, -0.005358, 0.002628, 1.163314]),
    "k1only": (700.0, 700.0, 320.0, 240.0, [-0.1, 0.0, 0.0, 0.0]),
}
for name, (fx, fy, cx, cy, d) in cams.items():
    w, h = (752, 480) if name == "euroc" else (640, 480)
    n = 3000
    # keypoint-like positions: integer level-0 pixels and integer pixels scaled by 1.2^l in float (pt *= scale)
    lv = rng.integers(0, 8, n)
    sc = np.float32(1.0) * np.power(np.float32(1.2), lv).astype(np.float32)
    xy = np.stack([rng.integers(16, w - 16, n) / sc, rng.integers(16, h - 16, n) / sc], 1).astype(np.int32).astype(np.float32)
    xy = (xy * sc[:, None]).astype(np.float32)
    xy[:4] = [[0, 0], [w - 1, 0], [0, h - 1], [w - 1, h - 1]]
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
    D = np.array(d, np.float32).reshape(-1, 1)
    und = cv2.undistortPoints(xy.reshape(-1, 1, 2), K, D, None, K).reshape(-1, 2)
    out[name + "_xy"], out[name + "_K"], out[name + "_D"], out[name + "_und"] = xy, K, D.reshape(-1), und.astype(np.float32)
np.savez_compressed(os.path.join(os.path.dirname(__file__), "cv413_undistort.npz"), **out)
print("written", {k: getattr(v, "shape", None) for k, v in out.items()})
