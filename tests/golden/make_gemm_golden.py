"""cv2 4.13 outputs of cv::gemm(Rcw, tcw, -1, GEMM_1_T) on CV_32F (the MatExpr `-Rcw.t()*tcw` of src/ORBmatcher.cc:2210)
for 256 random poses -> tests/golden/cv413_gemm_t.npz.  Run in the build container (cv2 is not needed by the tests)."""
import os

import cv2
import numpy as np

rng = np.random.default_rng(20260101)
N = 256
R = rng.standard_normal((N, 3, 3)).astype(np.float32)
t = (rng.standard_normal((N, 3)) * 10).astype(np.float32)
Ow = np.zeros((N, 3), np.float32)
for i in range(N):
    Ow[i] = cv2.gemm(R[i], t[i].reshape(3, 1), -1.0, None, 0.0, flags=cv2.GEMM_1_T).reshape(3)
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "cv413_gemm_t.npz"), R=R, t=t, Ow=Ow, cv2=cv2.__version__)
print("wrote", N, "cases with cv2", cv2.__version__)
