"""cv2 4.13 outputs of the CV_32F matrix products the matchers evaluate through cv::MatExpr -> tests/golden/*.npz.
Run in the build container (cv2 is not needed by the tests).

cv413_gemm_small.npz (what the oracle and the kernels restate):
  * y  = cv::gemm(R, x, 1, t, 1)            -- `Rcw*x3Dw+tcw` (src/ORBmatcher.cc:169, :892, :1618, :1746, :2104, :2230,
                                               src/Frame.cc:174): MatOp_GEMM::add folds the sum into one gemm; no transpose
                                               flag and len = 3 -> cv::gemm's small-matrix path (float accumulation)
  * Ow = cv::gemm(R.t() materialised, t, -1) -- `-Rcw.t()*tcw` (:871, :1723, :2210): operator-(MatExpr) materialises the
                                               transpose and scales it by -1, MatOp::matmul turns that into a plain gemm
cv413_gemm_t.npz: cv::gemm(R, t, -1, GEMM_1_T), the general (double-accumulating) path -- kept to show that the two paths
differ; nothing on the path uses it."""
import os

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
rng = np.random.default_rng(20260101)
N = 256
R = rng.standard_normal((N, 3, 3)).astype(np.float32)
t = (rng.standard_normal((N, 3)) * 10).astype(np.float32)
Ow = np.zeros((N, 3), np.float32)
for i in range(N):
    Ow[i] = cv2.gemm(R[i], t[i].reshape(3, 1), -1.0, None, 0.0, flags=cv2.GEMM_1_T).reshape(3)
np.savez_compressed(os.path.join(HERE, "cv413_gemm_t.npz"), R=R, t=t, Ow=Ow, cv2=cv2.__version__)

rng = np.random.default_rng(20260202)
N = 2048
R = rng.standard_normal((N, 3, 3)).astype(np.float32)
x = (rng.standard_normal((N, 3)) * 5).astype(np.float32)
t = (rng.standard_normal((N, 3)) * 3).astype(np.float32)
y = np.zeros((N, 3), np.float32); Ow = np.zeros((N, 3), np.float32)
for i in range(N):
    y[i] = cv2.gemm(R[i], x[i].reshape(3, 1), 1.0, t[i].reshape(3, 1), 1.0).reshape(3)
    Ow[i] = cv2.gemm(np.ascontiguousarray(R[i].T), t[i].reshape(3, 1), -1.0, None, 0.0).reshape(3)
np.savez_compressed(os.path.join(HERE, "cv413_gemm_small.npz"), R=R, x=x, t=t, y=y, Ow=Ow, cv2=cv2.__version__)
print("wrote 256 + 2048 cases with cv2", cv2.__version__)
