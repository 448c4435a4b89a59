// Shared device helpers for the gforb kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <float.h>
#include "../../include/gforb.h"
#include "glibc_sincosf.h"
#include "introselect.h"

namespace gforb {

#define GF_CUDA_OK(expr)                                                                     \
    do {                                                                                     \
        cudaError_t _e = (expr);                                                             \
        if (_e != cudaSuccess) {                                                             \
            fprintf(stderr, "[gforb] CUDA error %s at %s:%d: %s\n", cudaGetErrorName(_e), __FILE__, __LINE__, \
                    cudaGetErrorString(_e));                                                 \
            return GFORB_ERR_CUDA;                                                           \
        }                                                                                    \
    } while (0)

static const int EDGE = 16;          // EDGE_THRESHOLD (reference src/ORBextractor.cc:84)
static const int HALF_PATCH = 15;    // HALF_PATCH_SIZE
static const int MAX_LEVELS = 16;

// cvRound(float): round-half-even
__device__ __forceinline__ int cv_round(float v) { return __float2int_rn(v); }

// cv::fastAtan2 (degrees, [0,360)); every op explicitly un-fused so it matches the scalar CPU sequence
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
    const float s = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * s, p3 = -0.3258083974640975f * s;
    const float p5 = 0.1555786518463281f * s, p7 = -0.04432655554792128f * s;
    float ax = fabsf(x), ay = fabsf(y), a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, (float)DBL_EPSILON));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, (float)DBL_EPSILON));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

__device__ __forceinline__ int popc256(const uint4 a0, const uint4 a1, const uint4 b0, const uint4 b1) {
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// cv::Mat arithmetic of the matchers on CV_32F 3x3 / 3x1 operands (restated from OpenCV's matop.cpp / matmul.cpp; pinned
// against cv2 4.13 in tests/golden/cv413_gemm_small.npz).  `R*x + t` is ONE cv::gemm(R, x, 1, t, 1) (MatOp_GEMM::add folds the
// sum in); with no transpose flag and len = 3 cv::gemm takes its small-matrix path: the row's dot product accumulates in
// float, left to right, and d = (float)(t0*alpha + c*beta) is evaluated in double.
__device__ __forceinline__ float cv_gemm_row(float a0, float a1, float a2, float x0, float x1, float x2, float c) {
    const float t0 = __fadd_rn(__fadd_rn(__fmul_rn(a0, x0), __fmul_rn(a1, x1)), __fmul_rn(a2, x2));
    return (float)((double)t0 * 1.0 + (double)c * 1.0);
}
// one entry of `-R.t()*t`: the MatExpr materialises the transpose, scales it by -1 and multiplies: cv::gemm(Rt, t, -1), the
// same small-matrix path with c = 0.  a0..a2 = column i of R.
__device__ __forceinline__ float cv_neg_rt_t_row(float a0, float a1, float a2, float t0, float t1, float t2) {
    const float s = __fadd_rn(__fadd_rn(__fmul_rn(a0, t0), __fmul_rn(a1, t1)), __fmul_rn(a2, t2));
    return (float)((double)s * -1.0 + 0.0);
}

__device__ __forceinline__ uint4 ldg_nc_u4(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

}  // namespace gforb
