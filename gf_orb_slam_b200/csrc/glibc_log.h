// Bit-compatible double log() for the good-feature selection.
//
// logDet (include/Util.hpp:714-731) returns 2*log(prod diag L) and the greedy loops compare those doubles with strict `>`
// (src/Observability.cc:2950-2960: first drawn wins ties), so a 1-ulp difference between CUDA's log and the host libm's
// could turn a CPU tie into a GPU non-tie and change the selected subset.  This restates the algorithm glibc >= 2.28 uses
// for `log` (sysdeps/ieee754/dbl-64/e_log.c; the ARM optimized-routines design: 128-entry {1/c, log c} table, degree-6
// polynomial in r = z/c - 1, a degree-12 polynomial for 1 - 2^-4 <= x < 1 + 0x1.09p-4) in exactly the operation order of the
// x86-64 FMA build (`__log_fma`, the ifunc variant every FMA-capable host selects; sequence read off the disassembly of
// libm 2.39: r = fma(z, invc, -1), w = fma(k, Ln2hi, logc), lo = fma(k, Ln2lo, w - hi + r), ...).  The table is
// glibc_log_data.inc (tools/gen_glibc_log_table.py).  tests/cpp/test_glibc_log.cpp checks the host build against libm on
// > 10^7 inputs bit for bit; tests/test_gpu_obs.py checks the device build against the oracle's libm log.
#pragma once
#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#define GFL_HD __host__ __device__ __forceinline__
#else
#define GFL_HD inline
#endif

namespace gforb {

#ifdef __CUDACC__
static __device__ __constant__ unsigned long long d_glibc_log_data[18 + 256] = {
#include "glibc_log_data.inc"
};
#endif
#if defined(__CUDA_ARCH__)
#define GFL_FMA(a, b, c) __fma_rn((a), (b), (c))
#define GFL_MUL(a, b) __dmul_rn((a), (b))
#define GFL_ADD(a, b) __dadd_rn((a), (b))
#define GFL_SUB(a, b) __dsub_rn((a), (b))
#define GFL_TAB(i) __longlong_as_double((long long)d_glibc_log_data[(i)])
#define GFL_BITS(x) ((uint64_t)__double_as_longlong(x))
#define GFL_FROM(u) __longlong_as_double((long long)(u))
#else
// host: compiled with -ffp-contract=off; fma() of libm / the FMA unit is correctly rounded either way
#define GFL_FMA(a, b, c) __builtin_fma((a), (b), (c))
#define GFL_MUL(a, b) ((a) * (b))
#define GFL_ADD(a, b) ((a) + (b))
#define GFL_SUB(a, b) ((a) - (b))
static const uint64_t h_glibc_log_data[18 + 256] = {
#include "glibc_log_data.inc"
};
inline double gfl_from(uint64_t u) { double d; memcpy(&d, &u, 8); return d; }
inline uint64_t gfl_bits(double d) { uint64_t u; memcpy(&u, &d, 8); return u; }
#define GFL_TAB(i) gfl_from(h_glibc_log_data[(i)])
#define GFL_BITS(x) gfl_bits(x)
#define GFL_FROM(u) gfl_from(u)
#endif

GFL_HD double glibc_log(double x) {
    uint64_t ix = GFL_BITS(x);
    const uint64_t LO = 0x3FEE000000000000ull, HI = 0x3FF1090000000000ull;       // 1 - 2^-4, 1 + 0x1.09p-4
    if (ix - LO < HI - LO) {
        if (ix == 0x3FF0000000000000ull) return 0.0;
        const double r = GFL_SUB(x, 1.0);
        const double r2 = GFL_MUL(r, r), r3 = GFL_MUL(r, r2);
#define B_(k) GFL_TAB(7 + (k))
        const double p2 = GFL_FMA(r2, B_(3), GFL_FMA(r, B_(2), B_(1)));
        const double p3 = GFL_FMA(r2, B_(6), GFL_FMA(r, B_(5), B_(4)));
        double p1 = GFL_FMA(r, B_(8), B_(7));
        p1 = GFL_FMA(r2, B_(9), p1);
        p1 = GFL_FMA(r3, B_(10), p1);
        double q = GFL_FMA(p1, r3, p3);
        q = GFL_FMA(q, r3, p2);
        const double t = GFL_FMA(r, 0x1p27, r);
        const double rhi = GFL_FMA(-0x1p27, r, t);
        const double rlo = GFL_SUB(r, rhi);
        const double rhi2 = GFL_MUL(rhi, rhi);
        const double hi = GFL_FMA(rhi2, B_(0), r);
        double lo = GFL_FMA(rhi2, B_(0), GFL_SUB(r, hi));
        lo = GFL_FMA(GFL_MUL(B_(0), rlo), GFL_ADD(r, rhi), lo);
        const double y = GFL_FMA(q, r3, lo);
#undef B_
        return GFL_ADD(hi, y);
    }
    const uint32_t top = (uint32_t)(ix >> 48);
    if (top - 0x0010u >= 0x7ff0u - 0x0010u) {
        if (ix * 2 == 0) return -1.0 / 0.0;                                     // log(+-0) = -inf
        if (ix == 0x7FF0000000000000ull) return x;                              // log(inf) = inf
        if ((top & 0x8000u) || (top & 0x7ff0u) == 0x7ff0u) return GFL_FROM(0x7FF8000000000000ull);   // x < 0 or NaN
        ix = GFL_BITS(GFL_MUL(x, 0x1p52));                                       // subnormal: normalise
        ix -= 52ull << 52;
    }
    const uint64_t tmp = ix - 0x3FE6000000000000ull;
    const int i = (int)((tmp >> 45) & 127);
    const int k = (int)((int64_t)tmp >> 52);
    const uint64_t iz = ix - (tmp & (0xFFFull << 52));
    const double invc = GFL_TAB(18 + 2 * i), logc = GFL_TAB(19 + 2 * i);
    const double z = GFL_FROM(iz);
    const double kd = (double)k;
    const double w = GFL_FMA(kd, GFL_TAB(0), logc);
    const double r = GFL_FMA(z, invc, -1.0);
    const double t1 = GFL_FMA(r, GFL_TAB(4), GFL_TAB(3));                        // A2 r + A1
    const double hi = GFL_ADD(r, w);
    const double r2 = GFL_MUL(r, r);
    double lo = GFL_ADD(GFL_SUB(w, hi), r);
    lo = GFL_FMA(kd, GFL_TAB(1), lo);
    const double r3 = GFL_MUL(r, r2);
    const double t2 = GFL_FMA(r, GFL_TAB(6), GFL_TAB(5));                        // A4 r + A3
    lo = GFL_FMA(r2, GFL_TAB(2), lo);                                            // + A0 r^2
    const double t3 = GFL_FMA(t2, r2, t1);
    const double y = GFL_FMA(r3, t3, lo);
    return GFL_ADD(y, hi);
}

}  // namespace gforb
