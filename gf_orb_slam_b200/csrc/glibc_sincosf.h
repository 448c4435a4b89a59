// Bit-compatible sinf/cosf for the descriptor steering step.
//
// The reference computes `a = (float)cos(angle), b = (float)sin(angle)` with a *float* argument under
// `using namespace std` (src/ORBextractor.cc:75,166-167), i.e. libm's cosf/sinf.  Their results decide
// cvRound(x*b + y*a) and therefore descriptor *bits*, so the device has to reproduce glibc's float routines, not
// CUDA's (which differ in the last ulp on a few percent of inputs).  This is a restatement of the algorithm glibc
// >= 2.28 uses for sinf/cosf/sincosf (the ARM optimized-routines design: argument promoted to double, reduction by
// pi/2 with a 2^24-scaled reciprocal, degree-8/9 double polynomials, one final rounding to float), valid for
// |x| < 120 -- the steering angle is always in [0, 2*pi).  Double operations are written with explicit
// non-contracting intrinsics on the device so host and device agree exactly.
// tests/test_host_logic.py checks it against the platform libm over millions of angles (host build), and
// tests/test_gpu_extractor.py checks the device build against the same values.
#pragma once
#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#define GFS_HD __host__ __device__ __forceinline__
#else
#define GFS_HD inline
#endif

namespace gforb {

#if defined(__CUDA_ARCH__)
#define GFS_MUL(a, b) __dmul_rn((a), (b))
#define GFS_ADD(a, b) __dadd_rn((a), (b))
#else
// host: compiled with -ffp-contract=off (see Makefiles); volatile-free plain ops
#define GFS_MUL(a, b) ((a) * (b))
#define GFS_ADD(a, b) ((a) + (b))
#endif

struct SinCosTab { double sign[4]; double hpi_inv, hpi, c0, c1, c2, c3, c4, s1, s2, s3; };

GFS_HD uint32_t gfs_abstop12(float x) {
    uint32_t u;
#if defined(__CUDA_ARCH__)
    u = __float_as_uint(x);
#else
    memcpy(&u, &x, 4);
#endif
    return (u >> 20) & 0x7ff;
}

// n odd -> cosine polynomial, n even -> sine polynomial (of the reduced argument)
GFS_HD float gfs_poly(double x, double x2, int tab, int n) {
    // table 0: quadrants with positive cosine sign convention, table 1: negated cosine coefficients
    const double c0 = tab ? -0x1p0 : 0x1p0;
    const double c1 = tab ? 0x1.ffffffd0c621cp-2 : -0x1.ffffffd0c621cp-2;
    const double c2 = tab ? -0x1.55553e1068f19p-5 : 0x1.55553e1068f19p-5;
    const double c3 = tab ? 0x1.6c087e89a359dp-10 : -0x1.6c087e89a359dp-10;
    const double c4 = tab ? -0x1.99343027bf8c3p-16 : 0x1.99343027bf8c3p-16;
    const double s1 = -0x1.555545995a603p-3, s2 = 0x1.1107605230bc4p-7, s3 = -0x1.994eb3774cf24p-13;
    if ((n & 1) == 0) {
        double x3 = GFS_MUL(x, x2);
        double s1_ = GFS_ADD(s2, GFS_MUL(x2, s3));
        double x7 = GFS_MUL(x3, x2);
        double s = GFS_ADD(x, GFS_MUL(x3, s1));
        return (float)GFS_ADD(s, GFS_MUL(x7, s1_));
    } else {
        double x4 = GFS_MUL(x2, x2);
        double c2_ = GFS_ADD(c3, GFS_MUL(x2, c4));
        double c1_ = GFS_ADD(c0, GFS_MUL(x2, c1));
        double x6 = GFS_MUL(x4, x2);
        double c = GFS_ADD(c1_, GFS_MUL(x4, c2));
        return (float)GFS_ADD(c, GFS_MUL(x6, c2_));
    }
}

// want_cos = 0 -> sinf(y), 1 -> cosf(y).  Domain: |y| < 120.
GFS_HD float gfs_sincos1(float y, int want_cos) {
    double x = (double)y;
    const uint32_t top = gfs_abstop12(y);
    if (top < 0x3f4u) {                               // abstop12(pi/4 as float) == 0x3f4  -> |y| < 0.75
        double x2 = GFS_MUL(x, x);
        if (top < 0x398u)                             // |y| < 2^-12
            return want_cos ? 1.0f : y;
        return gfs_poly(x, x2, 0, want_cos);
    }
    // reduce_fast (non-TOINT variant): n = round(x * 2/pi), x -= n * pi/2
    const double hpi_inv = 0x1.45F306DC9C883p+23, hpi = 0x1.921FB54442D18p0;
    double r = GFS_MUL(x, hpi_inv);
    int n = ((int32_t)r + 0x800000) >> 24;
    x = GFS_ADD(x, -GFS_MUL((double)n, hpi));
    const double sgn = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;   // sign[] = {1,-1,-1,1}
    const int tab = (n & 2) ? 1 : 0;
    return gfs_poly(GFS_MUL(x, sgn), GFS_MUL(x, x), tab, want_cos ? (n ^ 1) : n);
}

GFS_HD float glibc_sinf(float y) { return gfs_sincos1(y, 0); }
GFS_HD float glibc_cosf(float y) { return gfs_sincos1(y, 1); }

}  // namespace gforb
