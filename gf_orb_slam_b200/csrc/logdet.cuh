// Shared device code of the good-feature kernels: logDet (include/Util.hpp:714-731) as a register-resident 7x7
// Cholesky with the LU fallback, and the warp-parallel glibc TYPE_3 rand() stream the selection loops draw from.
// Included by observability.cu (k_select, k_logdet) and matcher.cu (k_active_match); both TUs are built with
// --fmad=false so every double operation rounds like the scalar CPU sequence, and log() is glibc's (glibc_log.h): the
// log-dets are bit-identical to the host's, ties in the greedy arg-max included.
#pragma once
#include <cstdint>
#include <cmath>
#include "glibc_log.h"

namespace gforb {

// ------------------------------------------------------------------------------------------------ logDet
#define LT(i, j) ((i) * ((i) + 1) / 2 + (j))

// left-looking unblocked Cholesky of the lower triangle A (packed, 28 doubles) -> product of the diagonal of L.
// Every loop has the fixed bound 7 and a compile-time predicate: loops whose bound is an outer (unrolled) index were
// left rolled by nvcc 12.9, which put A and L in local memory behind runtime indices (3x slower per factorization).
__device__ __forceinline__ bool chol7_prod(const double (&A)[28], double& prod) {
    double L[28];
#pragma unroll
    for (int j = 0; j < 7; j++) {
        double ajj = A[LT(j, j)];
#pragma unroll
        for (int k = 0; k < 7; k++) if (k < j) ajj -= L[LT(j, k)] * L[LT(j, k)];
        if (!(ajj > 0.0)) return false;
        ajj = sqrt(ajj);
        L[LT(j, j)] = ajj;
        const double rinv = 1.0 / ajj;
#pragma unroll
        for (int i = 0; i < 7; i++) if (i > j) {
            double s = A[LT(i, j)];
#pragma unroll
            for (int k = 0; k < 7; k++) if (k < j) s -= L[LT(i, k)] * L[LT(j, k)];
            L[LT(i, j)] = s * rinv;
        }
    }
    double p = 1.0;
#pragma unroll
    for (int j = 0; j < 7; j++) p *= L[LT(j, j)];
    prod = p;
    return true;
}

// arma::log_det fallback: LU with partial pivoting on the full matrix, sum log|u_ii| (rare path, local memory)
static __device__ __noinline__ double lu7_logabsdet(const double* __restrict__ cur, const double* __restrict__ blk, double sgn) {
    double A[49];
    for (int i = 0; i < 49; i++) A[i] = cur[i] + sgn * blk[i];
    double ld = 0;
    for (int k = 0; k < 7; k++) {
        int p = k;
        for (int i = k + 1; i < 7; i++) if (fabs(A[i * 7 + k]) > fabs(A[p * 7 + k])) p = i;
        if (p != k) for (int c = 0; c < 7; c++) { double t = A[k * 7 + c]; A[k * 7 + c] = A[p * 7 + c]; A[p * 7 + c] = t; }
        const double piv = A[k * 7 + k];
        ld += glibc_log(fabs(piv));
        if (piv == 0) continue;
        for (int i = k + 1; i < 7; i++) {
            const double f = A[i * 7 + k] / piv;
            for (int c = k + 1; c < 7; c++) A[i * 7 + c] -= f * A[k * 7 + c];
        }
    }
    return ld;
}

// logDet(cur + sgn * blk), both 7x7 row-major (cur in shared memory, blk in global memory)
__device__ __forceinline__ double logdet_sum(const double* __restrict__ cur, const double* __restrict__ blk, double sgn) {
    double A[28];
#pragma unroll
    for (int i = 0; i < 7; i++)
#pragma unroll
        for (int j = 0; j < 7; j++) if (j <= i) A[LT(i, j)] = cur[i * 7 + j] + sgn * blk[i * 7 + j];
    double p;
    if (chol7_prod(A, p)) return 2 * glibc_log(p);
    return lu7_logabsdet(cur, blk, sgn);
}

// Packed form of a block for the selection loops: the 28 entries of the lower triangle, LT order, 224 bytes = 7 whole
// 32-byte sectors.  A candidate evaluation then moves 224 bytes through 14 128-bit loads instead of 28 sector-sized
// fetches out of a 392-byte row-major block (the loops are bound by the SM's L2 bandwidth: ~64 B/clk).
__device__ __forceinline__ void pack_block(const double* __restrict__ blk, double* __restrict__ pk) {
#pragma unroll
    for (int i = 0; i < 7; i++)
#pragma unroll
        for (int j = 0; j < 7; j++) if (j <= i) pk[LT(i, j)] = blk[i * 7 + j];
}

// logDet(cur + sgn * B) with the lower triangle of B read from its packed copy; bit-identical to logdet_sum (same
// additions, same factorization).  The rare LU fallback reads the full row-major block like the row-major path.
__device__ __forceinline__ double logdet_sum_packed(const double* __restrict__ cur, const double* __restrict__ pk,
                                                    const double* __restrict__ blkRow, double sgn) {
    double A[28];
    const double2* p2 = reinterpret_cast<const double2*>(pk);
    double B[28];
#pragma unroll
    for (int e = 0; e < 14; e++) { const double2 v = p2[e]; B[2 * e] = v.x; B[2 * e + 1] = v.y; }
#pragma unroll
    for (int i = 0; i < 7; i++)
#pragma unroll
        for (int j = 0; j < 7; j++) if (j <= i) A[LT(i, j)] = cur[i * 7 + j] + sgn * B[LT(i, j)];
    double p;
    if (chol7_prod(A, p)) return 2 * glibc_log(p);
    return lu7_logabsdet(cur, blkRow, sgn);
}

static const int RAW_RING = 1024;    // pre-generated rand() outputs (power of two)

// Warp-parallel glibc rand(): the additive-feedback recurrence r[n] = r[n-31] + r[n-3] splits into three independent
// chains (n mod 3), so lanes 0..2 produce 30 new states per call; outputs (state >> 1) go to a ring the sampler reads.
struct WarpRand {
    uint32_t* rs;      // 64-entry state ring (shared)
    uint32_t* raw;     // RAW_RING-entry output ring (shared), indexed by output number
    int n;             // next state index to generate (uniform across the warp); output number = n - 344
    __device__ void seed(uint32_t s, int lane) {
        if (lane == 0) {
            if (s == 0) s = 1;
            int32_t word = (int32_t)s;
            rs[0] = s;
            for (int k = 1; k < 31; k++) {
                const int32_t hi = word / 127773, lo = word % 127773;
                word = 16807 * lo - 2836 * hi;
                if (word < 0) word += 2147483647;
                rs[k] = (uint32_t)word;
            }
            for (int k = 31; k < 34; k++) rs[k] = rs[k - 31];
        }
        n = 34;
        __syncwarp();
    }
    // 30 new states per call.  Along each chain (lane mod 3) the recurrence is a running sum of the r[n-31] terms on top
    // of the chain's last state, so the 30 values are one stride-3 segmented scan (4 shuffle steps, uint32 wrap-around
    // addition is associative -> bit-identical to the sequential generator).
    __device__ __forceinline__ void gen30(int lane) {
        uint32_t a = (lane < 30) ? rs[(n + lane - 31) & 63] : 0u;
#pragma unroll
        for (int d = 3; d <= 24; d <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, a, d);
            if (lane >= d) a += t;
        }
        const uint32_t seed0 = rs[(n + (lane % 3) - 3) & 63];
        __syncwarp();
        if (lane < 30) {
            const uint32_t v = a + seed0;
            const int id = n + lane;
            rs[id & 63] = v;
            if (id >= 344) raw[(id - 344) & (RAW_RING - 1)] = v >> 1;
        }
        n += 30;
        __syncwarp();
    }
    // make outputs [rp, rp + 32) available
    __device__ __forceinline__ void ensure(int rp, int lane) {
        while (n - 344 < rp + 32) gen30(lane);
    }
};

}  // namespace gforb
