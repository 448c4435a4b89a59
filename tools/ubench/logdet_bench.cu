// Micro-benchmark of the selection kernels' inner function logdet_sum (gf_orb_slam_b200/csrc/logdet.cuh): cycles per
// call per thread at 1 / 8 / 16 warps per SM, with the block in global memory (L2 resident) or reused from registers.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 --fmad=false -I../../gf_orb_slam_b200/csrc -o logdet_bench.bin logdet_bench.cu
#include <cstdio>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>
#include "logdet.cuh"
using namespace gforb;

// VARIANT 0: logdet_sum on a different block per call; 1: same but without the final log (product only);
// 2: loads only (sum of the 28 entries)
template <int VARIANT, bool SMEM_BLOCK = false>
__global__ void __launch_bounds__(256) k(const double* __restrict__ blocks, int nblk, double* out, long long* cyc, int reps) {
    __shared__ double cur[49];
    for (int i = threadIdx.x; i < 49; i += blockDim.x) cur[i] = (i % 8 == 0) ? 5.0e5 : 1.0e3;
    __shared__ double sblk[32 * 49];      // one block per lane: the latency runs are not warp-uniform
    for (int i = threadIdx.x; i < 32 * 49; i += blockDim.x) sblk[i] = blocks[i];
    __syncthreads();
    double acc = 0.0;
    unsigned h = threadIdx.x * 2654435761u + blockIdx.x * 40503u;
    const long long t0 = clock64();
    for (int r = 0; r < reps; r++) {
        h = h * 1664525u + 1013904223u + (unsigned)(acc != 12345.0);     // next block depends on the previous result
        const double* b = blocks + (size_t)(h % (unsigned)nblk) * 49;
        if (SMEM_BLOCK) { b = sblk + (threadIdx.x & 31) * 49; if (acc == 12345.0) sblk[0] = acc; __syncwarp(); }
        if (VARIANT == 0) acc += logdet_sum(cur, b, 1.0);
        if (VARIANT == 1) {
            double A[28];
#pragma unroll
            for (int i = 0; i < 7; i++)
#pragma unroll
                for (int j = 0; j < 7; j++) if (j <= i) A[LT(i, j)] = cur[i * 7 + j] + b[i * 7 + j];
            double p; chol7_prod(A, p); acc += p;
        }
        if (VARIANT == 3) {      // division-free elimination with power-of-two row scaling: one division + one log at the end
            double A[28];
#pragma unroll
            for (int i = 0; i < 7; i++)
#pragma unroll
                for (int j = 0; j < 7; j++) if (j <= i) A[LT(i, j)] = cur[i * 7 + j] + b[i * 7 + j];
            double num = 1.0, den = 1.0; bool ok = true;
#pragma unroll
            for (int j = 0; j < 7; j++) {
                const double d = A[LT(j, j)];
                ok = ok && (d > 0.0);
                num *= d;
                if (j < 6) {
                    const int hi = __double2hiint(d);
                    const double s2 = __hiloint2double((2046 - ((hi >> 20) & 0x7ff)) << 20, 0);   // 2^-e
                    const double mj = d * s2;
#pragma unroll
                    for (int e = 0; e < 6; e++) if (e < 5 - j) den *= mj;
                    double as[7];
#pragma unroll
                    for (int i = 0; i < 7; i++) if (i > j) as[i] = A[LT(i, j)] * s2;
#pragma unroll
                    for (int i = 0; i < 7; i++)
#pragma unroll
                        for (int k2 = 0; k2 < 7; k2++) if (i > j && k2 > j && k2 <= i) A[LT(i, k2)] = mj * A[LT(i, k2)] - as[i] * A[LT(k2, j)];
                }
            }
            acc += ok ? log(num / den) : 0.0;
        }
        if (VARIANT == 2) {
            double s = 0;
#pragma unroll
            for (int i = 0; i < 7; i++)
#pragma unroll
                for (int j = 0; j < 7; j++) if (j <= i) s += cur[i * 7 + j] + b[i * 7 + j];
            acc += s;
        }
    }
    const long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int V, bool SB = false>
void run(const char* name, const double* d_blk, int nblk, int blocksPerSM, int threads) {
    double* d; long long* c; cudaMalloc(&d, 148 * 16 * 256 * 8); cudaMalloc(&c, 8);
    const int reps = 64;
    k<V, SB><<<148 * blocksPerSM, threads>>>(d_blk, nblk, d, c, reps);
    k<V, SB><<<148 * blocksPerSM, threads>>>(d_blk, nblk, d, c, reps);
    long long h; cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
    double o0; cudaMemcpy(&o0, d, 8, cudaMemcpyDeviceToHost);
    printf("%-28s warps/SM=%3d : %8.1f cycles per call   (mean result %.15g)\n", name, blocksPerSM * threads / 32, (double)h / reps, o0 / reps);
    cudaFree(d); cudaFree(c);
}

int main() {
    const int nblk = 64 * 1000;                      // 64 problems x 1000 blocks = 25 MB (L2 resident)
    std::vector<double> hb((size_t)nblk * 49);
    uint64_t s = 88172645463325252ull;
    for (int b = 0; b < nblk; b++) {
        double H[14];
        for (int i = 0; i < 14; i++) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; H[i] = ((double)(s >> 11) / 9007199254740992.0 - 0.5) * 600.0; }
        for (int r = 0; r < 7; r++) for (int c = 0; c < 7; c++) hb[(size_t)b * 49 + r * 7 + c] = H[r] * H[c] + H[7 + r] * H[7 + c];
    }
    double* d_blk; cudaMalloc(&d_blk, hb.size() * 8);
    cudaMemcpy(d_blk, hb.data(), hb.size() * 8, cudaMemcpyHostToDevice);
    run<0, true>("smem: chol + log (latency)", d_blk, nblk, 1, 32);
    run<1, true>("smem: chol (latency)", d_blk, nblk, 1, 32);
    run<2, true>("smem: loads+adds (latency)", d_blk, nblk, 1, 32);
    run<3, true>("smem: div-free + log (latency)", d_blk, nblk, 1, 32);
    for (int cfg = 0; cfg < 3; cfg++) {
        const int bps = cfg == 2 ? 2 : 1, th = cfg == 0 ? 32 : 256;
        run<0>("logdet_sum (load+chol+log)", d_blk, nblk, bps, th);
        run<1>("load + chol (no log)", d_blk, nblk, bps, th);
        run<2>("load only", d_blk, nblk, bps, th);
        run<3>("load + div-free elim + log", d_blk, nblk, bps, th);
    }
    return 0;
}
