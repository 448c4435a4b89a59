// Micro-benchmark: FP64 dependent-issue latency and per-SM throughput on sm_100a (DADD / DMUL / DFMA, IEEE division and
// sqrt), plus one register-resident 7x7 Cholesky per thread -- the building blocks of the good-feature selection kernel.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 --fmad=false -o fp64.bin fp64.cu ; run: ./fp64.bin
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 2048
// OP 0: dependent DADD chain, 1: DMUL+DADD pairs (un-fused multiply-add), 2: 1/x chain, 3: sqrt chain, 4: DFMA chain
template <int OP, int ILP>
__global__ void __launch_bounds__(256) k(double* out, double seed, long long* cyc) {
    double a[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) a[i] = seed + threadIdx.x * 1e-3 + i;
    const double b = seed * 0.5 + 1.0, c = 1.0 + seed * 1e-9;
    const long long t0 = clock64();
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) {
            if (OP == 0) a[i] = a[i] + b;
            if (OP == 1) a[i] = a[i] - a[i] * c;
            if (OP == 2) a[i] = 1.0 / a[i] + b;
            if (OP == 3) a[i] = sqrt(a[i]) + b;
            if (OP == 4) a[i] = __fma_rn(a[i], c, b);
        }
    }
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

#define LT(i, j) ((i) * ((i) + 1) / 2 + (j))
__global__ void __launch_bounds__(256) k_chol(double* out, double seed, long long* cyc, int reps) {
    double acc = 0;
    const long long t0 = clock64();
    for (int r = 0; r < reps; r++) {
        double A[28];
#pragma unroll
        for (int i = 0; i < 7; i++)
#pragma unroll
            for (int j = 0; j <= i; j++) A[LT(i, j)] = (i == j ? 10.0 + seed + acc * 1e-9 : 0.1 * (i + j) + threadIdx.x * 1e-4);
#pragma unroll
        for (int j = 0; j < 7; j++) {
            double ajj = A[LT(j, j)];
#pragma unroll
            for (int kk = 0; kk < j; kk++) ajj -= A[LT(j, kk)] * A[LT(j, kk)];
            ajj = sqrt(ajj);
            A[LT(j, j)] = ajj;
            const double rinv = 1.0 / ajj;
#pragma unroll
            for (int i = j + 1; i < 7; i++) {
                double s = A[LT(i, j)];
#pragma unroll
                for (int kk = 0; kk < j; kk++) s -= A[LT(i, kk)] * A[LT(j, kk)];
                A[LT(i, j)] = s * rinv;
            }
        }
        double p = 1.0;
#pragma unroll
        for (int j = 0; j < 7; j++) p *= A[LT(j, j)];
        acc += 2 * log(p);
    }
    const long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int OP, int ILP>
void run(const char* name, int blocksPerSM, int threads) {
    double* d; long long* c; cudaMalloc(&d, 148 * 16 * 256 * 8); cudaMalloc(&c, 8);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<OP, ILP><<<148 * blocksPerSM, threads>>>(d, 3.0, c);
    cudaEventRecord(e0);
    k<OP, ILP><<<148 * blocksPerSM, threads>>>(d, 3.0, c);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long h; cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
    printf("%-22s ilp=%d warps/SM=%3d : %7.1f cycles per dependent step, %7.2f steps/clk/SM\n", name, ILP, blocksPerSM * threads / 32,
           (double)h / ITERS, (double)ITERS * ILP * blocksPerSM * threads / (double)h);
    cudaFree(d); cudaFree(c);
}

void run_chol(int blocksPerSM, int threads) {
    double* d; long long* c; cudaMalloc(&d, 148 * 16 * 256 * 8); cudaMalloc(&c, 8);
    k_chol<<<148 * blocksPerSM, threads>>>(d, 3.0, c, 64);
    k_chol<<<148 * blocksPerSM, threads>>>(d, 3.0, c, 64);
    long long h; cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
    printf("chol7+log per thread  warps/SM=%3d : %8.1f cycles per logdet\n", blocksPerSM * threads / 32, (double)h / 64);
    cudaFree(d); cudaFree(c);
}

int main() {
    run<0, 1>("DADD", 1, 32);   run<0, 1>("DADD", 2, 256);  run<0, 4>("DADD", 2, 256);  run<0, 4>("DADD", 4, 256);
    run<1, 1>("DMUL+DADD", 1, 32); run<1, 4>("DMUL+DADD", 2, 256);
    run<4, 1>("DFMA", 1, 32);   run<4, 4>("DFMA", 2, 256);  run<4, 8>("DFMA", 4, 256);
    run<2, 1>("1/x + add", 1, 32); run<2, 1>("1/x + add", 2, 256); run<2, 4>("1/x + add", 2, 256);
    run<3, 1>("sqrt + add", 1, 32); run<3, 1>("sqrt + add", 2, 256); run<3, 4>("sqrt + add", 2, 256);
    run_chol(1, 32); run_chol(1, 256); run_chol(2, 256);
    return 0;
}
