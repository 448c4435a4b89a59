// Micro-benchmark: per-SM throughput of the packed min/max candidates for the FAST arc network (sm_100a).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes.bin pipes.cu ; run: ./pipes.bin
#include <cstdio>
#include <cstdint>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#define ITERS 4096
template <int OP>
__global__ void __launch_bounds__(256) k(uint32_t* out, uint32_t seed) {
    uint32_t a[8];
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = seed * (threadIdx.x + 1) + i * 0x01010101u;
    uint32_t b = seed ^ 0x00330055u, c = seed + 0x00110022u;
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (OP == 0) a[i] = __vimax3_u16x2(a[i], b, c);
            if (OP == 1) a[i] = __vmaxu2(a[i], b);
            if (OP == 2) { __half2 x = *reinterpret_cast<__half2*>(&a[i]); __half2 y = *reinterpret_cast<__half2*>(&b); x = __hmax2(x, y); a[i] = *reinterpret_cast<uint32_t*>(&x); }
            if (OP == 3) a[i] = __byte_perm(a[i], b, c);
            if (OP == 4) a[i] = a[i] * b + c;
            if (OP == 5) { a[i] = __vimax3_u16x2(a[i], b, c); }
            if (OP == 6) a[i] = __vimin3_u16x2(a[i], b, c);
            if (OP == 7) a[i] = max(max(a[i], b), c);   // 32-bit VIMNMX3?
            if (OP == 8) a[i] = __funnelshift_r(a[i], b, 16);
            if (OP == 9) a[i] = (a[i] & c) | (b & ~c);
            if (OP == 10) a[i] = a[i] + b;
            if (OP == 11) a[i] = (a[i] >> 16) | (b & 0xffff0000u);
            if (OP == 12) a[i] = __vminu2(a[i], b);
            if (OP == 13) a[i] = __vminu2(__vmaxu2(a[i], b), c);
            if (OP == 14) { __half2 x = *reinterpret_cast<__half2*>(&a[i]); x = __hmin2(__hmax2(x, *reinterpret_cast<__half2*>(&b)), *reinterpret_cast<__half2*>(&c)); a[i] = *reinterpret_cast<uint32_t*>(&x); }
            if (OP == 15) { if (i < 4) a[i] = __vimin3_u16x2(a[i], b, c); else { __half2 x = *reinterpret_cast<__half2*>(&a[i]); x = __hmin2(__hmax2(x, *reinterpret_cast<__half2*>(&b)), *reinterpret_cast<__half2*>(&c)); a[i] = *reinterpret_cast<uint32_t*>(&x); } }
            if (OP == 17) a[i] = __dp4a(a[i], b, c);
            if (OP == 18) a[i] = __dp2a_lo(a[i], b, c);
            if (OP == 19) a[i] = __vsadu4(a[i], b) + c;
            if (OP == 20) a[i] = __popc(a[i] ^ b) + c;           // LOP3 + POPC + IADD: POPC-bound if POPC is the slow pipe
            if (OP == 21) a[i] = __popc(a[i]) ^ b;              // POPC + LOP3
            if (OP == 16) { if (i < 4) a[i] = a[i] * b + c; else a[i] = __vimin3_u16x2(a[i], b, c); }
        }
        if (OP == 5) {   // interleave with half2 max on a second set
#pragma unroll
            for (int i = 0; i < 8; i++) { __half2 x = *reinterpret_cast<__half2*>(&a[i]); __half2 y = *reinterpret_cast<__half2*>(&c); x = __hmin2(x, y); a[i] = *reinterpret_cast<uint32_t*>(&x); }
        }
        b += 0x00010001u; c ^= it;
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int OP>
void run(const char* name, int opsPerIter) {
    uint32_t* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<OP><<<148 * 8, 256>>>(d, 12345);
    cudaEventRecord(e0);
    k<OP><<<148 * 8, 256>>>(d, 12345);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double ops = 148.0 * 8 * 256 * (double)ITERS * opsPerIter;
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("{\"op\": \"%s\", \"ms\": %.3f, \"tera_ops_per_s\": %.2f, \"lane_ops_per_clk_per_sm\": %.1f, \"clock_mhz\": %d}\n", name, ms, ops / ms / 1e9, ops / (ms * 1e-3) / 148 / (clk * 1e3), clk / 1000);
    cudaFree(d);
}
int main() {
    run<0>("vimax3_u16x2", 8); run<6>("vimin3_u16x2", 8); run<1>("vmaxu2 (2-input)", 8); run<2>("hmax2 on raw bits", 8);
    run<3>("prmt", 8); run<4>("imad", 8); run<5>("vimax3 + hmin2 interleaved", 16); run<7>("max3 u32", 8); run<8>("shf funnel", 8); run<9>("lop3 select", 8); run<10>("iadd", 8); run<11>("shr+lop3", 8); run<12>("vminu2", 8); run<13>("vmin(vmax) 2x2-input int", 16); run<14>("hmin(hmax) 2x2-input half", 16); run<15>("4 vimin3 + 4x2 half (indep)", 12); run<16>("4 imad + 4 vimin3 (indep)", 8);
    run<20>("popc(xor)+add (1 POPC counted)", 8); run<21>("popc^b (1 POPC counted)", 8);
    run<17>("dp4a (IDP.4A)", 8); run<18>("dp2a_lo (IDP.2A)", 8); run<19>("vsadu4 + add", 8);
    return 0;
}
