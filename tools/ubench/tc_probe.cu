// tc_probe.cu -- probe of the tcgen05 path used by the tensor-core all-pairs Hamming kernel (k_allpairs_tc, matcher_tc.cuh):
//   1. correctness of ONE 128 x 256 x 256 tile (queries x database rows x descriptor bits) for the shared-memory
//      descriptor encodings under test and for both operand encodings (kind::i8: A = +-1 s8, B = 0/1 u8;
//      kind::f8f6f4: A = +-2^-e e5m2, B = single-bit e4m3 patterns), checked against popcounts on the host;
//   2. tcgen05.mma throughput (M128 N256 K32) of both kinds on all SMs, tcgen05.ld throughput per SM.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tc_probe.bin tc_probe.cu ; run: ./tc_probe.bin
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(2); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ bool mbar_wait(uint64_t* bar, uint32_t parity) {
    for (int i = 0; i < (1 << 22); i++)
        if (mbar_try_wait(bar, parity)) return true;
    return false;      // bounded: a wrong descriptor must not hang the box
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t* dst, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void mma_i8(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void mma_f8(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,"
                 "%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                   "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
                   "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
                   "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// shared-memory matrix descriptor (K-major, no swizzle): start >> 4 | LBO >> 4 << 16 | SBO >> 4 << 32 | version 1 << 46
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((saddr & 0x3FFFF) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46);
}
// instruction descriptor: c_format [4,6), a_format [7,10), b_format [10,13), K-major both, N >> 3 [17,23), M >> 4 [24,29)
__host__ __device__ constexpr uint32_t make_idesc(int cfmt, int afmt, int bfmt, int M, int N) {
    return (uint32_t)cfmt << 4 | (uint32_t)afmt << 7 | (uint32_t)bfmt << 10 | (uint32_t)(N >> 3) << 17 | (uint32_t)(M >> 4) << 24;
}

// ---- operand expansion: descriptor bit (word i, bit j + 8 m) -> k index i*32 + j*4 + m; layout [chunk c = k/16][row][16 B]
// mode 0: kind::i8, A = +1 / -1 (s8), B = 1 / 0 (u8).   mode 1: kind::f8f6f4, B = e4m3 single-bit patterns, A = e5m2 +-2^-e
__device__ __forceinline__ void expand_b_row(const uint32_t (&w)[8], uint8_t* base, int rows, int r, int mode) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
        uint32_t x[8];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if (mode == 0 || j == 7) x[j] = (w[i] >> j) & 0x01010101u;
            else x[j] = w[i] & (0x01010101u << j);
        }
        *reinterpret_cast<uint4*>(base + (size_t)(2 * i) * rows * 16 + r * 16) = make_uint4(x[0], x[1], x[2], x[3]);
        *reinterpret_cast<uint4*>(base + (size_t)(2 * i + 1) * rows * 16 + r * 16) = make_uint4(x[4], x[5], x[6], x[7]);
    }
}
__device__ __forceinline__ void expand_a_row(const uint32_t (&w)[8], uint8_t* base, int rows, int r, int mode) {
    // e5m2 codes of 2^-e(j): e4m3 patterns 0x01,02,04,08,10,20,40 are 2^-9,-8,-7,-6,-5,-3,+1; bit 7 is shifted to 0x01
    const uint32_t code[8] = {0x60, 0x5C, 0x58, 0x54, 0x50, 0x48, 0x38, 0x60};
#pragma unroll
    for (int i = 0; i < 8; i++) {
        uint32_t y[8];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const uint32_t x = (w[i] >> j) & 0x01010101u;
            if (mode == 0) y[j] = (x * 0xFEu) ^ 0xFFFFFFFFu;                              // 1 -> 0x01, 0 -> 0xFF
            else y[j] = (code[j] * 0x01010101u) | ((x ^ 0x01010101u) << 7);               // sign bit when the bit is 0
        }
        *reinterpret_cast<uint4*>(base + (size_t)(2 * i) * rows * 16 + r * 16) = make_uint4(y[0], y[1], y[2], y[3]);
        *reinterpret_cast<uint4*>(base + (size_t)(2 * i + 1) * rows * 16 + r * 16) = make_uint4(y[4], y[5], y[6], y[7]);
    }
}

struct ProbeCfg {
    uint32_t lboA, sboA, lboB, sboB, kstepA, kstepB;   // bytes
    int mode;
};

// ---------------------------------------------------------------------------------------------- 1. one tile, checked
__global__ void __launch_bounds__(128) k_tile(const uint32_t* __restrict__ q, const uint32_t* __restrict__ db, ProbeCfg cfg,
                                              int32_t* __restrict__ out, int* __restrict__ err) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sA = smem;                  // 128 rows x 256 B
    uint8_t* sB = smem + 32768;          // 256 rows x 256 B
    __shared__ uint64_t bar;
    __shared__ uint32_t tmemBase;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) { mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    if (warp == 0) tmem_alloc(&tmemBase, 256);
    {
        uint32_t w[8];
#pragma unroll
        for (int i = 0; i < 8; i++) w[i] = q[tid * 8 + i];
        expand_a_row(w, sA, 128, tid, cfg.mode);
        for (int rr = 0; rr < 2; rr++) {
            const int r = tid + rr * 128;
#pragma unroll
            for (int i = 0; i < 8; i++) w[i] = db[r * 8 + i];
            expand_b_row(w, sB, 256, r, cfg.mode);
        }
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = tmemBase;
    if (tid == 0) {
        const uint32_t idesc = cfg.mode == 0 ? make_idesc(2, 1, 0, 128, 256) : make_idesc(1, 1, 0, 128, 256);
#pragma unroll 1
        for (int kb = 0; kb < 8; kb++) {
            const uint64_t da = make_desc(smem_u32(sA) + kb * cfg.kstepA, cfg.lboA, cfg.sboA);
            const uint64_t dbd = make_desc(smem_u32(sB) + kb * cfg.kstepB, cfg.lboB, cfg.sboB);
            if (cfg.mode == 0) mma_i8(tb, da, dbd, idesc, kb > 0);
            else mma_f8(tb, da, dbd, idesc, kb > 0);
        }
        mma_commit(&bar);
    }
    const bool ok = mbar_wait(&bar, 0);
    if (!ok && tid == 0) atomicExch(err, 1);
    tc_fence_after();
    if (ok) {
        for (int c = 0; c < 8; c++) {
            uint32_t r[32];
            tmem_ld32(tb + ((uint32_t)(warp * 32) << 16) + c * 32, r);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; j++) {
                int v = cfg.mode == 0 ? (int)r[j] : (int)__uint_as_float(r[j]);
                out[(warp * 32 + lane) * 256 + c * 32 + j] = v;
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tb, 256);
}

// ---------------------------------------------------------------------------------------------- 2. MMA throughput
__global__ void __launch_bounds__(128) k_mma_rate(int iters, int mode, int nTile, int* __restrict__ err) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sA = smem;
    uint8_t* sB = smem + 32768;
    __shared__ uint64_t bar[2];
    __shared__ uint32_t tmemBase;
    const int tid = threadIdx.x, warp = tid >> 5;
    if (tid == 0) { mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    if (warp == 0) tmem_alloc(&tmemBase, 512);
    for (int i = tid; i < (32768 + 65536) / 16; i += 128) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0x01010101u, 0x01010101u, 0x01010101u, 0x01010101u);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = tmemBase;
    if (tid == 0) {
        const uint32_t idesc = mode == 0 ? make_idesc(2, 1, 0, 128, nTile) : make_idesc(1, 1, 0, 128, nTile);
        uint32_t ph[2] = {0, 0};
        for (int it = 0; it < iters; it++) {
            const int a = it & 1;
            if (it >= 2) { if (!mbar_wait(&bar[a], ph[a])) { atomicExch(err, 2); break; } ph[a] ^= 1; }
#pragma unroll 1
            for (int kb = 0; kb < 8; kb++) {
                const uint64_t da = make_desc(smem_u32(sA) + kb * 4096, 2048, 128);
                const uint64_t dbd = make_desc(smem_u32(sB) + kb * 8192, 4096, 128);
                if (mode == 0) mma_i8(tb + a * 256, da, dbd, idesc, kb > 0);
                else mma_f8(tb + a * 256, da, dbd, idesc, kb > 0);
            }
            mma_commit(&bar[a]);
        }
        for (int a = 0; a < 2; a++)
            if (iters > a) mbar_wait(&bar[(iters - 1 - a) & 1], ph[(iters - 1 - a) & 1]);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tb, 512);
}

// ---------------------------------------------------------------------------------------------- 3. tcgen05.ld throughput
__global__ void __launch_bounds__(256) k_ld_rate(int iters, int nwarps, uint32_t* __restrict__ sink) {
    __shared__ uint32_t tmemBase;
    const int tid = threadIdx.x, warp = tid >> 5;
    if (warp == 0) tmem_alloc(&tmemBase, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = tmemBase;
    uint32_t m = 0;
    if (warp < nwarps) {
        const uint32_t lanebase = (uint32_t)((warp & 3) * 32) << 16;
        for (int it = 0; it < iters; it++) {
#pragma unroll 1
            for (int c = 0; c < 16; c += 2) {
                uint32_t r0[32], r1[32];
                tmem_ld32(tb + lanebase + c * 32, r0);
                tmem_ld32(tb + lanebase + c * 32 + 32, r1);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 32; j += 2) m = max(m, max(r0[j], r0[j + 1]));
#pragma unroll
                for (int j = 0; j < 32; j += 2) m = max(m, max(r1[j], r1[j + 1]));
            }
        }
    }
    if (m == 0x12345678u) sink[tid] = m;
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tb, 512);
}

static int popc32(uint32_t x) { return __builtin_popcount(x); }

int main() {
    int dev = 0; CK(cudaSetDevice(dev));
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, dev));
    int clk; CK(cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, dev));
    printf("{\"device\": \"%s\", \"sms\": %d, \"clock_mhz\": %d}\n", prop.name, prop.multiProcessorCount, clk / 1000);
    const int smemBytes = 32768 + 65536;
    CK(cudaFuncSetAttribute(k_tile, cudaFuncAttributeMaxDynamicSharedMemorySize, smemBytes));
    CK(cudaFuncSetAttribute(k_mma_rate, cudaFuncAttributeMaxDynamicSharedMemorySize, smemBytes));
    std::vector<uint32_t> q(128 * 8), db(256 * 8);
    uint64_t s = 0x9E3779B97F4A7C15ull;
    auto rnd = [&]() { s += 0x9E3779B97F4A7C15ull; uint64_t z = s; z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; return (uint32_t)((z ^ (z >> 31)) >> 16); };
    for (auto& v : q) v = rnd();
    for (auto& v : db) v = rnd();
    for (int i = 0; i < 8; i++) { db[5 * 8 + i] = q[3 * 8 + i]; db[6 * 8 + i] = ~q[3 * 8 + i]; }       // distance 0 and 256
    uint32_t *dq, *ddb; int32_t* dout; int* derr;
    CK(cudaMalloc(&dq, q.size() * 4)); CK(cudaMalloc(&ddb, db.size() * 4)); CK(cudaMalloc(&dout, 128 * 256 * 4)); CK(cudaMalloc(&derr, 4));
    CK(cudaMemcpy(dq, q.data(), q.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(ddb, db.data(), db.size() * 4, cudaMemcpyHostToDevice));
    std::vector<int32_t> ref(128 * 256), got(128 * 256);
    for (int m = 0; m < 128; m++)
        for (int n = 0; n < 256; n++) {
            int pand = 0, pd = 0;
            for (int i = 0; i < 8; i++) { pand += popc32(q[m * 8 + i] & db[n * 8 + i]); pd += popc32(db[n * 8 + i]); }
            ref[m * 256 + n] = 2 * pand - pd;        // = popc(q) - hamming(q, d)
        }
    struct V { const char* name; ProbeCfg c; };
    const V variants[] = {
        {"i8  LBO=K-chunk stride, SBO=128", {2048, 128, 4096, 128, 4096, 8192, 0}},
        {"f8  LBO=K-chunk stride, SBO=128", {2048, 128, 4096, 128, 4096, 8192, 1}},
    };
    int goodI8 = -1, goodF8 = -1;
    for (int v = 0; v < 2; v++) {   // (the swapped LBO/SBO encoding faults with an illegal address: measured once, removed)
        CK(cudaMemset(dout, 0xAB, 128 * 256 * 4)); CK(cudaMemset(derr, 0, 4));
        k_tile<<<1, 128, smemBytes>>>(dq, ddb, variants[v].c, dout, derr);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("{\"variant\": \"%s\", \"cuda_error\": \"%s\"}\n", variants[v].name, cudaGetErrorString(e)); return 3; }
        int herr; CK(cudaMemcpy(&herr, derr, 4, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(got.data(), dout, got.size() * 4, cudaMemcpyDeviceToHost));
        int bad = 0, first = -1;
        for (int i = 0; i < 128 * 256; i++) if (got[i] != ref[i]) { if (first < 0) first = i; bad++; }
        printf("{\"variant\": \"%s\", \"timeout\": %d, \"mismatches\": %d, \"first\": %d, \"got\": %d, \"ref\": %d, \"d0_check\": [%d, %d]}\n",
               variants[v].name, herr, bad, first, first >= 0 ? got[first] : 0, first >= 0 ? ref[first] : 0, got[3 * 256 + 5], got[3 * 256 + 6]);
        if (bad == 0 && !herr) { if (variants[v].c.mode == 0 && goodI8 < 0) goodI8 = v; if (variants[v].c.mode == 1 && goodF8 < 0) goodF8 = v; }
    }
    printf("{\"good_i8_variant\": %d, \"good_f8_variant\": %d}\n", goodI8, goodF8);
    // throughput (only meaningful with the LBO/SBO encoding that passed: variant 0 layout is hard-wired in k_mma_rate)
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const int nsm = prop.multiProcessorCount;
    for (int mode = 0; mode < 2; mode++)
        for (int nT : {256, 128}) {
            const int iters = 4000;
            CK(cudaMemset(derr, 0, 4));
            k_mma_rate<<<nsm, 128, smemBytes>>>(200, mode, nT, derr);
            CK(cudaEventRecord(e0));
            k_mma_rate<<<nsm, 128, smemBytes>>>(iters, mode, nT, derr);
            CK(cudaEventRecord(e1));
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("{\"mma_rate_error\": \"%s\"}\n", cudaGetErrorString(e)); return 3; }
            float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
            int herr; CK(cudaMemcpy(&herr, derr, 4, cudaMemcpyDeviceToHost));
            const double macs = (double)nsm * iters * 128.0 * nT * 256.0;
            printf("{\"mma_rate\": \"%s M128 N%d K32 x8\", \"ms\": %.3f, \"tera_ops\": %.1f, \"macs_per_clk_per_sm_at_max_clock\": %.0f, \"timeout\": %d}\n",
                   mode == 0 ? "kind::i8" : "kind::f8f6f4", nT, ms, 2 * macs / ms / 1e9, macs / (ms * 1e-3) / nsm / (clk * 1e3), herr);
        }
    uint32_t* dsink; CK(cudaMalloc(&dsink, 256 * 4));
    for (int nw : {4, 8}) {
        const int iters = 2000;
        k_ld_rate<<<nsm, 256>>>(50, nw, dsink);
        CK(cudaEventRecord(e0));
        k_ld_rate<<<nsm, 256>>>(iters, nw, dsink);
        CK(cudaEventRecord(e1));
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("{\"ld_rate_error\": \"%s\"}\n", cudaGetErrorString(e)); return 3; }
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        const double bytes = (double)iters * nw * 16 * 32 * 32 * 4;      // per SM
        printf("{\"tmem_ld_rate_warps\": %d, \"ms\": %.3f, \"bytes_per_clk_per_sm_at_max_clock\": %.1f, \"note\": \"includes a 2-op/value max reduction\"}\n",
               nw, ms, bytes / (ms * 1e-3) / (clk * 1e3));
    }
    return 0;
}
