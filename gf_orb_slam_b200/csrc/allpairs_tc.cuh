// Tensor-core all-pairs Hamming top-2 (BASELINE.json config 5; SURVEY section 7 hard part 8): the brute-force
// DescriptorDistance loop (src/ORBmatcher.cc:2384-2400 applied n_q x n_db times) as a dense contraction on the
// 5th-generation tensor cores.
//
//   query bit  -> A operand, s8:  +1 (bit set) / -1 (bit clear)          128 queries x 256 k per CTA, resident in SMEM
//   db row bit -> B operand, u8:   1 (bit set) /  0 (bit clear)          256 rows x 256 k per tile, double-buffered
//   D[q][r] = sum_k A[q][k] B[r][k] = 2 popc(q & r) - popc(r)   =>   hamming(q, r) = popc(q) - D[q][r]      (exact, s32)
//
// so per query the best database row is the one with the LARGEST accumulator; no per-row correction term is needed
// (only the per-query constant popc(q), applied once at the end).  tcgen05.mma kind::i8, M128 N256 K32 x 8 per tile,
// accumulators in TMEM (2 x 256 columns), operands in the no-swizzle K-major canonical layout
// [k-chunk of 16 B][row][16 B] (LBO = rows * 16, SBO = 128; encoding verified on B200 by tools/ubench/tc_probe.cu).
//
// Warp roles (448 threads, one CTA per SM):
//   warp 0      producer: cp.async.bulk (TMA bulk copy, mbarrier complete_tx) of the next tiles' raw 32-byte rows
//   warp 1      TMEM allocation + the single MMA-issuing thread (tcgen05.mma / tcgen05.commit)
//   warps 2-5   epilogue: tcgen05.ld 32 columns at a time; a chunk is examined exactly only if its maximum beats the
//               query's current second best (0.5 ALU op per accumulator otherwise); packed (D, row) top-2 keys
//   warps 6-13  expanders: one database row per thread, 256 bits -> 256 bytes (SHF + LOP3 per 4 bytes), 16 STS.128,
//               fence.proxy.async, mbarrier arrive
// Pipelines: raw (3 stages, producer -> expanders), B (2 stages, expanders -> MMA, released by tcgen05.commit),
// accumulator (2 stages, MMA -> epilogue, released by the epilogue warps).
#pragma once
#include <cstdint>

namespace gforb {
namespace tc {

static const int TC_THREADS = 448, TC_TILE = 256, TC_QTILE = 128, TC_RAW_STAGES = 3, TC_B_STAGES = 2;
static const int TC_SMEM_A = TC_QTILE * 256;                 // 32 KB
static const int TC_SMEM_B = TC_TILE * 256;                  // 64 KB per stage
static const int TC_SMEM_RAW = TC_TILE * 32;                 // 8 KB per stage
static const int TC_SMEM_BYTES = TC_SMEM_A + TC_B_STAGES * TC_SMEM_B + TC_RAW_STAGES * TC_SMEM_RAW + 1024;

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
// bounded wait: a protocol error must surface as an error code, never as a hung GPU
__device__ __forceinline__ bool mbar_wait(uint64_t* bar, uint32_t parity) {
    for (int i = 0; i < (1 << 24); i++)
        if (mbar_try_wait(bar, parity)) return true;
    return false;
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
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
// instruction descriptor: c_format [4,6) = S32, a_format [7,10) = s8, b_format [10,13) = u8, K-major, N >> 3 [17,23), M >> 4 [24,29)
__host__ __device__ constexpr uint32_t make_idesc_i8(int M, int N) {
    return 2u << 4 | 1u << 7 | 0u << 10 | (uint32_t)(N >> 3) << 17 | (uint32_t)(M >> 4) << 24;
}

// descriptor bit (word i, bit j + 8 m) -> k index i*32 + j*4 + m (the same permutation for A and B, so the dot product
// is over matching bits); operand layout [chunk c = k / 16][row][16 B]: chunk 2i = bits j 0..3, chunk 2i+1 = bits j 4..7
__device__ __forceinline__ void expand_b_row(const uint32_t (&w)[8], uint8_t* base, int rows, int r) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
        uint32_t x[8];
#pragma unroll
        for (int j = 0; j < 8; j++) x[j] = (w[i] >> j) & 0x01010101u;
        *reinterpret_cast<uint4*>(base + (size_t)(2 * i) * rows * 16 + r * 16) = make_uint4(x[0], x[1], x[2], x[3]);
        *reinterpret_cast<uint4*>(base + (size_t)(2 * i + 1) * rows * 16 + r * 16) = make_uint4(x[4], x[5], x[6], x[7]);
    }
}
__device__ __forceinline__ void expand_a_row(const uint32_t (&w)[8], uint8_t* base, int rows, int r, bool valid) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
        uint32_t y[8];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const uint32_t x = (w[i] >> j) & 0x01010101u;
            y[j] = valid ? ((x * 0xFEu) ^ 0xFFFFFFFFu) : 0u;              // 1 -> 0x01 (+1), 0 -> 0xFF (-1); padding rows 0
        }
        *reinterpret_cast<uint4*>(base + (size_t)(2 * i) * rows * 16 + r * 16) = make_uint4(y[0], y[1], y[2], y[3]);
        *reinterpret_cast<uint4*>(base + (size_t)(2 * i + 1) * rows * 16 + r * 16) = make_uint4(y[4], y[5], y[6], y[7]);
    }
}

struct TcParams {
    const uint4* q;          // n_q x 32 B
    int nq;
    const uint4* db;         // n_db x 32 B
    long long nDb;
    int tilesPerSplit;       // database tiles (256 rows) per blockIdx.x
    long long idxBase;
    uint64_t* partial;       // gridDim.x x n_q x 2 keys (dist << 32 | global row), merged by k_top2_merge
    int* err;
};

__device__ __forceinline__ bool mbar_wait_t(uint64_t* bar, uint32_t parity) {
    if (mbar_try_wait(bar, parity)) return true;
    const long long t0 = clock64();
    while (!mbar_try_wait(bar, parity))
        if (clock64() - t0 > 4000000000ll) return false;          // ~2 s: protocol error, reported through TcParams::err
    return true;
}

// running top-2 of one query over the columns of a 32-column accumulator chunk.  Keys: (D + 512) << 21 | (0x1FFFFF - local
// row): larger is better (higher accumulator = smaller distance, then lower row).  `thr` = accumulator of the second best.
template <bool MASK>
__device__ __forceinline__ void top2_chunk(const uint32_t (&r)[32], uint32_t invBase, int validCols, uint32_t& k1, uint32_t& k2, int& thr) {
    int m = (int)r[31];
#pragma unroll
    for (int j = 0; j < 31; j += 2) m = __vimax3_s32(m, (int)r[j], (int)r[j + 1]);      // VIMNMX3: two accumulators per issue slot
    if (MASK || m > thr) {
#pragma unroll
        for (int j = 0; j < 32; j++) {
            uint32_t key = ((uint32_t)((int)r[j] + 512) << 21) | (invBase - (uint32_t)j);
            if (MASK && j >= validCols) key = 0;
            k2 = max(k2, min(k1, key));
            k1 = max(k1, key);
        }
        thr = (int)(k2 >> 21) - 512;
    }
}

__global__ void __launch_bounds__(TC_THREADS, 1) k_allpairs_tc(TcParams P) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sA = smem;
    uint8_t* sB = smem + TC_SMEM_A;
    uint8_t* sRaw = sB + TC_B_STAGES * TC_SMEM_B;
    __shared__ uint64_t rawFull[TC_RAW_STAGES], rawEmpty[TC_RAW_STAGES], bFull[TC_B_STAGES], bEmpty[TC_B_STAGES], dFull[2], dEmpty[2];
    __shared__ uint32_t tmemBase;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long long nTilesAll = (P.nDb + TC_TILE - 1) / TC_TILE;
    const long long tile0 = (long long)blockIdx.x * P.tilesPerSplit;
    const int nT = (int)max(0ll, min((long long)P.tilesPerSplit, nTilesAll - tile0));
    const long long rowStart = tile0 * TC_TILE;
    const long long rangeRows = min((long long)nT * TC_TILE, P.nDb - rowStart);
    const int q0 = blockIdx.y * TC_QTILE;

    if (tid == 0) {
        for (int i = 0; i < TC_RAW_STAGES; i++) { mbar_init(&rawFull[i], 1); mbar_init(&rawEmpty[i], 8); }
        for (int i = 0; i < TC_B_STAGES; i++) { mbar_init(&bFull[i], 8); mbar_init(&bEmpty[i], 1); }
        for (int i = 0; i < 2; i++) { mbar_init(&dFull[i], 1); mbar_init(&dEmpty[i], 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) tmem_alloc(&tmemBase, 512);
    if (tid < TC_QTILE) {                                          // the CTA's queries: expanded once, resident
        const int q = q0 + tid;
        uint32_t w[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (q < P.nq) {
            const uint4 a0 = __ldg(P.q + 2 * (size_t)q), a1 = __ldg(P.q + 2 * (size_t)q + 1);
            w[0] = a0.x; w[1] = a0.y; w[2] = a0.z; w[3] = a0.w; w[4] = a1.x; w[5] = a1.y; w[6] = a1.z; w[7] = a1.w;
        }
        expand_a_row(w, sA, TC_QTILE, tid, q < P.nq);
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = tmemBase;

    if (warp == 0) {
        // ------------------------------------------------------------------ producer: raw rows, TMA bulk copies
        if (lane == 0) {
            for (int i = 0; i < nT; i++) {
                const int r = i % TC_RAW_STAGES;
                if (!mbar_wait_t(&rawEmpty[r], ((i / TC_RAW_STAGES) & 1) ^ 1)) { atomicExch(P.err, 11); break; }
                const int rows = (int)min((long long)TC_TILE, rangeRows - (long long)i * TC_TILE);
                const uint32_t bytes = (uint32_t)rows * 32u;
                mbar_arrive_expect_tx(&rawFull[r], bytes);
                bulk_g2s(sRaw + r * TC_SMEM_RAW, reinterpret_cast<const uint8_t*>(P.db) + (size_t)(rowStart + (long long)i * TC_TILE) * 32,
                         bytes, &rawFull[r]);
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer (one thread)
        if (lane == 0) {
            const uint32_t idesc = make_idesc_i8(TC_QTILE, TC_TILE);
            const uint32_t aBase = smem_u32(sA), bBase = smem_u32(sB);
            for (int i = 0; i < nT; i++) {
                const int s = i % TC_B_STAGES, a = i & 1;
                if (!mbar_wait_t(&dEmpty[a], ((i >> 1) & 1) ^ 1)) { atomicExch(P.err, 12); break; }
                if (!mbar_wait_t(&bFull[s], (i / TC_B_STAGES) & 1)) { atomicExch(P.err, 13); break; }
                tc_fence_after();
#pragma unroll
                for (int kb = 0; kb < 8; kb++) {
                    const uint64_t da = make_desc(aBase + kb * (2 * TC_QTILE * 16), TC_QTILE * 16, 128);
                    const uint64_t dbd = make_desc(bBase + s * TC_SMEM_B + kb * (2 * TC_TILE * 16), TC_TILE * 16, 128);
                    mma_i8(tb + a * TC_TILE, da, dbd, idesc, kb > 0);
                }
                mma_commit(&bEmpty[s]);
                mma_commit(&dFull[a]);
            }
        }
    } else if (warp < 6) {
        // ------------------------------------------------------------------ epilogue: top-2 per query out of TMEM
        const int quarter = warp & 3;
        const int q = q0 + quarter * 32 + lane;
        uint32_t k1 = 0, k2 = 0;
        int thr = -512;
        const uint32_t laneAddr = tb + ((uint32_t)(quarter * 32) << 16);
        bool ok = true;
        for (int i = 0; i < nT && ok; i++) {
            const int a = i & 1;
            if (!mbar_wait_t(&dFull[a], (i >> 1) & 1)) { if (lane == 0) atomicExch(P.err, 14); ok = false; break; }
            tc_fence_after();
            const int rows = (int)min((long long)TC_TILE, rangeRows - (long long)i * TC_TILE);
            const uint32_t inv0 = 0x1FFFFFu - (uint32_t)(i * TC_TILE);
            const uint32_t taddr = laneAddr + a * TC_TILE;
            uint32_t ra[32], rb[32];
            tmem_ld32(taddr, ra);
            if (rows == TC_TILE) {
#pragma unroll 1
                for (int c = 0; c < 8; c += 2) {
                    tmem_ld_wait();
                    tmem_ld32(taddr + (c + 1) * 32, rb);
                    top2_chunk<false>(ra, inv0 - c * 32, 32, k1, k2, thr);
                    tmem_ld_wait();
                    if (c + 2 < 8) tmem_ld32(taddr + (c + 2) * 32, ra);
                    top2_chunk<false>(rb, inv0 - (c + 1) * 32, 32, k1, k2, thr);
                }
            } else {
#pragma unroll 1
                for (int c = 0; c < 8; c++) {
                    tmem_ld_wait();
                    top2_chunk<true>(ra, inv0 - c * 32, rows - c * 32, k1, k2, thr);
                    if (c + 1 < 8) tmem_ld32(taddr + (c + 1) * 32, ra);
                }
                tmem_ld_wait();
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&dEmpty[a]);
        }
        if (q < P.nq) {
            const uint4 a0 = __ldg(P.q + 2 * (size_t)q), a1 = __ldg(P.q + 2 * (size_t)q + 1);
            const int pq = __popc(a0.x) + __popc(a0.y) + __popc(a0.z) + __popc(a0.w) + __popc(a1.x) + __popc(a1.y) + __popc(a1.z) + __popc(a1.w);
            uint64_t o1 = ~0ull, o2 = ~0ull;
            if (k1) o1 = ((uint64_t)(uint32_t)(pq - ((int)(k1 >> 21) - 512)) << 32) | (uint64_t)(P.idxBase + rowStart + (0x1FFFFFu - (k1 & 0x1FFFFFu)));
            if (k2) o2 = ((uint64_t)(uint32_t)(pq - ((int)(k2 >> 21) - 512)) << 32) | (uint64_t)(P.idxBase + rowStart + (0x1FFFFFu - (k2 & 0x1FFFFFu)));
            uint64_t* o = P.partial + ((size_t)blockIdx.x * P.nq + q) * 2;
            o[0] = o1; o[1] = o2;
        }
    } else {
        // ------------------------------------------------------------------ expanders: bits -> bytes, one row per thread
        const int e = tid - 6 * 32;
        for (int i = 0; i < nT; i++) {
            const int r = i % TC_RAW_STAGES, s = i % TC_B_STAGES;
            if (!mbar_wait_t(&rawFull[r], (i / TC_RAW_STAGES) & 1)) { if (lane == 0) atomicExch(P.err, 15); break; }
            const uint4* src = reinterpret_cast<const uint4*>(sRaw + r * TC_SMEM_RAW) + 2 * e;
            const uint4 a0 = src[0], a1 = src[1];
            const uint32_t w[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            __syncwarp();
            if (lane == 0) mbar_arrive(&rawEmpty[r]);
            if (!mbar_wait_t(&bEmpty[s], ((i / TC_B_STAGES) & 1) ^ 1)) { if (lane == 0) atomicExch(P.err, 16); break; }
            expand_b_row(w, sB + s * TC_SMEM_B, TC_TILE, e);
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bFull[s]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tb, 512);
}

}  // namespace tc
}  // namespace gforb
