// B200-native ORB extraction: batch-first CUDA implementation of ORB_SLAM::ORBextractor::operator()
// (reference src/ORBextractor.cc:769-875).  Stages (each a kernel over the whole batch):
//   K1a border0      level-0 copy + reflect-101 ring                      (ComputePyramid :990)
//   K1b resize       level l-1 -> l, 11-bit fixed-point bilinear + ring    (ComputePyramid :943,:982)
//   K2a fast_nms     dense FAST-9/16 score + cell-confined 3x3 NMS, register-rolling strips (detect.cuh)
//   K2b cell_emit    per cell: ordered candidate list, counts at iniThFAST and minThFAST (ComputeKeyPoints :573-628)
//                    (k_detect = first-generation fused per-cell kernel, kept for cell grids that do not tile the level)
//   K3b quota        per (frame, level): threshold fallback + quota redistribution (:673-721)
//   K3c cell_retain  per cell: retainBest emulation                        (:733-745)
//   K3d level_retain per (frame, level): retainBest(nDesired)              (:748-752)
//   K4  orient       warp per keypoint: IC_Angle + KeyPoint emission       (:131-158, :738-745, :865-871)
//   K5  blur2        7x7 Gaussian of every level ROI, register-rolling strips (blur.cuh)   (:842)
//   K6  describe     warp per keypoint: steered BRIEF                      (:161-201)
// Data layout in HBM: per frame one "pyramid block" holding all bordered level planes (pitch multiple of 128 B,
// ROI at (16,16)), a second block with the blurred ROIs, a candidate buffer (packed u32 score|y|x), per-level
// keypoint buffers.  All integer work is bit-exact by construction; float steps use un-fused IEEE ops.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <type_traits>
#include <vector>
#include "common.cuh"
#include "host_util.h"

namespace gforb {

// ------------------------------------------------------------------------------------------------ plans
struct LevelPlan {
    int w, h, pitch, rows;        // ROI size; bordered plane = rows x pitch bytes (rows = h + 32)
    unsigned long long off;       // byte offset of the plane inside a frame's pyramid block
    int nDesired, levelCols, levelRows, cellW, cellH, nCells, cellBase, nfeaturesCell;
    int kpBase, kpCap;            // slice of the per-frame level-keypoint buffer
    int xtab, ytab;               // offsets (in int2 units) of the resize tables for l-1 -> l
    int qtab, ybtab;              // k_resize2: per column quad {base, selA | selB << 16, 4 weight pairs} (4 int2), per bordered row {src row, b0 | b1 << 16}
    int scaledPatchSize;
    float scale;
    int fnStripsX, fnRows[2], fnTaskBase[2];   // k_fast_nms strips (detect.cuh): [0] rows <= 64 per strip (small batches), [1] <= 128
    int blStripsX, blRows[2], blTaskBase[2];   // k_blur2 strips (blur.cuh), same two strip heights
};
// per-level constants that travel as a kernel parameter (constant bank, indexed by a warp-uniform level) instead of a
// dependent global load of the LevelPlan: the row-mapped keypoint kernels and k_cell_emit (<= KP_MAX_LEVELS levels)
static const int KP_MAX_LEVELS = MAX_LEVELS;   // = the create-time limit on nlevels (common.cuh), so the global-memory fallbacks are never taken today
struct KpLevel { unsigned long long off; int pitch, kpBase, w, h, scaledPatchSize; float scale; };
struct KpLevels { KpLevel lv[KP_MAX_LEVELS]; };
struct CellPlan {
    int level, x0, y0, x1, y1;    // interior rectangle in level ROI coordinates [x0,x1) x [y0,y1) (may be empty)
    int skipped;                  // 1: the reference `continue`s before FAST (hX<=0 / hY<=0)
    int candBase, candCap;
};

}  // namespace gforb
#include "detect.cuh"
#include "blur.cuh"
#include "quadtree.cuh"
namespace gforb {

// bit_pattern_31_ (src/ORBextractor.cc:204-462) as 256 x {x0, y0, x1, y1}; lane j of k_describe reads its 32 bytes with
// two 128-bit loads (a per-lane index into __constant__ memory would serialise on the ADU pipe: ncu r1_v1, adu=97%)
__device__ __align__(16) signed char g_pattern[256 * 4];
__constant__ int c_umax[16];
// second-generation keypoint kernels (k_orient2 / k_describe2): the pattern as floats (no I2F on the XU pipe) and the
// IC_Angle task weights of k_orient's shared table, built once on the host
__device__ __align__(16) float g_patternF[256 * 4];
__device__ __align__(8) int2 g_orientW[256];
__constant__ int c_gauss[2][7] = {{18, 34, 48, 56, 48, 34, 18}, {18, 34, 49, 55, 49, 34, 18}};

static const signed char h_pattern[256][4] = {
#include "brief_pattern.inc"
};

__device__ __forceinline__ int reflect101(int p, int len) {
    if (len == 1) return 0;
    while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p;
    return p;
}

// ------------------------------------------------------------------------------------------------ K1a
// level 0 = copyMakeBorder(image, 16, BORDER_REFLECT_101) (ComputePyramid :990-991).  Task = one 16-byte chunk of a
// bordered row; a CTA owns B0_ROWS bordered rows and runs the chunks in two groups so that warps do not diverge: the interior
// chunks (one aligned 128-bit load each) first, then the few chunks per row that touch the reflected ring or the pitch padding.
static const int B0_ROWS = 16;
__global__ void __launch_bounds__(256) k_border0(const uint8_t* __restrict__ img, int w, int h, int stride,
                                                 size_t frameStride, uint8_t* __restrict__ pyr, size_t framePyr,
                                                 LevelPlan L) {
    const int Y0 = blockIdx.x * B0_ROWS, b = blockIdx.y, tid = threadIdx.x;
    const int nRows = min(B0_ROWS, L.rows - Y0);
    const uint8_t* src = img + (size_t)b * frameStride;
    uint8_t* dst = pyr + (size_t)b * framePyr + L.off;
    const int nC = L.pitch >> 4;
    // chunk c (bytes 16c .. 16c + 15 of the bordered row) lies inside the image iff 1 <= c <= w / 16
    const bool aligned = (((reinterpret_cast<uintptr_t>(src)) | (uintptr_t)stride) & 15) == 0;
    const int cLo = aligned ? 1 : nC, cHi = aligned ? min(w / 16 + 1, nC) : nC;
    const int nFast = max(cHi - cLo, 0), nSlow = nC - nFast;
    for (int t = tid; t < nFast * nRows; t += 256) {
        const int r = t / nFast, c = cLo + (t - r * nFast), Y = Y0 + r;
        const uint8_t* srow = src + (size_t)reflect101(Y - EDGE, h) * stride;
        *reinterpret_cast<uint4*>(dst + (size_t)Y * L.pitch + (c << 4)) = ldg_nc_u4(srow + ((c << 4) - EDGE));
    }
    for (int t = tid; t < nSlow * nRows; t += 256) {
        const int r = t / nSlow, k = t - r * nSlow, Y = Y0 + r;
        const int c = (nFast > 0 && k >= cLo) ? k + nFast : k;
        const int X0 = c << 4;
        const uint8_t* srow = src + (size_t)reflect101(Y - EDGE, h) * stride;
        uint32_t wv[4];
#pragma unroll
        for (int q = 0; q < 4; q++) {
            uint32_t v = 0;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int X = X0 + q * 4 + j;
                const uint32_t px = (X < w + 2 * EDGE) ? srow[reflect101(X - EDGE, w)] : 0u;
                v |= px << (8 * j);
            }
            wv[q] = v;
        }
        *reinterpret_cast<uint4*>(dst + (size_t)Y * L.pitch + X0) = make_uint4(wv[0], wv[1], wv[2], wv[3]);
    }
}

// ------------------------------------------------------------------------------------------------ K1b
// level l = resize(level l-1, INTER_LINEAR) + reflect-101 ring (ComputePyramid :943, :982-983); cv::resize 8U fixed
// point with 11-bit coefficients: out = ((b0 * (h0 >> 4) >> 16) + (b1 * (h1 >> 4) >> 16) + 2) >> 2,
// h = s[sx] * a0 + s[sx + 1] * a1.  tabs: int2 {src index, a0 | a1 << 16} per destination column / row.
// Thread = 4 adjacent bordered columns x a band of RS_BAND bordered rows; the column taps live in registers.
static const int RS_BAND = 4;
__global__ void __launch_bounds__(256, 6) k_resize(const uint8_t* __restrict__ spyr, uint8_t* __restrict__ pyr, size_t framePyr,
                                                LevelPlan S, LevelPlan D, const int2* __restrict__ tabs, int quadsPerRow,
                                                int nBands) {
    const int t = blockIdx.x * 256 + threadIdx.x, b = blockIdx.y;
    if (t >= quadsPerRow * nBands) return;
    const int band = t / quadsPerRow, q = t - band * quadsPerRow;
    const int2* __restrict__ xt = tabs + D.xtab;
    const int2* __restrict__ yt = tabs + D.ytab;
    int sx[4], a0[4], a1[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int X = (q << 2) + k;
        const int x = reflect101(min(X, D.w + 2 * EDGE - 1) - EDGE, D.w);   // columns past w+32 are pitch padding
        const int2 tx = __ldg(&xt[x]);
        sx[k] = tx.x; a0[k] = (short)(tx.y & 0xffff); a1[k] = (short)(tx.y >> 16);
    }
    const uint8_t* __restrict__ sroi = spyr + (size_t)b * framePyr + S.off + (size_t)EDGE * S.pitch + EDGE;
    uint8_t* __restrict__ dst = pyr + (size_t)b * framePyr + D.off + (q << 2);
    // The kernel is bound by the latency of its gathers, not by their number: every output row fetches its two source
    // rows itself (no carried state, no branches), the band is fully unrolled and all table and pixel loads of the
    // band are independent of each other.
    const int Y0 = band * RS_BAND;
    int2 ty[RS_BAND];
#pragma unroll
    for (int i = 0; i < RS_BAND; i++) ty[i] = __ldg(&yt[reflect101(min(Y0 + i, D.rows - 1) - EDGE, D.h)]);
#pragma unroll
    for (int i = 0; i < RS_BAND; i++) {
        const int Y = Y0 + i;
        const int b0 = (short)(ty[i].y & 0xffff), b1 = (short)(ty[i].y >> 16);
        const uint8_t* r0 = sroi + (ptrdiff_t)ty[i].x * S.pitch;     // row sy+1 == S.h is the ring (b1 == 0 there)
        const uint8_t* r1 = r0 + S.pitch;
        uint32_t outv = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int h0 = (__ldg(r0 + sx[k]) * a0[k] + __ldg(r0 + sx[k] + 1) * a1[k]) >> 4;
            const int h1 = (__ldg(r1 + sx[k]) * a0[k] + __ldg(r1 + sx[k] + 1) * a1[k]) >> 4;
            const int v = (((b0 * h0) >> 16) + ((b1 * h1) >> 16) + 2) >> 2;
            outv |= (uint32_t)min(max(v, 0), 255) << (8 * k);
        }
        if (Y < D.rows) *reinterpret_cast<uint32_t*>(dst + (size_t)Y * D.pitch) = outv;
    }
}

// K1b, word-load variant (GFORB_RESIZE=words): same thread / band layout and the same fixed-point arithmetic as k_resize, but
// the 8 byte gathers per source row (two taps x four columns, each warp-wide gather = 2-3 L1 wavefronts) become three aligned
// 32-bit loads: the four columns of a thread read source bytes base .. base + 6 (base = their smallest tap; adjacent
// destination columns -- reflected ring columns included -- lie <= scale apart, the host checks the span), two funnel shifts
// bring bytes base .. base + 7 into a register pair, one PRMT per column pair picks {s[sx], s[sx+1]} of both columns with a
// per-thread selector, and the horizontal pass of a column is ONE dp2a (16-bit weights x 8-bit pixels) instead of two IMAD.
__device__ __forceinline__ int dp2a_lo_su(uint32_t w, uint32_t px) {
    int d;
    asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(w), "r"(px), "r"(0));
    return d;
}
__device__ __forceinline__ int dp2a_hi_su(uint32_t w, uint32_t px) {
    int d;
    asm("dp2a.hi.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(w), "r"(px), "r"(0));
    return d;
}
// Everything that depends only on the column quad (smallest tap, the two PRMT selectors, the four weight pairs) and on the
// bordered output row (source row, weight pair; reflect-101 already applied) comes from host-built tables (LevelPlan::qtab,
// ybtab): the per-thread prologue of k_resize (four reflect-101 loops + table lookups per axis) was 40 % of its instructions.
// The saturation of cv::resize is dropped: (b0 h0 >> 16) + (b1 h1 >> 16) <= 32656 * 2049 / 65536 < 1021, so the result
// (.. + 2) >> 2 lies in [0, 255] for every input.
template <int BAND>
__global__ void __launch_bounds__(256, 6) k_resize2(const uint8_t* __restrict__ spyr, uint8_t* __restrict__ pyr, size_t framePyr,
                                                 LevelPlan S, LevelPlan D, const int2* __restrict__ tabs, int quadsPerRow,
                                                 int nBands) {
    const int t = blockIdx.x * 256 + threadIdx.x, b = blockIdx.y;
    if (t >= quadsPerRow * nBands) return;
    const int band = t / quadsPerRow, q = t - band * quadsPerRow;
    const int4* __restrict__ qt = reinterpret_cast<const int4*>(tabs + D.qtab) + 2 * q;
    const int2* __restrict__ ybt = tabs + D.ybtab;
    const int4 qa = __ldg(qt), qb = __ldg(qt + 1);
    const int base = qa.x;
    const uint32_t selA = (uint32_t)qa.y & 0xffffu, selB = (uint32_t)qa.y >> 16;
    const uint32_t wg0 = (uint32_t)qa.z, wg1 = (uint32_t)qa.w, wg2 = (uint32_t)qb.x, wg3 = (uint32_t)qb.y;
    const int sh = (base & 3) * 8;
    const uint8_t* __restrict__ sroi = spyr + (size_t)b * framePyr + S.off + (size_t)EDGE * S.pitch + EDGE + (base & ~3);
    uint8_t* __restrict__ dst = pyr + (size_t)b * framePyr + D.off + (q << 2);
    const int Y0 = band * BAND, spw = S.pitch >> 2;
    int2 ty[BAND];
#pragma unroll
    for (int i = 0; i < BAND; i++) ty[i] = __ldg(&ybt[min(Y0 + i, D.rows - 1)]);
#pragma unroll
    for (int i = 0; i < BAND; i++) {
        const int Y = Y0 + i;
        const int b0 = ty[i].y & 0xffff, b1 = (int)((uint32_t)ty[i].y >> 16);
        const uint32_t* r0 = reinterpret_cast<const uint32_t*>(sroi + (ptrdiff_t)ty[i].x * S.pitch);   // row sy+1 == S.h is the ring (b1 == 0 there)
        const uint32_t* r1 = r0 + spw;
        const uint32_t u0 = __ldg(r0), u1 = __ldg(r0 + 1), u2 = __ldg(r0 + 2);
        const uint32_t v0 = __ldg(r1), v1 = __ldg(r1 + 1), v2 = __ldg(r1 + 2);
        const uint32_t U0 = __funnelshift_r(u0, u1, sh), U1 = __funnelshift_r(u1, u2, sh);
        const uint32_t V0 = __funnelshift_r(v0, v1, sh), V1 = __funnelshift_r(v1, v2, sh);
        const uint32_t uA = __byte_perm(U0, U1, selA), uB = __byte_perm(U0, U1, selB);
        const uint32_t vA = __byte_perm(V0, V1, selA), vB = __byte_perm(V0, V1, selB);
        const int h0[4] = {dp2a_lo_su(wg0, uA) >> 4, dp2a_hi_su(wg1, uA) >> 4, dp2a_lo_su(wg2, uB) >> 4, dp2a_hi_su(wg3, uB) >> 4};
        const int h1[4] = {dp2a_lo_su(wg0, vA) >> 4, dp2a_hi_su(wg1, vA) >> 4, dp2a_lo_su(wg2, vB) >> 4, dp2a_hi_su(wg3, vB) >> 4};
        uint32_t outv = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int v = (((b0 * h0[k]) >> 16) + ((b1 * h1[k]) >> 16) + 2) >> 2;
            outv |= (uint32_t)v << (8 * k);
        }
        if (Y < D.rows) *reinterpret_cast<uint32_t*>(dst + (size_t)Y * D.pitch) = outv;
    }
}

// K1b, TMA variant (A/B partner of k_resize, GFORB_RESIZE=tma): CTA = RT_ROWS bordered output rows x the whole width.  The
// source rows those output rows read (<= ceil(RT_ROWS * scale) + 3 consecutive rows of level l-1, contiguous in the bordered
// plane) arrive in shared memory through ONE cp.async.bulk (TMA bulk copy, mbarrier complete_tx); the byte gathers of the
// bilinear taps then hit shared memory instead of L1/L2.  Same arithmetic, same output bytes.
static const int RT_ROWS = 8, RT_MAXSRC = 14;
__global__ void __launch_bounds__(256) k_resize_tma(const uint8_t* __restrict__ spyr, uint8_t* __restrict__ pyr, size_t framePyr,
                                                    LevelPlan S, LevelPlan D, const int2* __restrict__ tabs, int quadsPerRow) {
    extern __shared__ __align__(128) uint8_t rsm[];
    __shared__ uint64_t bar;
    __shared__ int s_sy0;
    const int tid = threadIdx.x, b = blockIdx.y;
    const int Y0 = blockIdx.x * RT_ROWS;
    const int2* __restrict__ xt = tabs + D.xtab;
    const int2* __restrict__ yt = tabs + D.ytab;
    if (tid == 0) {
        int lo = INT_MAX, hi = -1;
        for (int i = 0; i < RT_ROWS; i++) {
            const int sy = __ldg(&yt[reflect101(min(Y0 + i, D.rows - 1) - EDGE, D.h)]).x;
            lo = min(lo, sy); hi = max(hi, sy + 1);
        }
        const int nrows = min(hi - lo + 1, RT_MAXSRC);
        s_sy0 = lo;
        const uint32_t bytes = (uint32_t)nrows * (uint32_t)S.pitch;
        const uint8_t* src = spyr + (size_t)b * framePyr + S.off + (size_t)(EDGE + lo) * S.pitch;
        const uint32_t barA = (uint32_t)__cvta_generic_to_shared(&bar), dstA = (uint32_t)__cvta_generic_to_shared(rsm);
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(barA));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(barA), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dstA), "l"(src), "r"(bytes), "r"(barA) : "memory");
    }
    __syncthreads();
    {
        const uint32_t barA = (uint32_t)__cvta_generic_to_shared(&bar);
        uint32_t ok = 0;
        while (!ok)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(ok) : "r"(barA) : "memory");
    }
    const int sy0 = s_sy0;
    const uint8_t* __restrict__ sroi = rsm + EDGE;                      // ROI column 0 of source row sy0
    uint8_t* __restrict__ dplane = pyr + (size_t)b * framePyr + D.off;
    const int items = RT_ROWS * quadsPerRow;
    for (int it = tid; it < items; it += 256) {
        const int r = it / quadsPerRow, q = it - r * quadsPerRow;
        const int Y = Y0 + r;
        if (Y >= D.rows) break;
        const int2 ty = __ldg(&yt[reflect101(Y - EDGE, D.h)]);
        const int b0 = (short)(ty.y & 0xffff), b1 = (short)(ty.y >> 16);
        const uint8_t* r0 = sroi + (ptrdiff_t)(ty.x - sy0) * S.pitch;
        const uint8_t* r1 = r0 + S.pitch;
        uint32_t outv = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int X = (q << 2) + k;
            const int x = reflect101(min(X, D.w + 2 * EDGE - 1) - EDGE, D.w);
            const int2 tx = __ldg(&xt[x]);
            const int sx = tx.x, a0 = (short)(tx.y & 0xffff), a1 = (short)(tx.y >> 16);
            const int h0 = (r0[sx] * a0 + r0[sx + 1] * a1) >> 4;
            const int h1 = (r1[sx] * a0 + r1[sx + 1] * a1) >> 4;
            const int v = (((b0 * h0) >> 16) + ((b1 * h1) >> 16) + 2) >> 2;
            outv |= (uint32_t)min(max(v, 0), 255) << (8 * k);
        }
        *reinterpret_cast<uint32_t*>(dplane + (size_t)Y * D.pitch + (q << 2)) = outv;
    }
}

// ------------------------------------------------------------------------------------------------ K2 detect
// FAST-9/16 corner score in the pixel domain: with M1 = min over the 16 arcs of (max of the arc's 9 pixels) and
// M2 = max over arcs of (min of the arc), cornerScore = max(v - M1, M2 - v) - 1 and corner(t) <=> score >= t
// (OpenCV's cornerScore<16> restated).  Packed evaluation: two horizontally adjacent pixels share
// one u16x2 register and the arc windows are reduced with the native 3-input VIMNMX3.U16x2.
static const int DET_THREADS = 256;
static const int DET_BAND = 16;           // interior rows per band

__device__ __forceinline__ uint32_t pair_at(const uint8_t* row, int x) {
    // bytes row[x], row[x+1] -> u16x2 {lo = row[x], hi = row[x+1]}
    return (uint32_t)row[x] | ((uint32_t)row[x + 1] << 16);
}

// scores for the pixel pair (x, x+1) on smem row pointer `c` (centre row), row stride ps
__device__ __forceinline__ uint32_t fast_score_pair(const uint8_t* c, int ps, int minTh) {
    uint32_t p[16];
    p[0] = pair_at(c + 3 * ps, 0);   p[1] = pair_at(c + 3 * ps, 1);   p[2] = pair_at(c + 2 * ps, 2);
    p[3] = pair_at(c + ps, 3);       p[4] = pair_at(c, 3);            p[5] = pair_at(c - ps, 3);
    p[6] = pair_at(c - 2 * ps, 2);   p[7] = pair_at(c - 3 * ps, 1);   p[8] = pair_at(c - 3 * ps, 0);
    p[9] = pair_at(c - 3 * ps, -1);  p[10] = pair_at(c - 2 * ps, -2); p[11] = pair_at(c - ps, -3);
    p[12] = pair_at(c, -3);          p[13] = pair_at(c + ps, -3);     p[14] = pair_at(c + 2 * ps, -2);
    p[15] = pair_at(c + 3 * ps, -1);
    const uint32_t v = pair_at(c, 0);
    uint32_t mn3[16], mx3[16];
#pragma unroll
    for (int i = 0; i < 16; i++) {
        mn3[i] = __vimin3_u16x2(p[i], p[(i + 1) & 15], p[(i + 2) & 15]);
        mx3[i] = __vimax3_u16x2(p[i], p[(i + 1) & 15], p[(i + 2) & 15]);
    }
    uint32_t M1 = 0xffffffffu, M2 = 0u;   // M1 = min over arcs of max9, M2 = max over arcs of min9
#pragma unroll
    for (int i = 0; i < 16; i++) {
        uint32_t mx9 = __vimax3_u16x2(mx3[i], mx3[(i + 3) & 15], mx3[(i + 6) & 15]);
        uint32_t mn9 = __vimin3_u16x2(mn3[i], mn3[(i + 3) & 15], mn3[(i + 6) & 15]);
        M1 = __vminu2(M1, mx9);
        M2 = __vmaxu2(M2, mn9);
    }
    // score = max(v - M1, M2 - v) - 1, clamped at 0 (values below minTh are irrelevant and zeroed)
    uint32_t a = __vsubus2(v, M1), bq = __vsubus2(M2, v);
    uint32_t s = __vsubus2(__vmaxu2(a, bq), 0x00010001u);
    uint32_t lo = s & 0xffffu, hi = s >> 16;
    lo = lo >= (uint32_t)minTh ? lo : 0u;
    hi = hi >= (uint32_t)minTh ? hi : 0u;
    return lo | (hi << 16);
}

__global__ void __launch_bounds__(DET_THREADS) k_detect(const uint8_t* __restrict__ pyr, size_t framePyr,
                                                        const LevelPlan* __restrict__ levels,
                                                        const CellPlan* __restrict__ cells, int nCellsTotal,
                                                        uint32_t* __restrict__ cand, size_t frameCand,
                                                        int* __restrict__ cnt7, int* __restrict__ cnt20,
                                                        int iniTh, int minTh, int pixPitch, int* __restrict__ errflag) {
    extern __shared__ __align__(16) uint8_t smem[];
    const int cellId = blockIdx.x, b = blockIdx.y;
    const CellPlan C = cells[cellId];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    __shared__ int s_rowCnt[DET_BAND];
    __shared__ int s_n20, s_base;
    if (tid == 0) { s_n20 = 0; s_base = 0; }
    const int Wc = C.x1 - C.x0, Hc = C.y1 - C.y0;
    if (C.skipped || Wc <= 0 || Hc <= 0) {
        if (tid == 0) { cnt7[(size_t)b * nCellsTotal + cellId] = 0; cnt20[(size_t)b * nCellsTotal + cellId] = 0; }
        return;
    }
    const LevelPlan L = levels[C.level];
    const uint8_t* roi = pyr + (size_t)b * framePyr + L.off + (size_t)EDGE * L.pitch + EDGE;
    // smem: pixel tile [DET_BAND + 8][pixPitch] then score tile [DET_BAND + 2][scPitch], scPitch = pixPitch
    uint8_t* pix = smem;
    uint8_t* sc = smem + (size_t)(DET_BAND + 8) * pixPitch;
    uint32_t* masks = reinterpret_cast<uint32_t*>(sc + (size_t)(DET_BAND + 2) * pixPitch);
    const int maskWords = (Wc + 31) >> 5;
    const int gx0 = (C.x0 - 4) & ~3;                 // global x of smem column 0 (4-aligned, >= x0-7)
    const int spanW = ((C.x1 + 4 + 3) & ~3) - gx0;    // columns staged (multiple of 4)
    uint32_t* candOut = cand + (size_t)b * frameCand + C.candBase;
    __syncthreads();

    for (int yb = C.y0; yb < C.y1; yb += DET_BAND) {
        const int rowsB = min(DET_BAND, C.y1 - yb);
        // ---- stage pixel rows [yb-4, yb+rowsB+4) x [gx0, gx0+spanW) (always inside the bordered plane: the
        //      interior lies >= 16 px inside the level)
        const int nRowsPix = rowsB + 8, wordsPerRow = spanW >> 2;
        for (int i = tid; i < nRowsPix * wordsPerRow; i += DET_THREADS) {
            int r = i / wordsPerRow, wq = i - r * wordsPerRow;
            const uint8_t* g = roi + (ptrdiff_t)(yb - 4 + r) * L.pitch + gx0 + (wq << 2);
            *reinterpret_cast<uint32_t*>(pix + r * pixPitch + (wq << 2)) = __ldg(reinterpret_cast<const uint32_t*>(g));
        }
        // zero the score tile (rows yb-1 .. yb+rowsB, cols x0-1 .. x1)
        for (int i = tid; i < ((rowsB + 2) * pixPitch) >> 2; i += DET_THREADS) reinterpret_cast<uint32_t*>(sc)[i] = 0u;
        __syncthreads();
        // ---- scores for interior rows max(yb-1,y0) .. min(yb+rowsB, y1-1), pixel pairs along x
        const int ys0 = max(yb - 1, C.y0), ys1 = min(yb + rowsB + 1, C.y1);
        const int pairsPerRow = (Wc + 1) >> 1;
        for (int i = tid; i < (ys1 - ys0) * pairsPerRow; i += DET_THREADS) {
            int r = i / pairsPerRow, pq = i - r * pairsPerRow;
            int y = ys0 + r, x = C.x0 + (pq << 1);
            const uint8_t* c = pix + (y - (yb - 4)) * pixPitch + (x - gx0);
            uint32_t s2 = fast_score_pair(c, pixPitch, minTh);
            uint8_t* so = sc + (y - (yb - 1)) * pixPitch + (x - C.x0 + 1);
            so[0] = (uint8_t)(s2 & 0xff);
            if (x + 1 < C.x1) so[1] = (uint8_t)(s2 >> 16);
        }
        if (tid < DET_BAND) s_rowCnt[tid] = 0;
        __syncthreads();
        // ---- NMS (8-neighbourhood, strict, neighbours outside the cell interior are 0) -> per-row bit masks
        for (int r = warp; r < rowsB; r += DET_THREADS / 32) {
            const uint8_t* s1 = sc + (r + 1) * pixPitch + 1;    // score row of y = yb + r, col index = x - x0
            int rowCount = 0, n20 = 0;
            for (int wq = 0; wq < maskWords; wq++) {
                int xr = (wq << 5) + lane;
                bool keep = false;
                if (xr < Wc) {
                    int s = s1[xr];
                    if (s) {
                        const uint8_t* u = s1 - pixPitch + xr;
                        const uint8_t* d = s1 + pixPitch + xr;
                        keep = s > s1[xr - 1] && s > s1[xr + 1] && s > u[-1] && s > u[0] && s > u[1] && s > d[-1] &&
                               s > d[0] && s > d[1];
                        if (keep && s >= iniTh) n20++;
                    }
                }
                uint32_t m = __ballot_sync(0xffffffffu, keep);
                if (lane == 0) masks[r * maskWords + wq] = m;
                rowCount += __popc(m);
            }
            n20 = __reduce_add_sync(0xffffffffu, n20);
            if (lane == 0) { s_rowCnt[r] = rowCount; if (n20) atomicAdd(&s_n20, n20); }
        }
        __syncthreads();
        // ---- ordered (row-major) emission
        const int base0 = s_base;
        for (int r = warp; r < rowsB; r += DET_THREADS / 32) {
            int off = base0;
            for (int q = 0; q < r; q++) off += s_rowCnt[q];
            const uint8_t* s1 = sc + (r + 1) * pixPitch + 1;
            for (int wq = 0; wq < maskWords; wq++) {
                uint32_t m = masks[r * maskWords + wq];
                if ((m >> lane) & 1u) {
                    int xr = (wq << 5) + lane;
                    int pos = off + __popc(m & ((1u << lane) - 1u));
                    if (pos < C.candCap)
                        candOut[pos] = ((uint32_t)s1[xr] << 24) | ((uint32_t)(yb + r) << 12) | (uint32_t)(C.x0 + xr);
                    else
                        atomicExch(errflag, GFORB_ERR_OVERFLOW);
                }
                off += __popc(m);
            }
        }
        __syncthreads();
        if (tid == 0) {
            int t = 0;
            for (int q = 0; q < rowsB; q++) t += s_rowCnt[q];
            s_base = base0 + t;
        }
        __syncthreads();
    }
    if (tid == 0) {
        cnt7[(size_t)b * nCellsTotal + cellId] = s_base;
        cnt20[(size_t)b * nCellsTotal + cellId] = s_n20;
    }
}

// ------------------------------------------------------------------------------------------------ K3b quota
// src/ORBextractor.cc:621-628 (threshold fallback), :673-721 (quota redistribution).  One warp per (frame, level): the cells of
// the level are spread over the lanes, the loop-carried totals (all integers) are warp reductions, the output offsets a warp scan.
__global__ void __launch_bounds__(128) k_quota(const LevelPlan* __restrict__ levels, const CellPlan* __restrict__ cells, int nLevels,
                        int nCellsTotal, const int* __restrict__ cnt7, const int* __restrict__ cnt20,
                        int* __restrict__ nTotalOut, int* __restrict__ nToRetainOut, int* __restrict__ cellOutOff,
                        int* __restrict__ lvlPre, int batch, int* __restrict__ errflag) {
    const unsigned full = 0xffffffffu;
    const int gi = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (gi >= batch * nLevels) return;
    const int b = gi / nLevels, l = gi - b * nLevels;
    const LevelPlan L = levels[l];
    const size_t cb = (size_t)b * nCellsTotal + L.cellBase;
    int* nTotal = nTotalOut + cb;
    int* nToRetain = nToRetainOut + cb;
    // bNoMore is kept in cellOutOff during the loop (scratch), rewritten as offsets at the end
    int* noMore = cellOutOff + cb;
    int nNoMore = 0, nToDistribute = 0;
    for (int c = lane; c < L.nCells; c += 32) {
        if (cells[L.cellBase + c].skipped) { nTotal[c] = 0; nToRetain[c] = 0; noMore[c] = 0; continue; }
        int n20 = cnt20[cb + c], n7 = cnt7[cb + c];
        int nKeys = (n20 <= 3) ? n7 : n20;          // "if size()<=3 -> clear; FAST(cell, 7)"
        nTotal[c] = (n20 <= 3) ? nKeys : -nKeys - 1; // sign encodes which threshold produced the list (>=0: minTh)
        if (nKeys > L.nfeaturesCell) { nToRetain[c] = L.nfeaturesCell; noMore[c] = 0; }
        else { nToRetain[c] = nKeys; nToDistribute += L.nfeaturesCell - nKeys; noMore[c] = 1; nNoMore++; }
    }
    nNoMore = __reduce_add_sync(full, nNoMore);
    nToDistribute = __reduce_add_sync(full, nToDistribute);
    while (nToDistribute > 0 && nNoMore < L.nCells) {
        const int nNew = L.nfeaturesCell + (int)ceilf(__fdiv_rn((float)nToDistribute, (float)(L.nCells - nNoMore)));
        int dist = 0, more = 0;
        for (int c = lane; c < L.nCells; c += 32) {
            if (!noMore[c]) {
                int tot = nTotal[c] >= 0 ? nTotal[c] : -nTotal[c] - 1;
                if (tot > nNew) { nToRetain[c] = nNew; }
                else { nToRetain[c] = tot; dist += nNew - tot; noMore[c] = 1; more++; }
            }
        }
        nToDistribute = __reduce_add_sync(full, dist);
        nNoMore += __reduce_add_sync(full, more);
    }
    int off = 0;
    for (int c0 = 0; c0 < L.nCells; c0 += 32) {
        const int c = c0 + lane;
        const int r = c < L.nCells ? nToRetain[c] : 0;
        int inc = r;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(full, inc, d); if (lane >= d) inc += t; }
        if (c < L.nCells) noMore[c] = off + inc - r;
        off += __shfl_sync(full, inc, 31);
    }
    if (lane == 0) {
        lvlPre[(size_t)b * nLevels + l] = off;
        if (off > L.kpCap) atomicExch(errflag, GFORB_ERR_OVERFLOW);
    }
}

// ------------------------------------------------------------------------------------------------ warp-parallel nth_element
// The permutation of nth_element_libstdcxx (introselect.h), every Hoare partition pass done by the 32 lanes of a warp.
// __unguarded_partition swaps the k-th element from the left that is not "before" the pivot (a left stop: !comp(a, p)) with
// the k-th from the right that is not "after" it (a right stop: !comp(p, a)) while the left one lies left of the right one.
// Both pointers only ever see unmodified elements before they meet, so the pairs follow from the ORIGINAL array:
//   left stop i (rank = left stops before i) is swapped  iff  right stops after i  > rank,
//   right stop j (rank = right stops after j) is swapped  iff  left stops before j  > rank,
// and the returned cut is min(first unswapped left stop, last swapped right stop) -- the place where the scanning pointer
// is halted by an element the previous swap put there.  tmp: size / 2 entries, info: size entries of scratch.
template <class T, class C>
__device__ void nth_element_warp(T* a, int size, int nthIdx, C comp, T* tmp, uint16_t* info, int lane) {
    const unsigned full = 0xffffffffu;
    if (size <= 0 || nthIdx >= size) return;
    const unsigned lt = (1u << lane) - 1u, le = lt | (1u << lane);
    int first = 0, last = size;
    int depth = 2 * (31 - __clz(size));
    while (last - first > 3) {
        if (depth == 0) {                                  // introsort's heap-select fallback: sequential (never seen on images)
            if (lane == 0) { is_heap_select(a + first, a + nthIdx + 1, a + last, comp); gf_swap(a[first], a[nthIdx]); }
            __syncwarp();
            return;
        }
        --depth;
        if (lane == 0) is_move_median_to_first(a + first, a + first + 1, a + first + (last - first) / 2, a + last - 1, comp);
        __syncwarp();
        const T p = a[first];
        const int lo = first + 1, hi = last;
        int totalR = 0;
        for (int c0 = lo; c0 < hi; c0 += 32) {
            const int i = c0 + lane;
            const bool isR = i < hi && !comp(p, a[i]);
            totalR += __popc(__ballot_sync(full, isR));
        }
        int baseL = 0, baseR = 0, cut = INT_MAX;
        for (int c0 = lo; c0 < hi; c0 += 32) {
            const int i = c0 + lane;
            const bool in = i < hi;
            const T v = in ? a[i] : p;
            const bool isL = in && !comp(v, p), isR = in && !comp(p, v);
            const unsigned mL = __ballot_sync(full, isL), mR = __ballot_sync(full, isR);
            const int prefL = baseL + __popc(mL & lt);                    // left stops before i
            const int sufR = totalR - (baseR + __popc(mR & le));          // right stops after i
            const bool swL = isL && sufR > prefL, swR = isR && prefL > sufR;
            if (swR) tmp[sufR] = v;
            if (in) info[i] = swL ? (uint16_t)(0x8000 | prefL) : swR ? (uint16_t)(0x4000 | sufR) : (uint16_t)0;
            if ((isL && !swL) || swR) cut = min(cut, i);
            baseL += __popc(mL); baseR += __popc(mR);
        }
        __syncwarp();
        for (int i = lo + lane; i < hi; i += 32) {
            const uint16_t f = info[i];
            if (f & 0x8000) { const int k = f & 0x3fff; const T v = tmp[k]; tmp[k] = a[i]; a[i] = v; }
        }
        __syncwarp();
        for (int i = lo + lane; i < hi; i += 32) {
            const uint16_t f = info[i];
            if (f & 0x4000) a[i] = tmp[f & 0x3fff];
        }
        __syncwarp();
        cut = __reduce_min_sync(full, cut);
        if (cut <= nthIdx) first = cut; else last = cut;
    }
    if (lane == 0) is_insertion_sort(a + first, a + last, comp);
    __syncwarp();
}

// retain_best_first_n (introselect.h) with the warp-parallel nth_element; all 32 lanes call it
template <class T, class C>
__device__ int retain_best_first_n_warp(T* v, int size, int n, int variant, C comp, T* tmp, uint16_t* info, int lane) {
    if (n < 0 || size <= n) return size;
    if (n == 0) return 0;
    nth_element_warp(v, size, variant ? n - 1 : n, comp, tmp, info, lane);
    return n;
}

// ------------------------------------------------------------------------------------------------ K3c
// per cell: filter the candidate list to the threshold in force (order preserved), retainBest, append at the
// cell's offset of the level buffer.  One warp per cell; the selection itself is sequential (lane 0).
static const int RET_WARPS = 4;
static const int RET_SMEM = 768;    // entries per warp kept in shared memory (3 KB: 16 CTAs = 64 warps per SM)
__global__ void __launch_bounds__(RET_WARPS * 32) k_cell_retain(const CellPlan* __restrict__ cells,
                                                                const LevelPlan* __restrict__ levels, int nCellsTotal,
                                                                uint32_t* __restrict__ cand, size_t frameCand,
                                                                const int* __restrict__ cnt7,
                                                                const int* __restrict__ nTotalArr,
                                                                const int* __restrict__ nToRetainArr,
                                                                const int* __restrict__ cellOutOff,
                                                                uint32_t* __restrict__ lvlKp, size_t frameKp, int iniTh,
                                                                int retainVariant) {
    __shared__ uint32_t s_buf[RET_WARPS][RET_SMEM];
    __shared__ uint32_t s_tmp[RET_WARPS][RET_SMEM / 2];      // scratch of the warp-parallel selection
    __shared__ uint16_t s_info[RET_WARPS][RET_SMEM];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int cellId = blockIdx.x * RET_WARPS + warp, b = blockIdx.y;
    if (cellId >= nCellsTotal) return;
    const CellPlan C = cells[cellId];
    const size_t ci = (size_t)b * nCellsTotal + cellId;
    const int nKeep = nToRetainArr[ci];
    if (nKeep <= 0) return;
    const int enc = nTotalArr[ci];
    const bool useIni = enc < 0;
    const int nTot = useIni ? -enc - 1 : enc;
    const int nAll = cnt7[ci];
    uint32_t* list = cand + (size_t)b * frameCand + C.candBase;
    uint32_t* work = (nTot <= RET_SMEM) ? s_buf[warp] : list;     // large lists: compact in place in global memory
    // ordered filter (in place is safe: writes never pass reads within a 32-chunk processed after a warp sync)
    int outN = 0;
    for (int i0 = 0; i0 < nAll; i0 += 32) {
        int i = i0 + lane;
        uint32_t v = (i < nAll) ? list[i] : 0u;
        bool keep = (i < nAll) && (!useIni || (int)(v >> 24) >= iniTh);
        uint32_t m = __ballot_sync(0xffffffffu, keep);
        __syncwarp();
        if (keep) work[outN + __popc(m & ((1u << lane) - 1u))] = v;
        outN += __popc(m);
        __syncwarp();
    }
    __syncwarp();
    if (outN > nKeep) {
        if (outN <= RET_SMEM && !(retainVariant & 2)) retain_best_first_n_warp(work, outN, nKeep, retainVariant & 1, ScoreGreater(), s_tmp[warp], s_info[warp], lane);
        else if (lane == 0) retain_best_first_n(work, outN, nKeep, retainVariant & 1, ScoreGreater());    // huge lists / probe switch: sequential
    }
    __syncwarp();
    const int dstOff = cellOutOff[ci];
    if (dstOff + nKeep > levels[C.level].kpCap) return;      // overflow already flagged by k_quota
    uint32_t* dst = lvlKp + (size_t)b * frameKp + levels[C.level].kpBase + dstOff;
    for (int i = lane; i < nKeep; i += 32) dst[i] = work[i];
}

// ------------------------------------------------------------------------------------------------ K3d
__global__ void __launch_bounds__(128) k_level_retain(const LevelPlan* __restrict__ levels, int nLevels,
                                                      uint32_t* __restrict__ lvlKp, size_t frameKp,
                                                      const int* __restrict__ lvlPre, int* __restrict__ lvlCount,
                                                      int batch, int retainVariant, int scratchCap) {
    extern __shared__ __align__(16) unsigned char lr_smem[];        // per warp: scratchCap / 2 u32 + scratchCap u16 (0: sequential)
    const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (gw >= batch * nLevels) return;
    const int b = gw / nLevels, l = gw - b * nLevels;
    const LevelPlan L = levels[l];
    int n = lvlPre[gw];
    if (n > L.kpCap) n = L.kpCap;
    uint32_t* v = lvlKp + (size_t)b * frameKp + L.kpBase;
    if (n > L.nDesired) {
        if (n <= scratchCap && !(retainVariant & 2)) {
            uint32_t* tmp = reinterpret_cast<uint32_t*>(lr_smem + (size_t)warp * scratchCap * 4);
            uint16_t* info = reinterpret_cast<uint16_t*>(tmp + scratchCap / 2);
            n = retain_best_first_n_warp(v, n, L.nDesired, retainVariant & 1, ScoreGreater(), tmp, info, lane);
        } else {
            if (lane == 0) retain_best_first_n(v, n, L.nDesired, retainVariant & 1, ScoreGreater());
            n = L.nDesired;
        }
    }
    if (lane == 0) lvlCount[gw] = n;
}

// ------------------------------------------------------------------------------------------------ K3e HARRIS_SCORE
// scoreType == HARRIS_SCORE (src/ORBextractor.cc:667-671): the FAST keypoints of a cell are re-scored with
// HarrisResponses (:86-127: 7x7 block, Sobel-like integer gradients, float combination) before the quota loop, and both
// retainBest passes then order by that float response.  The candidate lists stay as FAST wrote them; the retain kernels
// of this mode work on 64-bit entries {response bits, packed FAST entry} in a global scratch list and keep the
// response of every survivor in lvlResp (parallel to lvlKp), which k_orient copies into KeyPoint::response.
struct HarrisGreater {
    GF_HD bool operator()(uint64_t a, uint64_t b) const {
        const uint32_t x = (uint32_t)(a >> 32), y = (uint32_t)(b >> 32);
        float fa, fb;
        memcpy(&fa, &x, 4); memcpy(&fb, &y, 4);
        return fa > fb;
    }
};

// p = the keypoint's pixel on the (unblurred) level plane
__device__ float harris_response(const uint8_t* __restrict__ p, int step) {
    int a = 0, b = 0, c = 0;
    for (int i = -3; i <= 3; i++) {
        const uint8_t* r = p + i * step;
        for (int j = -3; j <= 3; j++) {
            const uint8_t* q = r + j;
            const int Ix = ((int)q[1] - (int)q[-1]) * 2 + ((int)q[-step + 1] - (int)q[-step - 1]) + ((int)q[step + 1] - (int)q[step - 1]);
            const int Iy = ((int)q[step] - (int)q[-step]) * 2 + ((int)q[step - 1] - (int)q[-step - 1]) + ((int)q[step + 1] - (int)q[-step + 1]);
            a += Ix * Ix; b += Iy * Iy; c += Ix * Iy;
        }
    }
    float scale = __fmul_rn((float)((1 << 2) * 7), 255.0f);
    scale = __fdiv_rn(1.0f, scale);
    const float s4 = __fmul_rn(__fmul_rn(__fmul_rn(scale, scale), scale), scale);
    const float fa = (float)a, fb = (float)b, fc = (float)c;
    const float sum = __fadd_rn(fa, fb);
    const float t = __fsub_rn(__fsub_rn(__fmul_rn(fa, fb), __fmul_rn(fc, fc)), __fmul_rn(__fmul_rn(0.04f, sum), sum));
    return __fmul_rn(t, s4);
}

__global__ void __launch_bounds__(RET_WARPS * 32) k_cell_retain_harris(
    const CellPlan* __restrict__ cells, const LevelPlan* __restrict__ levels, int nCellsTotal, const uint32_t* __restrict__ cand,
    size_t frameCand, const int* __restrict__ cnt7, const int* __restrict__ nTotalArr, const int* __restrict__ nToRetainArr,
    const int* __restrict__ cellOutOff, uint32_t* __restrict__ lvlKp, float* __restrict__ lvlResp, size_t frameKp, int iniTh,
    int retainVariant, const uint8_t* __restrict__ pyr, size_t framePyr, uint64_t* __restrict__ work64, size_t frameWork) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int cellId = blockIdx.x * RET_WARPS + warp, b = blockIdx.y;
    if (cellId >= nCellsTotal) return;
    const CellPlan C = cells[cellId];
    const LevelPlan L = levels[C.level];
    const size_t ci = (size_t)b * nCellsTotal + cellId;
    const int nKeep = nToRetainArr[ci];
    if (nKeep <= 0) return;
    const int enc = nTotalArr[ci];
    const bool useIni = enc < 0;
    const int nAll = cnt7[ci];
    const uint32_t* list = cand + (size_t)b * frameCand + C.candBase;
    uint64_t* work = work64 + (size_t)b * frameWork + C.candBase;
    const uint8_t* roi = pyr + (size_t)b * framePyr + L.off + (size_t)EDGE * L.pitch + EDGE;
    int outN = 0;
    for (int i0 = 0; i0 < nAll; i0 += 32) {
        const int i = i0 + lane;
        const uint32_t v = (i < nAll) ? list[i] : 0u;
        const bool keep = (i < nAll) && (!useIni || (int)(v >> 24) >= iniTh);
        const uint32_t m = __ballot_sync(0xffffffffu, keep);
        if (keep) {
            const int x = v & 0xfff, y = (v >> 12) & 0xfff;
            const float r = harris_response(roi + (size_t)y * L.pitch + x, L.pitch);
            work[outN + __popc(m & ((1u << lane) - 1u))] = ((uint64_t)__float_as_uint(r) << 32) | v;
        }
        outN += __popc(m);
    }
    __syncwarp();
    if (lane == 0 && outN > nKeep) retain_best_first_n(work, outN, nKeep, retainVariant, HarrisGreater());
    __syncwarp();
    const int dstOff = cellOutOff[ci];
    if (dstOff + nKeep > L.kpCap) return;                     // overflow already flagged by k_quota
    const size_t d0 = (size_t)b * frameKp + L.kpBase + dstOff;
    for (int i = lane; i < nKeep; i += 32) {
        const uint64_t e = work[i];
        lvlKp[d0 + i] = (uint32_t)e;
        lvlResp[d0 + i] = __uint_as_float((uint32_t)(e >> 32));
    }
}

__global__ void __launch_bounds__(128) k_level_retain_harris(const LevelPlan* __restrict__ levels, int nLevels,
                                                             uint32_t* __restrict__ lvlKp, float* __restrict__ lvlResp,
                                                             size_t frameKp, const int* __restrict__ lvlPre,
                                                             int* __restrict__ lvlCount, int batch, int retainVariant,
                                                             uint64_t* __restrict__ work64, size_t frameWork) {
    const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (gw >= batch * nLevels) return;
    const int b = gw / nLevels, l = gw - b * nLevels;
    const LevelPlan L = levels[l];
    int n = lvlPre[gw];
    if (n > L.kpCap) n = L.kpCap;
    if (n > L.nDesired) {
        const size_t k0 = (size_t)b * frameKp + L.kpBase;
        uint64_t* w = work64 + (size_t)b * frameWork + L.kpBase;
        for (int i = lane; i < n; i += 32) w[i] = ((uint64_t)__float_as_uint(lvlResp[k0 + i]) << 32) | lvlKp[k0 + i];
        __syncwarp();
        int m = n;
        if (lane == 0) m = retain_best_first_n(w, n, L.nDesired, retainVariant, HarrisGreater());
        m = __shfl_sync(0xffffffffu, m, 0);
        __syncwarp();
        for (int i = lane; i < m; i += 32) {
            const uint64_t e = w[i];
            lvlKp[k0 + i] = (uint32_t)e;
            lvlResp[k0 + i] = __uint_as_float((uint32_t)(e >> 32));
        }
        n = m;
    }
    if (lane == 0) lvlCount[gw] = n;
}

// ------------------------------------------------------------------------------------------------ K4 orient
// IC_Angle on the unblurred level (src/ORBextractor.cc:131-158) + KeyPoint emission (:738-745, :865-871).
// One warp per keypoint slot.  The circular patch (31 rows, |u| <= umax[|v|]) is cut into 31 x 8 tasks of four
// horizontally adjacent pixels (u = -16 + 4c .. -13 + 4c); a lane takes 8 tasks, fetches the four bytes as two aligned
// words + one funnel shift and reduces them with two IDP.4A: sum u * I with the masked column offsets as signed byte
// weights, and the masked row sum (-> v * row sum).  All sums are exact int32.
__device__ __forceinline__ int dp4a_s8u8(int w, uint32_t px, int acc) {
    int d;
    asm("dp4a.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(w), "r"(px), "r"(acc));
    return d;
}

__global__ void __launch_bounds__(256) k_orient(const uint8_t* __restrict__ pyr, size_t framePyr,
                                                const LevelPlan* __restrict__ levels, int nLevels,
                                                const uint32_t* __restrict__ lvlKp, size_t frameKp, int kpCapTotal,
                                                const int* __restrict__ lvlCount, GforbKeyPoint* __restrict__ kps,
                                                int32_t* __restrict__ counts, int nfeatures,
                                                const float* __restrict__ lvlResp) {
    __shared__ int2 s_w[256];                 // task (row * 8 + chunk) -> {u weights (4 x s8), membership (4 x 0/1)}
    {
        const int row = threadIdx.x >> 3, c = threadIdx.x & 7, v = row - HALF_PATCH;
        const int d = row < 2 * HALF_PATCH + 1 ? c_umax[v < 0 ? -v : v] : -1;
        uint32_t wu = 0, w1 = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int u = -16 + 4 * c + k;
            if ((u < 0 ? -u : u) <= d) { wu |= (uint32_t)(u & 0xff) << (8 * k); w1 |= 1u << (8 * k); }
        }
        s_w[threadIdx.x] = make_int2((int)wu, (int)w1);
    }
    __syncthreads();
    const int slot = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), b = blockIdx.y, lane = threadIdx.x & 31;
    if (slot >= kpCapTotal) return;
    int l = 0, base = 0;
    while (l + 1 < nLevels && slot >= levels[l + 1].kpBase) { base += lvlCount[b * nLevels + l]; l++; }
    const LevelPlan L = levels[l];
    const int i = slot - L.kpBase;
    const int n = lvlCount[b * nLevels + l];
    if (slot == 0 && lane == 0) {
        int tot = 0;
        for (int q = 0; q < nLevels; q++) tot += lvlCount[b * nLevels + q];
        counts[b] = tot;
    }
    if (i >= n) return;
    const uint32_t e = lvlKp[(size_t)b * frameKp + slot];
    const int x = e & 0xfff, y = (e >> 12) & 0xfff, score = e >> 24;
    // plane column of u = -16 is EDGE + x - 16 = x; planes, pitches and plane offsets are multiples of 4 bytes
    const uint32_t* w00 = reinterpret_cast<const uint32_t*>(pyr + (size_t)b * framePyr + L.off + (size_t)(EDGE + y - HALF_PATCH) * L.pitch) +
                          (x >> 2) + (lane & 7);
    const int sh = (x & 3) * 8, pw = L.pitch >> 2;
    int m10 = 0, m01 = 0;
#pragma unroll
    for (int it = 0; it < 8; it++) {
        const int row = it * 4 + (lane >> 3);                     // 0..31, row 31 has empty weights
        const int2 w = s_w[it * 32 + lane];
        if (row < 2 * HALF_PATCH + 1) {
            const uint32_t* wp = w00 + row * pw;
            const uint32_t px = __funnelshift_r(__ldg(wp), __ldg(wp + 1), sh);
            m10 = dp4a_s8u8(w.x, px, m10);
            m01 += (row - HALF_PATCH) * (int)__dp4a((uint32_t)w.y, px, 0u);
        }
    }
    m10 = __reduce_add_sync(0xffffffffu, m10);
    m01 = __reduce_add_sync(0xffffffffu, m01);
    if (lane == 0) {
        GforbKeyPoint k;
        float fx = (float)x, fy = (float)y;
        if (l != 0) { fx = __fmul_rn(fx, L.scale); fy = __fmul_rn(fy, L.scale); }
        k.x = fx; k.y = fy;
        k.size = (float)L.scaledPatchSize;
        k.angle = fast_atan2_deg((float)m01, (float)m10);
        k.response = lvlResp ? lvlResp[(size_t)b * frameKp + slot] : (float)score;   // HARRIS_SCORE : FAST_SCORE
        k.octave = l;
        k.class_id = -1;
        kps[(size_t)b * nfeatures + base + i] = k;
    }
}

// ------------------------------------------------------------------------------------------------ K6 describe
// src/ORBextractor.cc:161-201.  Warp per keypoint, lane j builds descriptor byte j (8 tests, 16 taps).  Taps inside
// the level ROI read the blurred plane, taps in the 16-px ring read the unblurred reflect-101 ring (the reference
// blurs the ROI in place and leaves the ring untouched).
// ROWMAP (default): warp per output row j < nfeatures (kp_row_to_slot); otherwise warp per slot of the per-level lists
// (kpCapTotal slots, ~3.5 x nfeatures, most of which exit right after the level search: GFORB_KP=v1).
// The per-level constants the row-mapped keypoint kernels need travel as a kernel parameter (constant bank, indexed by the
// warp-uniform level) and the output row -> (level, list index) map is ONE load round: lane l reads the count of level l, a
// shuffle scan gives the prefix, a ballot the level.  (The first row-mapped build walked the counts with an early exit --
// up to nLevels dependent loads -- and then fetched the LevelPlan from global memory: k_orient2 ran at 19.6 long-scoreboard
// stalls per issued instruction.)
__device__ __forceinline__ bool kp_row_to_slot(int j, int b, int nLevels, const int* __restrict__ lvlCount, int lane, int& l, int& i, int& total) {
    const unsigned full = 0xffffffffu;
    const int c = lane < nLevels ? lvlCount[b * nLevels + lane] : 0;
    int inc = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(full, inc, d); if (lane >= d) inc += t; }
    total = __shfl_sync(full, inc, 31);
    const unsigned m = __ballot_sync(full, lane < nLevels && j < inc);
    if (!m) return false;
    l = __ffs(m) - 1;
    i = j - (__shfl_sync(full, inc, l) - __shfl_sync(full, c, l));
    return true;
}
template <bool ROWMAP, bool MAGIC>
__global__ void __launch_bounds__(256) k_describe(const uint8_t* __restrict__ pyr, const uint8_t* __restrict__ blr,
                                                  size_t framePyr, const LevelPlan* __restrict__ levels, int nLevels,
                                                  const uint32_t* __restrict__ lvlKp, size_t frameKp, int kpCapTotal,
                                                  const int* __restrict__ lvlCount,
                                                  const GforbKeyPoint* __restrict__ kps, uint8_t* __restrict__ desc,
                                                  int nfeatures, const __grid_constant__ KpLevels KL) {
    const int slot0 = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), b = blockIdx.y, lane = threadIdx.x & 31;
    int l = 0, base = 0, i, slot = slot0;
    KpLevel L;
    if (ROWMAP) {
        int total;
        if (slot0 >= nfeatures || !kp_row_to_slot(slot0, b, nLevels, lvlCount, lane, l, i, total)) return;
        base = slot0 - i;
        L = KL.lv[l];
        slot = L.kpBase + i;
    } else {
        if (slot >= kpCapTotal) return;
        while (l + 1 < nLevels && slot >= levels[l + 1].kpBase) { base += lvlCount[b * nLevels + l]; l++; }
        const LevelPlan G = levels[l];
        L.off = G.off; L.pitch = G.pitch; L.kpBase = G.kpBase; L.w = G.w; L.h = G.h;
        i = slot - L.kpBase;
        if (i >= lvlCount[b * nLevels + l]) return;
    }
    const uint32_t e = lvlKp[(size_t)b * frameKp + slot];
    const int cx = e & 0xfff, cy = (e >> 12) & 0xfff;
    const size_t oi = (size_t)b * nfeatures + base + i;
    const float factorPI = (float)(3.14159265358979323846 / 180.f);
    const float angle = __fmul_rn(kps[oi].angle, factorPI);
    const float a = glibc_cosf(angle), bb = glibc_sinf(angle);
    const uint8_t* P = pyr + (size_t)b * framePyr + L.off + (size_t)EDGE * L.pitch + EDGE;
    const uint8_t* B = blr + (size_t)b * framePyr + L.off + (size_t)EDGE * L.pitch + EDGE;
    const uint4 pw0 = reinterpret_cast<const uint4*>(g_pattern)[lane * 2], pw1 = reinterpret_cast<const uint4*>(g_pattern)[lane * 2 + 1];
    const uint32_t pw[8] = {pw0.x, pw0.y, pw0.z, pw0.w, pw1.x, pw1.y, pw1.z, pw1.w};
    int val = 0;
#pragma unroll
    for (int t = 0; t < 8; t++) {
        int tv[2];
#pragma unroll
        for (int q = 0; q < 2; q++) {
            const float px = (float)(signed char)(pw[t] >> (16 * q)), py = (float)(signed char)(pw[t] >> (16 * q + 8));
            const float fy = __fadd_rn(__fmul_rn(px, bb), __fmul_rn(py, a)), fx = __fsub_rn(__fmul_rn(px, a), __fmul_rn(py, bb));
            // MAGIC: cvRound by adding 1.5 * 2^23 (the float add rounds to nearest-even and leaves the integer in the mantissa,
            // bit image 0x4B400000 + n; identical to __float2int_rn for |v| < 2^22) -- keeps 32 conversions per lane off the XU pipe
            const int iy = MAGIC ? __float_as_int(__fadd_rn(fy, 12582912.f)) - 0x4B400000 : cv_round(fy);
            const int ix = MAGIC ? __float_as_int(__fadd_rn(fx, 12582912.f)) - 0x4B400000 : cv_round(fx);
            int x = cx + ix, y = cy + iy;
            const uint8_t* src = (x >= 0 && x < L.w && y >= 0 && y < L.h) ? B : P;
            tv[q] = src[(ptrdiff_t)y * L.pitch + x];
        }
        val |= (tv[0] < tv[1]) << t;
    }
    desc[oi * 32 + lane] = (uint8_t)val;
}

// ------------------------------------------------------------------------------------------------ K4b / K6b
// Second generation of the two keypoint kernels.  Both take one warp per OUTPUT row j < nfeatures (the level and the slot
// follow from the per-level counts) instead of one warp per slot of the over-provisioned per-level lists (kpCapTotal ~ 3.5 x
// nfeatures at the reference's defaults).  k_orient2 is the default (GFORB_KP=v1: k_orient + slot-mapped k_describe);
// k_describe2 is an A/B LOSER kept selectable (GFORB_DESCRIBE=smem) -- see its comment.

// IC_Angle as in k_orient; the task weights come from g_orientW (coalesced 8-byte loads into registers) instead of a
// table that every CTA rebuilt in shared memory through divergent __constant__ reads.
__global__ void __launch_bounds__(256) k_orient2(const uint8_t* __restrict__ pyr, size_t framePyr,
                                                 const LevelPlan* __restrict__ levels, int nLevels,
                                                 const uint32_t* __restrict__ lvlKp, size_t frameKp,
                                                 const int* __restrict__ lvlCount, GforbKeyPoint* __restrict__ kps,
                                                 int32_t* __restrict__ counts, int nfeatures,
                                                 const float* __restrict__ lvlResp, const __grid_constant__ KpLevels KL) {
    const int j = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), b = blockIdx.y, lane = threadIdx.x & 31;
    if (j >= nfeatures) return;
    int l, i, total;
    const bool live = kp_row_to_slot(j, b, nLevels, lvlCount, lane, l, i, total);
    if (j == 0 && lane == 0) counts[b] = total;
    if (!live) return;
    const KpLevel L = KL.lv[l];
    const int slot = L.kpBase + i;
    const uint32_t e = lvlKp[(size_t)b * frameKp + slot];
    const int x = e & 0xfff, y = (e >> 12) & 0xfff, score = e >> 24;
    const uint32_t* w00 = reinterpret_cast<const uint32_t*>(pyr + (size_t)b * framePyr + L.off + (size_t)(EDGE + y - HALF_PATCH) * L.pitch) +
                          (x >> 2) + (lane & 7);
    const int sh = (x & 3) * 8, pw = L.pitch >> 2;
    int m10 = 0, m01 = 0;
#pragma unroll
    for (int it = 0; it < 8; it++) {
        const int row = it * 4 + (lane >> 3);                     // 0..31, row 31 has empty weights
        const int2 w = g_orientW[it * 32 + lane];
        if (row < 2 * HALF_PATCH + 1) {
            const uint32_t* wp = w00 + row * pw;
            const uint32_t px = __funnelshift_r(__ldg(wp), __ldg(wp + 1), sh);
            m10 = dp4a_s8u8(w.x, px, m10);
            m01 += (row - HALF_PATCH) * (int)__dp4a((uint32_t)w.y, px, 0u);
        }
    }
    m10 = __reduce_add_sync(0xffffffffu, m10);
    m01 = __reduce_add_sync(0xffffffffu, m01);
    if (lane == 0) {
        GforbKeyPoint k;
        float fx = (float)x, fy = (float)y;
        if (l != 0) { fx = __fmul_rn(fx, L.scale); fy = __fmul_rn(fy, L.scale); }
        k.x = fx; k.y = fy;
        k.size = (float)L.scaledPatchSize;
        k.angle = fast_atan2_deg((float)m01, (float)m10);
        k.response = lvlResp ? lvlResp[(size_t)b * frameKp + slot] : (float)score;   // HARRIS_SCORE : FAST_SCORE
        k.octave = l;
        k.class_id = -1;
        kps[(size_t)b * nfeatures + j] = k;
    }
}

// A/B loser (B200: 0.369 ms per 250 frames against 0.2375 ms for k_describe; ncu: L1 data-pipe wavefronts 399 per keypoint
// against 186 -- the 128-bit staging and float-pattern loads cost one data-stage wavefront per 32-byte sector, 317 per
// keypoint, where a 32-lane byte gather costs ~9; shared-memory gathers add 49 + 31 for the stores).
// computeOrbDescriptor with the keypoint's 37 x 37 tap window staged in shared memory.  The rotated pattern reaches
// |ix|, |iy| <= 18 (largest pattern radius 18.38), so the window is plane rows cy - 2 .. cy + 34 and the 16-byte chunks
// that cover plane columns cx - 2 .. cx + 34 (3 or 4 per row): ~120 coalesced 128-bit loads per warp replace 512
// byte gathers that each touched up to 32 cache lines (the L1 tag stage was the limiter of k_describe: ncu l1 71 %).
// A chunk inside the level ROI comes from the blurred plane, one in the 16-px ring from the unblurred pyramid plane,
// a chunk that straddles the ROI edge is merged byte by byte -- the same source rule as k_describe, tap for tap.
// cvRound of the rotated coordinates is taken by adding 1.5 * 2^23 (round-to-nearest-even in the float add, the
// integer sits in the mantissa), which keeps the conversions off the XU pipe; identical to __float2int_rn for |v| < 2^22.
static const int DS_ROWS = 37, DS_PITCH = 80, DS_R = 18;
__global__ void __launch_bounds__(256) k_describe2(const uint8_t* __restrict__ pyr, const uint8_t* __restrict__ blr,
                                                   size_t framePyr, const LevelPlan* __restrict__ levels, int nLevels,
                                                   const uint32_t* __restrict__ lvlKp, size_t frameKp,
                                                   const int* __restrict__ lvlCount,
                                                   const GforbKeyPoint* __restrict__ kps, uint8_t* __restrict__ desc,
                                                   int nfeatures) {
    __shared__ __align__(16) uint8_t s_patch[8][DS_ROWS * DS_PITCH];
    const int j = blockIdx.x * 8 + (threadIdx.x >> 5), b = blockIdx.y, lane = threadIdx.x & 31;
    if (j >= nfeatures) return;
    int l, i, total;
    if (!kp_row_to_slot(j, b, nLevels, lvlCount, lane, l, i, total)) return;
    const LevelPlan L = levels[l];
    const uint32_t e = lvlKp[(size_t)b * frameKp + L.kpBase + i];
    const int cx = e & 0xfff, cy = (e >> 12) & 0xfff;
    const size_t oi = (size_t)b * nfeatures + j;
    uint8_t* sp = s_patch[threadIdx.x >> 5];
    // ---- stage the window (plane coordinates = ROI coordinates + EDGE)
    const int X0 = cx + EDGE - DS_R, Y0 = cy + EDGE - DS_R;
    const int a0 = X0 & ~15, o = X0 - a0;
    const int nch = o + 2 * DS_R >= 48 ? 4 : 3;
    const uint8_t* Pp = pyr + (size_t)b * framePyr + L.off;
    const uint8_t* Bp = blr + (size_t)b * framePyr + L.off;
    const int planeH = L.h + 2 * EDGE;
    for (int t = lane; t < DS_ROWS * nch; t += 32) {
        const int r = nch == 4 ? t >> 2 : t / 3, c = t - r * nch;
        const int Y = Y0 + r, X = a0 + 16 * c;
        uint4 v = make_uint4(0, 0, 0, 0);
        if ((unsigned)Y < (unsigned)planeH && X >= 0 && X + 16 <= L.pitch) {
            const size_t off = (size_t)Y * L.pitch + X;
            const bool rowIn = Y >= EDGE && Y < EDGE + L.h;
            if (rowIn && X >= EDGE && X + 16 <= EDGE + L.w) {
                v = __ldg(reinterpret_cast<const uint4*>(Bp + off));
            } else {
                v = __ldg(reinterpret_cast<const uint4*>(Pp + off));
                if (rowIn && X + 16 > EDGE && X < EDGE + L.w) {          // straddles the left or right ROI edge
                    const uint4 q = __ldg(reinterpret_cast<const uint4*>(Bp + off));
                    uint32_t pv[4] = {v.x, v.y, v.z, v.w};
                    const uint32_t qv[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
                    for (int k = 0; k < 16; k++) {
                        const int xx = X + k;
                        if (xx >= EDGE && xx < EDGE + L.w) {
                            const uint32_t m = 0xffu << (8 * (k & 3));
                            pv[k >> 2] = (pv[k >> 2] & ~m) | (qv[k >> 2] & m);
                        }
                    }
                    v = make_uint4(pv[0], pv[1], pv[2], pv[3]);
                }
            }
        }
        *reinterpret_cast<uint4*>(sp + r * DS_PITCH + 16 * c) = v;
    }
    const float factorPI = (float)(3.14159265358979323846 / 180.f);
    const float angle = __fmul_rn(kps[oi].angle, factorPI);
    const float a = glibc_cosf(angle), bb = glibc_sinf(angle);
    __syncwarp();
    // ---- 8 tests of this lane's descriptor byte
    const float MAGIC = 12582912.f;                               // 1.5 * 2^23; its bit image is 0x4B400000
    const uint32_t bias = (uint32_t)(DS_R * DS_PITCH + DS_R + o) - 0x4B400000u * (uint32_t)(DS_PITCH + 1);   // mod 2^32
    const float4* pf = reinterpret_cast<const float4*>(g_patternF) + lane * 8;
    int val = 0;
#pragma unroll
    for (int t = 0; t < 8; t++) {
        const float4 p4 = __ldg(pf + t);
        const float pxs[2] = {p4.x, p4.z}, pys[2] = {p4.y, p4.w};
        int tv[2];
#pragma unroll
        for (int q = 0; q < 2; q++) {
            const float fy = __fadd_rn(__fmul_rn(pxs[q], bb), __fmul_rn(pys[q], a));
            const float fx = __fsub_rn(__fmul_rn(pxs[q], a), __fmul_rn(pys[q], bb));
            const int iy = __float_as_int(__fadd_rn(fy, MAGIC)), ix = __float_as_int(__fadd_rn(fx, MAGIC));
            tv[q] = sp[(uint32_t)iy * DS_PITCH + (uint32_t)ix + bias];
        }
        val |= (tv[0] < tv[1]) << t;
    }
    desc[oi * 32 + lane] = (uint8_t)val;
}

// ================================================================================================ host side
struct Stage { const char* name; cudaEvent_t e0, e1; float ms; };
struct StageRec { int stage; cudaEvent_t e0, e1; };   // one (stage, call) pair recorded while profiling is on

}  // namespace gforb

using namespace gforb;

struct GforbExtractor {
    GforbExtractorConfig cfg;
    int device;
    cudaStream_t ownStream;
    // ctor tables (reference src/ORBextractor.cc:469-494)
    std::vector<float> scale, invScale;
    std::vector<int> featPerLevel;
    // plan for the current image size
    int planW = -1, planH = -1;
    std::vector<LevelPlan> levels;
    std::vector<CellPlan> cells;
    size_t framePyr = 0, frameCand = 0, frameKp = 0;
    int kpCapTotal = 0, nCellsTotal = 0, detSmem = 0, detPixPitch = 0, fnTasks[2] = {0, 0}, blTasks[2] = {0, 0};
    int blurQD = 2;                     // GFORB_BLUR_QD: input rows in flight ahead of the filtered one in k_blur3 (1, 2, 3; B200: 0.346 / 0.309 / 0.302 ms per 250 frames, pipelined 83.6k / 84.7k / 84.3k frames/s)
    int blurV2 = 0;                     // GFORB_BLUR=v2: second-generation blur (vertical pass on unpacked 32-bit lane-ops) instead of k_blur3 (IDP.2A row pairs)
    int fnTall = -1;                    // GFORB_FN_TALL=0/1 pins the strip height of k_fast_nms; -1: tall strips once the batch fills the GPU several times
    bool tiled = true;
    int detectVersion = 2;              // GFORB_DETECT=1 selects the first-generation per-cell kernel (A/B tool)
    int blurFork = 1;                   // the blur of a call runs on ownStream beside the quota / retainBest / orientation kernels (gforb_extractor_set_concurrency)
    cudaEvent_t evFork = nullptr, evJoin = nullptr;
    int distribution = 0;               // 0: the reference's cell quota; 1: ORB-SLAM2 quadtree (quadtree.cuh, gforb_extractor_set_distribution)
    std::vector<QtLevel> qlevels;
    QtLevel* d_qlv = nullptr;
    uint32_t *d_qtKeys = nullptr, *d_qtTmp = nullptr;
    QtNode* d_qtNodes = nullptr;
    int2* d_qtVsz = nullptr;
    size_t frameQtKeys = 0, frameQtNodes = 0, capQtKeys = 0, capQtTmp = 0, capQtNodes = 0, capQtVsz = 0;
    int retainSeq = 0;                  // GFORB_RETAIN=seq: retainBest's nth_element on one lane (the first-generation path)
    KpLevels kpLevels;                  // per-level constants of the row-mapped keypoint kernels (kernel parameter), filled by build_plan
    int describeSmem = 0;               // GFORB_DESCRIBE=smem: k_describe2 (tap window staged in shared memory; A/B loser)
    int kpV1 = 0;                       // GFORB_KP=v1: first-generation k_orient / k_describe (warp per list slot, byte gathers from L1)
    int resizeWords = 2;                // k_resize2 (aligned word loads + dp2a) with 8-row bands (default; GFORB_RESIZE=words: 4-row bands, =direct: the byte-gather k_resize);
                                        // needs every quad's tap span <= 6 (resizeWordsOk, true up to scaleFactor 2), else k_resize runs
    bool resizeWordsOk = true;
    int resizeTma = 0;                  // GFORB_RESIZE=tma / direct: source rows staged by cp.async.bulk (k_resize_tma) or gathered from L1 (k_resize)
    // device buffers
    uint8_t *d_img = nullptr, *d_pyr = nullptr, *d_blr = nullptr, *d_desc = nullptr;
    GforbKeyPoint* d_kps = nullptr;
    int32_t* d_counts = nullptr;
    uint32_t *d_cand = nullptr, *d_lvlKp = nullptr;
    uint64_t* d_work64 = nullptr;       // HARRIS_SCORE only: 64-bit retain scratch and the survivors' responses
    float* d_lvlResp = nullptr;
    size_t frameWork = 0, capWork = 0, capResp = 0;
    int *d_cnt7 = nullptr, *d_cnt20 = nullptr, *d_nTotal = nullptr, *d_nToRetain = nullptr, *d_cellOff = nullptr;
    int *d_lvlPre = nullptr, *d_lvlCount = nullptr, *d_err = nullptr;
    LevelPlan* d_levels = nullptr;
    CellPlan* d_cells = nullptr;
    int2* d_tabs = nullptr;
    size_t capPyr = 0, capCand = 0, capKp = 0, capCells = 0, capTabs = 0, capImg = 0;
    int lastBatch = 0, lastLaunches = 0, numSMs = 148;
    bool profiling = false;
    std::vector<Stage> stages;          // per-stage name + average ms over the recorded calls
    std::vector<StageRec> recs;         // raw event pairs since profiling was switched on
    std::vector<cudaEvent_t> evPool;    // recycled events
};

static int build_ctor_tables(GforbExtractor* E) {
    // reference src/ORBextractor.cc:469-494; `scaleFactor` member is a double holding the float argument
    const GforbExtractorConfig& c = E->cfg;
    const double scaleFactor = (double)c.scaleFactor;
    E->scale.assign(c.nlevels, 1.f);
    E->invScale.assign(c.nlevels, 1.f);
    for (int i = 1; i < c.nlevels; i++) E->scale[i] = (float)(E->scale[i - 1] * scaleFactor);
    float invScaleFactor = (float)(1.0f / scaleFactor);
    for (int i = 1; i < c.nlevels; i++) E->invScale[i] = E->invScale[i - 1] * invScaleFactor;
    E->featPerLevel.assign(c.nlevels, 0);
    float factor = (float)(1.0 / scaleFactor);
    float nDesired = c.nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)c.nlevels));
    int sum = 0;
    for (int l = 0; l < c.nlevels - 1; l++) {
        E->featPerLevel[l] = (int)lrintf(nDesired);
        sum += E->featPerLevel[l];
        nDesired *= factor;
    }
    E->featPerLevel[c.nlevels - 1] = std::max(c.nfeatures - sum, 0);
    return GFORB_OK;
}

static void linear_table(int ssize, int dsize, std::vector<int2>& out) {
    // cv::resize INTER_LINEAR coefficient generation
    double scale = 1.0 / ((double)dsize / ssize);
    for (int d = 0; d < dsize; d++) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)floorf(f);
        f -= s;
        if (s < 0) { s = 0; f = 0; }
        if (s >= ssize - 1) { s = ssize - 1; f = 0; }
        int a0 = (short)lrintf((1.f - f) * 2048.f), a1 = (short)lrintf(f * 2048.f);
        out.push_back(make_int2(s, (a0 & 0xffff) | (a1 << 16)));
    }
}

template <class T>
static int ensure(T** p, size_t* cap, size_t need) {
    if (need <= *cap && *p) return GFORB_OK;
    if (*p) cudaFree(*p);
    *p = nullptr;
    GF_CUDA_OK(cudaMalloc((void**)p, need * sizeof(T)));
    *cap = need;
    return GFORB_OK;
}

// Level / cell geometry exactly as ComputePyramid (:926-927) and ComputeKeyPoints (:540-612) derive it.
static int build_plan(GforbExtractor* E, int w, int h) {
    const GforbExtractorConfig& c = E->cfg;
    const int nl = c.nlevels;
    E->levels.assign(nl, LevelPlan());
    E->cells.clear();
    std::vector<int2> tabs;
    size_t off = 0;
    int kpBase = 0, candBase = 0, maxCellW = 0, fnTasks[2] = {0, 0}, blTasks[2] = {0, 0};
    const bool qt = E->distribution == 1;
    E->qlevels.assign(nl, QtLevel());
    int qtKeyBase = 0, qtNodeBase = 0;
    bool wordsOk = true;    // k_resize2 precondition (tap span of every column quad)
    bool tiled = true;      // cell interiors tile [16, w-16) x [16, h-16) (k_fast_nms precondition)
    const float imageRatio = (float)w / h;
    for (int l = 0; l < nl; l++) {
        LevelPlan& L = E->levels[l];
        L.w = (int)lrintf((float)w * E->invScale[l]);
        L.h = (int)lrintf((float)h * E->invScale[l]);
        if (L.w <= 2 * EDGE || L.h <= 2 * EDGE || L.w > 4095 || L.h > 4095) return GFORB_ERR_GEOMETRY;
        L.pitch = (L.w + 2 * EDGE + 127) & ~127;
        L.rows = L.h + 2 * EDGE;
        L.off = off;
        off += (size_t)L.pitch * L.rows;
        off = (off + 255) & ~(size_t)255;
        L.scale = E->scale[l];
        L.scaledPatchSize = (int)(31 * E->scale[l]);
        L.nDesired = E->featPerLevel[l];
        L.levelCols = (int)sqrtf((float)L.nDesired / (5 * imageRatio));
        L.levelRows = (int)(imageRatio * L.levelCols);
        const int minB = EDGE, maxBX = L.w - EDGE, maxBY = L.h - EDGE;
        const int W = maxBX - minB, H = maxBY - minB;
        if (qt) {
            // ORB-SLAM2 ComputeKeyPointsOctTree: cells of ~30 px over [13, w - 13) x [13, h - 13); a cell image is its interior
            // + 3 px on every side, so the interiors tile [16, w - 16) x [16, h - 16) with period (wCell, hCell)
            const float width = (float)(L.w - 2 * QT_MIN_BORDER), height = (float)(L.h - 2 * QT_MIN_BORDER);
            L.levelCols = (int)(width / 30.f);
            L.levelRows = (int)(height / 30.f);
            if (L.levelCols <= 0 || L.levelRows <= 0) return GFORB_ERR_GEOMETRY;
            L.cellW = (int)ceilf(width / L.levelCols);
            L.cellH = (int)ceilf(height / L.levelRows);
        } else {
            if (L.levelCols <= 0 || L.levelRows <= 0) return GFORB_ERR_GEOMETRY;
            L.cellW = (int)ceilf((float)W / L.levelCols);
            L.cellH = (int)ceilf((float)H / L.levelRows);
        }
        L.nCells = L.levelRows * L.levelCols;
        L.nfeaturesCell = (int)ceilf((float)L.nDesired / L.nCells);
        L.cellBase = (int)E->cells.size();
        L.fnStripsX = (W + FN_OUTW - 1) / FN_OUTW;
        for (int v = 0; v < 2; v++) {                        // equal strips per level: no ragged last strip
            const int sy = (H + FN_ROWS[v] - 1) / FN_ROWS[v];
            L.fnRows[v] = (H + sy - 1) / sy;
            L.fnTaskBase[v] = fnTasks[v];
            fnTasks[v] += L.fnStripsX * sy;
        }
        L.blStripsX = (L.w + 127) / 128;
        for (int v = 0; v < 2; v++) {
            const int sy = (L.h + BL2_ROWS[v] - 1) / BL2_ROWS[v];
            L.blRows[v] = (L.h + sy - 1) / sy;
            L.blTaskBase[v] = blTasks[v];
            blTasks[v] += L.blStripsX * sy;
        }
        if ((long long)L.cellW * (L.levelCols - 1) >= W || (long long)L.cellH * (L.levelRows - 1) >= H) tiled = false;
        L.kpBase = kpBase;
        L.kpCap = 2 * L.nDesired + L.nCells + 64;
        kpBase += L.kpCap;
        if (l > 0) {
            L.xtab = (int)tabs.size(); linear_table(E->levels[l - 1].w, L.w, tabs);
            L.ytab = (int)tabs.size(); linear_table(E->levels[l - 1].h, L.h, tabs);
            const auto refl = [](int p, int len) { if (len == 1) return 0; while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p; return p; };
            if (tabs.size() & 1) tabs.push_back(make_int2(0, 0));        // qtab is read with 128-bit loads
            L.qtab = (int)tabs.size();
            for (int q = 0; q < (L.w + 2 * EDGE + 3) / 4; q++) {
                int2 tx[4];
                int lo = INT_MAX, hi = 0;
                for (int k = 0; k < 4; k++) {
                    tx[k] = tabs[L.xtab + refl(std::min(4 * q + k, L.w + 2 * EDGE - 1) - EDGE, L.w)];   // columns past w+32 are pitch padding
                    lo = std::min(lo, tx[k].x); hi = std::max(hi, tx[k].x);
                }
                if (hi - lo > 6) wordsOk = false;                        // the taps of a quad must fit the 8-byte window
                unsigned sel[2];
                for (int pr = 0; pr < 2; pr++) {
                    const int da = (tx[2 * pr].x - lo) & 7, db = (tx[2 * pr + 1].x - lo) & 7;
                    sel[pr] = (unsigned)da | (unsigned)((da + 1) & 7) << 4 | (unsigned)db << 8 | (unsigned)((db + 1) & 7) << 12;
                }
                tabs.push_back(make_int2(lo, (int)(sel[0] | sel[1] << 16)));
                tabs.push_back(make_int2(tx[0].y, tx[1].y));
                tabs.push_back(make_int2(tx[2].y, tx[3].y));
                tabs.push_back(make_int2(0, 0));
            }
            L.ybtab = (int)tabs.size();
            for (int Y = 0; Y < L.rows; Y++) { const int2 e = tabs[L.ytab + refl(Y - EDGE, L.h)]; tabs.push_back(e); }
        } else { L.xtab = L.ytab = L.qtab = L.ybtab = 0; }
        for (int i = 0; i < L.levelRows; i++) {
            const int iniY = minB + i * L.cellH - 3;
            float hY = (float)(L.cellH + 6);
            bool rowSkipped = false;
            if (i == L.levelRows - 1) { hY = (float)(maxBY + 3 - iniY); if (hY <= 0) rowSkipped = true; }
            for (int j = 0; j < L.levelCols; j++) {
                const int iniX = minB + j * L.cellW - 3;
                float hX = (float)(L.cellW + 6);
                bool skipped = rowSkipped;
                if (j == L.levelCols - 1) { hX = (float)(maxBX + 3 - iniX); if (hX <= 0) skipped = true; }
                CellPlan C;
                C.level = l; C.skipped = skipped ? 1 : 0;
                C.x0 = C.y0 = C.x1 = C.y1 = 0;
                if (qt) {
                    C.x0 = EDGE + j * L.cellW; C.x1 = std::min(C.x0 + L.cellW, maxBX);
                    C.y0 = EDGE + i * L.cellH; C.y1 = std::min(C.y0 + L.cellH, maxBY);
                    C.skipped = (C.x1 <= C.x0 || C.y1 <= C.y0) ? 1 : 0;      // "iniX >= maxBorderX - 6" / a cell image too low for FAST
                    if (C.skipped) C.x0 = C.y0 = C.x1 = C.y1 = 0;
                } else if (!skipped) {
                    int x1 = iniX + (int)hX, y1 = iniY + (int)hY;            // cell image = [iniX,x1) x [iniY,y1)
                    if (iniX < 0 || iniY < 0 || x1 > L.w || y1 > L.h) return GFORB_ERR_GEOMETRY;  // cv::Mat range assert
                    C.x0 = iniX + 3; C.y0 = iniY + 3; C.x1 = x1 - 3; C.y1 = y1 - 3;                // FAST's 3-px margin
                    if (C.x1 < C.x0) C.x1 = C.x0;
                    if (C.y1 < C.y0) C.y1 = C.y0;
                }
                const int Wc = C.x1 - C.x0, Hc = C.y1 - C.y0;
                C.candBase = candBase;
                C.candCap = ((Wc + 1) / 2) * ((Hc + 1) / 2);
                candBase += C.candCap;
                maxCellW = std::max(maxCellW, Wc);
                E->cells.push_back(C);
            }
        }
    }
    if (qt) {
        for (int l = 0; l < nl; l++) {
            const LevelPlan& L = E->levels[l];
            QtLevel& Q = E->qlevels[l];
            Q.rootW = L.w - 2 * QT_MIN_BORDER; Q.rootH = L.h - 2 * QT_MIN_BORDER;
            Q.nIni = (int)roundf((float)Q.rootW / (float)Q.rootH);
            if (Q.nIni <= 0 || Q.nIni > 4) return GFORB_ERR_GEOMETRY;         // aspect ratios beyond 4.5 : 1 (or portrait below 1 : 2)
            Q.hX = (float)Q.rootW / Q.nIni;
            for (int i = 0; i <= Q.nIni; i++) Q.iniUL[i] = (int)(Q.hX * (float)i);
            Q.N = L.nDesired;
            long long sum = 0;
            for (int cI = 0; cI < L.nCells; cI++) sum += E->cells[L.cellBase + cI].candCap;
            Q.keyBase = qtKeyBase; Q.keyCap = (int)std::min<long long>(sum, 32768);
            qtKeyBase += (Q.keyCap + 63) & ~63;
            Q.nodeBase = qtNodeBase; Q.nodeCap = 8 * Q.N + 64;
            if (Q.nodeCap > 32767) return GFORB_ERR_CAPACITY;
            qtNodeBase += Q.nodeCap;
        }
        E->frameQtKeys = (size_t)qtKeyBase; E->frameQtNodes = (size_t)qtNodeBase;
    }
    E->framePyr = off;
    E->frameCand = (size_t)candBase + 4;
    E->frameKp = (size_t)kpBase;
    E->kpCapTotal = kpBase;
    E->nCellsTotal = (int)E->cells.size();
    E->fnTasks[0] = fnTasks[0]; E->fnTasks[1] = fnTasks[1];
    E->resizeWordsOk = wordsOk;
    memset(&E->kpLevels, 0, sizeof(E->kpLevels));
    for (int l = 0; l < (int)E->levels.size() && l < KP_MAX_LEVELS; l++) {
        const LevelPlan& P = E->levels[l];
        KpLevel& K = E->kpLevels.lv[l];
        K.off = P.off; K.pitch = P.pitch; K.kpBase = P.kpBase; K.w = P.w; K.h = P.h; K.scaledPatchSize = P.scaledPatchSize; K.scale = P.scale;
    }
    E->blTasks[0] = blTasks[0]; E->blTasks[1] = blTasks[1];
    E->tiled = tiled;
    E->detPixPitch = (maxCellW + 16 + 3) & ~3;
    E->detSmem = (DET_BAND + 8) * E->detPixPitch + (DET_BAND + 2) * E->detPixPitch + DET_BAND * ((maxCellW + 31) / 32) * 4 + 16;
    if (E->detSmem > 200 * 1024) return GFORB_ERR_CAPACITY;
    const size_t B = (size_t)c.max_batch;
    int rc;
    if ((rc = ensure(&E->d_pyr, &E->capPyr, E->framePyr * B))) return rc;
    { size_t cap2 = 0; uint8_t* old = E->d_blr; if (old) cudaFree(old); E->d_blr = nullptr; if ((rc = ensure(&E->d_blr, &cap2, E->framePyr * B))) return rc; }
    if ((rc = ensure(&E->d_cand, &E->capCand, E->frameCand * B))) return rc;
    if ((rc = ensure(&E->d_lvlKp, &E->capKp, E->frameKp * B))) return rc;
    if (c.scoreType == 0) {
        E->frameWork = std::max(E->frameCand, E->frameKp);
        if ((rc = ensure(&E->d_work64, &E->capWork, E->frameWork * B))) return rc;
        if ((rc = ensure(&E->d_lvlResp, &E->capResp, E->frameKp * B))) return rc;
    }
    if (qt) {
        if ((rc = ensure(&E->d_qtKeys, &E->capQtKeys, E->frameQtKeys * B))) return rc;
        if ((rc = ensure(&E->d_qtTmp, &E->capQtTmp, E->frameQtKeys * B))) return rc;
        if ((rc = ensure(&E->d_qtNodes, &E->capQtNodes, E->frameQtNodes * B))) return rc;
        if ((rc = ensure(&E->d_qtVsz, &E->capQtVsz, E->frameQtNodes * B * 2))) return rc;
        if (E->d_qlv) cudaFree(E->d_qlv);
        E->d_qlv = nullptr;
        GF_CUDA_OK(cudaMalloc((void**)&E->d_qlv, sizeof(QtLevel) * nl));
        GF_CUDA_OK(cudaMemcpy(E->d_qlv, E->qlevels.data(), sizeof(QtLevel) * nl, cudaMemcpyHostToDevice));
    }
    size_t nc = (size_t)E->nCellsTotal * B;
    if (nc > E->capCells) {
        for (int** p : {&E->d_cnt7, &E->d_cnt20, &E->d_nTotal, &E->d_nToRetain, &E->d_cellOff}) { if (*p) cudaFree(*p); *p = nullptr; }
        for (int** p : {&E->d_cnt7, &E->d_cnt20, &E->d_nTotal, &E->d_nToRetain, &E->d_cellOff}) GF_CUDA_OK(cudaMalloc((void**)p, nc * sizeof(int)));
        E->capCells = nc;
    }
    if (E->d_levels) cudaFree(E->d_levels);
    if (E->d_cells) cudaFree(E->d_cells);
    if (E->d_tabs) cudaFree(E->d_tabs);
    GF_CUDA_OK(cudaMalloc((void**)&E->d_levels, sizeof(LevelPlan) * nl));
    GF_CUDA_OK(cudaMalloc((void**)&E->d_cells, sizeof(CellPlan) * E->cells.size()));
    GF_CUDA_OK(cudaMalloc((void**)&E->d_tabs, sizeof(int2) * std::max<size_t>(tabs.size(), 1)));
    GF_CUDA_OK(cudaMemcpy(E->d_levels, E->levels.data(), sizeof(LevelPlan) * nl, cudaMemcpyHostToDevice));
    GF_CUDA_OK(cudaMemcpy(E->d_cells, E->cells.data(), sizeof(CellPlan) * E->cells.size(), cudaMemcpyHostToDevice));
    if (!tabs.empty()) GF_CUDA_OK(cudaMemcpy(E->d_tabs, tabs.data(), sizeof(int2) * tabs.size(), cudaMemcpyHostToDevice));
    GF_CUDA_OK(raise_dyn_smem(k_detect, E->detSmem, E->device));
    E->planW = w; E->planH = h;
    return GFORB_OK;
}

extern "C" {

const char* gforb_version(void) { return "gforb-b200 0.1 (sm_100a)"; }

int gforb_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

int gforb_extractor_create(const GforbExtractorConfig* cfg, GforbExtractor** out) {
    if (!cfg || !out) return GFORB_ERR_ARG;
    *out = nullptr;
    if (cfg->nfeatures <= 0 || cfg->nlevels <= 0 || cfg->nlevels > MAX_LEVELS || !(cfg->scaleFactor > 1.0f) ||
        cfg->scoreType < 0 || cfg->scoreType > 1 || cfg->max_batch <= 0 || cfg->max_width <= 0 || cfg->max_height <= 0 ||
        cfg->blur_variant < 0 || cfg->blur_variant > 1 || cfg->retain_variant < 0 || cfg->retain_variant > 1 ||
        cfg->minThFAST < 1 || cfg->iniThFAST < cfg->minThFAST || cfg->iniThFAST > 254)
        return GFORB_ERR_ARG;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || cfg->device < 0 || cfg->device >= ndev) {
        fprintf(stderr, "[gforb] no usable CUDA device (requested %d of %d): the B200 path has no CPU fallback\n", cfg->device, ndev);
        return GFORB_ERR_CUDA;
    }
    GF_CUDA_OK(cudaSetDevice(cfg->device));
    GforbExtractor* E = new GforbExtractor();
    const auto init = [&]() -> int {
    E->cfg = *cfg;
    E->device = cfg->device;
    GF_CUDA_OK(cudaStreamCreateWithFlags(&E->ownStream, cudaStreamNonBlocking));
    GF_CUDA_OK(cudaEventCreateWithFlags(&E->evFork, cudaEventDisableTiming));
    GF_CUDA_OK(cudaEventCreateWithFlags(&E->evJoin, cudaEventDisableTiming));
    if (const char* bf = getenv("GFORB_BLUR_FORK")) E->blurFork = atoi(bf);
    if (const char* rs = getenv("GFORB_RETAIN")) E->retainSeq = strcmp(rs, "seq") == 0;
    cudaDeviceGetAttribute(&E->numSMs, cudaDevAttrMultiProcessorCount, cfg->device);
    build_ctor_tables(E);
    if (const char* dv = getenv("GFORB_DETECT")) E->detectVersion = atoi(dv);
    if (const char* rv = getenv("GFORB_RESIZE")) { E->resizeTma = !strcmp(rv, "tma") ? 1 : 0; E->resizeWords = !strcmp(rv, "words") ? 1 : (!strcmp(rv, "direct") || E->resizeTma) ? 0 : 2; }
    if (const char* kv = getenv("GFORB_KP")) E->kpV1 = !strcmp(kv, "v1") ? 1 : 0;
    if (cfg->nlevels > KP_MAX_LEVELS) E->kpV1 = 1;          // the row-mapped kernels carry the level constants of <= 16 levels as a parameter
    if (const char* dv2 = getenv("GFORB_DESCRIBE")) E->describeSmem = !strcmp(dv2, "smem") ? 1 : !strcmp(dv2, "magic") ? 2 : 0;
    if (const char* tv = getenv("GFORB_FN_TALL")) E->fnTall = atoi(tv) ? 1 : 0;
    if (const char* qd = getenv("GFORB_BLUR_QD")) E->blurQD = atoi(qd);
    if (const char* bv2 = getenv("GFORB_BLUR")) E->blurV2 = !strcmp(bv2, "v2") ? 1 : 0;

    GF_CUDA_OK(cudaMemcpyToSymbol(g_pattern, h_pattern, sizeof(h_pattern)));
    {   // umax (src/ORBextractor.cc:500-517)
        int umax[16];
        int v, v0, vmax = (int)floorf(HALF_PATCH * sqrtf(2.f) / 2 + 1), vmin = (int)ceilf(HALF_PATCH * sqrtf(2.f) / 2);
        const double hp2 = HALF_PATCH * HALF_PATCH;
        for (v = 0; v <= vmax; ++v) umax[v] = (int)lrint(sqrt(hp2 - v * v));
        for (v = HALF_PATCH, v0 = 0; v >= vmin; --v) {
            while (umax[v0] == umax[v0 + 1]) ++v0;
            umax[v] = v0;
            ++v0;
        }
        GF_CUDA_OK(cudaMemcpyToSymbol(c_umax, umax, sizeof(umax)));
        int2 ow[256];                                            // k_orient2's task table: (row * 8 + chunk) -> {u weights, membership}
        for (int t = 0; t < 256; t++) {
            const int row = t >> 3, c = t & 7, vv = row - HALF_PATCH;
            const int d = row < 2 * HALF_PATCH + 1 ? umax[vv < 0 ? -vv : vv] : -1;
            uint32_t wu = 0, w1 = 0;
            for (int k = 0; k < 4; k++) {
                const int u = -16 + 4 * c + k;
                if ((u < 0 ? -u : u) <= d) { wu |= (uint32_t)(u & 0xff) << (8 * k); w1 |= 1u << (8 * k); }
            }
            ow[t] = make_int2((int)wu, (int)w1);
        }
        GF_CUDA_OK(cudaMemcpyToSymbol(g_orientW, ow, sizeof(ow)));
        float pf[1024];
        for (int t = 0; t < 1024; t++) pf[t] = (float)h_pattern[t >> 2][t & 3];
        GF_CUDA_OK(cudaMemcpyToSymbol(g_patternF, pf, sizeof(pf)));
    }
    size_t B = cfg->max_batch;
    GF_CUDA_OK(cudaMalloc((void**)&E->d_kps, B * cfg->nfeatures * sizeof(GforbKeyPoint)));
    GF_CUDA_OK(cudaMalloc((void**)&E->d_desc, B * cfg->nfeatures * 32));
    GF_CUDA_OK(cudaMalloc((void**)&E->d_counts, B * sizeof(int32_t)));
    GF_CUDA_OK(cudaMalloc((void**)&E->d_lvlPre, B * cfg->nlevels * sizeof(int)));
    GF_CUDA_OK(cudaMalloc((void**)&E->d_lvlCount, B * cfg->nlevels * sizeof(int)));
    GF_CUDA_OK(cudaMalloc((void**)&E->d_err, sizeof(int)));
    GF_CUDA_OK(cudaMemset(E->d_err, 0, sizeof(int)));
    E->capImg = (size_t)cfg->max_width * cfg->max_height * B;
    GF_CUDA_OK(cudaMalloc((void**)&E->d_img, E->capImg));
    if (const int rc = build_plan(E, cfg->max_width, cfg->max_height)) return rc;
        return GFORB_OK;
    };
    if (const int rc = init()) { gforb_extractor_destroy(E); return rc; }   // a failure part-way frees what was allocated
    *out = E;
    return GFORB_OK;
}

int gforb_extractor_destroy(GforbExtractor* E) {
    if (!E) return GFORB_ERR_ARG;
    cudaSetDevice(E->device);
    void* ptrs[] = {E->d_img, E->d_pyr, E->d_blr, E->d_desc, E->d_kps, E->d_counts, E->d_cand, E->d_lvlKp, E->d_cnt7,
                    E->d_cnt20, E->d_nTotal, E->d_nToRetain, E->d_cellOff, E->d_lvlPre, E->d_lvlCount, E->d_err,
                    E->d_levels, E->d_cells, E->d_tabs, E->d_work64, E->d_lvlResp, E->d_qlv, E->d_qtKeys, E->d_qtTmp, E->d_qtNodes,
                    E->d_qtVsz};
    for (void* p : ptrs) if (p) cudaFree(p);
    for (auto& r : E->recs) { cudaEventDestroy(r.e0); cudaEventDestroy(r.e1); }
    for (auto& e : E->evPool) cudaEventDestroy(e);
    if (E->ownStream) cudaStreamDestroy(E->ownStream);
    if (E->evFork) cudaEventDestroy(E->evFork);
    if (E->evJoin) cudaEventDestroy(E->evJoin);
    delete E;
    return GFORB_OK;
}

int gforb_extractor_levels(const GforbExtractor* E) { return E ? E->cfg.nlevels : GFORB_ERR_ARG; }
float gforb_extractor_scale_factor(const GforbExtractor* E) { return E ? E->cfg.scaleFactor : 0.f; }
int gforb_extractor_tables(const GforbExtractor* E, float* sf, int32_t* fpl) {
    if (!E) return GFORB_ERR_ARG;
    for (int i = 0; i < E->cfg.nlevels; i++) { if (sf) sf[i] = E->scale[i]; if (fpl) fpl[i] = E->featPerLevel[i]; }
    return GFORB_OK;
}
int gforb_extractor_last_launches(const GforbExtractor* E) { return E ? E->lastLaunches : GFORB_ERR_ARG; }
int gforb_extractor_set_profiling(GforbExtractor* E, int on) {
    if (!E) return GFORB_ERR_ARG;
    E->profiling = on != 0;
    if (on) {        // restart the accumulation
        for (auto& r : E->recs) { E->evPool.push_back(r.e0); E->evPool.push_back(r.e1); }
        E->recs.clear();
        for (auto& s : E->stages) s.ms = 0.f;
    }
    return GFORB_OK;
}
// average device time of stage i over every call recorded since set_profiling(h, 1); the caller must have
// synchronised the launch stream.  Returns the number of stages.
int gforb_extractor_stage_time(const GforbExtractor* Ec, int i, const char** name, float* ms) {
    GforbExtractor* E = const_cast<GforbExtractor*>(Ec);
    if (!E) return GFORB_ERR_ARG;
    if (i >= 0 && i < (int)E->stages.size()) {
        double sum = 0; int n = 0;
        for (auto& r : E->recs) {
            if (r.stage != i) continue;
            float t = 0.f;
            if (cudaEventElapsedTime(&t, r.e0, r.e1) == cudaSuccess) { sum += t; n++; }
        }
        E->stages[i].ms = n ? (float)(sum / n) : 0.f;
        if (name) *name = E->stages[i].name;
        if (ms) *ms = E->stages[i].ms;
    }
    return (int)E->stages.size();
}

int gforb_extract(GforbExtractor* E, const uint8_t* images, int w, int h, int stride, size_t frame_stride, int batch,
                  GforbKeyPoint* keypoints, uint8_t* descriptors, int32_t* counts, int flags, void* stream_) {
    if (!E) return GFORB_ERR_ARG;
    if (batch == 0 || images == nullptr || w <= 0 || h <= 0) return GFORB_OK;     // empty image: silent return (:772)
    if (batch < 0 || stride < w) return GFORB_ERR_ARG;
    if (batch > E->cfg.max_batch || (size_t)w * h > (size_t)E->cfg.max_width * E->cfg.max_height) return GFORB_ERR_CAPACITY;
    GF_CUDA_OK(cudaSetDevice(E->device));
    cudaStream_t st = (cudaStream_t)stream_;   // NULL = the CUDA default stream
    if (w != E->planW || h != E->planH) {
        GF_CUDA_OK(cudaStreamSynchronize(st));
        int rc = build_plan(E, w, h);
        if (rc != GFORB_OK) { E->planW = E->planH = -1; return rc; }
    }
    const GforbExtractorConfig& c = E->cfg;
    const int nl = c.nlevels;
    const uint8_t* dimg = images;
    int dstride = stride;
    size_t dfs = frame_stride;
    if (!(flags & GFORB_IN_DEVICE)) {
        dstride = w; dfs = (size_t)w * h;
        if (frame_stride == (size_t)stride * h)
            GF_CUDA_OK(cudaMemcpy2DAsync(E->d_img, w, images, stride, w, (size_t)h * batch, cudaMemcpyHostToDevice, st));
        else
            for (int b = 0; b < batch; b++)
                GF_CUDA_OK(cudaMemcpy2DAsync(E->d_img + (size_t)b * dfs, w, images + (size_t)b * frame_stride, stride, w, h, cudaMemcpyHostToDevice, st));
        dimg = E->d_img;
    }
    GforbKeyPoint* dk = (flags & GFORB_OUT_DEVICE) ? keypoints : E->d_kps;
    uint8_t* dd = (flags & GFORB_OUT_DEVICE) ? descriptors : E->d_desc;
    int32_t* dc = (flags & GFORB_OUT_DEVICE) ? counts : E->d_counts;
    if (!dk || !dd || !dc) return GFORB_ERR_ARG;

    int launches = 0;
    size_t si = 0;
    auto getEv = [&]() {
        cudaEvent_t e;
        if (!E->evPool.empty()) { e = E->evPool.back(); E->evPool.pop_back(); } else cudaEventCreate(&e);
        return e;
    };
    const bool prof = E->profiling && E->recs.size() < 65536;
    cudaStream_t stageStream = st;
    auto begin = [&](const char* name) {
        if (!prof) return;
        if (si >= E->stages.size()) { Stage s; s.name = name; s.ms = 0; s.e0 = s.e1 = nullptr; E->stages.push_back(s); }
        E->stages[si].name = name;
        StageRec r; r.stage = (int)si; r.e0 = getEv(); r.e1 = getEv();
        E->recs.push_back(r);
        cudaEventRecord(r.e0, stageStream);
    };
    auto end = [&]() { if (!prof) return; cudaEventRecord(E->recs.back().e1, stageStream); si++; };

    begin("pyramid");
    {
        const LevelPlan& L0 = E->levels[0];
        k_border0<<<dim3((L0.rows + B0_ROWS - 1) / B0_ROWS, batch), 256, 0, st>>>(dimg, w, h, dstride, dfs, E->d_pyr, E->framePyr, L0); launches++;
    }
    const uint8_t* spyr = E->d_pyr;     // source level l-1 and destination level l are disjoint planes of the same block
    for (int l = 1; l < nl; l++) {
        const LevelPlan& D = E->levels[l];
        const int quads = (D.w + 2 * EDGE + 3) >> 2, nBands = (D.rows + RS_BAND - 1) / RS_BAND;
        if (E->resizeTma && (size_t)RT_MAXSRC * E->levels[l - 1].pitch <= 48 * 1024)
            k_resize_tma<<<dim3((D.rows + RT_ROWS - 1) / RT_ROWS, batch), 256, (size_t)RT_MAXSRC * E->levels[l - 1].pitch, st>>>(
                spyr, E->d_pyr, E->framePyr, E->levels[l - 1], D, E->d_tabs, quads);
        else if (E->resizeWords == 1 && E->resizeWordsOk)
            k_resize2<4><<<dim3((quads * nBands + 255) / 256, batch), 256, 0, st>>>(spyr, E->d_pyr, E->framePyr, E->levels[l - 1], D, E->d_tabs, quads, nBands);
        else if (E->resizeWords == 2 && E->resizeWordsOk) {
            const int nB8 = (D.rows + 7) / 8;
            k_resize2<8><<<dim3((quads * nB8 + 255) / 256, batch), 256, 0, st>>>(spyr, E->d_pyr, E->framePyr, E->levels[l - 1], D, E->d_tabs, quads, nB8);
        }
        else
            k_resize<<<dim3((quads * nBands + 255) / 256, batch), 256, 0, st>>>(spyr, E->d_pyr, E->framePyr, E->levels[l - 1], D, E->d_tabs, quads, nBands);
        launches++;
    }
    end();
    begin("detect");
    if (E->detectVersion == 2 && E->tiled) {
        // strip height: every strip re-reads 7 rows of run-in, so tall strips (<= 128 rows) do ~6 % less work; short ones (<= 64)
        // keep the latency of a small batch low.  Tall once the short tasks of the batch would fill the resident warps 4 times over.
        const int fv = E->fnTall >= 0 ? E->fnTall : ((long long)E->fnTasks[0] * batch >= 4LL * E->numSMs * 16 ? 1 : 0);
        k_fast_nms<<<dim3((E->fnTasks[fv] + FN_WARPS - 1) / FN_WARPS, batch), FN_WARPS * 32, 0, st>>>(
            E->d_pyr, E->d_blr, E->framePyr, E->d_levels, nl, E->fnTasks[fv], c.minThFAST, fv); launches++;
        k_cell_emit<<<dim3((E->nCellsTotal + EM_WARPS - 1) / EM_WARPS, batch), EM_WARPS * 32, 0, st>>>(
            E->d_blr, E->framePyr, E->d_levels, E->d_cells, E->nCellsTotal, E->d_cand, E->frameCand, E->d_cnt7, E->d_cnt20,
            c.iniThFAST, E->d_err, E->kpLevels, nl <= KP_MAX_LEVELS ? 1 : 0); launches++;
    } else {
        k_detect<<<dim3(E->nCellsTotal, batch), DET_THREADS, E->detSmem, st>>>(
            E->d_pyr, E->framePyr, E->d_levels, E->d_cells, E->nCellsTotal, E->d_cand, E->frameCand, E->d_cnt7, E->d_cnt20,
            c.iniThFAST, c.minThFAST, E->detPixPitch, E->d_err); launches++;
    }
    end();
    // The blur needs only the pyramid and the score map's storage (free once k_cell_emit has read it); the quota / retainBest /
    // orientation kernels are latency-bound and leave the SMs almost empty: the blur runs beside them on the handle's own stream
    // and joins before the descriptors (issued first, so its blocks fill the SMs the short kernels leave idle).
    const bool fork = E->blurFork && st != E->ownStream;
    if (fork) {
        GF_CUDA_OK(cudaEventRecord(E->evFork, st));
        GF_CUDA_OK(cudaStreamWaitEvent(E->ownStream, E->evFork, 0));
    }
    if (fork) stageStream = E->ownStream;
    begin("blur");
    const int bv = E->fnTall >= 0 ? E->fnTall : ((long long)E->blTasks[0] * batch >= 4LL * E->numSMs * 16 ? 1 : 0);
    if (E->blurV2)
        k_blur2<<<dim3((E->blTasks[bv] + BL2_WARPS - 1) / BL2_WARPS, batch), BL2_WARPS * 32, 0, fork ? E->ownStream : st>>>(
            E->d_pyr, E->d_blr, E->framePyr, E->d_levels, nl, E->blTasks[bv], c.blur_variant, bv);
    else
    {
        const dim3 bg((E->blTasks[bv] + BL2_WARPS - 1) / BL2_WARPS, batch);
        cudaStream_t bs = fork ? E->ownStream : st;
        if (E->blurQD == 2) k_blur3<2><<<bg, BL2_WARPS * 32, 0, bs>>>(E->d_pyr, E->d_blr, E->framePyr, E->d_levels, nl, E->blTasks[bv], c.blur_variant, bv);
        else if (E->blurQD == 3) k_blur3<3><<<bg, BL2_WARPS * 32, 0, bs>>>(E->d_pyr, E->d_blr, E->framePyr, E->d_levels, nl, E->blTasks[bv], c.blur_variant, bv);
        else k_blur3<1><<<bg, BL2_WARPS * 32, 0, bs>>>(E->d_pyr, E->d_blr, E->framePyr, E->d_levels, nl, E->blTasks[bv], c.blur_variant, bv);
    }
    launches++;
    end();
    stageStream = st;
    if (fork) GF_CUDA_OK(cudaEventRecord(E->evJoin, E->ownStream));
    begin("select");
    const bool harris = c.scoreType == 0;
    if (E->distribution == 1) {
        k_quadtree<<<batch, 256, 0, st>>>(E->d_levels, E->d_qlv, nl, E->d_cells, E->nCellsTotal, E->d_cand, E->frameCand, E->d_cnt7,
                                          E->d_cnt20, c.iniThFAST, E->d_qtKeys, E->d_qtTmp, E->frameQtKeys, E->d_qtNodes, E->d_qtVsz,
                                          E->frameQtNodes, E->d_lvlKp, E->frameKp, E->d_lvlCount, c.nfeatures, E->d_err); launches++;
    } else {
    k_quota<<<(batch * nl * 32 + 127) / 128, 128, 0, st>>>(E->d_levels, E->d_cells, nl, E->nCellsTotal, E->d_cnt7, E->d_cnt20,
                                                     E->d_nTotal, E->d_nToRetain, E->d_cellOff, E->d_lvlPre, batch, E->d_err); launches++;
    if (!harris) {
        k_cell_retain<<<dim3((E->nCellsTotal + RET_WARPS - 1) / RET_WARPS, batch), RET_WARPS * 32, 0, st>>>(
            E->d_cells, E->d_levels, E->nCellsTotal, E->d_cand, E->frameCand, E->d_cnt7, E->d_nTotal, E->d_nToRetain,
            E->d_cellOff, E->d_lvlKp, E->frameKp, c.iniThFAST, c.retain_variant | (E->retainSeq ? 2 : 0)); launches++;
        // scratch of the warp-parallel selection: 4 bytes per level-buffer entry and warp (sequential when it would not fit)
        int cap = 0;
        for (int l = 0; l < nl; l++) cap = std::max(cap, E->levels[l].kpCap);
        cap = (cap + 7) & ~7;
        if ((size_t)cap * 16 > 48 * 1024 || cap > 0x3fff) cap = 0;
        k_level_retain<<<(batch * nl * 32 + 127) / 128, 128, (size_t)cap * 16, st>>>(E->d_levels, nl, E->d_lvlKp, E->frameKp, E->d_lvlPre,
                                                                                     E->d_lvlCount, batch, c.retain_variant | (E->retainSeq ? 2 : 0), cap); launches++;
    } else {
        k_cell_retain_harris<<<dim3((E->nCellsTotal + RET_WARPS - 1) / RET_WARPS, batch), RET_WARPS * 32, 0, st>>>(
            E->d_cells, E->d_levels, E->nCellsTotal, E->d_cand, E->frameCand, E->d_cnt7, E->d_nTotal, E->d_nToRetain,
            E->d_cellOff, E->d_lvlKp, E->d_lvlResp, E->frameKp, c.iniThFAST, c.retain_variant, E->d_pyr, E->framePyr,
            E->d_work64, E->frameWork); launches++;
        k_level_retain_harris<<<(batch * nl * 32 + 127) / 128, 128, 0, st>>>(
            E->d_levels, nl, E->d_lvlKp, E->d_lvlResp, E->frameKp, E->d_lvlPre, E->d_lvlCount, batch, c.retain_variant,
            E->d_work64, E->frameWork); launches++;
    }
    }
    end();
    begin("orient");
    if (E->kpV1)
        k_orient<<<dim3((E->kpCapTotal + 7) / 8, batch), 256, 0, st>>>(E->d_pyr, E->framePyr, E->d_levels, nl, E->d_lvlKp,
                                                                        E->frameKp, E->kpCapTotal, E->d_lvlCount, dk, dc, c.nfeatures,
                                                                        harris ? E->d_lvlResp : nullptr);
    else
        k_orient2<<<dim3((c.nfeatures + 7) / 8, batch), 256, 0, st>>>(E->d_pyr, E->framePyr, E->d_levels, nl, E->d_lvlKp,
                                                                       E->frameKp, E->d_lvlCount, dk, dc, c.nfeatures,
                                                                       harris ? E->d_lvlResp : nullptr, E->kpLevels);
    launches++;
    end();
    if (fork) GF_CUDA_OK(cudaStreamWaitEvent(st, E->evJoin, 0));
    begin("describe");
    if (E->kpV1)
        k_describe<false, false><<<dim3((E->kpCapTotal + 7) / 8, batch), 256, 0, st>>>(E->d_pyr, E->d_blr, E->framePyr, E->d_levels, nl, E->d_lvlKp,
                                                                                 E->frameKp, E->kpCapTotal, E->d_lvlCount, dk, dd, c.nfeatures, E->kpLevels);
    else if (E->describeSmem == 0)
        k_describe<true, false><<<dim3((c.nfeatures + 7) / 8, batch), 256, 0, st>>>(E->d_pyr, E->d_blr, E->framePyr, E->d_levels, nl, E->d_lvlKp,
                                                                              E->frameKp, E->kpCapTotal, E->d_lvlCount, dk, dd, c.nfeatures, E->kpLevels);
    else if (E->describeSmem == 2)
        k_describe<true, true><<<dim3((c.nfeatures + 7) / 8, batch), 256, 0, st>>>(E->d_pyr, E->d_blr, E->framePyr, E->d_levels, nl, E->d_lvlKp,
                                                                              E->frameKp, E->kpCapTotal, E->d_lvlCount, dk, dd, c.nfeatures, E->kpLevels);
    else
        k_describe2<<<dim3((c.nfeatures + 7) / 8, batch), 256, 0, st>>>(E->d_pyr, E->d_blr, E->framePyr, E->d_levels, nl, E->d_lvlKp,
                                                                         E->frameKp, E->d_lvlCount, dk, dd, c.nfeatures);
    launches++;
    end();
    GF_CUDA_OK(cudaGetLastError());
    E->lastLaunches = launches;
    E->lastBatch = batch;
    if (!(flags & GFORB_OUT_DEVICE)) {
        if (!keypoints || !descriptors || !counts) return GFORB_ERR_ARG;
        GF_CUDA_OK(cudaMemcpyAsync(keypoints, dk, (size_t)batch * c.nfeatures * sizeof(GforbKeyPoint), cudaMemcpyDeviceToHost, st));
        GF_CUDA_OK(cudaMemcpyAsync(descriptors, dd, (size_t)batch * c.nfeatures * 32, cudaMemcpyDeviceToHost, st));
        GF_CUDA_OK(cudaMemcpyAsync(counts, dc, (size_t)batch * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    }
    if (!(flags & GFORB_ASYNC) || !(flags & GFORB_OUT_DEVICE)) {
        GF_CUDA_OK(cudaStreamSynchronize(st));
        int err = 0;
        GF_CUDA_OK(cudaMemcpy(&err, E->d_err, sizeof(int), cudaMemcpyDeviceToHost));
        if (err) { cudaMemset(E->d_err, 0, sizeof(int)); return err; }
    }
    return GFORB_OK;
}

// error word of GFORB_ASYNC calls (cell-list overflow): synchronises `stream`, returns and clears the recorded error
int gforb_extractor_poll_error(GforbExtractor* E, void* stream) {
    if (!E) return GFORB_ERR_ARG;
    GF_CUDA_OK(cudaSetDevice(E->device));
    GF_CUDA_OK(cudaStreamSynchronize((cudaStream_t)stream));
    int err = 0;
    GF_CUDA_OK(cudaMemcpy(&err, E->d_err, sizeof(int), cudaMemcpyDeviceToHost));
    if (err) { cudaMemset(E->d_err, 0, sizeof(int)); return err; }
    return GFORB_OK;
}

// 1 (default): the Gaussian blur of a call runs on the handle's own stream beside the quota / retainBest / orientation
// kernels; 0: every kernel of a call on the caller's stream, one after the other (serial per-kernel timing)
int gforb_extractor_set_concurrency(GforbExtractor* E, int on) {
    if (!E) return GFORB_ERR_ARG;
    E->blurFork = on != 0;
    return GFORB_OK;
}

// 0 (default): the reference's cell-quota distribution (ComputeKeyPoints, src/ORBextractor.cc:535-758); 1: ORB-SLAM2's
// quadtree (quadtree.cuh).  Re-plans the cell grid for the sizes given at create().
int gforb_extractor_set_distribution(GforbExtractor* E, int mode) {
    if (!E || (mode != 0 && mode != 1)) return GFORB_ERR_ARG;
    if (mode == 1 && E->cfg.scoreType == 0) return GFORB_ERR_ARG;          // the quadtree mode ranks by the FAST score
    if (mode == E->distribution) return GFORB_OK;
    GF_CUDA_OK(cudaSetDevice(E->device));
    GF_CUDA_OK(cudaDeviceSynchronize());
    E->distribution = mode;
    const int w = E->planW > 0 ? E->planW : E->cfg.max_width, h = E->planH > 0 ? E->planH : E->cfg.max_height;
    const int rc = build_plan(E, w, h);
    if (rc != GFORB_OK) { E->planW = E->planH = -1; }
    return rc;
}

int gforb_extractor_debug_level(GforbExtractor* E, int b, int level, int which, uint8_t* dst, size_t dst_bytes, int* w,
                                int* hgt, int* pitch) {
    if (!E || level < 0 || level >= E->cfg.nlevels || b < 0 || b >= E->lastBatch) return GFORB_ERR_ARG;
    const LevelPlan& L = E->levels[level];
    if (w) *w = L.w;
    if (hgt) *hgt = L.h;
    if (pitch) *pitch = L.pitch;
    size_t bytes = (size_t)L.pitch * L.rows;
    if (dst) {
        if (dst_bytes < bytes) return GFORB_ERR_CAPACITY;
        GF_CUDA_OK(cudaSetDevice(E->device));
        GF_CUDA_OK(cudaDeviceSynchronize());
        GF_CUDA_OK(cudaMemcpy(dst, (which ? E->d_blr : E->d_pyr) + (size_t)b * E->framePyr + L.off, bytes, cudaMemcpyDeviceToHost));
    }
    return GFORB_OK;
}

}  // extern "C"
