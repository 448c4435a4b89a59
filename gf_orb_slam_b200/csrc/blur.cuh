// K5 blur, second generation: GaussianBlur(level, 7x7, sigma 2, BORDER_REFLECT_101) with OpenCV's 8U fixed-point rounding
// (src/ORBextractor.cc:842) as a register-rolling separable stencil.  Included by extractor.cu.
//
// One warp per strip of 128 columns x up to 64 or 128 rows (LevelPlan::blRows); lane l owns pixels x = X0 + 4 l .. +3.  Each input row is
// read once (coalesced 32-bit word per lane + neighbours' words by shuffle) and split into even / odd bytes, i.e. into
// the stride-2 pixel pairs {b0, b2} and {b1, b3}.  The horizontal pass runs on those packed u16x2 pairs (taps sum to
// 256 or 257, so 255 * sum fits 16 bits and a plain 32-bit IMAD never carries across halves); its result
// {h(x), h(x+2)}, {h(x+1), h(x+3)} enters a 7-row register window and the vertical pass emits one output row per input
// row: out = (sum_q g[q] * h[y-3+q] + 2^15) >> 16, saturated.  The ring of the source plane supplies the reflect-101
// border (it holds the unblurred reflection; verified identical to an isolated reflect-101 blur).
#pragma once

namespace gforb {

static const int BL2_ROWS[2] = {64, 128};   // most rows per strip, variant 0 / 1 (LevelPlan::blRows holds the per-level value)
static const int BL2_WARPS = 4;

// stride-2 pair {byte J, byte J+2} of the 12-byte window (J = 0..9) from the even/odd split words
template <int J>
__device__ __forceinline__ uint32_t bl_pair2(const uint32_t (&E)[3], const uint32_t (&O)[3]) {
    constexpr int j = J >> 2, q = J & 3;
    if (q == 0) return E[j];
    if (q == 1) return O[j];
    if (q == 2) return __byte_perm(E[j], E[j < 2 ? j + 1 : 2], 0x5432);
    return __byte_perm(O[j], O[j < 2 ? j + 1 : 2], 0x5432);
}

// horizontal 7-tap of the pair {pixel P, pixel P+2} (window bytes P+4-3 .. ), symmetric taps g0..g3
template <int P>
__device__ __forceinline__ uint32_t bl_hpass(const uint32_t (&E)[3], const uint32_t (&O)[3], uint32_t g0, uint32_t g1,
                                             uint32_t g2, uint32_t g3) {
    constexpr int C = 4 + P;     // window byte of pixel P
    const uint32_t s0 = bl_pair2<C - 3>(E, O) + bl_pair2<C + 3>(E, O);
    const uint32_t s1 = bl_pair2<C - 2>(E, O) + bl_pair2<C + 2>(E, O);
    const uint32_t s2 = bl_pair2<C - 1>(E, O) + bl_pair2<C + 1>(E, O);
    return g0 * s0 + g1 * s1 + g2 * s2 + g3 * bl_pair2<C>(E, O);
}

__global__ void __launch_bounds__(BL2_WARPS * 32) k_blur2(const uint8_t* __restrict__ pyr, uint8_t* __restrict__ blr,
                                                          size_t framePyr, const LevelPlan* __restrict__ levels, int nLevels,
                                                          int tasksPerFrame, int variant, int bv) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int task = blockIdx.x * BL2_WARPS + warp;
    const int b = blockIdx.y;
    if (task >= tasksPerFrame) return;
    int l = 0;
    while (l + 1 < nLevels && task >= (bv ? levels[l + 1].blTaskBase[1] : levels[l + 1].blTaskBase[0])) l++;
    const LevelPlan L = levels[l];
    task -= bv ? L.blTaskBase[1] : L.blTaskBase[0];
    const int stripRows = bv ? L.blRows[1] : L.blRows[0];
    const int sy = task / L.blStripsX, sx = task - sy * L.blStripsX;
    const int Y0 = stripRows * sy, Y1 = min(Y0 + stripRows, L.h);
    const int xl = 128 * sx + 4 * lane;
    const uint8_t* roi = pyr + (size_t)b * framePyr + L.off + (size_t)EDGE * L.pitch + EDGE;
    uint8_t* oroi = blr + (size_t)b * framePyr + L.off + (size_t)EDGE * L.pitch + EDGE;
    const uint32_t g0 = 18, g1 = 34, g2 = variant ? 49 : 48, g3 = variant ? 55 : 56;
    const unsigned full = 0xffffffffu;
    const int hoff = lane < 16 ? -4 : 4;
    const uint8_t* lp = roi + (ptrdiff_t)(Y0 - 3) * L.pitch + xl;
    uint32_t nw1 = __ldg(reinterpret_cast<const uint32_t*>(lp));
    uint32_t nh = __ldg(reinterpret_cast<const uint32_t*>(lp + hoff));
    uint32_t HA[7], HB[7];       // horizontal results {h(x), h(x+2)} / {h(x+1), h(x+3)} of the last 7 input rows
    const int Rend = Y1 + 2;

    auto phase = [&](auto PHc, int R) {
        constexpr int PH = decltype(PHc)::value;
        const uint32_t w1 = nw1, hw = nh;
        lp += L.pitch;
        nw1 = __ldg(reinterpret_cast<const uint32_t*>(lp));      // row Rend + 1 <= h + 3 is inside the bordered plane
        nh = __ldg(reinterpret_cast<const uint32_t*>(lp + hoff));
        uint32_t w0 = __shfl_up_sync(full, w1, 1), w2 = __shfl_down_sync(full, w1, 1);
        if (lane == 0) w0 = hw;
        if (lane == 31) w2 = hw;
        uint32_t E[3], O[3];
        E[0] = w0 & 0x00ff00ffu; O[0] = __byte_perm(w0, 0, 0x4341);
        E[1] = w1 & 0x00ff00ffu; O[1] = __byte_perm(w1, 0, 0x4341);
        E[2] = w2 & 0x00ff00ffu; O[2] = __byte_perm(w2, 0, 0x4341);
        HA[PH] = bl_hpass<0>(E, O, g0, g1, g2, g3);
        HB[PH] = bl_hpass<1>(E, O, g0, g1, g2, g3);
        if (R < Y0 + 3) return;
        // ---- output row y = R - 3: input rows y-3 .. y+3 sit in slots (PH+1) .. (PH+7) mod 7
        const int y = R - 3;
#define BL_S(k) ((PH + 1 + (k)) % 7)
        uint32_t o[4];
#pragma unroll
        for (int half = 0; half < 2; half++) {
            // pixel x + 2*half (from HA) and x + 1 + 2*half (from HB)
            const int sh = 16 * half;
            const uint32_t a0 = ((HA[BL_S(0)] >> sh) & 0xffffu) + ((HA[BL_S(6)] >> sh) & 0xffffu);
            const uint32_t a1 = ((HA[BL_S(1)] >> sh) & 0xffffu) + ((HA[BL_S(5)] >> sh) & 0xffffu);
            const uint32_t a2 = ((HA[BL_S(2)] >> sh) & 0xffffu) + ((HA[BL_S(4)] >> sh) & 0xffffu);
            const uint32_t a3 = (HA[BL_S(3)] >> sh) & 0xffffu;
            const uint32_t b0 = ((HB[BL_S(0)] >> sh) & 0xffffu) + ((HB[BL_S(6)] >> sh) & 0xffffu);
            const uint32_t b1 = ((HB[BL_S(1)] >> sh) & 0xffffu) + ((HB[BL_S(5)] >> sh) & 0xffffu);
            const uint32_t b2 = ((HB[BL_S(2)] >> sh) & 0xffffu) + ((HB[BL_S(4)] >> sh) & 0xffffu);
            const uint32_t b3 = (HB[BL_S(3)] >> sh) & 0xffffu;
            o[2 * half] = min((g0 * a0 + g1 * a1 + g2 * a2 + g3 * a3 + 32768u) >> 16, 255u);
            o[2 * half + 1] = min((g0 * b0 + g1 * b1 + g2 * b2 + g3 * b3 + 32768u) >> 16, 255u);
        }
#undef BL_S
        const uint32_t word = o[0] | (o[1] << 8) | (o[2] << 16) | (o[3] << 24);
        uint8_t* op = oroi + (ptrdiff_t)y * L.pitch + xl;
        if (xl + 3 < L.w) *reinterpret_cast<uint32_t*>(op) = word;
        else
            for (int k = 0; k < 4 && xl + k < L.w; k++) op[k] = (uint8_t)(word >> (8 * k));
    };
    for (int R = Y0 - 3; R <= Rend; R += 7) {
        phase(std::integral_constant<int, 0>(), R);
        if (R + 1 <= Rend) phase(std::integral_constant<int, 1>(), R + 1);
        if (R + 2 <= Rend) phase(std::integral_constant<int, 2>(), R + 2);
        if (R + 3 <= Rend) phase(std::integral_constant<int, 3>(), R + 3);
        if (R + 4 <= Rend) phase(std::integral_constant<int, 4>(), R + 4);
        if (R + 5 <= Rend) phase(std::integral_constant<int, 5>(), R + 5);
        if (R + 6 <= Rend) phase(std::integral_constant<int, 6>(), R + 6);
    }
}

// K5, third generation (default; GFORB_BLUR=v2 keeps k_blur2 as the A/B partner): same strips, loads and horizontal pass; the
// VERTICAL pass runs on IDP.2A.  k_blur2 unpacks every 16-bit horizontal result into its own 32-bit lane-op chain
// (7 extractions + 3 adds + 4 IMAD per pixel).  Here each new row is zipped with the previous one into per-pixel row pairs
// {h(r-1, x), h(r, x)} (one PRMT per pixel), and an output pixel is four dp2a (16-bit x 8-bit two-way dot products with a
// 32-bit accumulator) over the pairs (y-3,y-2) (y-1,y) (y+1,y+2) (y+2,y+3) with tap pairs (g0,g1) (g2,g3) (g2,g1) (0,g0),
// accumulator preloaded with the rounding constant 2^15: the same exact integer sum, 7 instead of ~17 lane-ops per pixel.
__device__ __forceinline__ uint32_t dp2a_lo_u32(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("dp2a.lo.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// QD = rows in flight ahead of the one being filtered (1, 2, 3 or 6: must divide the 6-phase unroll so that every queue slot is a
// fixed register that is consumed directly, never copied while its load is in flight)
template <int QD>
__global__ void __launch_bounds__(BL2_WARPS * 32) k_blur3(const uint8_t* __restrict__ pyr, uint8_t* __restrict__ blr,
                                                          size_t framePyr, const LevelPlan* __restrict__ levels, int nLevels,
                                                          int tasksPerFrame, int variant, int bv) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int task = blockIdx.x * BL2_WARPS + warp;
    const int b = blockIdx.y;
    if (task >= tasksPerFrame) return;
    int l = 0;
    while (l + 1 < nLevels && task >= (bv ? levels[l + 1].blTaskBase[1] : levels[l + 1].blTaskBase[0])) l++;
    const LevelPlan L = levels[l];
    task -= bv ? L.blTaskBase[1] : L.blTaskBase[0];
    const int stripRows = bv ? L.blRows[1] : L.blRows[0];
    const int sy = task / L.blStripsX, sx = task - sy * L.blStripsX;
    const int Y0 = stripRows * sy, Y1 = min(Y0 + stripRows, L.h);
    const int xl = 128 * sx + 4 * lane;
    const uint8_t* roi = pyr + (size_t)b * framePyr + L.off + (size_t)EDGE * L.pitch + EDGE;
    uint8_t* oroi = blr + (size_t)b * framePyr + L.off + (size_t)EDGE * L.pitch + EDGE;
    const uint32_t g0 = 18, g1 = 34, g2 = variant ? 49 : 48, g3 = variant ? 55 : 56;
    const uint32_t t01 = g0 | (g1 << 8), t23 = g2 | (g3 << 8), t21 = g2 | (g1 << 8), t_0 = g0 << 8;
    const unsigned full = 0xffffffffu;
    const int hoff = lane < 16 ? -4 : 4;
    const uint8_t* lp = roi + (ptrdiff_t)(Y0 - 3) * L.pitch + xl;
    uint32_t qw[QD], qh[QD];
#pragma unroll
    for (int i = 0; i < QD; i++) {
        qw[i] = __ldg(reinterpret_cast<const uint32_t*>(lp + (ptrdiff_t)i * L.pitch));
        qh[i] = __ldg(reinterpret_cast<const uint32_t*>(lp + (ptrdiff_t)i * L.pitch + hoff));
    }
    lp += (ptrdiff_t)(QD - 1) * L.pitch;
    uint32_t Q[6][4];            // row pairs {h(r-1), h(r)} of pixels x .. x+3 for the last 6 rows r, slot = phase
    uint32_t pA = 0, pB = 0;     // horizontal results of the previous row
    const int Rend = Y1 + 2;

    auto phase = [&](auto PHc, int R) {
        constexpr int PH = decltype(PHc)::value;
        const uint32_t w1 = qw[PH % QD], hw = qh[PH % QD];
        lp += L.pitch;
        qw[PH % QD] = __ldg(reinterpret_cast<const uint32_t*>(lp));      // rows up to Rend + QD <= h + 2 + 6 are inside the bordered plane (16 ring rows)
        qh[PH % QD] = __ldg(reinterpret_cast<const uint32_t*>(lp + hoff));
        uint32_t w0 = __shfl_up_sync(full, w1, 1), w2 = __shfl_down_sync(full, w1, 1);
        if (lane == 0) w0 = hw;
        if (lane == 31) w2 = hw;
        uint32_t E[3], O[3];
        E[0] = w0 & 0x00ff00ffu; O[0] = __byte_perm(w0, 0, 0x4341);
        E[1] = w1 & 0x00ff00ffu; O[1] = __byte_perm(w1, 0, 0x4341);
        E[2] = w2 & 0x00ff00ffu; O[2] = __byte_perm(w2, 0, 0x4341);
        const uint32_t hA = bl_hpass<0>(E, O, g0, g1, g2, g3);    // {h(x), h(x+2)}
        const uint32_t hB = bl_hpass<1>(E, O, g0, g1, g2, g3);    // {h(x+1), h(x+3)}
        Q[PH][0] = __byte_perm(pA, hA, 0x5410);
        Q[PH][1] = __byte_perm(pB, hB, 0x5410);
        Q[PH][2] = __byte_perm(pA, hA, 0x7632);
        Q[PH][3] = __byte_perm(pB, hB, 0x7632);
        pA = hA; pB = hB;
        if (R < Y0 + 3) return;
        // ---- output row y = R - 3: pairs ending at rows R-5, R-3, R-1, R sit in slots PH+1, PH+3, PH+5, PH (mod 6)
        uint32_t o[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            uint32_t acc = dp2a_lo_u32(Q[(PH + 1) % 6][k], t01, 32768u);
            acc = dp2a_lo_u32(Q[(PH + 3) % 6][k], t23, acc);
            acc = dp2a_lo_u32(Q[(PH + 5) % 6][k], t21, acc);
            acc = dp2a_lo_u32(Q[PH][k], t_0, acc);
            o[k] = min(acc >> 16, 255u);
        }
        const uint32_t word = __byte_perm(__byte_perm(o[0], o[1], 0x0040), __byte_perm(o[2], o[3], 0x0040), 0x5410);
        uint8_t* op = oroi + (ptrdiff_t)(R - 3) * L.pitch + xl;
        if (xl + 3 < L.w) *reinterpret_cast<uint32_t*>(op) = word;
        else
            for (int k = 0; k < 4 && xl + k < L.w; k++) op[k] = (uint8_t)(word >> (8 * k));
    };
    for (int R = Y0 - 3; R <= Rend; R += 6) {
        phase(std::integral_constant<int, 0>(), R);
        if (R + 1 <= Rend) phase(std::integral_constant<int, 1>(), R + 1);
        if (R + 2 <= Rend) phase(std::integral_constant<int, 2>(), R + 2);
        if (R + 3 <= Rend) phase(std::integral_constant<int, 3>(), R + 3);
        if (R + 4 <= Rend) phase(std::integral_constant<int, 4>(), R + 4);
        if (R + 5 <= Rend) phase(std::integral_constant<int, 5>(), R + 5);
    }
}

}  // namespace gforb
