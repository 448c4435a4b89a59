// TEST INFRASTRUCTURE ONLY (see oracle.hpp).  Restatements of the un-vendored OpenCV primitives the
// reference's ORBextractor calls (src/ORBextractor.cc:943 resize, :982/:990 copyMakeBorder, :621/:627 FAST,
// :734/:750 KeyPointsFilter::retainBest, :842 GaussianBlur, :157 fastAtan2, cvRound everywhere).
// OpenCV is a third-party dependency absent from /root/reference (pinned 2.4.13 or 3.4.1 by
// CMakeLists.txt:42-49 / build_dep.sh:37-45); what is restated is its published algorithm, cross-checked
// against cv2 4.13 in tests/test_oracle_cv.py.
#include "oracle.hpp"
#if defined(__SSE2__)
#include <emmintrin.h>
#endif
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>

namespace oracle {

int cvRound(double v) { return (int)lrint(v); }     // round-half-even under the default FP environment
int cvRoundf(float v) { return (int)lrintf(v); }
static inline int cvFloor(double v) { int i = (int)v; return i - (i > v); }

// ---------------------------------------------------------------------------------------------------------------
// cv::fastAtan2 (float polynomial, degrees in [0,360)); no FMA contraction (build with -ffp-contract=off)
float fastAtan2(float y, float x) {
    const float s = (float)(180.0 / M_PI);
    const float p1 = 0.9997878412794807f * s, p3 = -0.3258083974640975f * s;
    const float p5 = 0.1555786518463281f * s, p7 = -0.04432655554792128f * s;
    float ax = std::fabs(x), ay = std::fabs(y), a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

// ---------------------------------------------------------------------------------------------------------------
// cv::resize(..., INTER_LINEAR) for 8UC1: 11-bit fixed-point coefficients, HResize into int32, VResize with the
// ((b*(r>>4))>>16 ... +2)>>2 rounding.
static void linearCoeffs(int ssize, int dsize, std::vector<int>& ofs, std::vector<short>& coef) {
    double scale = 1.0 / ((double)dsize / ssize);
    ofs.resize(dsize);
    coef.resize(2 * dsize);
    for (int d = 0; d < dsize; d++) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = cvFloor(f);
        f -= s;
        if (s < 0) { s = 0; f = 0; }
        if (s >= ssize - 1) { s = ssize - 1; f = 0; }
        ofs[d] = s;
        coef[2 * d] = (short)cvRoundf((1.f - f) * 2048.f);
        coef[2 * d + 1] = (short)cvRoundf(f * 2048.f);
    }
}

void resizeLinear8u(const uint8_t* src, int sw, int sh, int sstep, uint8_t* dst, int dw, int dh, int dstep) {
    std::vector<int> xofs, yofs;
    std::vector<short> alpha, beta;
    linearCoeffs(sw, dw, xofs, alpha);
    linearCoeffs(sh, dh, yofs, beta);
    std::vector<int> r0(dw), r1(dw);
    for (int y = 0; y < dh; y++) {
        int sy = yofs[y];
        const uint8_t* S0 = src + (size_t)sy * sstep;
        const uint8_t* S1 = src + (size_t)std::min(sy + 1, sh - 1) * sstep;
        for (int x = 0; x < dw; x++) {
            int sx = xofs[x], sx1 = std::min(sx + 1, sw - 1);
            int a0 = alpha[2 * x], a1 = alpha[2 * x + 1];
            r0[x] = S0[sx] * a0 + S0[sx1] * a1;
            r1[x] = S1[sx] * a0 + S1[sx1] * a1;
        }
        int b0 = beta[2 * y], b1 = beta[2 * y + 1];
        uint8_t* D = dst + (size_t)y * dstep;
        for (int x = 0; x < dw; x++) {
            int v = (((b0 * (r0[x] >> 4)) >> 16) + ((b1 * (r1[x] >> 4)) >> 16) + 2) >> 2;
            D[x] = (uint8_t)std::min(std::max(v, 0), 255);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
static inline int reflect101(int p, int len) {
    if (len == 1) return 0;
    while (p < 0 || p >= len) {
        if (p < 0) p = -p;
        else p = 2 * (len - 1) - p;
    }
    return p;
}

void copyMakeBorder101(const uint8_t* src, int w, int h, int sstep, uint8_t* dst, int dstep, int b) {
    for (int y = -b; y < h + b; y++) {
        const uint8_t* S = src + (size_t)reflect101(y, h) * sstep;
        uint8_t* D = dst + (size_t)(y + b) * dstep;
        for (int x = -b; x < w + b; x++) D[x + b] = S[reflect101(x, w)];
    }
}

// ---------------------------------------------------------------------------------------------------------------
// GaussianBlur 7x7 sigma=2.  Variant 0: OpenCV >=3.4.x "bit-exact" fixed-point path: Q8 kernel
// {18,34,48,56,48,34,18} applied per axis in 8.8 / 16.16 fixed point, final (acc + 2^15) >> 16.
// Variant 1 (legacy 2.4.x integer path): kernel cvRound(k*256) = {18,34,49,55,49,34,18}, saturated.
static const int kGauss[2][7] = {{18, 34, 48, 56, 48, 34, 18}, {18, 34, 49, 55, 49, 34, 18}};

// Separable pass over rows that are directly addressable with a 3-pixel margin on every side (a reflect-101 padded copy,
// or the ROI of a bordered pyramid plane): a ring of 7 horizontally filtered int32 rows, then the vertical taps.  Same
// integer arithmetic as the textbook double loop ((sum_y k_y sum_x k_x p + 2^15) >> 16); written with plain pointer loops so
// that -O3 -march=native vectorises it like OpenCV's SIMD path.
static void blurRows(const uint8_t* src, ptrdiff_t sstep, int w, int h, uint8_t* dst, ptrdiff_t dstep, int variant) {
    const int* k = kGauss[variant];
    std::vector<int> ring((size_t)7 * w);
    auto hrow = [&](int y) {
        const uint8_t* __restrict p = src + (ptrdiff_t)y * sstep - 3;
        int* __restrict r = ring.data() + (size_t)((y + 3) % 7) * w;
        const int k0 = k[0], k1 = k[1], k2 = k[2], k3 = k[3], k4 = k[4], k5 = k[5], k6 = k[6];
        for (int x = 0; x < w; x++)
            r[x] = k0 * p[x] + k1 * p[x + 1] + k2 * p[x + 2] + k3 * p[x + 3] + k4 * p[x + 4] + k5 * p[x + 5] + k6 * p[x + 6];
    };
    for (int y = -3; y < 3; y++) hrow(y);
    for (int y = 0; y < h; y++) {
        hrow(y + 3);
        const int* __restrict r0 = ring.data() + (size_t)((y + 0) % 7) * w;       // row y - 3 lives in slot (y - 3 + 3) % 7
        const int* __restrict r1 = ring.data() + (size_t)((y + 1) % 7) * w;
        const int* __restrict r2 = ring.data() + (size_t)((y + 2) % 7) * w;
        const int* __restrict r3 = ring.data() + (size_t)((y + 3) % 7) * w;
        const int* __restrict r4 = ring.data() + (size_t)((y + 4) % 7) * w;
        const int* __restrict r5 = ring.data() + (size_t)((y + 5) % 7) * w;
        const int* __restrict r6 = ring.data() + (size_t)((y + 6) % 7) * w;
        uint8_t* __restrict d = dst + (ptrdiff_t)y * dstep;
        const int k0 = k[0], k1 = k[1], k2 = k[2], k3 = k[3], k4 = k[4], k5 = k[5], k6 = k[6];
        for (int x = 0; x < w; x++) {
            const int sum = k0 * r0[x] + k1 * r1[x] + k2 * r2[x] + k3 * r3[x] + k4 * r4[x] + k5 * r5[x] + k6 * r6[x];
            const int v = (sum + 32768) >> 16;
            d[x] = (uint8_t)std::min(std::max(v, 0), 255);
        }
    }
}

void gaussianBlur7x7(const uint8_t* src, int w, int h, int sstep, uint8_t* dst, int dstep, int variant) {
    // reflect-101 padded copy (3 pixels), then the shared row kernel; dst may alias src
    const int pw = w + 6;
    std::vector<uint8_t> pad((size_t)pw * (h + 6));
    for (int y = -3; y < h + 3; y++) {
        const uint8_t* srow = src + (size_t)reflect101(y, h) * sstep;
        uint8_t* prow = pad.data() + (size_t)(y + 3) * pw;
        for (int x = -3; x < w + 3; x++) prow[x + 3] = srow[reflect101(x, w)];
    }
    std::vector<uint8_t> tmp((size_t)w * h);
    blurRows(pad.data() + (size_t)3 * pw + 3, pw, w, h, tmp.data(), w, variant);
    for (int y = 0; y < h; y++) memcpy(dst + (size_t)y * dstep, tmp.data() + (size_t)y * w, w);
}

void gaussianBlur7x7InPlace(uint8_t* roi, int w, int h, int step, int variant) {
    // The ROI sits inside a bordered plane, so (without BORDER_ISOLATED) OpenCV reads the ring pixels, which
    // hold the reflect-101 copy of the unblurred level (src/ORBextractor.cc:982-983) => same as isolated blur.
    std::vector<uint8_t> tmp((size_t)w * h);
    blurRows(roi, step, w, h, tmp.data(), w, variant);
    for (int y = 0; y < h; y++) memcpy(roi + (size_t)y * step, tmp.data() + (size_t)y * w, w);
}

// ---------------------------------------------------------------------------------------------------------------
// cv::FAST TYPE_9_16 with non-max suppression.
static const int kCircle[16][2] = {{0, 3},  {1, 3},   {2, 2},   {3, 1},   {3, 0},  {3, -1}, {2, -2}, {1, -3},
                                   {0, -3}, {-1, -3}, {-2, -2}, {-3, -1}, {-3, 0}, {-3, 1}, {-2, 2}, {-1, 3}};

bool g_fast_scalar_only = false;      // tests: force the scalar FAST path (orc_set_fast_scalar)

static bool isCorner9(const uint8_t* p, const int* ofs, int t) {
    int v = p[0];
    // quick reject (result-neutral): every 9-arc of the 16-circle contains position 0 or 8, and 4 or 12
    if (std::abs(p[ofs[0]] - v) <= t && std::abs(p[ofs[8]] - v) <= t) return false;
    if (std::abs(p[ofs[4]] - v) <= t && std::abs(p[ofs[12]] - v) <= t) return false;
    int bright = 0, dark = 0;   // run lengths over the 16+8 wrapped circle
    for (int k = 0; k < 25; k++) {
        int x = p[ofs[k & 15]];
        if (x > v + t) { if (++bright > 8) return true; } else bright = 0;
        if (x < v - t) { if (++dark > 8) return true; } else dark = 0;
    }
    return false;
}

static int cornerScore9(const uint8_t* p, const int* ofs, int threshold) {
    int d[25];
    int v = p[0];
    for (int k = 0; k < 25; k++) d[k] = v - p[ofs[k & 15]];
    int a0 = threshold;
    for (int k = 0; k < 16; k += 2) {
        int a = std::min(d[k + 1], d[k + 2]);
        a = std::min(a, d[k + 3]);
        if (a <= a0) continue;
        a = std::min(a, d[k + 4]); a = std::min(a, d[k + 5]); a = std::min(a, d[k + 6]);
        a = std::min(a, d[k + 7]); a = std::min(a, d[k + 8]);
        a0 = std::max(a0, std::min(a, d[k]));
        a0 = std::max(a0, std::min(a, d[k + 9]));
    }
    int b0 = -a0;
    for (int k = 0; k < 16; k += 2) {
        int b = std::max(d[k + 1], d[k + 2]);
        b = std::max(b, d[k + 3]); b = std::max(b, d[k + 4]); b = std::max(b, d[k + 5]);
        if (b >= b0) continue;
        b = std::max(b, d[k + 6]); b = std::max(b, d[k + 7]); b = std::max(b, d[k + 8]);
        b0 = std::min(b0, std::max(b, d[k]));
        b0 = std::min(b0, std::max(b, d[k + 9]));
    }
    return -b0 - 1;
}

#if defined(__SSE2__)
// The same score with 16-bit lanes, one lane per arc start (the structure of OpenCV's SIMD cornerScore<16>): per lane k
// a = min(d[k+1..k+8]), b = max(d[k+1..k+8]); A = max_k max(min(a, d[k]), min(a, d[k+9])), B likewise with min/max swapped;
// result max(A, -B) - 1.  For a pixel that IS a corner at `threshold` this equals cornerScore9 (whose running maximum
// merely starts at the threshold).
static int cornerScore9_simd(const uint8_t* p, const int* ofs) {
    alignas(16) short d[32];
    const int v = p[0];
    for (int k = 0; k < 25; k++) d[k] = (short)(v - p[ofs[k & 15]]);
    __m128i q0 = _mm_set1_epi16(-1000), q1 = _mm_set1_epi16(1000);
    for (int k = 0; k < 16; k += 8) {
        __m128i v0 = _mm_loadu_si128((const __m128i*)(d + k + 1)), v1 = _mm_loadu_si128((const __m128i*)(d + k + 2));
        __m128i a = _mm_min_epi16(v0, v1), b = _mm_max_epi16(v0, v1);
        for (int t = 3; t <= 8; t++) {
            v0 = _mm_loadu_si128((const __m128i*)(d + k + t));
            a = _mm_min_epi16(a, v0); b = _mm_max_epi16(b, v0);
        }
        v0 = _mm_loadu_si128((const __m128i*)(d + k));
        q0 = _mm_max_epi16(q0, _mm_min_epi16(a, v0)); q1 = _mm_min_epi16(q1, _mm_max_epi16(b, v0));
        v0 = _mm_loadu_si128((const __m128i*)(d + k + 9));
        q0 = _mm_max_epi16(q0, _mm_min_epi16(a, v0)); q1 = _mm_min_epi16(q1, _mm_max_epi16(b, v0));
    }
    q0 = _mm_max_epi16(q0, _mm_sub_epi16(_mm_setzero_si128(), q1));
    q0 = _mm_max_epi16(q0, _mm_srli_si128(q0, 8));
    q0 = _mm_max_epi16(q0, _mm_srli_si128(q0, 4));
    q0 = _mm_max_epi16(q0, _mm_srli_si128(q0, 2));
    return (short)_mm_cvtsi128_si32(q0) - 1;
}
#endif

void FAST9_16(const uint8_t* img, int w, int h, int step, int threshold, bool nonmax, std::vector<KeyPoint>& out) {
    out.clear();
    if (w < 7 || h < 7) return;
    int ofs[16];
    for (int k = 0; k < 16; k++) ofs[k] = kCircle[k][0] + kCircle[k][1] * step;
    threshold = std::min(std::max(threshold, 0), 255);
    std::vector<uint8_t> buf((size_t)3 * w, 0);
    std::vector<int> cpbuf((size_t)3 * (w + 1), 0);
    for (int i = 3; i < h - 2; i++) {
        const uint8_t* ptr = img + (size_t)i * step;
        uint8_t* curr = buf.data() + (size_t)((i - 3) % 3) * w;
        int* cornerpos = cpbuf.data() + (size_t)((i - 3) % 3) * (w + 1) + 1;
        memset(curr, 0, w);
        int ncorners = 0;
        if (i < h - 3) {
            int j = 3;
#if defined(__SSE2__)
            if (!g_fast_scalar_only)
            // 16 pixels per step, the structure of OpenCV's own SIMD FAST: signed-domain compares against v +- t
            // (saturating), block-wide early out on the four compass points, then per-lane run-length counters over
            // the wrapped circle.  Decides exactly what isCorner9 decides (a run of > 8 brighter or darker pixels).
            {
                const __m128i delta = _mm_set1_epi8((char)-128), tt = _mm_set1_epi8((char)threshold), K8 = _mm_set1_epi8(8);
                for (; j < w - 18; j += 16) {
                    const uint8_t* q = ptr + j;
                    const __m128i v = _mm_loadu_si128((const __m128i*)q);
                    const __m128i v0 = _mm_xor_si128(_mm_adds_epu8(v, tt), delta), v1 = _mm_xor_si128(_mm_subs_epu8(v, tt), delta);
                    __m128i x0 = _mm_xor_si128(_mm_loadu_si128((const __m128i*)(q + ofs[0])), delta);
                    __m128i x1 = _mm_xor_si128(_mm_loadu_si128((const __m128i*)(q + ofs[4])), delta);
                    __m128i x2 = _mm_xor_si128(_mm_loadu_si128((const __m128i*)(q + ofs[8])), delta);
                    __m128i x3 = _mm_xor_si128(_mm_loadu_si128((const __m128i*)(q + ofs[12])), delta);
                    __m128i m0 = _mm_and_si128(_mm_cmpgt_epi8(x0, v0), _mm_cmpgt_epi8(x1, v0));
                    __m128i m1 = _mm_and_si128(_mm_cmpgt_epi8(v1, x0), _mm_cmpgt_epi8(v1, x1));
                    m0 = _mm_or_si128(m0, _mm_and_si128(_mm_cmpgt_epi8(x1, v0), _mm_cmpgt_epi8(x2, v0)));
                    m1 = _mm_or_si128(m1, _mm_and_si128(_mm_cmpgt_epi8(v1, x1), _mm_cmpgt_epi8(v1, x2)));
                    m0 = _mm_or_si128(m0, _mm_and_si128(_mm_cmpgt_epi8(x2, v0), _mm_cmpgt_epi8(x3, v0)));
                    m1 = _mm_or_si128(m1, _mm_and_si128(_mm_cmpgt_epi8(v1, x2), _mm_cmpgt_epi8(v1, x3)));
                    m0 = _mm_or_si128(m0, _mm_and_si128(_mm_cmpgt_epi8(x3, v0), _mm_cmpgt_epi8(x0, v0)));
                    m1 = _mm_or_si128(m1, _mm_and_si128(_mm_cmpgt_epi8(v1, x3), _mm_cmpgt_epi8(v1, x0)));
                    if (_mm_movemask_epi8(_mm_or_si128(m0, m1)) == 0) continue;      // no 9-arc can exist in these 16 pixels
                    __m128i c0 = _mm_setzero_si128(), c1 = c0, max0 = c0, max1 = c0;
                    for (int k = 0; k < 25; k++) {
                        const __m128i x = _mm_xor_si128(_mm_loadu_si128((const __m128i*)(q + ofs[k & 15])), delta);
                        m0 = _mm_cmpgt_epi8(x, v0);
                        m1 = _mm_cmpgt_epi8(v1, x);
                        c0 = _mm_and_si128(_mm_sub_epi8(c0, m0), m0);
                        c1 = _mm_and_si128(_mm_sub_epi8(c1, m1), m1);
                        max0 = _mm_max_epu8(max0, c0);
                        max1 = _mm_max_epu8(max1, c1);
                    }
                    int m = _mm_movemask_epi8(_mm_cmpgt_epi8(_mm_max_epu8(max0, max1), K8));
                    for (int l = 0; m; l++, m >>= 1)
                        if (m & 1) {
                            cornerpos[ncorners++] = j + l;
                            if (nonmax) curr[j + l] = (uint8_t)cornerScore9_simd(q + l, ofs);
                        }
                }
            }
#endif
            for (; j < w - 3; j++) {
                if (isCorner9(ptr + j, ofs, threshold)) {
                    cornerpos[ncorners++] = j;
                    if (nonmax) curr[j] = (uint8_t)cornerScore9(ptr + j, ofs, threshold);
                }
            }
        }
        cornerpos[-1] = ncorners;
        if (i == 3) continue;
        const uint8_t* prev = buf.data() + (size_t)((i - 4 + 3) % 3) * w;
        const uint8_t* pprev = buf.data() + (size_t)((i - 5 + 3) % 3) * w;
        cornerpos = cpbuf.data() + (size_t)((i - 4 + 3) % 3) * (w + 1) + 1;
        ncorners = cornerpos[-1];
        for (int k = 0; k < ncorners; k++) {
            int j = cornerpos[k];
            int score = prev[j];
            if (!nonmax || (score > prev[j + 1] && score > prev[j - 1] && score > pprev[j - 1] && score > pprev[j] &&
                            score > pprev[j + 1] && score > curr[j - 1] && score > curr[j] && score > curr[j + 1])) {
                KeyPoint kp{(float)j, (float)(i - 1), 7.f, -1.f, (float)score, 0, -1};
                out.push_back(kp);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// cv::KeyPointsFilter::retainBest -- uses the toolchain's own std::nth_element / std::partition (libstdc++),
// exactly as the reference binary would.
void retainBest(std::vector<KeyPoint>& kps, int n, int variant) {
    if (n >= 0 && kps.size() > (size_t)n) {
        if (n == 0) { kps.clear(); return; }
        auto greater = [](const KeyPoint& a, const KeyPoint& b) { return a.response > b.response; };
        int nth = (variant == RETAIN_NTH_AT_N_MINUS_1) ? n - 1 : n;
        std::nth_element(kps.begin(), kps.begin() + nth, kps.end(), greater);
        float amb = kps[n - 1].response;
        auto new_end = std::partition(kps.begin() + n, kps.end(), [amb](const KeyPoint& k) { return k.response >= amb; });
        kps.resize(new_end - kps.begin());
    }
}

}  // namespace oracle
