// TEST INFRASTRUCTURE ONLY (see oracle.hpp).  Restatements of the un-vendored OpenCV primitives the
// reference's ORBextractor calls (src/ORBextractor.cc:943 resize, :982/:990 copyMakeBorder, :621/:627 FAST,
// :734/:750 KeyPointsFilter::retainBest, :842 GaussianBlur, :157 fastAtan2, cvRound everywhere).
// OpenCV is a third-party dependency absent from /root/reference (pinned 2.4.13 or 3.4.1 by
// CMakeLists.txt:42-49 / build_dep.sh:37-45); what is restated is its published algorithm, cross-checked
// against cv2 4.13 in tests/test_oracle_cv.py.
#include "oracle.hpp"
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

template <class Fetch>
static void blurCore(Fetch fetch, int w, int h, uint8_t* dst, int dstep, int variant) {
    const int* k = kGauss[variant];
    std::vector<int> rows((size_t)(h + 6) * w);
    for (int y = -3; y < h + 3; y++)
        for (int x = 0; x < w; x++) {
            int s = 0;
            for (int i = 0; i < 7; i++) s += k[i] * fetch(x + i - 3, y);
            rows[(size_t)(y + 3) * w + x] = s;
        }
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            int s = 0;
            for (int i = 0; i < 7; i++) s += k[i] * rows[(size_t)(y + i) * w + x];
            int v = (s + 32768) >> 16;
            dst[(size_t)y * dstep + x] = (uint8_t)std::min(std::max(v, 0), 255);
        }
}

void gaussianBlur7x7(const uint8_t* src, int w, int h, int sstep, uint8_t* dst, int dstep, int variant) {
    std::vector<uint8_t> tmp((size_t)w * h);
    blurCore([&](int x, int y) { return (int)src[(size_t)reflect101(y, h) * sstep + reflect101(x, w)]; }, w, h,
             tmp.data(), w, variant);
    for (int y = 0; y < h; y++) memcpy(dst + (size_t)y * dstep, tmp.data() + (size_t)y * w, w);
}

void gaussianBlur7x7InPlace(uint8_t* roi, int w, int h, int step, int variant) {
    // The ROI sits inside a bordered plane, so (without BORDER_ISOLATED) OpenCV reads the ring pixels, which
    // hold the reflect-101 copy of the unblurred level (src/ORBextractor.cc:982-983) => same as isolated blur.
    std::vector<uint8_t> tmp((size_t)w * h);
    blurCore([&](int x, int y) { return (int)roi[(ptrdiff_t)y * step + x]; }, w, h, tmp.data(), w, variant);
    for (int y = 0; y < h; y++) memcpy(roi + (size_t)y * step, tmp.data() + (size_t)y * w, w);
}

// ---------------------------------------------------------------------------------------------------------------
// cv::FAST TYPE_9_16 with non-max suppression.
static const int kCircle[16][2] = {{0, 3},  {1, 3},   {2, 2},   {3, 1},   {3, 0},  {3, -1}, {2, -2}, {1, -3},
                                   {0, -3}, {-1, -3}, {-2, -2}, {-3, -1}, {-3, 0}, {-3, 1}, {-2, 2}, {-1, 3}};

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
            for (int j = 3; j < w - 3; j++) {
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
