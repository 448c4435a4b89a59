// B200-native Hamming matching: batch-first CUDA implementation of the per-frame tracking overloads of
// ORB_SLAM::ORBmatcher (reference src/ORBmatcher.cc) and of the Frame grid they search (src/Frame.cc).
//
//   K7  k_grid_build   one CTA per frame: PosInGrid (Frame.cc:367-377) -> CSR over the 64x48 grid, cell index =
//                      ix*48+iy (so one grid column range [iy0,iy1] is ONE contiguous CSR segment), entries in
//                      insertion (= keypoint index) order, 16 B records {x, y, octave, index}.
//   K8a k_candidates   one warp per query (map point / source keypoint): GetFeaturesInArea (Frame.cc:300-365) in
//                      the reference's order (ix outer, iy inner, insertion order), level + window filter, 256-bit
//                      Hamming distance (LOP3 xor + POPC on the integer pipe, 2 x 128-bit loads per descriptor);
//                      writes packed (dist | octave | index) candidate lists.
//   K8b k_resolve      one warp per problem: the reference's greedy loops are order dependent (claimed keypoints are
//                      skipped, ORBmatcher.cc:426; SearchForInitialization steals, :1211-1237).  32 consecutive
//                      queries are evaluated speculatively, one per lane, against the current claim state; the longest
//                      prefix of lanes whose candidate lists are untouched by an earlier lane's claim commits, the
//                      rest re-evaluate.  The result is identical to the sequential loop.  Rotation histogram +
//                      ComputeThreeMaxima (:2338-2379) run in the same kernel.
//   K10 k_allpairs     brute-force top-2 of n_q queries against a descriptor database shard.
// All integer results are exact; the float window geometry uses the same un-fused float operations as the reference
// (this TU is compiled with --fmad=false).
#include <climits>
#include <cstring>
#include <vector>
#include "common.cuh"
#include "host_util.h"

namespace gforb {

static const int GRID_COLS = 64, GRID_ROWS = 48, GRID_CELLS = GRID_COLS * GRID_ROWS;
static const int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;
enum { M_INIT = 0, M_WINDOW = 1, M_MAP = 2, M_LAST = 3, M_AREA = 4, M_PROJ2 = 5, M_BOW = 6, M_KF = 7,
       M_BOW2 = 8,      // SearchByBoW(KeyFrame, KeyFrame)
       M_TRI = 9,       // SearchForTriangulation
       M_PKF = 10 };    // map points projected into a keyframe: Fuse (both), SearchByProjection(KF, Scw, ..), SearchBySim3
#define GF_IS_BOW(M) ((M) == M_BOW || (M) == M_BOW2 || (M) == M_TRI)

struct FrameDev {
    const GforbKeyPoint* kps;
    const uint8_t* desc;
    const int32_t* counts;
    int stride;
};
struct GridGeom {
    float minX, minY, invW, invH;
    int iminX, imaxX, iminY, imaxY;
};
struct GridDev {
    const int* cellStart;     // batch x (GRID_CELLS + 1)
    const int4* entries;      // batch x maxKp : {x bits, y bits, octave, keypoint index}
    int maxKp;
};

// ------------------------------------------------------------------------------------------------ K7 grid
__global__ void __launch_bounds__(256) k_grid_build(FrameDev F, GridGeom G, int* __restrict__ cellStart,
                                                    int4* __restrict__ entries, int maxKp) {
    __shared__ int s_start[GRID_CELLS + 1];
    __shared__ int s_fill[GRID_CELLS];
    __shared__ int s_part[8];
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = min(F.counts[b], maxKp);
    const GforbKeyPoint* kp = F.kps + (size_t)b * F.stride;
    for (int c = tid; c < GRID_CELLS; c += 256) s_fill[c] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += 256) {
        // PosInGrid: posX = round((kp.pt.x - mnMinX) * mfGridElementWidthInv) (round half away from zero)
        int px = (int)roundf((kp[i].x - G.minX) * G.invW);
        int py = (int)roundf((kp[i].y - G.minY) * G.invH);
        if (px >= 0 && px < GRID_COLS && py >= 0 && py < GRID_ROWS) atomicAdd(&s_fill[px * GRID_ROWS + py], 1);
    }
    __syncthreads();
    // exclusive scan over 3072 cells: 12 consecutive cells per thread
    int loc[12], sum = 0;
#pragma unroll
    for (int k = 0; k < 12; k++) { loc[k] = sum; sum += s_fill[tid * 12 + k]; }
    int inc = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { int t = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += t; }
    if (lane == 31) s_part[warp] = inc;
    __syncthreads();
    int base = inc - sum;
    for (int w = 0; w < warp; w++) base += s_part[w];
#pragma unroll
    for (int k = 0; k < 12; k++) s_start[tid * 12 + k] = base + loc[k];
    if (tid == 255) s_start[GRID_CELLS] = base + sum;
    __syncthreads();
    int* cs = cellStart + (size_t)b * (GRID_CELLS + 1);
    for (int c = tid; c <= GRID_CELLS; c += 256) cs[c] = s_start[c];
    for (int c = tid; c < GRID_CELLS; c += 256) s_fill[c] = s_start[c];
    __syncthreads();
    int4* ent = entries + (size_t)b * maxKp;
    for (int i = tid; i < n; i += 256) {
        const float x = kp[i].x, y = kp[i].y;
        int px = (int)roundf((x - G.minX) * G.invW);
        int py = (int)roundf((y - G.minY) * G.invH);
        if (px >= 0 && px < GRID_COLS && py >= 0 && py < GRID_ROWS) {
            int pos = atomicAdd(&s_fill[px * GRID_ROWS + py], 1);
            ent[pos] = make_int4(__float_as_int(x), __float_as_int(y), kp[i].octave, i);
        }
    }
    __syncthreads();
    // restore insertion order inside each cell (the scatter above is unordered): insertion sort by keypoint index
    for (int c = tid; c < GRID_CELLS; c += 256) {
        const int beg = s_start[c], end = s_start[c + 1];
        for (int i = beg + 1; i < end; i++) {
            int4 v = ent[i];
            int j = i - 1;
            while (j >= beg && ent[j].w > v.w) { ent[j + 1] = ent[j]; j--; }
            ent[j + 1] = v;
        }
    }
}

// ------------------------------------------------------------------------------------------------ K8a candidates
struct QParams {
    FrameDev src;                 // INIT / WINDOW / LAST: F1 / LastFrame
    const float* prevMatched;     // INIT: batch x src.stride x 2
    const int32_t* srcMp;         // WINDOW / LAST: batch x src.stride, >= 0 valid
    const uint8_t* srcOutlier;    // LAST (may be null)
    const float* srcWorld;        // LAST: batch x src.stride x 3
    const float* Tcw;             // LAST: batch x 16
    float fx, fy, cx, cy;         // LAST
    float window;                 // INIT / WINDOW
    int minScaleLevel, maxScaleLevel;   // WINDOW
    float th;                     // MAP / LAST
    float thGood;                 // MAP / LAST, SearchByProjection_GoodFeature: window factor of the good features
    int goodMode;                 // 1: good features (MAP: mpFlags bit2, LAST: srcGood) use thGood, the rest th (= th_rest)
    const uint8_t* srcGood;       // LAST + goodMode: batch x src.stride
    const int32_t* tgtMp0;        // PROJ2: F2.mvpMapPoints (batch x T.stride) = spMapPointsAlreadyFound
    const int32_t *srcNode, *tgtNode;   // BOW: DBoW2 FeatureVector node of every KF / F feature (-1 = none)
    int32_t* order;               // BOW out: batch x maxQ, the queries in (node, index) order = the reference's walk
    const float* srcMinDist;      // KF: MapPoint::GetMinDistanceInvariance per KF keypoint
    const uint8_t* qDesc;         // KF: MapPoint::GetDescriptor per KF keypoint (batch x src.stride x 32)
    const float *mpU, *mpV, *mpCos;     // MAP
    const int32_t* mpLevel;
    const uint8_t *mpFlags, *mpDesc;
    const int32_t* mpCounts;
    int mpStride;
    const int32_t* tgtMp;         // BOW2 / TRI: pKF2->GetMapPointMatches() as ids (batch x T.stride)
    const float* F12;             // TRI: batch x 9
    const float *pWorld, *pNormal, *pMinD, *pMaxD;   // PKF: batch x mpStride (x 3); pNormal == nullptr: no viewing-angle gate
    const uint8_t* pSkip;         // PKF
    int pkfMode;                  // PKF: 0 pose = Tcw, 1 pose = Scw (decomposed), 2 / 3 SearchBySim3 direction 1 -> 2 / 2 -> 1
    const float *s12, *R12, *t12; // PKF, SearchBySim3: batch, batch x 9, batch x 3
    const float *ax, *ay, *ar;    // AREA
    const int32_t *aMin, *aMax;
    int aCount, aFrame;
    float scale[MAX_LEVELS];
    int nlevels;
};

// Warp-cooperative GetFeaturesInArea + distance.  Returns the number of candidates; with WRITE, stores
// (dist << 20 | octave << 16 | index) in the reference's order.
template <bool WRITE>
__device__ __forceinline__ int enum_candidates(const int* __restrict__ cs, const int4* __restrict__ ent,
                                               const uint8_t* __restrict__ tdesc, const GridGeom G, float x, float y,
                                               float r, int minL, int maxL, const uint32_t (&qd)[8],
                                               uint32_t* __restrict__ out, int lane) {
    const unsigned full = 0xffffffffu;
    int cx0 = (int)floorf((x - G.minX - r) * G.invW);
    cx0 = max(0, cx0);
    if (cx0 >= GRID_COLS) return 0;
    int cx1 = (int)ceilf((x - G.minX + r) * G.invW);
    cx1 = min(GRID_COLS - 1, cx1);
    if (cx1 < 0) return 0;
    int cy0 = (int)floorf((y - G.minY - r) * G.invH);
    cy0 = max(0, cy0);
    if (cy0 >= GRID_ROWS) return 0;
    int cy1 = (int)ceilf((y - G.minY + r) * G.invH);
    cy1 = min(GRID_ROWS - 1, cy1);
    if (cy1 < 0) return 0;
    if (cx1 < cx0 || cy1 < cy0) return 0;
    const bool checkLevels = !(minL == -1 && maxL == -1);
    const bool same = checkLevels && (minL == maxL);
    int total = 0;
    for (int c0 = cx0; c0 <= cx1; c0 += 32) {
        const int ix = c0 + lane;
        int beg = 0, len = 0;
        if (ix <= cx1) {
            beg = cs[ix * GRID_ROWS + cy0];
            len = cs[ix * GRID_ROWS + cy1 + 1] - beg;
        }
        int inc = len;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { int t = __shfl_up_sync(full, inc, d); if (lane >= d) inc += t; }
        const int T = __shfl_sync(full, inc, 31);
        const int exc = inc - len;
        for (int base = 0; base < T; base += 32) {
            const int flat = base + lane;
            int lo = 0;
#pragma unroll
            for (int step = 16; step >= 1; step >>= 1) {
                int v = __shfl_sync(full, inc, lo + step - 1);
                if (v <= flat) lo += step;
            }
            const int sbeg = __shfl_sync(full, beg, lo), sexc = __shfl_sync(full, exc, lo);
            bool valid = false;
            int4 e = make_int4(0, 0, 0, 0);
            if (flat < T) {
                e = ent[sbeg + flat - sexc];
                const float ex = __int_as_float(e.x), ey = __int_as_float(e.y);
                valid = true;
                if (checkLevels && !same) { if (e.z < minL || e.z > maxL) valid = false; }
                else if (same) { if (e.z != minL) valid = false; }
                if (valid && (fabsf(ex - x) > r || fabsf(ey - y) > r)) valid = false;
            }
            const unsigned m = __ballot_sync(full, valid);
            if (WRITE && valid) {
                const uint4* dp = reinterpret_cast<const uint4*>(tdesc + (size_t)e.w * 32);
                const uint4 d0 = __ldg(dp), d1 = __ldg(dp + 1);
                const int dist = __popc(qd[0] ^ d0.x) + __popc(qd[1] ^ d0.y) + __popc(qd[2] ^ d0.z) + __popc(qd[3] ^ d0.w) +
                                 __popc(qd[4] ^ d1.x) + __popc(qd[5] ^ d1.y) + __popc(qd[6] ^ d1.z) + __popc(qd[7] ^ d1.w);
                out[total + __popc(m & ((1u << lane) - 1u))] = ((uint32_t)dist << 20) | ((uint32_t)(e.z & 15) << 16) | (uint32_t)e.w;
            }
            total += __popc(m);
        }
    }
    return total;
}

// SearchByBoW / SearchForTriangulation: the features of the target frame that sit in the query's vocabulary node, in index
// order (vIndicesF / f2it->second).  FILT 0: every feature (SearchByBoW(KF, F): claims are the resolver's business);
// FILT 1: features that hold a map point (SearchByBoW(KF1, KF2), :1341-1347); FILT 2: SearchForTriangulation -- features without
// a map point whose distance is <= TH_LOW (:1485-1493); bit 31 of the entry = CheckDistEpipolarLine(kp1, kp2, F12) (:705-722).
template <bool WRITE, int FILT>
__device__ __forceinline__ int enum_node(const int32_t* __restrict__ tnode, const GforbKeyPoint* __restrict__ tkps,
                                         const uint8_t* __restrict__ tdesc, int N2, int node, const uint32_t (&qd)[8],
                                         uint32_t* __restrict__ out, int lane, const int32_t* __restrict__ tmp = nullptr,
                                         float k1x = 0.f, float k1y = 0.f, const float* __restrict__ F12 = nullptr,
                                         const float* scale = nullptr) {
    const unsigned full = 0xffffffffu;
    int total = 0;
    for (int t0 = 0; t0 < N2; t0 += 32) {
        const int t = t0 + lane;
        bool valid = t < N2 && tnode[t] == node;
        if (FILT == 1 && valid) valid = tmp[t] >= 0;
        if (FILT == 2 && valid) valid = tmp[t] < 0;
        int dist = 0;
        if ((WRITE || FILT == 2) && valid) {
            const uint4* dp = reinterpret_cast<const uint4*>(tdesc + (size_t)t * 32);
            const uint4 d0 = __ldg(dp), d1 = __ldg(dp + 1);
            dist = __popc(qd[0] ^ d0.x) + __popc(qd[1] ^ d0.y) + __popc(qd[2] ^ d0.z) + __popc(qd[3] ^ d0.w) +
                   __popc(qd[4] ^ d1.x) + __popc(qd[5] ^ d1.y) + __popc(qd[6] ^ d1.z) + __popc(qd[7] ^ d1.w);
            if (FILT == 2 && dist > TH_LOW) valid = false;
        }
        const unsigned m = __ballot_sync(full, valid);
        if (WRITE && valid) {
            const int oct = tkps[t].octave;
            uint32_t e = ((uint32_t)dist << 20) | ((uint32_t)(oct & 15) << 16) | (uint32_t)t;
            if (FILT == 2) {
                const float k2x = tkps[t].x, k2y = tkps[t].y;
                const float a = __fadd_rn(__fadd_rn(__fmul_rn(k1x, F12[0]), __fmul_rn(k1y, F12[3])), F12[6]);
                const float b = __fadd_rn(__fadd_rn(__fmul_rn(k1x, F12[1]), __fmul_rn(k1y, F12[4])), F12[7]);
                const float c = __fadd_rn(__fadd_rn(__fmul_rn(k1x, F12[2]), __fmul_rn(k1y, F12[5])), F12[8]);
                const float num = __fadd_rn(__fadd_rn(__fmul_rn(a, k2x), __fmul_rn(b, k2y)), c);
                const float den = __fadd_rn(__fmul_rn(a, a), __fmul_rn(b, b));
                bool ok = false;
                if (den != 0.0f) {
                    const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
                    const float sc = scale[min(max(oct, 0), MAX_LEVELS - 1)];
                    ok = (double)dsqr < 3.84 * (double)__fmul_rn(sc, sc);              // mvLevelSigma2[i] = scale[i]^2 (src/Frame.cc:110)
                }
                if (ok) e |= 0x80000000u;
            }
            out[total + __popc(m & ((1u << lane) - 1u))] = e;
        }
        total += __popc(m);
    }
    return total;
}

template <int MODE>
__global__ void __launch_bounds__(128) k_candidates(const QParams P, const FrameDev T, const GridDev Gd, const GridGeom G,
                                                    uint32_t* __restrict__ pool, int poolCap, int* __restrict__ poolCount,
                                                    int2* __restrict__ qMeta, int maxQ, int* __restrict__ err) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int q = blockIdx.x * 4 + warp, b = blockIdx.y;
    const int tb = (MODE == M_AREA) ? P.aFrame : b;      // target frame
    int Q;
    if (MODE == M_MAP || MODE == M_PKF) Q = min(P.mpCounts[b], P.mpStride);
    else if (MODE == M_AREA) Q = P.aCount;
    else Q = P.src.counts[b];
    Q = min(Q, maxQ);
    if (q >= Q) return;
    float x = 0.f, y = 0.f, r = 0.f;
    int minL = -1, maxL = -1;
    bool active = true;
    const uint8_t* qdp = nullptr;
    if (MODE == M_INIT) {
        const GforbKeyPoint* kp = P.src.kps + (size_t)b * P.src.stride + q;
        const int level1 = kp->octave;
        if (level1 > 0) active = false;                   // src/ORBmatcher.cc:1188-1190
        x = P.prevMatched[((size_t)b * P.src.stride + q) * 2];
        y = P.prevMatched[((size_t)b * P.src.stride + q) * 2 + 1];
        r = P.window; minL = maxL = level1;
        qdp = P.src.desc + ((size_t)b * P.src.stride + q) * 32;
    } else if (MODE == M_WINDOW) {
        const GforbKeyPoint* kp = P.src.kps + (size_t)b * P.src.stride + q;
        const int level1 = kp->octave;
        if (P.srcMp[(size_t)b * P.src.stride + q] < 0) active = false;
        if (P.minScaleLevel > 0 && level1 < P.minScaleLevel) active = false;
        if (P.maxScaleLevel < INT_MAX && level1 > P.maxScaleLevel) active = false;
        x = kp->x; y = kp->y; r = P.window; minL = maxL = level1;
        qdp = P.src.desc + ((size_t)b * P.src.stride + q) * 32;
    } else if (MODE == M_MAP) {
        const size_t i = (size_t)b * P.mpStride + q;
        const uint8_t fl = P.mpFlags[i];
        if (!(fl & 1) || (fl & 2)) active = false;        // !mbTrackInView || isBad()
        int lvl = P.mpLevel[i];
        lvl = min(max(lvl, 0), P.nlevels - 1);
        float rr = (P.mpCos[i] > 0.998) ? 2.5f : 4.0f;    // RadiusByViewingCos (float vs double literal)
        if (P.goodMode && (fl & 4)) rr *= P.thGood;       // src/ORBmatcher.cc:73-80 (goodAtFrameId == F.mnId)
        else if (P.th != 1.0f) rr *= P.th;
        x = P.mpU[i]; y = P.mpV[i];
        r = rr * P.scale[lvl];
        minL = P.mpLevel[i] - 1; maxL = P.mpLevel[i];
        qdp = P.mpDesc + i * 32;
    } else if (MODE == M_LAST) {
        const size_t i = (size_t)b * P.src.stride + q;
        if (P.srcMp[i] < 0) active = false;
        if (P.srcOutlier && P.srcOutlier[i]) active = false;
        if (active) {
            const float* X = P.srcWorld + i * 3;
            const float* Tm = P.Tcw + (size_t)b * 16;
            float xc3[3];                                   // Rcw*x3Dw+tcw (common.cuh: cv_gemm_row)
#pragma unroll
            for (int rw = 0; rw < 3; rw++) xc3[rw] = cv_gemm_row(Tm[rw * 4], Tm[rw * 4 + 1], Tm[rw * 4 + 2], X[0], X[1], X[2], Tm[rw * 4 + 3]);
            const float invzc = (float)(1.0 / (double)xc3[2]);
            const float u = P.fx * xc3[0] * invzc + P.cx;
            const float v = P.fy * xc3[1] * invzc + P.cy;
            if (u < (float)G.iminX || u > (float)G.imaxX) active = false;
            if (v < (float)G.iminY || v > (float)G.imaxY) active = false;
            const int oct = P.src.kps[i].octave;
            x = u; y = v;
            r = ((P.goodMode && P.srcGood[i]) ? P.thGood : P.th) * P.scale[min(max(oct, 0), P.nlevels - 1)];   // :46-53
            minL = oct - 1; maxL = oct + 1;
        }
        qdp = P.src.desc + i * 32;
    } else if (MODE == M_PROJ2) {
        // SearchByProjection(F1, F2, windowSize, ...) (src/ORBmatcher.cc:1089-1168): F1 map points not yet in F2, projected
        // with F2's pose; no bounds / depth test; fixed pixel window, same level
        const size_t i = (size_t)b * P.src.stride + q;
        const int id = P.srcMp[i];
        if (id < 0) active = false;
        if (active) {                                     // spMapPointsAlreadyFound.count(pMP1) (:1092, :1106)
            const int n2 = min(T.counts[b], T.stride);
            bool found = false;
            for (int t = lane; t < n2; t += 32) found |= P.tgtMp0[(size_t)b * T.stride + t] == id;
            if (__any_sync(0xffffffffu, found)) active = false;
        }
        if (active) {
            const float* X = P.srcWorld + i * 3;
            const float* Tm = P.Tcw + (size_t)b * 16;
            float xc3[3];                                   // Rcw*x3Dw+tcw (common.cuh: cv_gemm_row)
#pragma unroll
            for (int rw = 0; rw < 3; rw++) xc3[rw] = cv_gemm_row(Tm[rw * 4], Tm[rw * 4 + 1], Tm[rw * 4 + 2], X[0], X[1], X[2], Tm[rw * 4 + 3]);
            const float invzc = (float)(1.0 / (double)xc3[2]);
            x = P.fx * xc3[0] * invzc + P.cx;
            y = P.fy * xc3[1] * invzc + P.cy;
            r = P.window;
            minL = maxL = P.src.kps[i].octave;
            if (!(x == x) || !(y == y) || fabsf(x) > 1e30f || fabsf(y) > 1e30f) active = false;   // z = 0: no cell range exists
        }
        qdp = P.src.desc + i * 32;
    } else if (MODE == M_KF) {
        // SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (src/ORBmatcher.cc:2204-2336): the keyframe's map
        // points projected with the current pose, level predicted from the distance to the camera centre
        const size_t i = (size_t)b * P.src.stride + q;
        if (P.srcMp[i] < 0) active = false;
        if (P.srcOutlier && P.srcOutlier[i]) active = false;      // isBad() || sAlreadyFound.count(pMP)
        if (active) {
            const float* X = P.srcWorld + i * 3;
            const float* Tm = P.Tcw + (size_t)b * 16;
            float xc3[3], Ow[3];
#pragma unroll
            for (int rw = 0; rw < 3; rw++) {
                xc3[rw] = cv_gemm_row(Tm[rw * 4], Tm[rw * 4 + 1], Tm[rw * 4 + 2], X[0], X[1], X[2], Tm[rw * 4 + 3]);
                Ow[rw] = cv_neg_rt_t_row(Tm[rw], Tm[4 + rw], Tm[8 + rw], Tm[3], Tm[7], Tm[11]);    // Ow = -Rcw.t()*tcw (:2210)
            }
            const float invzc = (float)(1.0 / (double)xc3[2]);
            const float u = P.fx * xc3[0] * invzc + P.cx;
            const float v = P.fy * xc3[1] * invzc + P.cy;
            if (u < (float)G.iminX || u > (float)G.imaxX) active = false;
            if (v < (float)G.iminY || v > (float)G.imaxY) active = false;
            const float PO0 = X[0] - Ow[0], PO1 = X[1] - Ow[1], PO2 = X[2] - Ow[2];
            const double n2 = (double)PO0 * (double)PO0 + (double)PO1 * (double)PO1 + (double)PO2 * (double)PO2;
            const float dist3D = (float)sqrt(n2);
            const float ratio = dist3D / P.srcMinDist[i];
            int lvl = 0;                                          // std::lower_bound(mvScaleFactors, ratio)
            while (lvl < P.nlevels && P.scale[lvl] < ratio) lvl++;
            if (lvl >= P.nlevels) lvl = P.nlevels - 1;
            x = u; y = v;
            r = P.th * P.scale[lvl];
            minL = lvl - 1; maxL = lvl + 1;
        }
        qdp = P.qDesc + i * 32;
    } else if (MODE == M_BOW || MODE == M_BOW2) {
        const size_t i = (size_t)b * P.src.stride + q;
        if (P.srcMp[i] < 0 || P.srcNode[i] < 0) active = false;
        qdp = P.src.desc + i * 32;
    } else if (MODE == M_TRI) {
        const size_t i = (size_t)b * P.src.stride + q;
        if (P.srcMp[i] >= 0 || P.srcNode[i] < 0) active = false;     // "if there is already a MapPoint skip" (:1468-1470)
        x = P.src.kps[i].x; y = P.src.kps[i].y;                       // kp1 of the epipolar test
        qdp = P.src.desc + i * 32;
    } else if (MODE == M_PKF) {
        // the projection block of Fuse (:1617-1659, :1743-1788), SearchByProjection(pKF, Scw, ..) (:889-934) and one direction
        // of SearchBySim3 (:1903-1939, :1986-2024); every lane computes it (a few dozen operations)
        const size_t i = (size_t)b * P.mpStride + q;
        if (P.pSkip && P.pSkip[i]) active = false;
        if (active) {
            const float* S = P.Tcw + (size_t)b * 16;
            float R[9], t[3];
            if (P.pkfMode == 1) {
                // scw = sqrt(sRcw.row(0).dot(sRcw.row(0))) (Mat::dot: double products), Rcw = sRcw/scw, tcw = Scw.col(3)/scw:
                // the MatExpr scale converts with alpha = 1./scw -> every entry times (float)alpha
                const double d = (double)S[0] * (double)S[0] + (double)S[1] * (double)S[1] + (double)S[2] * (double)S[2];
                const float scw = (float)sqrt(d);
                const float al = (float)(1.0 / (double)scw);
#pragma unroll
                for (int rw = 0; rw < 3; rw++) {
#pragma unroll
                    for (int c = 0; c < 3; c++) R[rw * 3 + c] = __fadd_rn(__fmul_rn(S[rw * 4 + c], al), 0.0f);
                    t[rw] = __fadd_rn(__fmul_rn(S[rw * 4 + 3], al), 0.0f);
                }
            } else {
#pragma unroll
                for (int rw = 0; rw < 3; rw++) { R[rw * 3] = S[rw * 4]; R[rw * 3 + 1] = S[rw * 4 + 1]; R[rw * 3 + 2] = S[rw * 4 + 2]; t[rw] = S[rw * 4 + 3]; }
            }
            const float* X = P.pWorld + i * 3;
            float pc[3];
#pragma unroll
            for (int rw = 0; rw < 3; rw++) pc[rw] = cv_gemm_row(R[rw * 3], R[rw * 3 + 1], R[rw * 3 + 2], X[0], X[1], X[2], t[rw]);
            float dist3D;
            if (P.pkfMode >= 2) {
                // SearchBySim3: sR12 = s12*R12; sR21 = (1.0/s12)*R12.t(); t21 = -sR21*t12 (:1858-1860), then sR*p3Dc + t
                const float s12 = P.s12[b];
                const float* R12 = P.R12 + (size_t)b * 9;
                const float* t12 = P.t12 + (size_t)b * 3;
                float Qm[9], tq[3];
                if (P.pkfMode == 3) {
                    const float al = (float)(double)s12;
#pragma unroll
                    for (int k = 0; k < 9; k++) Qm[k] = __fadd_rn(__fmul_rn(R12[k], al), 0.0f);
                    tq[0] = t12[0]; tq[1] = t12[1]; tq[2] = t12[2];
                } else {
                    const float al = (float)(1.0 / (double)s12);
#pragma unroll
                    for (int rw = 0; rw < 3; rw++)
#pragma unroll
                        for (int c = 0; c < 3; c++) Qm[rw * 3 + c] = __fadd_rn(__fmul_rn(R12[c * 3 + rw], al), 0.0f);
#pragma unroll
                    for (int rw = 0; rw < 3; rw++) tq[rw] = cv_neg_rt_t_row(Qm[rw * 3], Qm[rw * 3 + 1], Qm[rw * 3 + 2], t12[0], t12[1], t12[2]);
                }
                float pd[3];
#pragma unroll
                for (int rw = 0; rw < 3; rw++) pd[rw] = cv_gemm_row(Qm[rw * 3], Qm[rw * 3 + 1], Qm[rw * 3 + 2], pc[0], pc[1], pc[2], tq[rw]);
                pc[0] = pd[0]; pc[1] = pd[1]; pc[2] = pd[2];
                dist3D = (float)sqrt((double)pc[0] * (double)pc[0] + (double)pc[1] * (double)pc[1] + (double)pc[2] * (double)pc[2]);   // cv::norm(p3Dc2)
            } else {
                float PO[3];
#pragma unroll
                for (int rw = 0; rw < 3; rw++) PO[rw] = X[rw] - cv_neg_rt_t_row(R[rw], R[3 + rw], R[6 + rw], t[0], t[1], t[2]);      // p3Dw - Ow
                dist3D = (float)sqrt((double)PO[0] * (double)PO[0] + (double)PO[1] * (double)PO[1] + (double)PO[2] * (double)PO[2]);
                if (P.pNormal) {                           // PO.dot(Pn) < 0.5*dist3D: viewing angle above 60 degrees
                    const float* Pn = P.pNormal + i * 3;
                    const double dot = (double)PO[0] * (double)Pn[0] + (double)PO[1] * (double)Pn[1] + (double)PO[2] * (double)Pn[2];
                    if (dot < 0.5 * (double)dist3D) active = false;
                }
            }
            if (pc[2] < 0.0f) active = false;              // depth must be positive
            const float invz = (float)(1.0 / (double)pc[2]);
            const float xn = __fmul_rn(pc[0], invz), yn = __fmul_rn(pc[1], invz);
            const float u = __fadd_rn(__fmul_rn(P.fx, xn), P.cx), v = __fadd_rn(__fmul_rn(P.fy, yn), P.cy);
            // KeyFrame::IsInImage (src/KeyFrame.cc:654-657); a NaN projection fails it
            if (!(u >= (float)G.iminX && u < (float)G.imaxX && v >= (float)G.iminY && v < (float)G.imaxY)) active = false;
            const float minD = P.pMinD[i];
            if (dist3D < minD || dist3D > P.pMaxD[i]) active = false;
            const float ratio = __fdiv_rn(dist3D, minD);
            int lvl = 0;                                   // std::lower_bound(vfScaleFactors, ratio), capped at nMaxLevel
            while (lvl < P.nlevels && P.scale[lvl] < ratio) lvl++;
            if (lvl >= P.nlevels) lvl = P.nlevels - 1;
            x = u; y = v;
            r = __fmul_rn(P.th, P.scale[lvl]);
            minL = lvl - 1; maxL = lvl;
        }
        qdp = P.mpDesc + i * 32;
    } else {
        x = P.ax[q]; y = P.ay[q]; r = P.ar[q]; minL = P.aMin[q]; maxL = P.aMax[q];
    }
    if (GF_IS_BOW(MODE)) {
        // position of this query in the reference's walk: FeatureVector nodes ascending, features of a node in index order
        const size_t so = (size_t)b * P.src.stride;
        const int myNode = P.srcNode[so + q];
        int rank = 0;
        for (int t = lane; t < Q; t += 32) {
            const int nd = P.srcNode[so + t];
            rank += (nd < myNode || (nd == myNode && t < q)) ? 1 : 0;
        }
        rank = __reduce_add_sync(0xffffffffu, rank);
        if (lane == 0) P.order[(size_t)b * maxQ + rank] = q;
    }
    int2 meta = make_int2(0, 0);
    if (active) {
        uint32_t qd[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (qdp) {
            const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(qdp)), a1 = __ldg(reinterpret_cast<const uint4*>(qdp) + 1);
            qd[0] = a0.x; qd[1] = a0.y; qd[2] = a0.z; qd[3] = a0.w; qd[4] = a1.x; qd[5] = a1.y; qd[6] = a1.z; qd[7] = a1.w;
        }
        const int* cs = Gd.cellStart + (size_t)tb * (GRID_CELLS + 1);
        const int4* ent = Gd.entries + (size_t)tb * Gd.maxKp;
        const uint8_t* tdesc = T.desc + (size_t)tb * T.stride * 32;
        const int N2 = min(T.counts[tb], T.stride);
        const int32_t* tnode = GF_IS_BOW(MODE) ? P.tgtNode + (size_t)tb * T.stride : nullptr;
        const int myNode = GF_IS_BOW(MODE) ? P.srcNode[(size_t)b * P.src.stride + q] : 0;
        const int32_t* tmp = (MODE == M_BOW2 || MODE == M_TRI) ? P.tgtMp + (size_t)tb * T.stride : nullptr;
        const float* F12 = (MODE == M_TRI) ? P.F12 + (size_t)b * 9 : nullptr;
        constexpr int FILT = MODE == M_BOW2 ? 1 : (MODE == M_TRI ? 2 : 0);
        int cnt = GF_IS_BOW(MODE) ? enum_node<false, FILT>(tnode, T.kps + (size_t)tb * T.stride, tdesc, N2, myNode, qd, nullptr, lane, tmp, x, y, F12, P.scale)
                                  : enum_candidates<false>(cs, ent, tdesc, G, x, y, r, minL, maxL, qd, nullptr, lane);
        if (cnt > 0) {
            int off = 0;
            if (lane == 0) off = atomicAdd(&poolCount[b], cnt);
            off = __shfl_sync(0xffffffffu, off, 0);
            if (off + cnt > poolCap) {
                if (lane == 0) atomicExch(err, GFORB_ERR_OVERFLOW);
            } else {
                if (GF_IS_BOW(MODE)) enum_node<true, FILT>(tnode, T.kps + (size_t)tb * T.stride, tdesc, N2, myNode, qd, pool + (size_t)b * poolCap + off, lane, tmp, x, y, F12, P.scale);
                else enum_candidates<true>(cs, ent, tdesc, G, x, y, r, minL, maxL, qd, pool + (size_t)b * poolCap + off, lane);
                meta = make_int2(off, cnt);
            }
        }
    }
    if (lane == 0) qMeta[(size_t)b * maxQ + q] = meta;
}

// ------------------------------------------------------------------------------------------------ K8b resolve
struct RParams {
    FrameDev src, tgt;
    const uint32_t* pool;
    int poolCap;
    const int2* qMeta;
    int maxQ, maxKp;
    const int32_t* qCounts;
    const int32_t* srcMp;
    float nnratio;
    int checkOri;
    int32_t* tgtMp;       // MAP / LAST: in/out batch x tgt.stride;  WINDOW: matches2 (out)
    int32_t* tgtScore;    // MAP / LAST
    int32_t* matches12;   // INIT: batch x src.stride
    float* prevMatched;   // INIT
    int32_t* nmatches;
    int* rotBinQ;         // INIT scratch: batch x maxQ
    const int32_t* order; // BOW / BOW2 / TRI: batch x maxQ, query index per position of the walk
    const int32_t* tgtMpIds;   // BOW2: pKF2->GetMapPointMatches() as ids (the value written to matches12)
    int thHigh;           // LAST / KF: acceptance threshold (TH_HIGH; ORBdist for the keyframe overload)
    int laneParallel;     // INIT: queries one after the other, the candidates of a query over the lanes (0: speculative rounds)
};

__device__ __forceinline__ int rot_bin(float a1, float a2) {
    float rot = a1 - a2;
    if (rot < 0.0f) rot += 360.0f;
    int bin = (int)roundf(rot * (1.0f / HISTO_LENGTH));
    if (bin == HISTO_LENGTH) bin = 0;
    return min(max(bin, 0), HISTO_LENGTH - 1);
}

static const int RS_STAGE = 4096;
template <int MODE>
__global__ void __launch_bounds__(32) k_resolve(const RParams R) {
    extern __shared__ int sm[];
    const unsigned full = 0xffffffffu;
    const int b = blockIdx.x, lane = threadIdx.x;
    int* state = sm;                          // MAP/LAST/WINDOW: id held by target keypoint (-1 free); INIT: vMatchedDistance
    int* owner = sm + R.maxKp;                // INIT: vnMatches21;  LAST/WINDOW: rotation bin of a claimed keypoint
    unsigned* rnd = reinterpret_cast<unsigned*>(sm + 2 * R.maxKp);
    __shared__ uint32_t stage[MODE == M_INIT ? RS_STAGE : 1];   // INIT, lane-parallel walk: the candidate lists of 32 queries
    __shared__ int hist[32];
    __shared__ int top3[3];
    const int N2 = min(R.tgt.counts[b], R.maxKp);
    const int Q = min(R.qCounts[b], R.maxQ);
    const size_t tOff = (size_t)b * R.tgt.stride, sOff = (size_t)b * R.src.stride;
    for (int i = lane; i < N2; i += 32) {
        if (MODE == M_MAP || MODE == M_LAST || MODE == M_PROJ2) state[i] = R.tgtMp[tOff + i];
        else if (MODE == M_WINDOW || GF_IS_BOW(MODE)) state[i] = -1;
        else state[i] = INT_MAX;
        owner[i] = -1;
        rnd[i] = 0xffffffffu;
    }
    hist[lane] = 0;
    if (MODE == M_INIT)
        for (int i = lane; i < R.src.stride; i += 32) {
            R.matches12[sOff + i] = -1;
            if (i < R.maxQ) R.rotBinQ[(size_t)b * R.maxQ + i] = -1;
        }
    if (MODE == M_BOW2 || MODE == M_TRI)
        for (int i = lane; i < R.src.stride; i += 32) R.matches12[sOff + i] = -1;
    __syncwarp();
    int accepted = 0;
    if (MODE == M_INIT && R.laneParallel) {
        // SearchForInitialization has few active queries (level-0 keypoints only) with long candidate lists (a 100-pixel
        // window): the queries are taken strictly in order and the 32 lanes share the candidate walk of one query.  The
        // sequential scan's outcome is the minimum of (dist, list position) and the second order statistic of dist over the
        // eligible candidates (vMatchedDistance > dist), which two warp reductions reproduce.
        const uint32_t* poolB = R.pool + (size_t)b * R.poolCap;
        for (int qb = 0; qb < Q; qb += 32) {
            int2 meta = make_int2(0, 0);
            if (qb + lane < Q) meta = R.qMeta[(size_t)b * R.maxQ + qb + lane];
            unsigned act = __ballot_sync(full, meta.y > 0);
            if (!act) continue;
            // the candidate lists of this block of 32 queries go to shared memory with every copy in flight at once
            // (cp.async, 4 bytes per element): one L2 round trip per block instead of one or two per query
            int inc = meta.y;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(full, inc, d); if (lane >= d) inc += t; }
            const int soff = inc - meta.y;
            const bool fits = inc <= RS_STAGE;
            for (unsigned m = act; m; m &= m - 1) {
                const int src = __ffs(m) - 1;
                const int c = __shfl_sync(full, meta.y, src), o = __shfl_sync(full, soff, src), g = __shfl_sync(full, meta.x, src);
                if (__shfl_sync(full, (int)fits, src))
                    for (int k = lane; k < c; k += 32) {
                        const unsigned sa = (unsigned)__cvta_generic_to_shared(&stage[o + k]);
                        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa), "l"(poolB + g + k) : "memory");
                    }
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
            asm volatile("cp.async.wait_group 0;" ::: "memory");
            __syncwarp();
            for (; act; act &= act - 1) {
                const int src = __ffs(act) - 1, q = qb + src;
                const int cnt = __shfl_sync(full, meta.y, src);
                const bool st = __shfl_sync(full, (int)fits, src) != 0;
                const uint32_t* lst = st ? stage + __shfl_sync(full, soff, src) : poolB + __shfl_sync(full, meta.x, src);
                // local best / second over this lane's candidates (list positions lane, lane + 32, ...)
                unsigned k1 = 0xffffffffu, d2 = 0x7fffffffu;      // k1 = dist << 16 | list position
                int myBi = -1;
                {   // the first two chunks (lists of up to 64 candidates) with both loads in flight
                    const bool h0 = lane < cnt, h1 = lane + 32 < cnt;
                    const uint32_t ea = h0 ? lst[lane] : 0u, eb = h1 ? lst[lane + 32] : 0u;
                    const int ia = ea & 0xffff, ib = eb & 0xffff, da = (int)(ea >> 20), db = (int)(eb >> 20);
                    const bool oka = h0 && state[ia] > da, okb = h1 && state[ib] > db;         // :1211
                    if (oka) { k1 = ((unsigned)da << 16) | (unsigned)lane; myBi = ia; }
                    if (okb) {
                        const unsigned key = ((unsigned)db << 16) | (unsigned)(lane + 32);
                        if (key < k1) { if (oka) d2 = (unsigned)da; k1 = key; myBi = ib; }
                        else d2 = (unsigned)db;
                    }
                }
                for (int c = lane + 64; c < cnt; c += 32) {
                    const uint32_t e = lst[c];
                    const int i2 = e & 0xffff, dist = (int)(e >> 20);
                    if (state[i2] <= dist) continue;                             // :1211
                    const unsigned key = ((unsigned)dist << 16) | (unsigned)c;
                    if (key < k1) { if (k1 != 0xffffffffu) d2 = min(d2, k1 >> 16); k1 = key; myBi = i2; }
                    else d2 = min(d2, (unsigned)dist);
                }
                const unsigned kb = __reduce_min_sync(full, k1);
                if (kb != 0xffffffffu) {
                    const unsigned other = (k1 == kb) ? d2 : (k1 == 0xffffffffu ? 0x7fffffffu : (k1 >> 16));
                    const int best2 = (int)__reduce_min_sync(full, other);
                    const int best = (int)(kb >> 16);
                    const int bi = __shfl_sync(full, myBi, (int)(kb & 31u));     // list position c sits on lane c & 31
                    if (best <= TH_LOW && (float)best < (float)best2 * R.nnratio) {     // :1219-1221
                        if (lane == 0) {
                            const int prev = owner[bi];
                            if (prev >= 0) R.matches12[sOff + prev] = -1;       // steal (:1230-1234)
                            R.matches12[sOff + q] = bi;
                            owner[bi] = q;
                            state[bi] = best;
                            // the rotation bin depends only on (q, bi): it is filled in after the walk; stale entries of
                            // stolen matches stay and count, as in the reference (:1249)
                            if (R.checkOri) R.rotBinQ[(size_t)b * R.maxQ + q] = bi;
                        }
                        __syncwarp();
                    }
                }
            }
            __syncwarp();
        }
        __threadfence_block();
        __syncwarp();
        if (R.checkOri)
            for (int q = lane; q < Q; q += 32) {
                const int bi = __ldcg(&R.rotBinQ[(size_t)b * R.maxQ + q]);
                if (bi >= 0) {
                    const int bin = rot_bin(R.src.kps[sOff + q].angle, R.tgt.kps[tOff + bi].angle);
                    R.rotBinQ[(size_t)b * R.maxQ + q] = bin;
                    atomicAdd(&hist[bin], 1);
                }
            }
    } else
    for (int qb = 0; qb < Q; qb += 32) {
        int q = qb + lane;
        if (GF_IS_BOW(MODE) && q < Q) q = R.order[(size_t)b * R.maxQ + q];     // (node, index) order of the reference's walk
        int2 meta = make_int2(0, 0);
        if (qb + lane < Q) meta = R.qMeta[(size_t)b * R.maxQ + q];
        const uint32_t* list = R.pool + (size_t)b * R.poolCap + meta.x;
        const int cnt = meta.y;
        bool done = cnt <= 0;
        while (__any_sync(full, !done)) {
            bool acc = false;
            int bi = -1, bi2 = -1, best = INT_MAX;     // bi2 = the candidate that defines the second-best distance
            if (MODE == M_TRI && !done) {
                // SearchForTriangulation (:1476-1533): the unmatched candidates sorted by (distance, index); the first one with
                // distance <= 2 * best that passes the epipolar test (bit 31, set by k_candidates) is taken.  bi2 = the entry
                // that defines `best`, bi = the chosen one: the outcome depends on the claim state only through these two.
                unsigned kmin = 0xffffffffu, kch = 0xffffffffu;
                for (int c = 0; c < cnt; c++) {
                    const uint32_t e = list[c];
                    const int i2 = e & 0xffff;
                    if (state[i2] >= 0) continue;                               // vbMatched2 (:1485)
                    kmin = min(kmin, (((e >> 20) & 0x1ffu) << 16) | (unsigned)i2);
                }
                if (kmin != 0xffffffffu) {
                    bi2 = (int)(kmin & 0xffffu);
                    const unsigned distTh = 2u * (kmin >> 16);                  // round(2*BestDist)
                    for (int c = 0; c < cnt; c++) {
                        const uint32_t e = list[c];
                        const int i2 = e & 0xffff;
                        const unsigned dist = (e >> 20) & 0x1ffu;
                        if (!(e >> 31) || dist > distTh || state[i2] >= 0) continue;
                        kch = min(kch, (dist << 16) | (unsigned)i2);
                    }
                    if (kch != 0xffffffffu) { bi = (int)(kch & 0xffffu); best = (int)(kch >> 16); acc = true; }
                }
            } else if (!done) {
                int best2 = INT_MAX, bl = -1, bl2 = -1;
                for (int c = 0; c < cnt; c++) {
                    const uint32_t e = list[c];
                    const int i2 = e & 0xffff, dist = (int)(e >> 20), oct = (e >> 16) & 15;
                    if (MODE == M_INIT) {
                        if (state[i2] <= dist) continue;                        // :1211
                        if (dist < best) { best2 = best; bi2 = bi; best = dist; bi = i2; }
                        else if (dist < best2) { best2 = dist; bi2 = i2; }
                    } else {
                        if (state[i2] >= 0) continue;                           // claimed (:426, :1040, :2150)
                        if (MODE == M_LAST) {
                            if (dist < best) { best = dist; bi = i2; }
                        } else if (dist < best) { best2 = best; bi2 = bi; best = dist; bl2 = bl; bl = oct; bi = i2; }
                        else if (dist < best2) { bl2 = oct; best2 = dist; bi2 = i2; }
                    }
                }
                if (MODE == M_INIT || MODE == M_BOW) acc = best <= TH_LOW && (float)best < (float)best2 * R.nnratio;   // :1219-1221, :790-792
                else if (MODE == M_BOW2) acc = best < TH_LOW && (float)best < R.nnratio * (float)best2;                // :1365-1367
                else if (MODE == M_WINDOW || MODE == M_PROJ2) acc = (float)best <= (float)best2 * R.nnratio && best <= TH_HIGH;
                else if (MODE == M_LAST) acc = best <= R.thHigh;
                else acc = best <= TH_HIGH && !(bl == bl2 && (float)best > R.nnratio * (float)best2);
                if (bi < 0) acc = false;
            }
            if (!done && acc) atomicMin(&rnd[bi], (unsigned)lane);
            __syncwarp();
            // A claim by an earlier lane only removes (MAP/LAST/WINDOW) or may remove (INIT: vMatchedDistance drops) that
            // one target from this lane's candidate walk.  The walk's outcome is a function of its best and second-best
            // entries alone, so it is unchanged unless one of those two was claimed this round by an earlier lane.
            bool conflict = false;
            if (!done) conflict = (bi >= 0 && rnd[bi] < (unsigned)lane) || (bi2 >= 0 && rnd[bi2] < (unsigned)lane);
            const unsigned cm = __ballot_sync(full, !done && conflict);
            const int first = cm ? (__ffs(cm) - 1) : 32;
            __syncwarp();
            if (!done && acc) rnd[bi] = 0xffffffffu;
            if (!done && lane < first) {
                done = true;
                if (acc) {
                    accepted++;
                    if (MODE == M_MAP) {
                        state[bi] = q;
                        R.tgtMp[tOff + bi] = q;
                        R.tgtScore[tOff + bi] = best;
                    } else if (MODE == M_PROJ2) {
                        const int id = R.srcMp[sOff + q];
                        state[bi] = id;
                        R.tgtMp[tOff + bi] = id;                                 // vpMapPointMatches2[bestIdx2] = pMP1
                        R.tgtScore[tOff + bi] = best;                            // F2.mvpMatchScore[bestIdx2] (:1158)
                    } else if (MODE == M_LAST) {
                        const int id = R.srcMp[sOff + q];
                        state[bi] = id;
                        R.tgtMp[tOff + bi] = id;
                        R.tgtScore[tOff + bi] = best;
                        if (R.checkOri) {
                            const int bin = rot_bin(R.src.kps[sOff + q].angle, R.tgt.kps[tOff + bi].angle);
                            owner[bi] = bin;
                            atomicAdd(&hist[bin], 1);
                        }
                    } else if (MODE == M_BOW2 || MODE == M_TRI) {
                        state[bi] = q;                                           // vbMatched2[bestIdx2] = true, by whom
                        R.matches12[sOff + q] = (MODE == M_BOW2) ? R.tgtMpIds[tOff + bi] : bi;   // :1369 / :1515
                        if (R.checkOri) {
                            const int bin = rot_bin(R.src.kps[sOff + q].angle, R.tgt.kps[tOff + bi].angle);
                            owner[bi] = bin;
                            atomicAdd(&hist[bin], 1);
                        }
                    } else if (MODE == M_WINDOW || MODE == M_BOW) {
                        state[bi] = R.srcMp[sOff + q];
                        if (MODE == M_WINDOW || R.checkOri) {                    // :1064 pushes always, :796 under mbCheckOrientation
                            const int bin = rot_bin(R.src.kps[sOff + q].angle, R.tgt.kps[tOff + bi].angle);
                            owner[bi] = bin;
                            atomicAdd(&hist[bin], 1);
                        }
                    } else {
                        const int prev = owner[bi];
                        if (prev >= 0) R.matches12[sOff + prev] = -1;           // steal (:1230-1234)
                        R.matches12[sOff + q] = bi;
                        owner[bi] = q;
                        state[bi] = best;
                        if (R.checkOri) {
                            const int bin = rot_bin(R.src.kps[sOff + q].angle, R.tgt.kps[tOff + bi].angle);
                            R.rotBinQ[(size_t)b * R.maxQ + q] = bin;             // stale entries stay and count (:1249)
                            atomicAdd(&hist[bin], 1);
                        }
                    }
                }
            }
            __syncwarp();
        }
    }
    __threadfence_block();
    __syncwarp();
    // ComputeThreeMaxima (:2338-2379)
    if (lane == 0) {
        int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
        for (int i = 0; i < HISTO_LENGTH; i++) {
            const int s = hist[i];
            if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
            else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
            else if (s > max3) { max3 = s; ind3 = i; }
        }
        if ((float)max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
        else if ((float)max3 < 0.1f * (float)max1) ind3 = -1;
        top3[0] = ind1; top3[1] = ind2; top3[2] = ind3;
    }
    __syncwarp();
    const int i1 = top3[0], i2m = top3[1], i3 = top3[2];
    int total = 0;
    if (MODE == M_INIT) {
        const int N1 = min(R.src.counts[b], R.maxQ);
        for (int i = lane; i < N1; i += 32) {
            int m = __ldcg(&R.matches12[sOff + i]);
            if (R.checkOri && m >= 0) {
                const int bin = __ldcg(&R.rotBinQ[(size_t)b * R.maxQ + i]);
                if (bin >= 0 && bin != i1 && bin != i2m && bin != i3) { m = -1; R.matches12[sOff + i] = -1; }
            }
            if (m >= 0) {
                total++;
                R.prevMatched[(sOff + i) * 2] = R.tgt.kps[tOff + m].x;          // :1282-1284
                R.prevMatched[(sOff + i) * 2 + 1] = R.tgt.kps[tOff + m].y;
            }
        }
        total = __reduce_add_sync(full, total);
    } else if (MODE == M_MAP || MODE == M_PROJ2) {
        total = __reduce_add_sync(full, accepted);
    } else if (MODE == M_BOW2 || MODE == M_TRI) {
        int removed = 0;
        if (R.checkOri)
            for (int i = lane; i < N2; i += 32) {
                const int bin = owner[i];
                if (bin >= 0 && bin != i1 && bin != i2m && bin != i3) { removed++; R.matches12[sOff + state[i]] = -1; }   // :1409-1419
            }
        total = __reduce_add_sync(full, accepted - removed);
    } else {
        int removed = 0;
        if (R.checkOri)
            for (int i = lane; i < N2; i += 32) {
                const int bin = owner[i];
                if (bin >= 0 && bin != i1 && bin != i2m && bin != i3) {
                    removed++;
                    state[i] = -1;
                    if (MODE == M_LAST) { R.tgtMp[tOff + i] = -1; R.tgtScore[tOff + i] = 999; }
                }
            }
        if (MODE == M_WINDOW || MODE == M_BOW)
            for (int i = lane; i < R.tgt.stride; i += 32) R.tgtMp[tOff + i] = (i < N2) ? state[i] : -1;
        total = __reduce_add_sync(full, accepted - removed);
    }
    if (lane == 0) R.nmatches[b] = total;
}

// Claim-free overloads (Fuse, the two directions of SearchBySim3): the best candidate of every query, first occurrence of the
// smallest distance (strict `<` over the list in the reference's order).  outIdx = the keypoint when dist <= thrIdx, else -1.
__global__ void __launch_bounds__(128) k_pick_best(const uint32_t* __restrict__ pool, int poolCap, const int2* __restrict__ qMeta,
                                                   int maxQ, const int32_t* __restrict__ qCounts, int outStride, int thrIdx,
                                                   int thrCount, int32_t* __restrict__ outIdx, int32_t* __restrict__ outDist,
                                                   int32_t* __restrict__ nAccepted) {
    const int b = blockIdx.y, q = blockIdx.x * 128 + threadIdx.x;
    const int Q = min(qCounts[b], maxQ);
    int bestDist = INT_MAX, bestIdx = -1;
    if (q < Q) {
        const int2 meta = qMeta[(size_t)b * maxQ + q];
        const uint32_t* list = pool + (size_t)b * poolCap + meta.x;
        for (int c = 0; c < meta.y; c++) {
            const uint32_t e = list[c];
            const int dist = (int)(e >> 20);
            if (dist < bestDist) { bestDist = dist; bestIdx = (int)(e & 0xffff); }
        }
    }
    if (q < outStride) {
        outIdx[(size_t)b * outStride + q] = (q < Q && bestDist <= thrIdx) ? bestIdx : -1;
        if (outDist) outDist[(size_t)b * outStride + q] = (q < Q) ? bestDist : INT_MAX;
    }
    if (nAccepted) {
        const unsigned m = __ballot_sync(0xffffffffu, q < Q && bestDist <= thrCount);
        if ((threadIdx.x & 31) == 0 && m) atomicAdd(&nAccepted[b], __popc(m));
    }
}

// SearchBySim3, "check agreement" (:2060-2076)
__global__ void __launch_bounds__(128) k_sim3_agree(const int32_t* __restrict__ vnMatch1, int stride1, const int32_t* __restrict__ vnMatch2,
                                                    int stride2, const int32_t* __restrict__ n1, const int32_t* __restrict__ mp2,
                                                    int32_t* __restrict__ matches12, int32_t* __restrict__ nFound) {
    const int b = blockIdx.y, i1 = blockIdx.x * 128 + threadIdx.x;
    bool ok = false;
    if (i1 < min(n1[b], stride1)) {
        const int idx2 = vnMatch1[(size_t)b * stride1 + i1];
        if (idx2 >= 0 && idx2 < stride2 && vnMatch2[(size_t)b * stride2 + idx2] == i1) {
            matches12[(size_t)b * stride1 + i1] = mp2[(size_t)b * stride2 + idx2];
            ok = true;
        }
    }
    const unsigned m = __ballot_sync(0xffffffffu, ok);
    if ((threadIdx.x & 31) == 0 && m) atomicAdd(&nFound[b], __popc(m));
}

}  // namespace gforb
#include "active.cuh"
#include "allpairs_tc.cuh"
namespace gforb {

// ------------------------------------------------------------------------------------------------ Frame::isInFrustum
// Frame::UndistortKeyPoints (src/Frame.cc:389-423) for the keypoints of `batch` frames: thread per keypoint.
// cvUndistortPoints restated (double arithmetic on CV_32F inputs, 5 fixed-point iterations, R = I, P = K); the keypoint
// is copied and its position replaced.  This TU is built with --fmad=false, so every operation rounds on its own.
struct UndistortParams {
    const GforbKeyPoint* keys;
    GforbKeyPoint* keysUn;
    const int32_t* counts;
    int stride;
    double fx, fy, cx, cy, k[8];
    int identity;                          // mDistCoef[0] == 0: mvKeysUn = mvKeys
};

__global__ void __launch_bounds__(128) k_undistort(const UndistortParams P) {
    const int b = blockIdx.y, i = blockIdx.x * 128 + threadIdx.x;
    if (i >= P.stride || i >= P.counts[b]) return;
    GforbKeyPoint kp = P.keys[(size_t)b * P.stride + i];
    if (!P.identity) {
        const double ifx = 1. / P.fx, ify = 1. / P.fy;
        double x = ((double)kp.x - P.cx) * ifx, y = ((double)kp.y - P.cy) * ify;
        const double x0 = x, y0 = y;
        for (int j = 0; j < 5; j++) {
            const double r2 = x * x + y * y;
            const double icdist = (1 + ((P.k[7] * r2 + P.k[6]) * r2 + P.k[5]) * r2) / (1 + ((P.k[4] * r2 + P.k[1]) * r2 + P.k[0]) * r2);
            const double deltaX = 2 * P.k[2] * x * y + P.k[3] * (r2 + 2 * x * x);
            const double deltaY = P.k[2] * (r2 + 2 * y * y) + 2 * P.k[3] * x * y;
            x = (x0 - deltaX) * icdist;
            y = (y0 - deltaY) * icdist;
        }
        const double xx = P.fx * x + 0. * y + P.cx;
        const double yy = 0. * x + P.fy * y + P.cy;
        const double ww = 1. / (0. * x + 0. * y + 1.);
        kp.x = (float)(xx * ww);
        kp.y = (float)(yy * ww);
    }
    P.keysUn[(size_t)b * P.stride + i] = kp;
}

// src/Frame.cc:166-227 for every map point of `batch` frames: thread per point.  cv::Mat arithmetic restated: Rcw * P + tcw
// is one small-matrix cv::gemm (common.cuh: cv_gemm_row); cv::norm / Mat::dot accumulate double products.
struct FrustumParams {
    const float *Rcw, *tcw, *Ow;          // batch x 9 / 3 / 3
    float fx, fy, cx, cy;
    int minX, maxX, minY, maxY;
    const float *world, *normal, *minDist, *maxDist;
    const int32_t* counts;
    int stride;
    float cosLimit;
    float scale[MAX_LEVELS];
    int nlevels;
    float *u, *v, *viewCos;
    int32_t* level;
    uint8_t* flags;
};

__global__ void __launch_bounds__(128) k_is_in_frustum(const FrustumParams F) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x, b = blockIdx.y;
    if (i >= min(F.counts[b], F.stride)) return;
    const size_t p = (size_t)b * F.stride + i;
    const float* R = F.Rcw + (size_t)b * 9;
    const float* t = F.tcw + (size_t)b * 3;
    const float* O = F.Ow + (size_t)b * 3;
    const float P0 = F.world[p * 3], P1 = F.world[p * 3 + 1], P2 = F.world[p * 3 + 2];
    uint8_t fl = F.flags[p] & ~1;
    bool in = true;
    float Pc[3];
#pragma unroll
    for (int r = 0; r < 3; r++) Pc[r] = cv_gemm_row(R[r * 3], R[r * 3 + 1], R[r * 3 + 2], P0, P1, P2, t[r]);   // mRcw*P+mtcw
    if (Pc[2] < 0.0f) in = false;
    const float invz = (float)(1.0 / (double)Pc[2]);
    const float u = F.fx * Pc[0] * invz + F.cx;
    const float v = F.fy * Pc[1] * invz + F.cy;
    if (u < (float)F.minX || u > (float)F.maxX) in = false;
    if (v < (float)F.minY || v > (float)F.maxY) in = false;
    const float PO0 = P0 - O[0], PO1 = P1 - O[1], PO2 = P2 - O[2];
    const double n2 = (double)PO0 * (double)PO0 + (double)PO1 * (double)PO1 + (double)PO2 * (double)PO2;
    const float dist = (float)sqrt(n2);
    const float minD = F.minDist[p], maxD = F.maxDist[p];
    if (dist < minD || dist > maxD) in = false;
    const double dot = (double)PO0 * (double)F.normal[p * 3] + (double)PO1 * (double)F.normal[p * 3 + 1] + (double)PO2 * (double)F.normal[p * 3 + 2];
    const float viewCos = (float)(dot / (double)dist);
    if (viewCos < F.cosLimit) in = false;
    if (in) {
        const float ratio = dist / minD;
        int lvl = 0;                                          // std::lower_bound(mvScaleFactors, ratio)
        while (lvl < F.nlevels && F.scale[lvl] < ratio) lvl++;
        if (lvl >= F.nlevels) lvl = F.nlevels - 1;
        F.u[p] = u; F.v[p] = v; F.level[p] = lvl; F.viewCos[p] = viewCos;
        fl |= 1;
    }
    F.flags[p] = fl;
}

// ------------------------------------------------------------------------------------------------ distinctive descriptor
// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:196-262) for many map points: the observation descriptor with
// the least median Hamming distance to the others (median = sorted row [int(0.5 (N - 1))], the row's own 0 included; the
// first row with the smallest median wins).  CTA of 4 warps per map point, a warp per row: distances go into a 257-bin
// histogram (they are integers 0..256) and the median is the bin where the running count passes the rank.
static const int DD_WARPS = 4;
__global__ void __launch_bounds__(DD_WARPS * 32) k_distinctive(const uint8_t* __restrict__ desc, const int32_t* __restrict__ offsets,
                                                              int nPoints, int32_t* __restrict__ bestIdx,
                                                              uint8_t* __restrict__ outDesc) {
    __shared__ int hist[DD_WARPS][264];
    __shared__ unsigned long long s_best[DD_WARPS];
    const int mp = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (mp >= nPoints) return;
    const int o0 = offsets[mp], N = offsets[mp + 1] - o0;
    if (N <= 0) {
        if (threadIdx.x == 0) bestIdx[mp] = -1;          // no observation: the reference returns and keeps mDescriptor
        return;
    }
    const uint4* D = reinterpret_cast<const uint4*>(desc + (size_t)o0 * 32);
    const int rank = (int)(0.5 * (N - 1));
    unsigned long long best = ~0ull;                      // (median << 32 | row), smallest wins
    for (int i = warp; i < N; i += DD_WARPS) {
        for (int q = lane; q < 264; q += 32) hist[warp][q] = 0;
        __syncwarp();
        const uint4 a0 = __ldg(D + 2 * i), a1 = __ldg(D + 2 * i + 1);
        for (int j = lane; j < N; j += 32) atomicAdd(&hist[warp][popc256(a0, a1, __ldg(D + 2 * j), __ldg(D + 2 * j + 1))], 1);
        __syncwarp();
        // lane l owns bins [9 l, 9 l + 9) (32 x 9 = 288 >= 257): inclusive scan of the lane sums, then walk the own bins
        int c[9], sum = 0;
#pragma unroll
        for (int q = 0; q < 9; q++) { const int bin = 9 * lane + q; c[q] = bin < 257 ? hist[warp][bin] : 0; sum += c[q]; }
        int incl = sum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += t; }
        int before = incl - sum, med = 1 << 30;
#pragma unroll
        for (int q = 0; q < 9; q++) { if (med == (1 << 30) && before + c[q] > rank && before <= rank) med = 9 * lane + q; before += c[q]; }
        med = __reduce_min_sync(0xffffffffu, med);
        const unsigned long long key = ((unsigned long long)(unsigned)med << 32) | (unsigned)i;
        if (key < best) best = key;
        __syncwarp();
    }
    if (lane == 0) s_best[warp] = best;
    __syncthreads();
    if (threadIdx.x < 32) {
        unsigned long long b = s_best[0];
        for (int w = 1; w < DD_WARPS; w++) b = s_best[w] < b ? s_best[w] : b;
        const int row = (int)(b & 0xffffffffu);
        if (lane == 0) bestIdx[mp] = row;
        if (outDesc && lane < 2) reinterpret_cast<uint4*>(outDesc + (size_t)mp * 32)[lane] = __ldg(D + 2 * row + lane);
    }
}

// ------------------------------------------------------------------------------------------------ B1 distance
__global__ void __launch_bounds__(256) k_distance(const uint4* __restrict__ a, const uint4* __restrict__ b, int n,
                                                  int32_t* __restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = popc256(__ldg(a + 2 * i), __ldg(a + 2 * i + 1), __ldg(b + 2 * i), __ldg(b + 2 * i + 1));
}

// ------------------------------------------------------------------------------------------------ K10 all-pairs
// Thread = AP_R queries held in registers; CTA streams its database chunk through shared memory (broadcast LDS.128).
// Per pair: 8 LOP3(xor) + 8 POPC + adds, then a 3-instruction top-2 update on packed (dist << 22 | row) keys.
static const int AP_R = 4, AP_THREADS = 128, AP_TILE = 256;
__global__ void __launch_bounds__(AP_THREADS) k_allpairs(const uint4* __restrict__ qv, int nq, const uint4* __restrict__ db,
                                                         long long nDb, int chunkRows, long long idxBase,
                                                         uint64_t* __restrict__ partial) {
    __shared__ uint4 s_db[2][AP_TILE * 2];
    const int tid = threadIdx.x;
    const int q0 = (blockIdx.x * AP_THREADS + tid) * AP_R;
    const long long row0 = (long long)blockIdx.y * chunkRows;
    const int rows = (int)min((long long)chunkRows, nDb - row0);
    uint32_t qd[AP_R][8];
    uint32_t k1[AP_R], k2[AP_R];
#pragma unroll
    for (int r = 0; r < AP_R; r++) {
        uint4 a0 = make_uint4(0, 0, 0, 0), a1 = a0;
        if (q0 + r < nq) { a0 = __ldg(qv + 2 * (size_t)(q0 + r)); a1 = __ldg(qv + 2 * (size_t)(q0 + r) + 1); }
        qd[r][0] = a0.x; qd[r][1] = a0.y; qd[r][2] = a0.z; qd[r][3] = a0.w;
        qd[r][4] = a1.x; qd[r][5] = a1.y; qd[r][6] = a1.z; qd[r][7] = a1.w;
        k1[r] = k2[r] = 0xffffffffu;
    }
    const uint4* src = db + 2 * (size_t)row0;
    const int nTiles = (rows + AP_TILE - 1) / AP_TILE;
    auto load_tile = [&](int t, int buf) {
        const int base = t * AP_TILE, n16 = min(AP_TILE, rows - base) * 2;
        for (int i = tid; i < AP_TILE * 2; i += AP_THREADS)
            s_db[buf][i] = (i < n16) ? ldg_nc_u4(src + (size_t)base * 2 + i) : make_uint4(0, 0, 0, 0);
    };
    if (nTiles > 0) load_tile(0, 0);
    __syncthreads();
    for (int t = 0; t < nTiles; t++) {
        const int buf = t & 1;
        if (t + 1 < nTiles) load_tile(t + 1, buf ^ 1);
        const int n = min(AP_TILE, rows - t * AP_TILE);
        const uint32_t rowBase = (uint32_t)(t * AP_TILE);
#pragma unroll 4
        for (int j = 0; j < n; j++) {
            const uint4 d0 = s_db[buf][2 * j], d1 = s_db[buf][2 * j + 1];
#pragma unroll
            for (int r = 0; r < AP_R; r++) {
                const int dist = __popc(qd[r][0] ^ d0.x) + __popc(qd[r][1] ^ d0.y) + __popc(qd[r][2] ^ d0.z) +
                                 __popc(qd[r][3] ^ d0.w) + __popc(qd[r][4] ^ d1.x) + __popc(qd[r][5] ^ d1.y) +
                                 __popc(qd[r][6] ^ d1.z) + __popc(qd[r][7] ^ d1.w);
                const uint32_t key = ((uint32_t)dist << 22) + (rowBase + (uint32_t)j);
                k2[r] = min(k2[r], max(k1[r], key));
                k1[r] = min(k1[r], key);
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int r = 0; r < AP_R; r++) {
        if (q0 + r >= nq) continue;
        uint64_t o1 = ~0ull, o2 = ~0ull;
        if (k1[r] != 0xffffffffu) o1 = ((uint64_t)(k1[r] >> 22) << 32) | (uint64_t)(idxBase + row0 + (k1[r] & 0x3fffff));
        if (k2[r] != 0xffffffffu) o2 = ((uint64_t)(k2[r] >> 22) << 32) | (uint64_t)(idxBase + row0 + (k2[r] & 0x3fffff));
        uint64_t* o = partial + ((size_t)blockIdx.y * nq + (q0 + r)) * 2;
        o[0] = o1; o[1] = o2;
    }
}

// merge n_parts partial top-2 lists per query (keys are unique: they carry the global row index)
__global__ void __launch_bounds__(256) k_top2_merge(const uint64_t* __restrict__ parts, int nParts, int nq,
                                                    uint64_t* __restrict__ out) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    uint64_t a = ~0ull, b2 = ~0ull;
    for (int p = 0; p < nParts; p++) {
        const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(parts + ((size_t)p * nq + q) * 2);
        uint64_t k = v.x;
        b2 = min(b2, max(a, k)); a = min(a, k);
        k = v.y;
        b2 = min(b2, max(a, k)); a = min(a, k);
    }
    out[(size_t)q * 2] = a;
    out[(size_t)q * 2 + 1] = b2;
}

}  // namespace gforb

using namespace gforb;

// ================================================================================================ host side
struct GforbMatcher {
    GforbMatcherConfig cfg;
    int device = 0, numSMs = 148, poolCap = 0, lastLaunches = 0;
    float scale[MAX_LEVELS];
    Arena arena;
    int* d_cellStart = nullptr;      // max_batch x 3073
    int4* d_entries = nullptr;       // max_batch x max_keypoints
    uint32_t* d_pool = nullptr;      // max_batch x poolCap
    int* d_poolCount = nullptr;      // max_batch
    int2* d_qMeta = nullptr;         // max_batch x max_queries
    int* d_rotBinQ = nullptr;        // max_batch x max_queries
    int initLaneParallel = 1;        // GFORB_RESOLVE=spec: SearchForInitialization through the speculative rounds of the other modes
    int* d_err = nullptr;
    uint64_t* d_partial = nullptr;
    size_t capPartial = 0;
    uint8_t* d_active = nullptr;     // k_active_match scratch (allocated by the first gforb_obs_active_match call)
    std::vector<OutCopy> outs;
};

namespace {

struct Ctx {
    GforbMatcher* M;
    cudaStream_t st;
    bool inDev, outDev;
    int rc = GFORB_OK;
    template <class T>
    const T* in(const T* p, size_t n) {           // pure input
        if (!p || rc) return p;
        if (inDev) return p;
        T* d = (T*)M->arena.take(n * sizeof(T));
        if (!d) { rc = GFORB_ERR_CAPACITY; return nullptr; }
        if (cudaMemcpyAsync(d, p, n * sizeof(T), cudaMemcpyHostToDevice, st) != cudaSuccess) rc = GFORB_ERR_CUDA;
        return d;
    }
    template <class T>
    T* out(T* p, size_t n, bool copyIn) {          // output or in/out
        if (!p || rc) return p;
        if (outDev) return p;
        T* d = (T*)M->arena.take(n * sizeof(T));
        if (!d) { rc = GFORB_ERR_CAPACITY; return nullptr; }
        if (copyIn && cudaMemcpyAsync(d, p, n * sizeof(T), cudaMemcpyHostToDevice, st) != cudaSuccess) rc = GFORB_ERR_CUDA;
        M->outs.push_back(OutCopy{p, d, n * sizeof(T)});
        return d;
    }
    FrameDev frame(const GforbFrameSet* F, int batch) {
        FrameDev f;
        f.stride = F->kp_stride;
        f.kps = in(F->keypoints, (size_t)batch * F->kp_stride);
        f.desc = in(F->descriptors, (size_t)batch * F->kp_stride * 32);
        f.counts = in(F->counts, (size_t)batch);
        return f;
    }
};

GridGeom geom_of(const GforbFrameSet* F) {
    GridGeom G;
    G.iminX = F->min_x; G.imaxX = F->max_x; G.iminY = F->min_y; G.imaxY = F->max_y;
    G.minX = (float)F->min_x; G.minY = (float)F->min_y;
    G.invW = (float)GRID_COLS / (float)(F->max_x - F->min_x);       // src/Frame.cc:97-98
    G.invH = (float)GRID_ROWS / (float)(F->max_y - F->min_y);
    return G;
}

int begin_call(GforbMatcher* M, int batch, int flags, void* stream, Ctx& c) {
    if (!M) return GFORB_ERR_ARG;
    if (batch < 0) return GFORB_ERR_ARG;
    if (batch > M->cfg.max_batch) return GFORB_ERR_CAPACITY;
    GF_CUDA_OK(cudaSetDevice(M->device));
    c.M = M; c.st = (cudaStream_t)stream;
    c.inDev = (flags & GFORB_IN_DEVICE) != 0; c.outDev = (flags & GFORB_OUT_DEVICE) != 0;
    M->arena.reset();
    M->outs.clear();
    M->lastLaunches = 0;
    return GFORB_OK;
}

int end_call(GforbMatcher* M, Ctx& c, int flags) {
    if (c.rc) return c.rc;
    GF_CUDA_OK(cudaGetLastError());
    for (auto& o : M->outs) GF_CUDA_OK(cudaMemcpyAsync(o.host, o.dev, o.bytes, cudaMemcpyDeviceToHost, c.st));
    if (!(flags & GFORB_ASYNC) || !M->outs.empty()) {
        GF_CUDA_OK(cudaStreamSynchronize(c.st));
        int err = 0;
        GF_CUDA_OK(cudaMemcpy(&err, M->d_err, sizeof(int), cudaMemcpyDeviceToHost));
        if (err) { cudaMemset(M->d_err, 0, sizeof(int)); return err; }
    }
    return GFORB_OK;
}

int check_frames(GforbMatcher* M, const GforbFrameSet* F) {
    if (!F || !F->keypoints || !F->descriptors || !F->counts) return GFORB_ERR_ARG;
    if (F->kp_stride <= 0 || F->max_x <= F->min_x || F->max_y <= F->min_y) return GFORB_ERR_ARG;
    if (F->kp_stride > M->cfg.max_keypoints) return GFORB_ERR_CAPACITY;
    return GFORB_OK;
}

void build_grid(GforbMatcher* M, Ctx& c, const FrameDev& T, const GridGeom& G, int batch) {
    k_grid_build<<<batch, 256, 0, c.st>>>(T, G, M->d_cellStart, M->d_entries, M->cfg.max_keypoints);
    M->lastLaunches++;
}

template <int MODE>
void run_candidates(GforbMatcher* M, Ctx& c, const QParams& P, const FrameDev& T, const GridGeom& G, int batch, int maxQueries) {
    cudaMemsetAsync(M->d_poolCount, 0, sizeof(int) * batch, c.st);
    GridDev Gd{M->d_cellStart, M->d_entries, M->cfg.max_keypoints};
    dim3 grid((maxQueries + 3) / 4, batch);
    k_candidates<MODE><<<grid, 128, 0, c.st>>>(P, T, Gd, G, M->d_pool, M->poolCap, M->d_poolCount, M->d_qMeta,
                                                M->cfg.max_queries, M->d_err);
    M->lastLaunches++;
}

template <int MODE>
void run_resolve(GforbMatcher* M, Ctx& c, const RParams& R, int batch) {
    const size_t smem = (size_t)3 * M->cfg.max_keypoints * sizeof(int);
    k_resolve<MODE><<<batch, 32, smem, c.st>>>(R);
    M->lastLaunches++;
}

void fill_scale(GforbMatcher* M, QParams& P) {
    for (int i = 0; i < MAX_LEVELS; i++) P.scale[i] = M->scale[i < M->cfg.nlevels ? i : M->cfg.nlevels - 1];
    P.nlevels = M->cfg.nlevels;
}

}  // namespace

extern "C" {

int gforb_matcher_create(const GforbMatcherConfig* cfg, GforbMatcher** out) {
    if (!cfg || !out) return GFORB_ERR_ARG;
    *out = nullptr;
    if (cfg->max_batch <= 0 || cfg->max_keypoints <= 0 || cfg->max_keypoints > 16384 || cfg->max_queries <= 0 ||
        cfg->nlevels <= 0 || cfg->nlevels > MAX_LEVELS || !(cfg->scaleFactor > 1.0f) || cfg->cand_pool < 0)
        return GFORB_ERR_ARG;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || cfg->device < 0 || cfg->device >= ndev) {
        fprintf(stderr, "[gforb] no usable CUDA device (requested %d of %d): the B200 path has no CPU fallback\n", cfg->device, ndev);
        return GFORB_ERR_CUDA;
    }
    GF_CUDA_OK(cudaSetDevice(cfg->device));
    GforbMatcher* M = new GforbMatcher();
    const auto init = [&]() -> int {
    M->cfg = *cfg;
    M->device = cfg->device;
    cudaDeviceGetAttribute(&M->numSMs, cudaDevAttrMultiProcessorCount, cfg->device);
    if (const char* e = getenv("GFORB_RESOLVE")) M->initLaneParallel = strcmp(e, "spec") != 0;
    // Frame::mvScaleFactors (src/Frame.cc:62-75): mvScaleFactors[i] = mvScaleFactors[i-1] * mfScaleFactor (float)
    M->scale[0] = 1.0f;
    for (int i = 1; i < MAX_LEVELS; i++) M->scale[i] = M->scale[i - 1] * cfg->scaleFactor;
    M->poolCap = cfg->cand_pool > 0 ? cfg->cand_pool : 64 * cfg->max_queries;
    const size_t B = cfg->max_batch, K = cfg->max_keypoints, Qn = cfg->max_queries;
    GF_CUDA_OK(cudaMalloc((void**)&M->d_cellStart, B * (GRID_CELLS + 1) * sizeof(int)));
    GF_CUDA_OK(cudaMalloc((void**)&M->d_entries, B * K * sizeof(int4)));
    GF_CUDA_OK(cudaMalloc((void**)&M->d_pool, B * (size_t)M->poolCap * sizeof(uint32_t)));
    GF_CUDA_OK(cudaMalloc((void**)&M->d_poolCount, B * sizeof(int)));
    GF_CUDA_OK(cudaMalloc((void**)&M->d_qMeta, B * Qn * sizeof(int2)));
    GF_CUDA_OK(cudaMalloc((void**)&M->d_rotBinQ, B * Qn * sizeof(int)));
    GF_CUDA_OK(cudaMalloc((void**)&M->d_err, sizeof(int)));
    GF_CUDA_OK(cudaMemset(M->d_err, 0, sizeof(int)));
    // staging arena for host-pointer calls: two frame sets + one query set + outputs
    const size_t arenaBytes = B * (K * 2 * (28 + 32 + 16) + Qn * (64 + 32)) + (4u << 20);
    GF_CUDA_OK(M->arena.init(arenaBytes));
    const int smem = 3 * cfg->max_keypoints * (int)sizeof(int);
    GF_CUDA_OK(raise_dyn_smem(k_resolve<M_INIT>, smem, cfg->device));
    GF_CUDA_OK(raise_dyn_smem(k_resolve<M_WINDOW>, smem, cfg->device));
    GF_CUDA_OK(raise_dyn_smem(k_resolve<M_MAP>, smem, cfg->device));
    GF_CUDA_OK(raise_dyn_smem(k_resolve<M_LAST>, smem, cfg->device));
    GF_CUDA_OK(raise_dyn_smem(k_resolve<M_PROJ2>, smem, cfg->device));
    GF_CUDA_OK(raise_dyn_smem(k_resolve<M_BOW>, smem, cfg->device));
    GF_CUDA_OK(raise_dyn_smem(k_resolve<M_BOW2>, smem, cfg->device));
    GF_CUDA_OK(raise_dyn_smem(k_resolve<M_TRI>, smem, cfg->device));
        return GFORB_OK;
    };
    if (const int rc = init()) { gforb_matcher_destroy(M); return rc; }   // a failure part-way frees what was allocated
    *out = M;
    return GFORB_OK;
}

int gforb_matcher_destroy(GforbMatcher* M) {
    if (!M) return GFORB_ERR_ARG;
    cudaSetDevice(M->device);
    void* ptrs[] = {M->d_cellStart, M->d_entries, M->d_pool, M->d_poolCount, M->d_qMeta, M->d_rotBinQ, M->d_err, M->d_partial,
                    M->d_active};
    for (void* p : ptrs) if (p) cudaFree(p);
    M->arena.release();
    delete M;
    return GFORB_OK;
}

int gforb_matcher_last_launches(const GforbMatcher* M) { return M ? M->lastLaunches : GFORB_ERR_ARG; }

int gforb_mappoint_distinctive_descriptors(GforbMatcher* M, int n_points, const uint8_t* descriptors, const int32_t* offsets,
                                           int32_t* best_idx, uint8_t* out_desc, int flags, void* stream) {
    Ctx c;
    int rc = begin_call(M, 0, flags, stream, c);
    if (rc) return rc;
    if (n_points == 0) return GFORB_OK;
    if (n_points < 0 || !offsets || !best_idx) return GFORB_ERR_ARG;
    // the total number of observation descriptors is offsets[n_points]: host-visible for host inputs; for device inputs
    // it is read back (one int) because the staging size is not needed then
    size_t total = 0;
    if (!(flags & GFORB_IN_DEVICE)) {
        total = (size_t)offsets[n_points];
        for (int i = 0; i < n_points; i++) if (offsets[i + 1] < offsets[i]) return GFORB_ERR_ARG;
        if (total > 0 && !descriptors) return GFORB_ERR_ARG;
    }
    const uint8_t* dd = (flags & GFORB_IN_DEVICE) ? descriptors : c.in(descriptors, total * 32);
    const int32_t* doff = c.in(offsets, (size_t)n_points + 1);
    int32_t* dbest = c.out(best_idx, (size_t)n_points, false);
    uint8_t* dout = out_desc ? c.out(out_desc, (size_t)n_points * 32, true) : nullptr;
    if (c.rc) return c.rc;
    k_distinctive<<<n_points, DD_WARPS * 32, 0, c.st>>>(dd, doff, n_points, dbest, dout);
    M->lastLaunches++;
    return end_call(M, c, flags);
}

int gforb_descriptor_distance(GforbMatcher* M, const uint8_t* a, const uint8_t* b, int n, int32_t* out, int flags, void* stream) {
    Ctx c;
    int rc = begin_call(M, 0, flags, stream, c);
    if (rc) return rc;
    if (n == 0) return GFORB_OK;
    if (!a || !b || !out || n < 0) return GFORB_ERR_ARG;
    const uint8_t* da = c.in(a, (size_t)n * 32);
    const uint8_t* db = c.in(b, (size_t)n * 32);
    int32_t* dout = c.out(out, (size_t)n, false);
    if (c.rc) return c.rc;
    k_distance<<<(n + 255) / 256, 256, 0, c.st>>>((const uint4*)da, (const uint4*)db, n, dout);
    M->lastLaunches++;
    return end_call(M, c, flags);
}

int gforb_features_in_area(GforbMatcher* M, const GforbFrameSet* frames, int batch, int frame, const float* x,
                           const float* y, const float* r, const int32_t* min_level, const int32_t* max_level,
                           int n_queries, int32_t* out_idx, int cap, int32_t* out_n, int flags, void* stream) {
    Ctx c;
    int rc = begin_call(M, batch, flags, stream, c);
    if (rc) return rc;
    if ((rc = check_frames(M, frames))) return rc;
    if (frame < 0 || frame >= batch || n_queries < 0 || cap <= 0 || !out_idx || !out_n) return GFORB_ERR_ARG;
    if (n_queries > M->cfg.max_queries) return GFORB_ERR_CAPACITY;
    if (n_queries == 0) return GFORB_OK;
    FrameDev T = c.frame(frames, batch);
    GridGeom G = geom_of(frames);
    QParams P;
    memset(&P, 0, sizeof(P));
    fill_scale(M, P);
    P.ax = c.in(x, n_queries); P.ay = c.in(y, n_queries); P.ar = c.in(r, n_queries);
    P.aMin = c.in(min_level, n_queries); P.aMax = c.in(max_level, n_queries);
    P.aCount = n_queries; P.aFrame = frame;
    if (c.rc) return c.rc;
    build_grid(M, c, T, G, batch);
    run_candidates<M_AREA>(M, c, P, T, G, 1, n_queries);
    GF_CUDA_OK(cudaStreamSynchronize(c.st));
    // unpack on the host (parity tap, not a hot path)
    std::vector<int2> meta(n_queries);
    GF_CUDA_OK(cudaMemcpy(meta.data(), M->d_qMeta, sizeof(int2) * n_queries, cudaMemcpyDeviceToHost));
    int used = 0;
    GF_CUDA_OK(cudaMemcpy(&used, M->d_poolCount, sizeof(int), cudaMemcpyDeviceToHost));
    std::vector<uint32_t> pool(std::max(used, 1));
    if (used > 0 && used <= M->poolCap) GF_CUDA_OK(cudaMemcpy(pool.data(), M->d_pool, sizeof(uint32_t) * used, cudaMemcpyDeviceToHost));
    std::vector<int32_t> idx((size_t)n_queries * cap, -1), cnt(n_queries, 0);
    for (int q = 0; q < n_queries; q++) {
        cnt[q] = meta[q].y;
        for (int k = 0; k < meta[q].y && k < cap; k++) idx[(size_t)q * cap + k] = (int32_t)(pool[meta[q].x + k] & 0xffff);
    }
    if (c.outDev) {
        GF_CUDA_OK(cudaMemcpy(out_idx, idx.data(), idx.size() * 4, cudaMemcpyHostToDevice));
        GF_CUDA_OK(cudaMemcpy(out_n, cnt.data(), cnt.size() * 4, cudaMemcpyHostToDevice));
    } else {
        memcpy(out_idx, idx.data(), idx.size() * 4);
        memcpy(out_n, cnt.data(), cnt.size() * 4);
    }
    return end_call(M, c, flags);
}

int gforb_search_for_initialization(GforbMatcher* M, const GforbFrameSet* F1, const GforbFrameSet* F2, int batch,
                                    float* prev_matched, int32_t* matches12, int32_t* nmatches, int window_size,
                                    float nnratio, int check_orientation, int flags, void* stream) {
    Ctx c;
    int rc = begin_call(M, batch, flags, stream, c);
    if (rc) return rc;
    if ((rc = check_frames(M, F1)) || (rc = check_frames(M, F2))) return rc;
    if (!prev_matched || !matches12 || !nmatches) return GFORB_ERR_ARG;
    if (F1->kp_stride > M->cfg.max_queries) return GFORB_ERR_CAPACITY;
    if (batch == 0) return GFORB_OK;
    FrameDev S = c.frame(F1, batch), T = c.frame(F2, batch);
    GridGeom G = geom_of(F2);
    QParams P;
    memset(&P, 0, sizeof(P));
    fill_scale(M, P);
    float* dPrev = c.out(prev_matched, (size_t)batch * F1->kp_stride * 2, true);
    int32_t* dM12 = c.out(matches12, (size_t)batch * F1->kp_stride, false);
    int32_t* dN = c.out(nmatches, (size_t)batch, false);
    if (c.rc) return c.rc;
    P.src = S; P.prevMatched = dPrev; P.window = (float)window_size;
    build_grid(M, c, T, G, batch);
    run_candidates<M_INIT>(M, c, P, T, G, batch, F1->kp_stride);
    RParams R;
    memset(&R, 0, sizeof(R));
    R.src = S; R.tgt = T; R.pool = M->d_pool; R.poolCap = M->poolCap; R.qMeta = M->d_qMeta; R.maxQ = M->cfg.max_queries;
    R.maxKp = M->cfg.max_keypoints; R.qCounts = S.counts; R.nnratio = nnratio; R.checkOri = check_orientation;
    R.matches12 = dM12; R.prevMatched = dPrev; R.nmatches = dN; R.rotBinQ = M->d_rotBinQ;
    R.laneParallel = M->initLaneParallel;
    run_resolve<M_INIT>(M, c, R, batch);
    return end_call(M, c, flags);
}

// shared body of SearchByProjection(F, MPs, th) and SearchByProjection_GoodFeature(F, MPs, th_good, th_rest); the latter
// does not store match scores (match_score == nullptr -> the resolver writes into scratch)
static int sbp_map(GforbMatcher* M, const GforbFrameSet* F, const GforbMapPointSet* mps, int batch, float th, float thGood,
                   int goodMode, float nnratio, int32_t* frame_mappoints, int32_t* match_score, int32_t* nmatches, int flags,
                   void* stream) {
    Ctx c;
    int rc = begin_call(M, batch, flags, stream, c);
    if (rc) return rc;
    if ((rc = check_frames(M, F))) return rc;
    if (!mps || !mps->proj_x || !mps->proj_y || !mps->scale_level || !mps->view_cos || !mps->flags || !mps->descriptors ||
        !mps->counts || mps->mp_stride <= 0 || !frame_mappoints || (!match_score && !goodMode) || !nmatches)
        return GFORB_ERR_ARG;
    if (mps->mp_stride > M->cfg.max_queries) return GFORB_ERR_CAPACITY;
    if (batch == 0) return GFORB_OK;
    FrameDev T = c.frame(F, batch);
    GridGeom G = geom_of(F);
    QParams P;
    memset(&P, 0, sizeof(P));
    fill_scale(M, P);
    const size_t nm = (size_t)batch * mps->mp_stride;
    P.mpU = c.in(mps->proj_x, nm); P.mpV = c.in(mps->proj_y, nm); P.mpCos = c.in(mps->view_cos, nm);
    P.mpLevel = c.in(mps->scale_level, nm); P.mpFlags = c.in(mps->flags, nm); P.mpDesc = c.in(mps->descriptors, nm * 32);
    P.mpCounts = c.in(mps->counts, (size_t)batch); P.mpStride = mps->mp_stride; P.th = th;
    P.thGood = thGood; P.goodMode = goodMode;
    int32_t* dMp = c.out(frame_mappoints, (size_t)batch * F->kp_stride, true);
    int32_t* dSc = match_score ? c.out(match_score, (size_t)batch * F->kp_stride, true)
                               : (int32_t*)M->arena.take((size_t)batch * F->kp_stride * sizeof(int32_t));
    if (!dSc) return GFORB_ERR_CAPACITY;
    int32_t* dN = c.out(nmatches, (size_t)batch, false);
    if (c.rc) return c.rc;
    build_grid(M, c, T, G, batch);
    run_candidates<M_MAP>(M, c, P, T, G, batch, mps->mp_stride);
    RParams R;
    memset(&R, 0, sizeof(R));
    R.src = T; R.tgt = T; R.pool = M->d_pool; R.poolCap = M->poolCap; R.qMeta = M->d_qMeta; R.maxQ = M->cfg.max_queries;
    R.maxKp = M->cfg.max_keypoints; R.qCounts = P.mpCounts; R.nnratio = nnratio; R.checkOri = 0;
    R.tgtMp = dMp; R.tgtScore = dSc; R.nmatches = dN; R.rotBinQ = M->d_rotBinQ;
    run_resolve<M_MAP>(M, c, R, batch);
    return end_call(M, c, flags);
}

int gforb_search_by_projection(GforbMatcher* M, const GforbFrameSet* F, const GforbMapPointSet* mps, int batch,
                               float th, float nnratio, int32_t* frame_mappoints, int32_t* match_score,
                               int32_t* nmatches, int flags, void* stream) {
    return sbp_map(M, F, mps, batch, th, 0.0f, 0, nnratio, frame_mappoints, match_score, nmatches, flags, stream);
}

int gforb_search_by_projection_good_feature(GforbMatcher* M, const GforbFrameSet* F, const GforbMapPointSet* mps, int batch,
                                            float th_good, float th_rest, float nnratio, int32_t* frame_mappoints,
                                            int32_t* nmatches, int flags, void* stream) {
    return sbp_map(M, F, mps, batch, th_rest, th_good, 1, nnratio, frame_mappoints, nullptr, nmatches, flags, stream);
}

int gforb_obs_active_match(GforbObs* O, GforbMatcher* M, const GforbFrameSet* F, const GforbMapPointSet* mps, int batch,
                           const double* obs_score, const double* obs_mat, const double* H_meas, const double* base_info,
                           const float* level_sigma2, int nlevels, int mat_type, float th, float nnratio, int num_to_match,
                           const uint32_t* seeds, int32_t* frame_mappoints, int32_t* match_score, int32_t* nmatched,
                           int32_t* left_idx, int32_t* left_n, double* out_info, int flags, void* stream) {
    Ctx c;
    int rc = begin_call(M, batch, flags, stream, c);
    if (rc) return rc;
    if (!O) return GFORB_ERR_ARG;
    if ((rc = check_frames(M, F))) return rc;
    if (!mps || !mps->proj_x || !mps->proj_y || !mps->scale_level || !mps->view_cos || !mps->flags || !mps->descriptors ||
        !mps->counts || mps->mp_stride <= 0 || !frame_mappoints || !match_score || !nmatched || !left_idx || !left_n ||
        !obs_mat || !H_meas || !base_info || !level_sigma2 || nlevels <= 0)
        return GFORB_ERR_ARG;
    // ORB_SLAM::MAP_INFO_MATRIX = 0, FRAME_INFO_MATRIX = 1; the hybrid types (2, 3) multiply the 2x7 Jacobian by the
    // 13x13 kinematic F and exist only for DIMENSION_OF_STATE_MODEL = 13 (include/Observability.h:45, 94-102)
    if (mat_type != 0 && mat_type != 1) return GFORB_ERR_ARG;
    if (mps->mp_stride > M->cfg.max_queries) return GFORB_ERR_CAPACITY;
    if (batch == 0) return GFORB_OK;
    FrameDev T = c.frame(F, batch);
    GridGeom G = geom_of(F);
    QParams P;
    memset(&P, 0, sizeof(P));
    fill_scale(M, P);
    const size_t nm = (size_t)batch * mps->mp_stride;
    P.mpU = c.in(mps->proj_x, nm); P.mpV = c.in(mps->proj_y, nm); P.mpCos = c.in(mps->view_cos, nm);
    P.mpLevel = c.in(mps->scale_level, nm); P.mpFlags = c.in(mps->flags, nm); P.mpDesc = c.in(mps->descriptors, nm * 32);
    P.mpCounts = c.in(mps->counts, (size_t)batch); P.mpStride = mps->mp_stride; P.th = th;
    AParams A;
    memset(&A, 0, sizeof(A));
    A.obsScore = c.in(obs_score, nm);
    A.blocks = c.in(obs_mat, nm * 49);
    A.Hmeas = c.in(H_meas, nm * 14);
    A.baseInfo = c.in(base_info, (size_t)batch * 49);
    A.levelSigma2 = c.in(level_sigma2, (size_t)nlevels);
    A.seeds = c.in(seeds, (size_t)batch);
    A.tgtMp = c.out(frame_mappoints, (size_t)batch * F->kp_stride, true);
    A.tgtScore = c.out(match_score, (size_t)batch * F->kp_stride, true);
    A.nmatched = c.out(nmatched, (size_t)batch, false);
    A.leftIdx = c.out(left_idx, nm, false);
    A.leftN = c.out(left_n, (size_t)batch, false);
    A.outInfo = c.out(out_info, (size_t)batch * 49, false);
    if (c.rc) return c.rc;
    const size_t S = (size_t)M->cfg.max_batch * M->cfg.max_queries;
    if (!M->d_active) GF_CUDA_OK(cudaMalloc((void**)&M->d_active, S * (5 * sizeof(int) + 2 * sizeof(double) + 1)));
    const size_t n = (size_t)M->cfg.max_batch * mps->mp_stride <= S ? S : 0;
    if (!n) return GFORB_ERR_CAPACITY;
    A.sHeapVal = (double*)M->d_active; A.sListVal = A.sHeapVal + S;
    A.sIdx = (int*)(A.sListVal + S); A.sVis = A.sIdx + S; A.sHeapIdx = A.sVis + S; A.sList = A.sHeapIdx + S;
    A.sRemove = A.sList + S; A.sMark = (uint8_t*)(A.sRemove + S);
    build_grid(M, c, T, G, batch);
    run_candidates<M_MAP>(M, c, P, T, G, batch, mps->mp_stride);
    A.tgt = T; A.pool = M->d_pool; A.poolCap = M->poolCap; A.qMeta = M->d_qMeta; A.maxQ = M->cfg.max_queries;
    A.mpFlags = P.mpFlags; A.mpCounts = P.mpCounts; A.mpStride = mps->mp_stride;
    A.nlevels = nlevels; A.nnratio = nnratio; A.numToMatch = num_to_match;
    k_active_match<<<batch, 32, 0, c.st>>>(A);
    M->lastLaunches++;
    return end_call(M, c, flags);
}

int gforb_frame_is_in_frustum(GforbMatcher* M, int batch, const float* Rcw, const float* tcw, const float* Ow, float fx,
                              float fy, float cx, float cy, int min_x, int max_x, int min_y, int max_y, const float* world,
                              const float* normal, const float* min_distance, const float* max_distance,
                              const int32_t* counts, int mp_stride, float viewing_cos_limit, float* proj_x, float* proj_y,
                              int32_t* scale_level, float* view_cos, uint8_t* mp_flags, int flags, void* stream) {
    Ctx c;
    int rc = begin_call(M, batch, flags, stream, c);
    if (rc) return rc;
    if (!Rcw || !tcw || !Ow || !world || !normal || !min_distance || !max_distance || !counts || mp_stride <= 0 || !proj_x ||
        !proj_y || !scale_level || !view_cos || !mp_flags)
        return GFORB_ERR_ARG;
    if (mp_stride > M->cfg.max_queries) return GFORB_ERR_CAPACITY;
    if (batch == 0) return GFORB_OK;
    const size_t nm = (size_t)batch * mp_stride;
    FrustumParams F;
    memset(&F, 0, sizeof(F));
    F.Rcw = c.in(Rcw, (size_t)batch * 9); F.tcw = c.in(tcw, (size_t)batch * 3); F.Ow = c.in(Ow, (size_t)batch * 3);
    F.fx = fx; F.fy = fy; F.cx = cx; F.cy = cy; F.minX = min_x; F.maxX = max_x; F.minY = min_y; F.maxY = max_y;
    F.world = c.in(world, nm * 3); F.normal = c.in(normal, nm * 3);
    F.minDist = c.in(min_distance, nm); F.maxDist = c.in(max_distance, nm);
    F.counts = c.in(counts, (size_t)batch); F.stride = mp_stride; F.cosLimit = viewing_cos_limit;
    for (int i = 0; i < MAX_LEVELS; i++) F.scale[i] = M->scale[i < M->cfg.nlevels ? i : M->cfg.nlevels - 1];
    F.nlevels = M->cfg.nlevels;
    F.u = c.out(proj_x, nm, true); F.v = c.out(proj_y, nm, true); F.level = c.out(scale_level, nm, true);
    F.viewCos = c.out(view_cos, nm, true); F.flags = c.out(mp_flags, nm, true);
    if (c.rc) return c.rc;
    k_is_in_frustum<<<dim3((mp_stride + 127) / 128, batch), 128, 0, c.st>>>(F);
    M->lastLaunches++;
    return end_call(M, c, flags);
}

int gforb_frame_undistort_keypoints(GforbMatcher* M, int batch, const GforbKeyPoint* keys, const int32_t* counts,
                                    int kp_stride, float fx, float fy, float cx, float cy, const float* dist_coef,
                                    int n_dist, GforbKeyPoint* keys_un, int flags, void* stream) {
    Ctx c;
    int rc = begin_call(M, batch, flags, stream, c);
    if (rc) return rc;
    if (!keys || !counts || !keys_un || kp_stride <= 0 || n_dist < 0 || n_dist > 8 || (n_dist > 0 && !dist_coef)) return GFORB_ERR_ARG;
    if (kp_stride > M->cfg.max_keypoints) return GFORB_ERR_CAPACITY;
    if (batch == 0) return GFORB_OK;
    const size_t nk = (size_t)batch * kp_stride;
    UndistortParams P;
    memset(&P, 0, sizeof(P));
    P.keys = c.in(keys, nk); P.counts = c.in(counts, (size_t)batch); P.keysUn = c.out(keys_un, nk, true);     // in/out: entries past counts[b] keep their contents
    P.stride = kp_stride;
    P.fx = fx; P.fy = fy; P.cx = cx; P.cy = cy;
    for (int i = 0; i < n_dist; i++) P.k[i] = dist_coef[i];          // host array (Frame::mDistCoef, CV_32F)
    P.identity = (n_dist == 0 || dist_coef[0] == 0.0f) ? 1 : 0;
    if (c.rc) return c.rc;
    k_undistort<<<dim3((kp_stride + 127) / 128, batch), 128, 0, c.st>>>(P);
    M->lastLaunches++;
    return end_call(M, c, flags);
}

// shared body of SearchByProjection(Cur, Last, th) and SearchByProjection_GoodFeature(Cur, Last, th_good, th_rest)
static int sbp_last(GforbMatcher* M, const GforbFrameSet* cur, const GforbFrameSet* last, int batch,
                    const int32_t* last_mappoints, const uint8_t* last_outlier, const uint8_t* last_good, const float* last_world,
                    const float* Tcw, float fx, float fy, float cx, float cy, float th, float thGood, int goodMode, float nnratio,
                    int check_orientation, int32_t* cur_mappoints, int32_t* cur_score, int32_t* nmatches, int flags,
                    void* stream) {
    Ctx c;
    int rc = begin_call(M, batch, flags, stream, c);
    if (rc) return rc;
    if ((rc = check_frames(M, cur)) || (rc = check_frames(M, last))) return rc;
    if (!last_mappoints || !last_world || !Tcw || !cur_mappoints || (!cur_score && !goodMode) || (goodMode && !last_good) || !nmatches)
        return GFORB_ERR_ARG;
    if (last->kp_stride > M->cfg.max_queries) return GFORB_ERR_CAPACITY;
    if (batch == 0) return GFORB_OK;
    FrameDev T = c.frame(cur, batch), S = c.frame(last, batch);
    GridGeom G = geom_of(cur);
    QParams P;
    memset(&P, 0, sizeof(P));
    fill_scale(M, P);
    const size_t ns = (size_t)batch * last->kp_stride;
    P.src = S; P.srcMp = c.in(last_mappoints, ns); P.srcOutlier = c.in(last_outlier, ns);
    P.srcWorld = c.in(last_world, ns * 3); P.Tcw = c.in(Tcw, (size_t)batch * 16);
    P.fx = fx; P.fy = fy; P.cx = cx; P.cy = cy; P.th = th;
    P.thGood = thGood; P.goodMode = goodMode; P.srcGood = goodMode ? c.in(last_good, ns) : nullptr;
    int32_t* dMp = c.out(cur_mappoints, (size_t)batch * cur->kp_stride, true);
    int32_t* dSc = cur_score ? c.out(cur_score, (size_t)batch * cur->kp_stride, true)
                             : (int32_t*)M->arena.take((size_t)batch * cur->kp_stride * sizeof(int32_t));
    if (!dSc) return GFORB_ERR_CAPACITY;
    int32_t* dN = c.out(nmatches, (size_t)batch, false);
    if (c.rc) return c.rc;
    build_grid(M, c, T, G, batch);
    run_candidates<M_LAST>(M, c, P, T, G, batch, last->kp_stride);
    RParams R;
    memset(&R, 0, sizeof(R));
    R.src = S; R.tgt = T; R.pool = M->d_pool; R.poolCap = M->poolCap; R.qMeta = M->d_qMeta; R.maxQ = M->cfg.max_queries;
    R.maxKp = M->cfg.max_keypoints; R.qCounts = S.counts; R.srcMp = P.srcMp; R.nnratio = nnratio;
    R.checkOri = check_orientation; R.tgtMp = dMp; R.tgtScore = dSc; R.nmatches = dN; R.rotBinQ = M->d_rotBinQ;
    R.thHigh = TH_HIGH;
    run_resolve<M_LAST>(M, c, R, batch);
    return end_call(M, c, flags);
}

int gforb_search_by_projection_frames(GforbMatcher* M, const GforbFrameSet* cur, const GforbFrameSet* last, int batch,
                                      const int32_t* last_mappoints, const uint8_t* last_outlier,
                                      const float* last_world, const float* Tcw, float fx, float fy, float cx, float cy,
                                      float th, float nnratio, int check_orientation, int32_t* cur_mappoints,
                                      int32_t* cur_score, int32_t* nmatches, int flags, void* stream) {
    return sbp_last(M, cur, last, batch, last_mappoints, last_outlier, nullptr, last_world, Tcw, fx, fy, cx, cy, th, 0.0f, 0, nnratio,
                    check_orientation, cur_mappoints, cur_score, nmatches, flags, stream);
}

int gforb_search_by_projection_frames_good_feature(GforbMatcher* M, const GforbFrameSet* cur, const GforbFrameSet* last, int batch,
                                                   const int32_t* last_mappoints, const uint8_t* last_outlier,
                                                   const uint8_t* last_good, const float* last_world, const float* Tcw, float fx,
                                                   float fy, float cx, float cy, float th_good, float th_rest, float nnratio,
                                                   int check_orientation, int32_t* cur_mappoints, int32_t* nmatches, int flags,
                                                   void* stream) {
    return sbp_last(M, cur, last, batch, last_mappoints, last_outlier, last_good, last_world, Tcw, fx, fy, cx, cy, th_rest, th_good, 1,
                    nnratio, check_orientation, cur_mappoints, nullptr, nmatches, flags, stream);
}

int gforb_window_search(GforbMatcher* M, const GforbFrameSet* F1, const GforbFrameSet* F2, int batch,
                        const int32_t* f1_mappoints, int window_size, int min_scale_level, int max_scale_level,
                        float nnratio, int check_orientation, int32_t* matches2, int32_t* nmatches, int flags,
                        void* stream) {
    Ctx c;
    int rc = begin_call(M, batch, flags, stream, c);
    if (rc) return rc;
    if ((rc = check_frames(M, F1)) || (rc = check_frames(M, F2))) return rc;
    if (!f1_mappoints || !matches2 || !nmatches) return GFORB_ERR_ARG;
    if (F1->kp_stride > M->cfg.max_queries) return GFORB_ERR_CAPACITY;
    if (batch == 0) return GFORB_OK;
    FrameDev S = c.frame(F1, batch), T = c.frame(F2, batch);
    GridGeom G = geom_of(F2);
    QParams P;
    memset(&P, 0, sizeof(P));
    fill_scale(M, P);
    P.src = S; P.srcMp = c.in(f1_mappoints, (size_t)batch * F1->kp_stride);
    P.window = (float)window_size; P.minScaleLevel = min_scale_level; P.maxScaleLevel = max_scale_level;
    int32_t* dM2 = c.out(matches2, (size_t)batch * F2->kp_stride, false);
    int32_t* dN = c.out(nmatches, (size_t)batch, false);
    if (c.rc) return c.rc;
    build_grid(M, c, T, G, batch);
    run_candidates<M_WINDOW>(M, c, P, T, G, batch, F1->kp_stride);
    RParams R;
    memset(&R, 0, sizeof(R));
    R.src = S; R.tgt = T; R.pool = M->d_pool; R.poolCap = M->poolCap; R.qMeta = M->d_qMeta; R.maxQ = M->cfg.max_queries;
    R.maxKp = M->cfg.max_keypoints; R.qCounts = S.counts; R.srcMp = P.srcMp; R.nnratio = nnratio;
    R.checkOri = check_orientation; R.tgtMp = dM2; R.nmatches = dN; R.rotBinQ = M->d_rotBinQ;
    run_resolve<M_WINDOW>(M, c, R, batch);
    return end_call(M, c, flags);
}

int gforb_search_by_projection_window(GforbMatcher* M, const GforbFrameSet* F1, const GforbFrameSet* F2, int batch,
                                      const int32_t* f1_mappoints, const float* f1_world, const float* Tcw2, float fx, float fy,
                                      float cx, float cy, int window_size, float nnratio, int32_t* matches2, int32_t* f2_score,
                                      int32_t* nmatches, int flags, void* stream) {
    Ctx c;
    int rc = begin_call(M, batch, flags, stream, c);
    if (rc) return rc;
    if ((rc = check_frames(M, F1)) || (rc = check_frames(M, F2))) return rc;
    if (!f1_mappoints || !f1_world || !Tcw2 || !matches2 || !f2_score || !nmatches) return GFORB_ERR_ARG;
    if (F1->kp_stride > M->cfg.max_queries) return GFORB_ERR_CAPACITY;
    if (batch == 0) return GFORB_OK;
    FrameDev S = c.frame(F1, batch), T = c.frame(F2, batch);
    GridGeom G = geom_of(F2);
    QParams P;
    memset(&P, 0, sizeof(P));
    fill_scale(M, P);
    P.src = S; P.srcMp = c.in(f1_mappoints, (size_t)batch * F1->kp_stride);
    P.srcWorld = c.in(f1_world, (size_t)batch * F1->kp_stride * 3);
    P.Tcw = c.in(Tcw2, (size_t)batch * 16);
    P.fx = fx; P.fy = fy; P.cx = cx; P.cy = cy; P.window = (float)window_size;
    int32_t* dM2 = c.out(matches2, (size_t)batch * F2->kp_stride, true);       // in: F2.mvpMapPoints
    int32_t* dSc = c.out(f2_score, (size_t)batch * F2->kp_stride, true);
    int32_t* dN = c.out(nmatches, (size_t)batch, false);
    if (c.rc) return c.rc;
    P.tgtMp0 = dM2;            // read by k_candidates before k_resolve (same stream) starts to write it
    build_grid(M, c, T, G, batch);
    run_candidates<M_PROJ2>(M, c, P, T, G, batch, F1->kp_stride);
    RParams R;
    memset(&R, 0, sizeof(R));
    R.src = S; R.tgt = T; R.pool = M->d_pool; R.poolCap = M->poolCap; R.qMeta = M->d_qMeta; R.maxQ = M->cfg.max_queries;
    R.maxKp = M->cfg.max_keypoints; R.qCounts = S.counts; R.srcMp = P.srcMp; R.nnratio = nnratio;
    R.tgtMp = dM2; R.tgtScore = dSc; R.nmatches = dN; R.rotBinQ = M->d_rotBinQ;
    run_resolve<M_PROJ2>(M, c, R, batch);
    return end_call(M, c, flags);
}

int gforb_search_by_bow(GforbMatcher* M, const GforbFrameSet* KF, const int32_t* kf_mappoints, const int32_t* kf_nodes,
                        const GforbFrameSet* F, const int32_t* f_nodes, int batch, float nnratio, int check_orientation,
                        int32_t* matches, int32_t* nmatches, int flags, void* stream) {
    Ctx c;
    int rc = begin_call(M, batch, flags, stream, c);
    if (rc) return rc;
    if ((rc = check_frames(M, KF)) || (rc = check_frames(M, F))) return rc;
    if (!kf_mappoints || !kf_nodes || !f_nodes || !matches || !nmatches) return GFORB_ERR_ARG;
    if (KF->kp_stride > M->cfg.max_queries) return GFORB_ERR_CAPACITY;
    if (batch == 0) return GFORB_OK;
    FrameDev S = c.frame(KF, batch), T = c.frame(F, batch);
    GridGeom G = geom_of(F);
    QParams P;
    memset(&P, 0, sizeof(P));
    fill_scale(M, P);
    P.src = S; P.srcMp = c.in(kf_mappoints, (size_t)batch * KF->kp_stride);
    P.srcNode = c.in(kf_nodes, (size_t)batch * KF->kp_stride);
    P.tgtNode = c.in(f_nodes, (size_t)batch * F->kp_stride);
    P.order = M->d_rotBinQ;                        // the INIT scratch doubles as the walk order of this call
    int32_t* dM = c.out(matches, (size_t)batch * F->kp_stride, false);
    int32_t* dN = c.out(nmatches, (size_t)batch, false);
    if (c.rc) return c.rc;
    run_candidates<M_BOW>(M, c, P, T, G, batch, KF->kp_stride);     // no grid: candidates come from the node lists
    RParams R;
    memset(&R, 0, sizeof(R));
    R.src = S; R.tgt = T; R.pool = M->d_pool; R.poolCap = M->poolCap; R.qMeta = M->d_qMeta; R.maxQ = M->cfg.max_queries;
    R.maxKp = M->cfg.max_keypoints; R.qCounts = S.counts; R.srcMp = P.srcMp; R.nnratio = nnratio;
    R.checkOri = check_orientation; R.tgtMp = dM; R.nmatches = dN; R.rotBinQ = M->d_rotBinQ; R.order = M->d_rotBinQ;
    run_resolve<M_BOW>(M, c, R, batch);
    return end_call(M, c, flags);
}

int gforb_search_by_projection_keyframe(GforbMatcher* M, const GforbFrameSet* cur, const GforbFrameSet* KF, int batch,
                                        const int32_t* kf_mappoints, const uint8_t* kf_skip, const float* kf_world,
                                        const float* kf_min_distance, const uint8_t* kf_mp_descriptors, const float* Tcw,
                                        float fx, float fy, float cx, float cy, float th, int orb_dist, int check_orientation,
                                        int32_t* cur_mappoints, int32_t* cur_score, int32_t* nmatches, int flags, void* stream) {
    Ctx c;
    int rc = begin_call(M, batch, flags, stream, c);
    if (rc) return rc;
    if ((rc = check_frames(M, cur)) || (rc = check_frames(M, KF))) return rc;
    if (!kf_mappoints || !kf_world || !kf_min_distance || !kf_mp_descriptors || !Tcw || !cur_mappoints || !cur_score || !nmatches)
        return GFORB_ERR_ARG;
    if (KF->kp_stride > M->cfg.max_queries) return GFORB_ERR_CAPACITY;
    if (batch == 0) return GFORB_OK;
    FrameDev T = c.frame(cur, batch), S = c.frame(KF, batch);
    GridGeom G = geom_of(cur);
    QParams P;
    memset(&P, 0, sizeof(P));
    fill_scale(M, P);
    const size_t ns = (size_t)batch * KF->kp_stride;
    P.src = S; P.srcMp = c.in(kf_mappoints, ns); P.srcOutlier = c.in(kf_skip, ns);
    P.srcWorld = c.in(kf_world, ns * 3); P.srcMinDist = c.in(kf_min_distance, ns); P.qDesc = c.in(kf_mp_descriptors, ns * 32);
    P.Tcw = c.in(Tcw, (size_t)batch * 16);
    P.fx = fx; P.fy = fy; P.cx = cx; P.cy = cy; P.th = th;
    int32_t* dMp = c.out(cur_mappoints, (size_t)batch * cur->kp_stride, true);
    int32_t* dSc = c.out(cur_score, (size_t)batch * cur->kp_stride, true);
    int32_t* dN = c.out(nmatches, (size_t)batch, false);
    if (c.rc) return c.rc;
    build_grid(M, c, T, G, batch);
    run_candidates<M_KF>(M, c, P, T, G, batch, KF->kp_stride);
    RParams R;                                     // the walk, the claims and the rotation histogram are those of (Cur, Last)
    memset(&R, 0, sizeof(R));
    R.src = S; R.tgt = T; R.pool = M->d_pool; R.poolCap = M->poolCap; R.qMeta = M->d_qMeta; R.maxQ = M->cfg.max_queries;
    R.maxKp = M->cfg.max_keypoints; R.qCounts = S.counts; R.srcMp = P.srcMp; R.nnratio = 0.0f;
    R.checkOri = check_orientation; R.tgtMp = dMp; R.tgtScore = dSc; R.nmatches = dN; R.rotBinQ = M->d_rotBinQ;
    R.thHigh = orb_dist;
    run_resolve<M_LAST>(M, c, R, batch);
    return end_call(M, c, flags);
}

}  // extern "C"

// ---- the overloads of the mapping / loop-closing threads
// shared body of SearchByBoW(pKF1, pKF2, vpMatches12) and SearchForTriangulation: the walk of gforb_search_by_bow with the
// keyframe-pair filters and outputs per KF1 feature
template <int MODE>
static int bow_pair(GforbMatcher* M, const GforbFrameSet* KF1, const int32_t* kf1_mappoints, const int32_t* kf1_nodes,
                    const GforbFrameSet* KF2, const int32_t* kf2_mappoints, const int32_t* kf2_nodes, int batch, const float* F12,
                    float nnratio, int check_orientation, int32_t* matches12, int32_t* nmatches, int flags, void* stream) {
    Ctx c;
    int rc = begin_call(M, batch, flags, stream, c);
    if (rc) return rc;
    if ((rc = check_frames(M, KF1)) || (rc = check_frames(M, KF2))) return rc;
    if (!kf1_mappoints || !kf1_nodes || !kf2_mappoints || !kf2_nodes || !matches12 || !nmatches || (MODE == M_TRI && !F12))
        return GFORB_ERR_ARG;
    if (KF1->kp_stride > M->cfg.max_queries) return GFORB_ERR_CAPACITY;
    if (batch == 0) return GFORB_OK;
    FrameDev S = c.frame(KF1, batch), T = c.frame(KF2, batch);
    GridGeom G = geom_of(KF2);
    QParams P;
    memset(&P, 0, sizeof(P));
    fill_scale(M, P);
    P.src = S; P.srcMp = c.in(kf1_mappoints, (size_t)batch * KF1->kp_stride);
    P.srcNode = c.in(kf1_nodes, (size_t)batch * KF1->kp_stride);
    P.tgtNode = c.in(kf2_nodes, (size_t)batch * KF2->kp_stride);
    P.tgtMp = c.in(kf2_mappoints, (size_t)batch * KF2->kp_stride);
    P.F12 = (MODE == M_TRI) ? c.in(F12, (size_t)batch * 9) : nullptr;
    P.order = M->d_rotBinQ;
    int32_t* dM = c.out(matches12, (size_t)batch * KF1->kp_stride, false);
    int32_t* dN = c.out(nmatches, (size_t)batch, false);
    if (c.rc) return c.rc;
    run_candidates<MODE>(M, c, P, T, G, batch, KF1->kp_stride);
    RParams R;
    memset(&R, 0, sizeof(R));
    R.src = S; R.tgt = T; R.pool = M->d_pool; R.poolCap = M->poolCap; R.qMeta = M->d_qMeta; R.maxQ = M->cfg.max_queries;
    R.maxKp = M->cfg.max_keypoints; R.qCounts = S.counts; R.srcMp = P.srcMp; R.nnratio = nnratio;
    R.checkOri = check_orientation; R.matches12 = dM; R.nmatches = dN; R.rotBinQ = M->d_rotBinQ; R.order = M->d_rotBinQ;
    R.tgtMpIds = P.tgtMp;
    run_resolve<MODE>(M, c, R, batch);
    return end_call(M, c, flags);
}

extern "C" {

int gforb_search_by_bow_keyframes(GforbMatcher* M, const GforbFrameSet* KF1, const int32_t* kf1_mappoints, const int32_t* kf1_nodes,
                                  const GforbFrameSet* KF2, const int32_t* kf2_mappoints, const int32_t* kf2_nodes, int batch,
                                  float nnratio, int check_orientation, int32_t* matches12, int32_t* nmatches, int flags,
                                  void* stream) {
    return bow_pair<M_BOW2>(M, KF1, kf1_mappoints, kf1_nodes, KF2, kf2_mappoints, kf2_nodes, batch, nullptr, nnratio, check_orientation,
                            matches12, nmatches, flags, stream);
}

int gforb_search_for_triangulation(GforbMatcher* M, const GforbFrameSet* KF1, const int32_t* kf1_mappoints, const int32_t* kf1_nodes,
                                   const GforbFrameSet* KF2, const int32_t* kf2_mappoints, const int32_t* kf2_nodes, int batch,
                                   const float* F12, int check_orientation, int32_t* matches12, int32_t* nmatches, int flags,
                                   void* stream) {
    return bow_pair<M_TRI>(M, KF1, kf1_mappoints, kf1_nodes, KF2, kf2_mappoints, kf2_nodes, batch, F12, 0.0f, check_orientation,
                           matches12, nmatches, flags, stream);
}

namespace {
int check_points(GforbMatcher* M, const GforbWorldPointSet* p, bool needNormal) {
    if (!p || !p->world || !p->min_distance || !p->max_distance || !p->descriptors || !p->counts || p->mp_stride <= 0) return GFORB_ERR_ARG;
    if (needNormal && !p->normal) return GFORB_ERR_ARG;
    if (p->mp_stride > M->cfg.max_queries) return GFORB_ERR_CAPACITY;
    return GFORB_OK;
}
void stage_points(Ctx& c, QParams& P, const GforbWorldPointSet* p, int batch, bool withNormal) {
    const size_t nm = (size_t)batch * p->mp_stride;
    P.pWorld = c.in(p->world, nm * 3);
    P.pNormal = withNormal ? c.in(p->normal, nm * 3) : nullptr;
    P.pMinD = c.in(p->min_distance, nm); P.pMaxD = c.in(p->max_distance, nm);
    P.mpDesc = c.in(p->descriptors, nm * 32);
    P.pSkip = c.in(p->skip, nm);
    P.mpCounts = c.in(p->counts, (size_t)batch); P.mpStride = p->mp_stride;
}
void run_pick(GforbMatcher* M, Ctx& c, const int32_t* qCounts, int batch, int outStride, int thrIdx, int thrCount, int32_t* outIdx,
              int32_t* outDist, int32_t* nAccepted) {
    k_pick_best<<<dim3((outStride + 127) / 128, batch), 128, 0, c.st>>>(M->d_pool, M->poolCap, M->d_qMeta, M->cfg.max_queries, qCounts,
                                                                         outStride, thrIdx, thrCount, outIdx, outDist, nAccepted);
    M->lastLaunches++;
}
}  // namespace

int gforb_fuse(GforbMatcher* M, const GforbFrameSet* KF, int batch, const float* pose, int is_sim3, float fx, float fy, float cx,
               float cy, const GforbWorldPointSet* points, float th, int32_t* best_idx, int32_t* best_dist, int32_t* nfused,
               int flags, void* stream) {
    Ctx c;
    int rc = begin_call(M, batch, flags, stream, c);
    if (rc) return rc;
    if ((rc = check_frames(M, KF)) || (rc = check_points(M, points, true))) return rc;
    if (!pose || !best_idx || !best_dist || !nfused) return GFORB_ERR_ARG;
    if (batch == 0) return GFORB_OK;
    FrameDev T = c.frame(KF, batch);
    GridGeom G = geom_of(KF);
    QParams P;
    memset(&P, 0, sizeof(P));
    fill_scale(M, P);
    stage_points(c, P, points, batch, true);
    P.Tcw = c.in(pose, (size_t)batch * 16); P.pkfMode = is_sim3 ? 1 : 0;
    P.fx = fx; P.fy = fy; P.cx = cx; P.cy = cy; P.th = th;
    const size_t nm = (size_t)batch * points->mp_stride;
    int32_t* dIdx = c.out(best_idx, nm, false);
    int32_t* dDist = c.out(best_dist, nm, false);
    int32_t* dN = c.out(nfused, (size_t)batch, false);
    if (c.rc) return c.rc;
    GF_CUDA_OK(cudaMemsetAsync(dN, 0, sizeof(int32_t) * batch, c.st));
    build_grid(M, c, T, G, batch);
    run_candidates<M_PKF>(M, c, P, T, G, batch, points->mp_stride);
    run_pick(M, c, P.mpCounts, batch, points->mp_stride, INT_MAX - 1, TH_LOW, dIdx, dDist, dN);
    return end_call(M, c, flags);
}

int gforb_search_by_projection_sim3(GforbMatcher* M, const GforbFrameSet* KF, int batch, const float* Scw, float fx, float fy,
                                    float cx, float cy, const GforbWorldPointSet* points, const int32_t* point_ids, float th,
                                    int32_t* matched, int32_t* nmatches, int flags, void* stream) {
    Ctx c;
    int rc = begin_call(M, batch, flags, stream, c);
    if (rc) return rc;
    if ((rc = check_frames(M, KF)) || (rc = check_points(M, points, true))) return rc;
    if (!Scw || !point_ids || !matched || !nmatches) return GFORB_ERR_ARG;
    if (batch == 0) return GFORB_OK;
    FrameDev T = c.frame(KF, batch);
    GridGeom G = geom_of(KF);
    QParams P;
    memset(&P, 0, sizeof(P));
    fill_scale(M, P);
    stage_points(c, P, points, batch, true);
    P.Tcw = c.in(Scw, (size_t)batch * 16); P.pkfMode = 1;
    P.fx = fx; P.fy = fy; P.cx = cx; P.cy = cy; P.th = th;
    const int32_t* dIds = c.in(point_ids, (size_t)batch * points->mp_stride);
    int32_t* dM = c.out(matched, (size_t)batch * KF->kp_stride, true);
    int32_t* dSc = (int32_t*)M->arena.take((size_t)batch * KF->kp_stride * sizeof(int32_t));     // distances of the claims (not returned)
    if (!dSc) return GFORB_ERR_CAPACITY;
    int32_t* dN = c.out(nmatches, (size_t)batch, false);
    if (c.rc) return c.rc;
    build_grid(M, c, T, G, batch);
    run_candidates<M_PKF>(M, c, P, T, G, batch, points->mp_stride);
    RParams R;                                     // claims as in (Cur, Last): best unclaimed candidate, threshold TH_LOW, no histogram
    memset(&R, 0, sizeof(R));
    R.src = T; R.src.stride = points->mp_stride; R.tgt = T; R.pool = M->d_pool; R.poolCap = M->poolCap; R.qMeta = M->d_qMeta;
    R.maxQ = M->cfg.max_queries; R.maxKp = M->cfg.max_keypoints; R.qCounts = P.mpCounts; R.srcMp = dIds; R.checkOri = 0;
    R.tgtMp = dM; R.tgtScore = dSc; R.nmatches = dN; R.rotBinQ = M->d_rotBinQ; R.thHigh = TH_LOW;
    run_resolve<M_LAST>(M, c, R, batch);
    return end_call(M, c, flags);
}

int gforb_search_by_sim3(GforbMatcher* M, const GforbFrameSet* KF1, const GforbFrameSet* KF2, int batch,
                         const GforbWorldPointSet* points1, const GforbWorldPointSet* points2, const int32_t* kf2_mappoints,
                         const float* T1w, const float* T2w, const float* s12, const float* R12, const float* t12, float fx, float fy,
                         float cx, float cy, float th, int32_t* matches12, int32_t* nfound, int flags, void* stream) {
    Ctx c;
    int rc = begin_call(M, batch, flags, stream, c);
    if (rc) return rc;
    if ((rc = check_frames(M, KF1)) || (rc = check_frames(M, KF2)) || (rc = check_points(M, points1, false)) ||
        (rc = check_points(M, points2, false)))
        return rc;
    if (!kf2_mappoints || !T1w || !T2w || !s12 || !R12 || !t12 || !matches12 || !nfound) return GFORB_ERR_ARG;
    if (points1->mp_stride != KF1->kp_stride || points2->mp_stride != KF2->kp_stride) return GFORB_ERR_ARG;
    if (batch == 0) return GFORB_OK;
    FrameDev F1 = c.frame(KF1, batch), F2 = c.frame(KF2, batch);
    const float* dT1 = c.in(T1w, (size_t)batch * 16);
    const float* dT2 = c.in(T2w, (size_t)batch * 16);
    const float* dS = c.in(s12, (size_t)batch);
    const float* dR = c.in(R12, (size_t)batch * 9);
    const float* dt = c.in(t12, (size_t)batch * 3);
    const int32_t* dMp2 = c.in(kf2_mappoints, (size_t)batch * KF2->kp_stride);
    int32_t* dM12 = c.out(matches12, (size_t)batch * KF1->kp_stride, true);
    int32_t* dN = c.out(nfound, (size_t)batch, false);
    int32_t* vn1 = (int32_t*)M->arena.take((size_t)batch * KF1->kp_stride * sizeof(int32_t));
    int32_t* vn2 = (int32_t*)M->arena.take((size_t)batch * KF2->kp_stride * sizeof(int32_t));
    if (!vn1 || !vn2) return GFORB_ERR_CAPACITY;
    QParams P1, P2;
    memset(&P1, 0, sizeof(P1)); memset(&P2, 0, sizeof(P2));
    fill_scale(M, P1); fill_scale(M, P2);
    stage_points(c, P1, points1, batch, false);
    stage_points(c, P2, points2, batch, false);
    if (c.rc) return c.rc;
    P1.Tcw = dT1; P1.pkfMode = 2; P2.Tcw = dT2; P2.pkfMode = 3;
    for (QParams* P : {&P1, &P2}) { P->s12 = dS; P->R12 = dR; P->t12 = dt; P->fx = fx; P->fy = fy; P->cx = cx; P->cy = cy; P->th = th; }
    GF_CUDA_OK(cudaMemsetAsync(dN, 0, sizeof(int32_t) * batch, c.st));
    // KF1's map points into KF2 (:1893-1973), then KF2's into KF1 (:1976-2058): the grid, pool and query scratch are reused
    // one direction after the other on the stream
    const GridGeom G2 = geom_of(KF2), G1 = geom_of(KF1);
    build_grid(M, c, F2, G2, batch);
    run_candidates<M_PKF>(M, c, P1, F2, G2, batch, KF1->kp_stride);
    run_pick(M, c, P1.mpCounts, batch, KF1->kp_stride, TH_HIGH, INT_MIN, vn1, nullptr, nullptr);
    build_grid(M, c, F1, G1, batch);
    run_candidates<M_PKF>(M, c, P2, F1, G1, batch, KF2->kp_stride);
    run_pick(M, c, P2.mpCounts, batch, KF2->kp_stride, TH_HIGH, INT_MIN, vn2, nullptr, nullptr);
    k_sim3_agree<<<dim3((KF1->kp_stride + 127) / 128, batch), 128, 0, c.st>>>(vn1, KF1->kp_stride, vn2, KF2->kp_stride, P1.mpCounts, dMp2,
                                                                              dM12, dN);
    M->lastLaunches++;
    return end_call(M, c, flags);
}

int gforb_allpairs_top2(GforbMatcher* M, const uint8_t* queries, int n_q, const uint8_t* db, int64_t n_db,
                        int64_t idx_base, uint64_t* out_keys, int flags, void* stream) {
    if (!M || !queries || !db || !out_keys || n_q < 0 || n_db < 0) return GFORB_ERR_ARG;
    if (!(flags & GFORB_IN_DEVICE) || !(flags & GFORB_OUT_DEVICE)) return GFORB_ERR_ARG;   // database lives in HBM
    if (n_q == 0) return GFORB_OK;
    GF_CUDA_OK(cudaSetDevice(M->device));
    cudaStream_t st = (cudaStream_t)stream;
    M->lastLaunches = 0;
    // GFORB_ALLPAIRS = "popc": the integer-pipe kernel (parity reference, A/B partner); default: tcgen05 kind::i8 kernel
    const char* sel = getenv("GFORB_ALLPAIRS");
    const bool usePopc = n_db == 0 || (sel && !strcmp(sel, "popc"));
    int nChunks;
    if (!usePopc) {
        const int qTiles = (n_q + tc::TC_QTILE - 1) / tc::TC_QTILE;
        const long long nTiles = (n_db + tc::TC_TILE - 1) / tc::TC_TILE;
        long long splits = std::max<long long>(1, M->numSMs / qTiles);             // one wave of (splits x qTiles) CTAs
        if (const char* e = getenv("GFORB_ALLPAIRS_SPLITS")) splits = std::max(1, atoi(e));
        long long tilesPer = std::min<long long>((nTiles + splits - 1) / splits, (1 << 21) / tc::TC_TILE);   // packed key: 2^21 rows
        tilesPer = std::max<long long>(tilesPer, 1);
        nChunks = (int)((nTiles + tilesPer - 1) / tilesPer);
        const size_t need = (size_t)nChunks * n_q * 2;
        if (need > M->capPartial) {
            if (M->d_partial) cudaFree(M->d_partial);
            M->d_partial = nullptr; M->capPartial = 0;
            GF_CUDA_OK(cudaMalloc((void**)&M->d_partial, need * sizeof(uint64_t)));
            M->capPartial = need;
        }
        GF_CUDA_OK(raise_dyn_smem(tc::k_allpairs_tc, tc::TC_SMEM_BYTES, M->device));
        tc::TcParams P;
        P.q = (const uint4*)queries; P.nq = n_q; P.db = (const uint4*)db; P.nDb = n_db; P.tilesPerSplit = (int)tilesPer;
        P.idxBase = idx_base; P.partial = M->d_partial; P.err = M->d_err;
        tc::k_allpairs_tc<<<dim3(nChunks, qTiles), tc::TC_THREADS, tc::TC_SMEM_BYTES, st>>>(P);
    } else {
        const int qTiles = (n_q + AP_THREADS * AP_R - 1) / (AP_THREADS * AP_R);
        // chunk the shard so that the grid covers every SM several times; chunk rows <= 2^22 (packed key)
        long long targetBlocks = (long long)M->numSMs * 8;
        long long chunks = std::max<long long>(1, targetBlocks / qTiles);
        long long chunkRows = (n_db + chunks - 1) / std::max<long long>(chunks, 1);
        chunkRows = std::max<long long>(AP_TILE, ((chunkRows + AP_TILE - 1) / AP_TILE) * AP_TILE);
        chunkRows = std::min<long long>(chunkRows, 1 << 22);
        nChunks = (int)std::max<long long>(1, (n_db + chunkRows - 1) / chunkRows);
        const size_t need = (size_t)nChunks * n_q * 2;
        if (need > M->capPartial) {
            if (M->d_partial) cudaFree(M->d_partial);
            M->d_partial = nullptr; M->capPartial = 0;
            GF_CUDA_OK(cudaMalloc((void**)&M->d_partial, need * sizeof(uint64_t)));
            M->capPartial = need;
        }
        k_allpairs<<<dim3(qTiles, nChunks), AP_THREADS, 0, st>>>((const uint4*)queries, n_q, (const uint4*)db, n_db,
                                                                 (int)chunkRows, idx_base, M->d_partial);
    }
    k_top2_merge<<<(n_q + 255) / 256, 256, 0, st>>>(M->d_partial, nChunks, n_q, out_keys);
    M->lastLaunches = 2;
    GF_CUDA_OK(cudaGetLastError());
    if (!(flags & GFORB_ASYNC)) {
        GF_CUDA_OK(cudaStreamSynchronize(st));
        int err = 0;
        GF_CUDA_OK(cudaMemcpy(&err, M->d_err, sizeof(int), cudaMemcpyDeviceToHost));
        if (err) { cudaMemset(M->d_err, 0, sizeof(int)); return err > 0 ? GFORB_ERR_CUDA : err; }
    }
    return GFORB_OK;
}

// error word of the asynchronous calls (GFORB_ASYNC): internal overflow / protocol errors raised by kernels since the last poll.
// Synchronises `stream`; returns GFORB_OK or the recorded GFORB_ERR_* and clears it.
int gforb_matcher_poll_error(GforbMatcher* M, void* stream) {
    if (!M) return GFORB_ERR_ARG;
    GF_CUDA_OK(cudaSetDevice(M->device));
    GF_CUDA_OK(cudaStreamSynchronize((cudaStream_t)stream));
    int err = 0;
    GF_CUDA_OK(cudaMemcpy(&err, M->d_err, sizeof(int), cudaMemcpyDeviceToHost));
    if (err) { cudaMemset(M->d_err, 0, sizeof(int)); return err > 0 ? GFORB_ERR_CUDA : err; }
    return GFORB_OK;
}

int gforb_top2_merge(GforbMatcher* M, const uint64_t* gathered, int n_shards, int n_q, uint64_t* out_keys, int flags,
                     void* stream) {
    if (!M || !gathered || !out_keys || n_shards <= 0 || n_q < 0) return GFORB_ERR_ARG;
    if (!(flags & GFORB_IN_DEVICE) || !(flags & GFORB_OUT_DEVICE)) return GFORB_ERR_ARG;
    if (n_q == 0) return GFORB_OK;
    GF_CUDA_OK(cudaSetDevice(M->device));
    cudaStream_t st = (cudaStream_t)stream;
    k_top2_merge<<<(n_q + 255) / 256, 256, 0, st>>>(gathered, n_shards, n_q, out_keys);
    M->lastLaunches = 1;
    GF_CUDA_OK(cudaGetLastError());
    if (!(flags & GFORB_ASYNC)) GF_CUDA_OK(cudaStreamSynchronize(st));
    return GFORB_OK;
}

// ------------------------------------------------------------------------------------------------ allocation-free all-pairs
// gforb_allpairs_top2_static: the tensor-core all-pairs top-2 search for callers that may not allocate (CUDA graph capture,
// serving loops, harnesses that forbid cudaMalloc).  No handle: the per-split partial keys and the kernel error word live in
// a file-scope device array that every GPU loading the library holds until exit.  The array's name is only a host-side
// stand-in in host code; the device address comes from cudaGetSymbolAddress, once per device, with that device current.
static const int SA_MAX_QUERIES = 4096;                  // queries per call
static const size_t SA_MAX_PAIRS = 65536;                // (database splits) x (queries) partial key pairs = 1 MiB
static const size_t SA_HEADER = 256;                     // error word of the kernel (the rest of the header is unused)
static __device__ __align__(256) unsigned char g_static_scratch[SA_HEADER + SA_MAX_PAIRS * 2 * sizeof(uint64_t)];

namespace {
struct StaticDev { unsigned char* ptr = nullptr; int numSMs = 0; bool ready = false; };
std::mutex g_staticMu;
StaticDev g_staticDev[64];
}  // namespace

int gforb_allpairs_top2_static(int device, const uint8_t* queries, int n_q, const uint8_t* db, int64_t n_db, int64_t idx_base,
                               uint64_t* out_keys, int flags, void* stream) {
    if (!queries || !db || !out_keys || n_q < 0 || n_db <= 0 || device < 0 || device >= 64) return GFORB_ERR_ARG;
    if (!(flags & GFORB_IN_DEVICE) || !(flags & GFORB_OUT_DEVICE)) return GFORB_ERR_ARG;   // queries, database and keys live in HBM
    if (n_q == 0) return GFORB_OK;
    if (n_q > SA_MAX_QUERIES) {
        fprintf(stderr, "[gforb] gforb_allpairs_top2_static: %d queries exceed the static limit of %d per call\n", n_q, SA_MAX_QUERIES);
        return GFORB_ERR_CAPACITY;
    }
    GF_CUDA_OK(cudaSetDevice(device));
    StaticDev* D = &g_staticDev[device];
    {
        std::lock_guard<std::mutex> g(g_staticMu);
        if (!D->ready) {                                 // set-up, once per device, with that device current
            void* p = nullptr;
            GF_CUDA_OK(cudaGetSymbolAddress(&p, g_static_scratch));
            GF_CUDA_OK(cudaDeviceGetAttribute(&D->numSMs, cudaDevAttrMultiProcessorCount, device));
            GF_CUDA_OK(raise_dyn_smem(tc::k_allpairs_tc, tc::TC_SMEM_BYTES, device));
            D->ptr = (unsigned char*)p;
            D->ready = true;
        }
    }
    cudaStream_t st = (cudaStream_t)stream;
    const int qTiles = (n_q + tc::TC_QTILE - 1) / tc::TC_QTILE;
    const long long nTiles = (n_db + tc::TC_TILE - 1) / tc::TC_TILE;
    const long long splits = std::max<long long>(1, D->numSMs / qTiles);            // one wave of (splits x qTiles) CTAs
    long long tilesPer = std::min<long long>((nTiles + splits - 1) / splits, (1 << 21) / tc::TC_TILE);   // packed key: 2^21 rows
    tilesPer = std::max<long long>(tilesPer, 1);
    const long long nChunks = (nTiles + tilesPer - 1) / tilesPer;
    if ((size_t)nChunks * (size_t)n_q > SA_MAX_PAIRS) {
        fprintf(stderr, "[gforb] gforb_allpairs_top2_static: %lld database splits x %d queries exceed the static scratch of %zu key pairs\n",
                nChunks, n_q, SA_MAX_PAIRS);
        return GFORB_ERR_CAPACITY;
    }
    int* dErr = reinterpret_cast<int*>(D->ptr);
    uint64_t* dPartial = reinterpret_cast<uint64_t*>(D->ptr + SA_HEADER);
    // the array keeps the previous call's contents: the error word must start at zero; the partial keys of this call's
    // (split, query) range are all written by the kernel before the merge reads them
    GF_CUDA_OK(cudaMemsetAsync(dErr, 0, sizeof(int), st));
    tc::TcParams P;
    P.q = (const uint4*)queries; P.nq = n_q; P.db = (const uint4*)db; P.nDb = n_db; P.tilesPerSplit = (int)tilesPer;
    P.idxBase = idx_base; P.partial = dPartial; P.err = dErr;
    tc::k_allpairs_tc<<<dim3((unsigned)nChunks, qTiles), tc::TC_THREADS, tc::TC_SMEM_BYTES, st>>>(P);
    k_top2_merge<<<(n_q + 255) / 256, 256, 0, st>>>(dPartial, (int)nChunks, n_q, out_keys);
    GF_CUDA_OK(cudaGetLastError());
    if (!(flags & GFORB_ASYNC)) {
        GF_CUDA_OK(cudaStreamSynchronize(st));
        int err = 0;
        GF_CUDA_OK(cudaMemcpy(&err, dErr, sizeof(int), cudaMemcpyDeviceToHost));
        if (err) return err > 0 ? GFORB_ERR_CUDA : err;
    }
    return GFORB_OK;
}

}  // extern "C"
