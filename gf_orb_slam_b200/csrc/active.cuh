// K13  k_active_match -- Observability::runActiveMapMatching (reference src/Observability.cc:1249-1524), the default
// good-feature mode: greedy log-det ranking of the visible local-map points interleaved with one-point matching.
//
// One warp per frame, persistent over all num_to_match rounds.  Per round (reference lines in brackets):
//   * szActualSubset distinct columns are drawn with the device glibc rand() stream by rejection [:1343-1362]
//     (lane 0 walks the pre-generated ring; the generator itself is warp-parallel, logdet.cuh);
//   * logDet(curMat + ObsMat) of the drawn points: one lane per point, register-resident Cholesky [:1374];
//   * std::priority_queue<SimplePoint> [:1331, include/Util.hpp:47-62] restated as libstdc++'s __push_heap /
//     __adjust_heap on a flat array, so equal scores leave the heap in the same order as the reference's;
//   * the heap top goes through ORBmatcher::SearchByProjection_OnePoint [include/ORBmatcher.h:71-145]: its candidate
//     list (GetFeaturesInArea order, Hamming distances) was produced for every map point by k_candidates<M_MAP>;
//     here only the claim-dependent walk runs (already matched keypoints are skipped);
//   * success: reWeightInfoMat with the matched keypoint's octave and curMat += H_rw^T H_rw [:1394-1417];
//     failure: pop, draw one more point [:1427-1433];
//   * the matched and the failed points leave lmkIdx at the end of the round, order preserved [:1448-1483].
// Time caps are infinite; INFORMATION_EFFICIENCY_SCORE is off (as compiled in the reference); hybrid matrix types
// need DIMENSION_OF_STATE_MODEL = 13 and are rejected by the host entry.
#pragma once
#include "logdet.cuh"

namespace gforb {

struct AParams {
    FrameDev tgt;
    const uint32_t* pool;         // candidate entries (dist << 20 | octave << 16 | keypoint), k_candidates<M_MAP>
    int poolCap;
    const int2* qMeta;            // per map point: (offset, count) into pool
    int maxQ;
    const uint8_t* mpFlags;       // bit0 mbTrackInView, bit1 isBad
    const int32_t* mpCounts;
    int mpStride;
    const double* obsScore;       // batch x mpStride: > 0 <=> updateAtFrameId == frame id (null: every point)
    const double* blocks;         // MapPoint::ObsMat, batch x mpStride x 49
    const double* Hmeas;          // MapPoint::H_meas, batch x mpStride x 14
    const double* baseInfo;       // batch x 49 (mBaseInfoMat)
    const float* levelSigma2;     // Frame::mvLevelSigma2
    int nlevels;
    float nnratio;
    int numToMatch;
    const uint32_t* seeds;
    int32_t* tgtMp;               // in/out Frame::mvpMapPoints (ids)
    int32_t* tgtScore;            // in/out Frame::mvpMatchScore
    int32_t* nmatched;
    int32_t* leftIdx;             // batch x mpStride: mLeftMapPoints
    int32_t* leftN;
    double* outInfo;              // batch x 49 or null: curMat after the last round
    // scratch, batch x mpStride each
    int *sIdx, *sVis, *sHeapIdx, *sList, *sRemove;
    double *sHeapVal, *sListVal;
    uint8_t* sMark;
};

// libstdc++ std::__push_heap with std::less<SimplePoint> (score only)
__device__ __forceinline__ void heap_push(int* hi, double* hv, int holeIndex, int topIndex, int idx, double val) {
    int parent = (holeIndex - 1) / 2;
    while (holeIndex > topIndex && hv[parent] < val) {
        hi[holeIndex] = hi[parent]; hv[holeIndex] = hv[parent];
        holeIndex = parent;
        parent = (holeIndex - 1) / 2;
    }
    hi[holeIndex] = idx; hv[holeIndex] = val;
}

// std::pop_heap + pop_back on a heap of `size` entries (libstdc++ __pop_heap -> __adjust_heap); returns the new size
__device__ __forceinline__ int heap_pop(int* hi, double* hv, int size) {
    if (size <= 1) return 0;
    const int len = size - 1;
    const int vi = hi[len]; const double vv = hv[len];
    int holeIndex = 0, secondChild = 0;
    while (secondChild < (len - 1) / 2) {
        secondChild = 2 * (secondChild + 1);
        if (hv[secondChild] < hv[secondChild - 1]) secondChild--;
        hi[holeIndex] = hi[secondChild]; hv[holeIndex] = hv[secondChild];
        holeIndex = secondChild;
    }
    if ((len & 1) == 0 && secondChild == (len - 2) / 2) {
        secondChild = 2 * (secondChild + 1);
        hi[holeIndex] = hi[secondChild - 1]; hv[holeIndex] = hv[secondChild - 1];
        holeIndex = secondChild - 1;
    }
    heap_push(hi, hv, holeIndex, 0, vi, vv);
    return len;
}

__global__ void __launch_bounds__(32) k_active_match(const AParams A) {
    const unsigned full = 0xffffffffu;
    const int b = blockIdx.x, lane = threadIdx.x;
    __shared__ double cur[49];
    __shared__ uint32_t s_rs[64];
    __shared__ uint32_t s_raw[RAW_RING];
    __shared__ uint32_t s_ent[32];
    __shared__ int s_claim[32];
    const size_t mOff = (size_t)b * A.mpStride, tOff = (size_t)b * A.tgt.stride;
    const int M = min(A.mpCounts[b], min(A.mpStride, A.maxQ));
    int* idx = A.sIdx + mOff;          // lmkIdx
    int* vis = A.sVis + mOff;          // lmkVisited (round stamps; a stale stamp is always < the current round)
    int* hi = A.sHeapIdx + mOff;       // heapSubset
    double* hv = A.sHeapVal + mOff;
    int* list = A.sList + mOff;        // points drawn in the fill phase of a round
    double* listVal = A.sListVal + mOff;
    int* rem = A.sRemove + mOff;       // removeIdx
    uint8_t* mark = A.sMark + mOff;
    const uint8_t* fl = A.mpFlags + mOff;
    const double* sc = A.obsScore ? A.obsScore + mOff : nullptr;
    const double* blk = A.blocks + mOff * 49;
    int32_t* left = A.leftIdx + mOff;
    const unsigned lt = (1u << lane) - 1u;
    if (lane < 32) { cur[lane] = A.baseInfo[(size_t)b * 49 + lane]; if (lane + 32 < 49) cur[lane + 32] = A.baseInfo[(size_t)b * 49 + lane + 32]; }
    // lmkIdx = visible, not bad, updated at this frame (:1290-1310); with nothing to do every visible point is "left"
    const bool trivial = M <= 0 || A.numToMatch <= 0;
    int N = 0;
    for (int base = 0; base < M; base += 32) {
        const int i = base + lane;
        bool ok = false;
        if (i < M) {
            ok = (fl[i] & 1) && !(fl[i] & 2);
            if (!trivial && sc && !(sc[i] > 0.0)) ok = false;
            mark[i] = 0;
        }
        const unsigned bal = __ballot_sync(full, ok);
        if (ok) {
            const int o = N + __popc(bal & lt);
            if (trivial) left[o] = i; else { idx[o] = i; vis[o] = -1; }
        }
        N += __popc(bal);
    }
    __syncwarp();
    if (trivial) {
        if (lane == 0) { A.leftN[b] = N; A.nmatched[b] = 0; }
        if (A.outInfo) { A.outInfo[(size_t)b * 49 + lane] = cur[lane]; if (lane + 32 < 49) A.outInfo[(size_t)b * 49 + lane + 32] = cur[lane + 32]; }
        return;
    }
    const int szLazier = (int)(size_t)((double)((float)N / (float)A.numToMatch) * 1.0);      // :1320
    WarpRand rng;
    rng.rs = s_rs; rng.raw = s_raw; rng.n = 34;
    rng.seed(A.seeds ? A.seeds[b] : 1u, lane);
    int rp = 0;                        // rand() calls consumed
    int m = N, nMatched = 0;
    const uint32_t* pool = A.pool + (size_t)b * A.poolCap;
    const int2* qMeta = A.qMeta + (size_t)b * A.maxQ;
    for (int i = 0; i < A.numToMatch; i++) {
        const int szActual = min(szLazier, m);
        int cnt = 0, tries = 0, nRem = 0, hs = 0;
        int failed = 0;
        // ---- fill: szActual distinct unvisited columns (:1343-1366)
        while (cnt < szActual && !failed) {
            rng.ensure(rp, lane);
            int used = 0;
            if (lane == 0) {
                while (used < 32 && cnt < szActual) {
                    const int j = (int)(s_raw[(rp + used) & (RAW_RING - 1)] % (uint32_t)m);
                    used++;
                    if (vis[j] < i) { vis[j] = i; list[cnt++] = idx[j]; tries = 0; }
                    else if (++tries >= 2000) { failed = 1; break; }
                }
            }
            used = __shfl_sync(full, used, 0); cnt = __shfl_sync(full, cnt, 0);
            tries = __shfl_sync(full, tries, 0); failed = __shfl_sync(full, failed, 0);
            rp += used;
        }
        __syncwarp();
        if (failed || szActual == 0) break;                 // numRndQue >= MAX or empty heap: "early termination" (:1437)
        // ---- potential gain of every drawn point, then the heap in draw order (:1374-1379)
        for (int t = lane; t < cnt; t += 32) listVal[t] = logdet_sum(cur, blk + (size_t)list[t] * 49, 1.0);
        __syncwarp();
        if (lane == 0) for (int t = 0; t < cnt; t++) heap_push(hi, hv, t, 0, list[t], listVal[t]);
        hs = cnt;
        __syncwarp();
        // ---- match the heap top until one succeeds (:1381-1434)
        while (true) {
            const int top = hi[0];
            const int2 meta = qMeta[top];
            int best = INT_MAX, best2 = INT_MAX, bl = -1, bl2 = -1, bi = -1;
            for (int c0 = 0; c0 < meta.y; c0 += 32) {
                const int c = c0 + lane;
                uint32_t e = 0; int cl = 1;
                if (c < meta.y) { e = pool[meta.x + c]; cl = A.tgtMp[tOff + (e & 0xffff)] >= 0; }
                s_ent[lane] = e; s_claim[lane] = cl;
                __syncwarp();
                if (lane == 0) {
                    const int lim = min(32, meta.y - c0);
                    for (int q = 0; q < lim; q++) {
                        if (s_claim[q]) continue;                                  // F.mvpMapPoints[idx] (ORBmatcher.h:105)
                        const uint32_t ee = s_ent[q];
                        const int dist = (int)(ee >> 20), oct = (ee >> 16) & 15, i2 = ee & 0xffff;
                        if (dist < best) { best2 = best; best = dist; bl2 = bl; bl = oct; bi = i2; }
                        else if (dist < best2) { bl2 = oct; best2 = dist; }
                    }
                }
                __syncwarp();
            }
            int ok = 0;
            if (lane == 0) {
                ok = bi >= 0 && best <= TH_HIGH && !(bl == bl2 && (float)best > A.nnratio * (float)best2);
                rem[nRem] = top;
                if (ok) { A.tgtMp[tOff + bi] = top; A.tgtScore[tOff + bi] = best; }
            }
            nRem++;
            ok = __shfl_sync(full, ok, 0); bi = __shfl_sync(full, bi, 0);
            if (ok) {
                // reWeightInfoMat with the matched keypoint's octave; curMat += H_rw^T H_rw (:1394-1417)
                const int oct = A.tgt.kps[tOff + bi].octave;
                const float Sigma2 = A.levelSigma2[min(max(oct, 0), A.nlevels - 1)];
                const double wi = 1.0 / sqrt((double)Sigma2);
                const double* H = A.Hmeas + (mOff + top) * 14;
                for (int e = lane; e < 49; e += 32) {
                    const int r = e / 7, c = e % 7;
                    const double a0 = wi * H[r], a1 = wi * H[c], b0 = wi * H[7 + r], b1 = wi * H[7 + c];
                    cur[e] = cur[e] + (a0 * a1 + b0 * b1);
                }
                nMatched++;
                __syncwarp();
                break;
            }
            // failed: drop the top, draw one more point (:1427-1433)
            if (lane == 0) hs = heap_pop(hi, hv, hs);
            hs = __shfl_sync(full, hs, 0);
            int que = -1;
            tries = 0;
            while (que < 0 && !failed) {
                rng.ensure(rp, lane);
                int used = 0;
                if (lane == 0) {
                    while (used < 32) {
                        const int j = (int)(s_raw[(rp + used) & (RAW_RING - 1)] % (uint32_t)m);
                        used++;
                        if (vis[j] < i) { vis[j] = i; que = idx[j]; break; }
                        else if (++tries >= 2000) { failed = 1; break; }
                    }
                }
                used = __shfl_sync(full, used, 0); que = __shfl_sync(full, que, 0);
                tries = __shfl_sync(full, tries, 0); failed = __shfl_sync(full, failed, 0);
                rp += used;
            }
            if (failed) break;
            const double val = logdet_sum(cur, blk + (size_t)que * 49, 1.0);
            if (lane == 0) heap_push(hi, hv, hs, 0, que, val);
            hs++;
            __syncwarp();
        }
        if (failed || hs == 0 || nRem == 0) break;           // :1437-1440
        if (m == nRem) break;                                // "went through all map points" (:1449-1452)
        // ---- the matched and the failed points leave lmkIdx, order preserved (:1455-1483)
        for (int r = lane; r < nRem; r += 32) mark[rem[r]] = 1;
        __syncwarp();
        int off = 0;
        for (int base = 0; base < m; base += 32) {
            const int j = base + lane;
            int v = -1;
            if (j < m) v = idx[j];
            const bool keep = v >= 0 && !mark[v];
            const unsigned bal = __ballot_sync(full, keep);
            __syncwarp();
            if (keep) idx[off + __popc(bal & lt)] = v;
            off += __popc(bal);
        }
        __syncwarp();
        for (int r = lane; r < nRem; r += 32) mark[rem[r]] = 0;
        m = off;
        __syncwarp();
    }
    for (int j = lane; j < m; j += 32) left[j] = idx[j];
    if (lane == 0) { A.leftN[b] = m; A.nmatched[b] = nMatched; }
    if (A.outInfo) { A.outInfo[(size_t)b * 49 + lane] = cur[lane]; if (lane + 32 < 49) A.outInfo[(size_t)b * 49 + lane + 32] = cur[lane + 32]; }
}

}  // namespace gforb
