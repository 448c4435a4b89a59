// B200-native good-feature path of ORB_SLAM::Observability (reference include/Observability.h, include/Util.hpp,
// src/Observability.cc), DIMENSION_OF_STATE_MODEL = 7 (position + quaternion).
//
//   host   updatePWLSVec / predictPWLSVec   13-state kinematics, a handful of 4x4 / 13x13 matrices per frame
//                                            (Observability.h:222-295, Util.hpp:787-1005): plain host C++.
//   K11    k_build_info     one thread per matched point: 2x7 measurement Jacobian (compute_H_subblock_simplied,
//                           Observability.h:460-515), octave re-weighting (reWeightInfoMat, :517-596) and the 7x7
//                           information block H^T H (src/Observability.cc:464-470), FP64.
//   K12    k_select         one CTA per selection problem, persistent over all rounds of maxVolSelection_LazierGreedy
//                           / maxVolDeletion_LazierGreedy / the full greedy (src/Observability.cc:2815-3029, 2339-2542,
//                           3031-3139): glibc-TYPE_3 rand() stream re-implemented on the device (the reference draws
//                           from the process-global rand()), one thread per sampled candidate evaluating
//                           logDet(cur +- block) with a register-resident 7x7 Cholesky (Util.hpp:714-731), block
//                           arg-max with "first drawn wins" ties, in-place update of the running matrix.
// All double arithmetic is un-fused (this TU is built with --fmad=false) and follows the CPU operation order, so the
// information blocks, Cholesky products and -- with glibc's log() restated in glibc_log.h -- the log-dets themselves are
// bit-identical to the scalar sequence (north_star tolerance on log-det scores: 1e-5 relative; they are in fact equal).
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "common.cuh"
#include "host_util.h"
#include "logdet.cuh"

namespace gforb {

struct CamDev { double fu, fv, Cx, Cy; };

// ------------------------------------------------------------------------------------------------ K11
__device__ __forceinline__ void h_subblock(const double* __restrict__ Xv, const CamDev cam, const double Y0, const double Y1,
                                           const double Y2, double (&H)[14], float& u, float& v, double& hz) {
    const double r = Xv[3], x = Xv[4], y = Xv[5], z = Xv[6];
    // q2r (Util.hpp:302-317)
    double R[9];
    R[0] = r * r + x * x - y * y - z * z; R[1] = 2.0 * (x * y - r * z);         R[2] = 2.0 * (z * x + r * y);
    R[3] = 2.0 * (x * y + r * z);         R[4] = r * r - x * x + y * y - z * z; R[5] = 2.0 * (y * z - r * x);
    R[6] = 2.0 * (z * x - r * y);         R[7] = 2.0 * (y * z + r * x);         R[8] = r * r - x * x - y * y + z * z;
    // R_rw = inv(R): 3x3 cofactor inverse (arma::inv tiny-matrix path)
    const double c00 = R[4] * R[8] - R[5] * R[7], c01 = R[5] * R[6] - R[3] * R[8], c02 = R[3] * R[7] - R[4] * R[6];
    const double det = R[0] * c00 + R[1] * c01 + R[2] * c02;
    const double id = 1.0 / det;
    double W[9];
    W[0] = c00 * id; W[1] = (R[2] * R[7] - R[1] * R[8]) * id; W[2] = (R[1] * R[5] - R[2] * R[4]) * id;
    W[3] = c01 * id; W[4] = (R[0] * R[8] - R[2] * R[6]) * id; W[5] = (R[2] * R[3] - R[0] * R[5]) * id;
    W[6] = c02 * id; W[7] = (R[1] * R[6] - R[0] * R[7]) * id; W[8] = (R[0] * R[4] - R[1] * R[3]) * id;
    const double t0 = Y0 - Xv[0], t1 = Y1 - Xv[1], t2 = Y2 - Xv[2];
    const double h0 = W[0] * t0 + W[1] * t1 + W[2] * t2;
    const double h1 = W[3] * t0 + W[4] * t1 + W[5] * t2;
    const double h2 = W[6] * t0 + W[7] * t1 + W[8] * t2;
    hz = h2;
    if (h2 > 0) {
        u = (float)((double)(float)cam.fu * h0 / h2 + (double)(float)cam.Cx);
        v = (float)((double)(float)cam.fv * h1 / h2 + (double)(float)cam.Cy);
    } else { u = FLT_MAX; v = FLT_MAX; }
    // dhu_dhrl (2x3 pinhole Jacobian; zeros when |hz| < 1e-6)
    double J[6] = {0, 0, 0, 0, 0, 0};
    if (!(fabs(h2) < 1e-6)) {
        J[0] = cam.fu / h2; J[2] = -h0 * cam.fu / (h2 * h2);
        J[4] = cam.fv / h2; J[5] = -h1 * cam.fv / (h2 * h2);
    }
    // H13 = -(dhu_dhrl * R_rw)
#pragma unroll
    for (int rr = 0; rr < 2; rr++)
#pragma unroll
        for (int c = 0; c < 3; c++) {
            double s = 0;
#pragma unroll
            for (int k = 0; k < 3; k++) s += J[rr * 3 + k] * W[k * 3 + c];
            H[rr * 7 + c] = -1.0 * s;
        }
    // H47 = dhu_dhrl * (dRq_times_a_by_dq(qconj(q), t_rw) * diag(1,-1,-1,-1))   (Util.hpp:391-438)
    const double q0 = r, qx = -x, qy = -y, qz = -z;
    const double d0[9] = {2 * q0, -2 * qz, 2 * qy, 2 * qz, 2 * q0, -2 * qx, -2 * qy, 2 * qx, 2 * q0};
    const double dx[9] = {2 * qx, 2 * qy, 2 * qz, 2 * qy, -2 * qx, -2 * q0, 2 * qz, 2 * q0, -2 * qx};
    const double dy[9] = {-2 * qy, 2 * qx, 2 * q0, 2 * qx, 2 * qy, 2 * qz, -2 * q0, 2 * qz, -2 * qy};
    const double dz[9] = {-2 * qz, -2 * q0, 2 * qx, 2 * q0, -2 * qz, 2 * qy, 2 * qx, 2 * qy, 2 * qz};
    double D[12];
#pragma unroll
    for (int rr = 0; rr < 3; rr++) {
        D[rr * 4 + 0] = (d0[rr * 3] * t0 + d0[rr * 3 + 1] * t1 + d0[rr * 3 + 2] * t2) * 1.0;
        D[rr * 4 + 1] = (dx[rr * 3] * t0 + dx[rr * 3 + 1] * t1 + dx[rr * 3 + 2] * t2) * -1.0;
        D[rr * 4 + 2] = (dy[rr * 3] * t0 + dy[rr * 3 + 1] * t1 + dy[rr * 3 + 2] * t2) * -1.0;
        D[rr * 4 + 3] = (dz[rr * 3] * t0 + dz[rr * 3 + 1] * t1 + dz[rr * 3 + 2] * t2) * -1.0;
    }
#pragma unroll
    for (int rr = 0; rr < 2; rr++)
#pragma unroll
        for (int c = 0; c < 4; c++) {
            double s = 0;
#pragma unroll
            for (int k = 0; k < 3; k++) s += J[rr * 3 + k] * D[k * 4 + c];
            H[rr * 7 + 3 + c] = s;
        }
}

// mode 0: batchInfoMat_Frame (valid[] gate, sigma2 = levelSigma2[octave], score = match score)
// mode 1: batchInfoMat_Map   (trackInView gate unless check_viz, identity noise, score 1 / -1 / -2 untouched)
// The 392-byte blocks leave through shared memory: a warp stages its 32 blocks (12.5 KB) and writes them as one contiguous
// run (every store instruction covers 256 contiguous bytes); written straight from the registers, each of a thread's 49 stores
// hit 32 different sectors (ncu: lg_throttle 46 stall cycles per issued instruction).
static const int BI_THREADS = 64;
__global__ void __launch_bounds__(BI_THREADS) k_build_info(int mode, const double* __restrict__ Xv, CamDev cam,
                                                    const float* __restrict__ world, const int32_t* __restrict__ octave,
                                                    const int32_t* __restrict__ score, const uint8_t* __restrict__ gate,
                                                    const int32_t* __restrict__ counts, int stride,
                                                    const float* __restrict__ levelSigma2, int nlevels, int checkViz,
                                                    int minX, int maxX, int minY, int maxY, double boundX, double boundY,
                                                    double boundDepth, double* __restrict__ blocks,
                                                    double* __restrict__ obsScore, double* __restrict__ Hmeas,
                                                    float* __restrict__ uv) {
    __shared__ double s_blk[BI_THREADS / 32][32 * 49];
    const int i = blockIdx.x * blockDim.x + threadIdx.x, b = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t p = (size_t)b * stride + i;
    double* sb = s_blk[warp] + lane * 49;
    const auto compute = [&]() -> bool {                     // true: this thread's block is in sb
        if (i >= min(counts[b], stride)) return false;
        if (mode == 0) {
            if (!gate[p]) { obsScore[p] = -1.0; return false; }
        } else {
            obsScore[p] = -2.0;
            if (!checkViz && !gate[p]) return false;
        }
        double H[14];
        float u, v;
        double hz;
        h_subblock(Xv + (size_t)b * 13, cam, (double)world[p * 3], (double)world[p * 3 + 1], (double)world[p * 3 + 2], H, u, v, hz);
        if (uv) { uv[p * 2] = u; uv[p * 2 + 1] = v; }
        if (mode == 1 && checkViz) {
            bool ok = true;
            if (hz < 0.0 + boundDepth) ok = false;
            if ((double)u < (double)minX - boundX || (double)u > (double)maxX + boundX) ok = false;
            if ((double)v < (double)minY - boundY || (double)v > (double)maxY + boundY) ok = false;
            if (!ok) { obsScore[p] = -1.0; return false; }
        }
        if (Hmeas)
#pragma unroll
            for (int k = 0; k < 14; k++) Hmeas[p * 14 + k] = H[k];
        double sigma2 = 1.0;
        if (mode == 0) sigma2 = (double)levelSigma2[min(max(octave[p], 0), nlevels - 1)];
        const double wi = 1.0 / sqrt(sigma2);
        double Hr[14];
#pragma unroll
        for (int k = 0; k < 14; k++) Hr[k] = wi * H[k];
#pragma unroll
        for (int rr = 0; rr < 7; rr++)
#pragma unroll
            for (int c = 0; c < 7; c++) sb[rr * 7 + c] = Hr[rr] * Hr[c] + Hr[7 + rr] * Hr[7 + c];
        obsScore[p] = (mode == 0) ? (double)score[p] : 1.0;
        return true;
    };
    const bool wr = compute();
    const unsigned m = __ballot_sync(0xffffffffu, wr);
    if (!m) return;
    __syncwarp();
    double* o = blocks + ((size_t)b * stride + (size_t)(i - lane)) * 49;      // block of the warp's first match
    const double* sw = s_blk[warp];
    if (m == 0xffffffffu) {
#pragma unroll 7
        for (int e = lane; e < 32 * 49; e += 32) o[e] = sw[e];
    } else {
        for (int e = lane; e < 32 * 49; e += 32)
            if ((m >> (e / 49)) & 1u) o[e] = sw[e];
    }
}

__global__ void __launch_bounds__(128) k_logdet(const double* __restrict__ mats, int n, double* __restrict__ out) {
    __shared__ double zero[49];
    if (threadIdx.x < 49) zero[threadIdx.x] = 0.0;
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = logdet_sum(zero, mats + (size_t)i * 49, 1.0);
}

// ------------------------------------------------------------------------------------------------ K12 selection
enum { MTD_GREEDY = 1, MTD_LAZIER_SELECT = 2, MTD_LAZIER_AUTO = 3, MTD_LAZIER_DELETE = 5 };
static const int MAX_RANDOM_QUERY_TIME = 2000;
static const int SEL_WIN = 512;      // rand() draws examined per sampling pass
static const int SA_DPL = 4, SA_WIN = 32 * SA_DPL;   // k_select_ahead: draws per lane / per pass of the sampler warp

struct SelParams {
    const double* blocks;      // batch x stride x 49
    const double* obsScore;    // batch x stride (may be null: all valid)
    const int32_t* counts;     // batch
    const uint32_t* seeds;     // batch (may be null: seed 1 == the never-seeded reference process)
    int stride, k, mtd, maxPoints;
    double sampleScale;
    int32_t* outIdx;           // batch x outStride
    double* outLogdet;         // batch x outStride (may be null)
    int32_t* outN;             // batch
    int outStride;
    long long* prof;           // optional phase-cycle counters of problem 0 (GFORB_SELECT_PROF=1), else null
    double* pack;              // batch x stride x 28: packed lower triangles (pack_block), scratch
    int split;                 // 1: problems whose pool entries are all valid belong to k_select_ahead, the rest to k_select
};

__global__ void k_rand_stream(uint32_t seed, int n, int32_t* out) {
    __shared__ uint32_t rs[64];
    __shared__ uint32_t raw[RAW_RING];
    const int lane = threadIdx.x;
    WarpRand g; g.rs = rs; g.raw = raw;
    g.seed(seed, lane);
    for (int base = 0; base < n; base += 32) {
        g.ensure(base, lane);
        if (base + lane < n) out[base + lane] = (int32_t)raw[(base + lane) & (RAW_RING - 1)];
        __syncwarp();
    }
}

// T threads per problem (64 / 128 / 256): the kernel is a 100-round dependency chain that leaves the issue slots
// mostly idle, so what it costs the rest of the pipeline is the register file it holds while it runs (T x 128
// registers per problem).  Small T trades a longer chain (each thread evaluates cnt / T candidates in turn) for a
// footprint that lets the extraction kernels of the next step co-reside on the same SMs.
template <int T>
__global__ void __launch_bounds__(T, 512 / T) k_select(const SelParams P) {
    constexpr int DPT = SEL_WIN / T;          // rand() draws examined per thread and sampling pass
    constexpr int NW = T / 32;
    extern __shared__ __align__(16) unsigned char smraw[];
    const unsigned full = 0xffffffffu;
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = min(P.counts[b], min(P.stride, P.maxPoints));
    int* idxA = reinterpret_cast<int*>(smraw);
    int* vis = idxA + P.maxPoints;           // lmkVisited: round stamp per column (never shifted: only stamps of the
                                              // current round are ever compared, and a round erases no column)
    int* idxB = vis + P.maxPoints;
    int* win = idxB + P.maxPoints;           // per column: (pass epoch << 9 | 511 - draw) of the first draw in a pass
    int* list = win + P.maxPoints;
    unsigned char* alive = reinterpret_cast<unsigned char*>(list + P.maxPoints);   // `selected` flag per pool entry
    unsigned char* okf = alive + P.maxPoints;                                      // obs_score >= 0
    __shared__ double cur[49];
    __shared__ int redT[NW];
    __shared__ uint32_t s_rs[64];
    __shared__ uint32_t s_raw[RAW_RING];
    __shared__ int s_cnt;
    __shared__ double redV[NW];
    __shared__ int redT2[NW];
    __shared__ unsigned s_new[16], s_valid[16];
    __shared__ int s_cut[16];
    const double* blk = P.blocks + (size_t)b * P.stride * 49;
    const double* sc = P.obsScore ? P.obsScore + (size_t)b * P.stride : nullptr;
    int32_t* outIdx = P.outIdx + (size_t)b * P.outStride;
    double* outLd = P.outLogdet ? P.outLogdet + (size_t)b * P.outStride : nullptr;
    const int k = P.k;
    if (n <= 0) { if (tid == 0) P.outN[b] = 0; return; }
    if (k >= n) {        // "if (mpLimit >= edIdx - stIdx) take all" (:2832-2840)
        for (int i = tid; i < n && i < P.outStride; i += T) { outIdx[i] = i; if (outLd) outLd[i] = 0.0; }
        if (tid == 0) P.outN[b] = min(n, P.outStride);
        return;
    }
    int mtd = P.mtd;
    if (mtd == MTD_LAZIER_AUTO) mtd = ((size_t)k * 2 > (size_t)n) ? MTD_LAZIER_DELETE : MTD_LAZIER_SELECT;   // :3141-3155
    const bool deletion = mtd == MTD_LAZIER_DELETE;
    const double sgn = deletion ? -1.0 : 1.0;
    const size_t szLazier = (size_t)((double)n / (double)k * P.sampleScale);
    // init
    if (tid < 49) cur[tid] = (tid % 8 == 0) ? 0.00001 : 0.0;
    int allOk = 1;
    for (int i = tid; i < n; i += T) {
        idxA[i] = i; vis[i] = -1; win[i] = 0; alive[i] = deletion ? 1 : 0;
        okf[i] = (sc && sc[i] < 0) ? 0 : 1;
        allOk &= okf[i];
    }
    allOk = __syncthreads_and(allOk);
    if (P.split && allOk && mtd != MTD_GREEDY) return;      // k_select_ahead owns this problem
    double* pk = P.pack + (size_t)b * P.stride * 28;
    for (int c = tid; c < n; c += T) pack_block(blk + (size_t)c * 49, pk + (size_t)c * 28);
    if (deletion && tid < 49) {       // curMat += every block, in pool order (:2370-2380)
        double s = cur[tid];
        for (int i = 0; i < n; i++) s = s + blk[(size_t)i * 49 + tid];
        cur[tid] = s;
    }
    WarpRand rng;
    rng.rs = s_rs; rng.raw = s_raw; rng.n = 34;
    if (warp == 0 && mtd != MTD_GREEDY) rng.seed(P.seeds ? P.seeds[b] : 1u, lane);
    int rp = 0;                       // rand() calls consumed so far (uniform)
    __syncthreads();
    int m = n;                        // live columns of lmkIdx / lmkVisited
    int* idx = idxA; int* idx2 = idxB;
    int epoch = 1;                    // sampling pass counter (uniform)
    const int rounds = deletion ? n - k : k;
    int nOut = 0;
    long long tprev = 0;
    const bool prof = P.prof && b == 0 && tid == 0;
#define SEL_MARK(slot) if (prof) { const long long t_ = clock64(); P.prof[slot] += t_ - tprev; tprev = t_; }
    if (prof) tprev = clock64();
    for (int i = 0; i < rounds; i++) {
        // ---- candidate list of this round, in evaluation order
        if (mtd == MTD_GREEDY) {
            // full greedy == the reference's lazy greedy with the sum-log-diag bound (:3031-3139): evaluate every
            // unselected valid entry (ties -> lowest pool index)
            if (tid == 0) s_cnt = 0;
            __syncthreads();
            for (int base = 0; base < m; base += T) {
                const int j = base + tid;
                const bool ok = j < m && okf[idx[j]];
                const unsigned bal = __ballot_sync(full, ok);
                if (lane == 0) redT[warp] = __popc(bal);
                __syncthreads();
                int off = s_cnt;
                for (int w = 0; w < warp; w++) off += redT[w];
                if (ok) list[off + __popc(bal & ((1u << lane) - 1u))] = idx[j];
                __syncthreads();
                if (tid == 0) { int t = 0; for (int w = 0; w < NW; w++) t += redT[w]; s_cnt += t; }
                __syncthreads();
            }
        } else {
            // RANDOM_ACCESS_LAZIER_GREEDY sampling (:2895-2921), block-parallel: each pass examines the next SEL_WIN
            // rand() draws (2 per thread, draw number d = h * 256 + tid).  A draw is "new" if its column was not visited
            // in this round (stamp in vis[] from earlier passes, first-in-pass decided by atomicMax on win[]); new + valid
            // draws count towards szActualSubset; 2000 consecutive rejections stop the round.  The pass is cut at the
            // draw where the sequential loop would have stopped; only draws up to the cut are consumed and committed.
            size_t szActual = szLazier;
            if ((size_t)m < szActual) szActual = (size_t)m;
            int numHit = 0, cntL = 0, rej = 0;
            bool stop = (size_t)numHit >= szActual;
            while (!stop) {
                if (warp == 0) while (rng.n - 344 < rp + SEL_WIN) rng.gen30(lane);
                __syncthreads();
                int jj[DPT]; int key[DPT]; bool seenPrev[DPT];
#pragma unroll
                for (int h = 0; h < DPT; h++) {
                    const int d = h * T + tid;
                    const uint32_t r = s_raw[(rp + d) & (RAW_RING - 1)];
                    jj[h] = (int)(r % (uint32_t)m);
                    seenPrev[h] = vis[jj[h]] >= i;
                    key[h] = (epoch << 9) | (511 - d);
                    if (!seenPrev[h]) atomicMax(&win[jj[h]], key[h]);
                }
                __syncthreads();
                bool isNew[DPT], valid[DPT]; int que[DPT]; unsigned nm[DPT], vm[DPT];
#pragma unroll
                for (int h = 0; h < DPT; h++) {
                    isNew[h] = !seenPrev[h] && win[jj[h]] == key[h];
                    que[h] = idx[jj[h]];
                    const bool sel = alive[que[h]] != 0;
                    valid[h] = isNew[h] && !((deletion ? !sel : sel) || !okf[que[h]]);
                    nm[h] = __ballot_sync(full, isNew[h]);
                    vm[h] = __ballot_sync(full, valid[h]);
                    if (lane == 0) { s_new[h * NW + warp] = nm[h]; s_valid[h * NW + warp] = vm[h]; }
                }
                __syncthreads();
                const unsigned lt = (1u << lane) - 1u;
                int vBefore[DPT];
#pragma unroll
                for (int h = 0; h < DPT; h++) {
                    const int g = h * NW + warp, d = h * T + tid;
                    int vb = numHit;
                    for (int q = 0; q < g; q++) vb += __popc(s_valid[q]);
                    vb += __popc(vm[h] & lt);
                    vBefore[h] = vb;
                    const bool hitA = valid[h] && (size_t)(vb + 1) == szActual;
                    // consecutive rejections ending at this draw (only meaningful for a rejected draw)
                    int run;
                    if (nm[h] & lt) run = lane - (31 - __clz(nm[h] & lt));
                    else {
                        int q = g - 1;
                        while (q >= 0 && s_new[q] == 0u) q--;
                        run = (q >= 0) ? d - (q * 32 + 31 - __clz(s_new[q])) : rej + d + 1;
                    }
                    const bool hitB = !isNew[h] && run >= MAX_RANDOM_QUERY_TIME;
                    const unsigned cm = __ballot_sync(full, hitA || hitB);
                    if (lane == 0) s_cut[g] = cm ? g * 32 + __ffs(cm) - 1 : INT_MAX;
                }
                __syncthreads();
                int cut = INT_MAX;
                for (int q = 0; q < 16; q++) if (s_cut[q] != INT_MAX) { cut = s_cut[q]; break; }
                stop = cut != INT_MAX;
                if (!stop) cut = SEL_WIN - 1;
                // commit the consumed draws
#pragma unroll
                for (int h = 0; h < DPT; h++) {
                    const int d = h * T + tid;
                    if (d <= cut) {
                        if (isNew[h]) vis[jj[h]] = i;
                        if (valid[h]) list[cntL + vBefore[h] - numHit] = que[h];
                    }
                }
                // uniform bookkeeping: valid draws taken, trailing rejection run, draws consumed
                int V = 0, lastNew = -1;
                for (int q = 0; q < 16; q++) {
                    const int lo = q * 32;
                    if (lo > cut) break;
                    const unsigned take = (cut - lo >= 31) ? full : ((2u << (cut - lo)) - 1u);
                    V += __popc(s_valid[q] & take);
                    const unsigned nw = s_new[q] & take;
                    if (nw) lastNew = lo + 31 - __clz(nw);
                }
                cntL += V; numHit += V;
                rej = (lastNew >= 0) ? cut - lastNew : rej + cut + 1;
                rp += cut + 1;
                epoch++;
                __syncthreads();
            }
            if (tid == 0) s_cnt = cntL;
        }
        __syncthreads();
        SEL_MARK(0)
        const int cnt = s_cnt;
        // ---- evaluate logDet(cur +- block) for every candidate; arg-max, first in list wins ties (strict >)
        double bestV = -DBL_MAX; int bestT = INT_MAX;
        for (int t = tid; t < cnt; t += T) {
            const int q = list[t];
            const double v = logdet_sum_packed(cur, pk + (size_t)q * 28, blk + (size_t)q * 49, sgn);
            if (v > bestV) { bestV = v; bestT = t; }
        }
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) {
            const double ov = __shfl_xor_sync(full, bestV, d);
            const int ot = __shfl_xor_sync(full, bestT, d);
            if (ov > bestV || (ov == bestV && ot < bestT)) { bestV = ov; bestT = ot; }
        }
        if (lane == 0) { redV[warp] = bestV; redT2[warp] = bestT; }
        __syncthreads();
        double winV = redV[0]; int winT = redT2[0];        // every thread scans the per-warp winners
#pragma unroll
        for (int w = 1; w < NW; w++)
            if (redV[w] > winV || (redV[w] == winV && redT2[w] < winT)) { winV = redV[w]; winT = redT2[w]; }
        SEL_MARK(1)
        const int pick = winT == INT_MAX ? -1 : list[winT];
        if (pick < 0) break;                                  // "if (maxLmk == -1) break"
        // ---- update the running matrix, record the pick, erase its column (order preserved)
        if (tid < 49) cur[tid] = cur[tid] + sgn * blk[(size_t)pick * 49 + tid];
        if (tid == 0) {
            alive[pick] = deletion ? 0 : 1;
            if (!deletion && i < P.outStride) { outIdx[i] = pick; if (outLd) outLd[i] = winV; }
        }
        if (!deletion) nOut = i + 1;
        // lmkIdx stays ascending (identity minus erased columns): every entry >= pick moves one column left
        for (int j = tid; j < m - 1; j += T) { const int a = idx[j]; idx2[j] = (a >= pick) ? idx[j + 1] : a; }
        m -= 1;
        { int* t = idx; idx = idx2; idx2 = t; }
        __syncthreads();
        SEL_MARK(2)
    }
    if (deletion) {
        // survivors in ascending pool order (:2530-2540); single warp, ordered compaction
        __syncthreads();
        if (warp == 0) {
            int off = 0;
            for (int base = 0; base < n; base += 32) {
                const int j = base + lane;
                const bool ok = j < n && alive[j];
                const unsigned bal = __ballot_sync(full, ok);
                const int o = off + __popc(bal & ((1u << lane) - 1u));
                if (ok && o < P.outStride) { outIdx[o] = j; if (outLd) outLd[o] = 0.0; }
                off += __popc(bal);
            }
            if (lane == 0) P.outN[b] = min(off, P.outStride);
        }
    } else if (tid == 0) P.outN[b] = min(nOut, P.outStride);
}


// ------------------------------------------------------------------------------------------------ K12b selection, sampler ahead
// When every pool entry is valid (obs_score >= 0, the normal case) the rejection sampler of the lazier-greedy loops is
// independent of the picks: round i draws from rand() % (n - i) column NUMBERS until szActualSubset distinct columns
// were hit, and which pool entry a column holds only matters when the list is evaluated.  So the sampler runs one
// round AHEAD in a warp of its own (producer), while T evaluator threads (consumers) map columns to pool entries,
// evaluate logDet(cur +- block), take the arg-max and erase the picked column.  The round's critical path shrinks to
// evaluate -> arg-max -> update.  Hand-off through two column lists and named barriers (bar.sync / bar.arrive):
//   FULL[p]  sampler arrives after filling list p          evaluators wait before reading it
//   EMPTY[p] evaluators arrive after their last read of p   sampler waits before refilling it
// The rand() stream is consumed in exactly the reference's order (rounds are sampled in order, each draw is either
// taken or rejected by the reference's rule), so picks, order and log-dets are those of k_select.
// barrier ids are immediates so that ptxas reserves exactly the six barriers in use
template <int ID, int N> __device__ __forceinline__ void bar_sync_i() { asm volatile("bar.sync %0, %1;" ::"n"(ID), "n"(N) : "memory"); }
template <int ID, int N> __device__ __forceinline__ void bar_arrive_i() { asm volatile("bar.arrive %0, %1;" ::"n"(ID), "n"(N) : "memory"); }
template <int ID, int N> __device__ __forceinline__ void bar_sync2(int p) { if (p) bar_sync_i<ID + 1, N>(); else bar_sync_i<ID, N>(); }
template <int ID, int N> __device__ __forceinline__ void bar_arrive2(int p) { if (p) bar_arrive_i<ID + 1, N>(); else bar_arrive_i<ID, N>(); }

// MINB = resident CTAs per SM the register budget is cut for: the evaluation wants ~170 registers (no spills of the two
// 28-entry triangles), which fits two CTAs of 160 threads per SM; batches above two problems per SM trade that for a third CTA.
template <int T, int MINB>
__global__ void __launch_bounds__(T + 32, MINB) k_select_ahead(const SelParams P) {
    constexpr int NW = T / 32;                // evaluator warps; warp NW is the sampler
    constexpr int ALL = T + 32;
    enum { B_FULL = 1, B_EMPTY = 3, B_EVAL = 5 };
    extern __shared__ __align__(16) unsigned char smraw[];
    const unsigned full = 0xffffffffu;
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = min(P.counts[b], min(P.stride, P.maxPoints));
    int* idxA = reinterpret_cast<int*>(smraw);
    int* idxB = idxA + P.maxPoints;
    int* vis = idxB + P.maxPoints;           // sampler only: round stamp per column
    int* col0 = vis + P.maxPoints;           // column lists of even / odd rounds
    int* col1 = col0 + P.maxPoints;
    int* win = col1 + P.maxPoints;           // sampler only: key of the first draw of a column within a pass
    __shared__ double cur[49];
    __shared__ uint32_t s_rs[64];
    __shared__ uint32_t s_raw[RAW_RING];
    __shared__ int s_cnt2[2];
    __shared__ double redV[2][NW];           // per-warp winners, double-buffered over rounds (one barrier per arg-max)
    __shared__ int redT[2][NW];
    const double* blk = P.blocks + (size_t)b * P.stride * 49;
    const double* sc = P.obsScore ? P.obsScore + (size_t)b * P.stride : nullptr;
    int32_t* outIdx = P.outIdx + (size_t)b * P.outStride;
    double* outLd = P.outLogdet ? P.outLogdet + (size_t)b * P.outStride : nullptr;
    const int k = P.k;
    if (n <= 0 || k >= n) return;            // trivial problems are written by k_select (launched alongside)
    int mtd = P.mtd;
    if (mtd == MTD_LAZIER_AUTO) mtd = ((size_t)k * 2 > (size_t)n) ? MTD_LAZIER_DELETE : MTD_LAZIER_SELECT;
    const bool deletion = mtd == MTD_LAZIER_DELETE;
    const double sgn = deletion ? -1.0 : 1.0;
    const int szLazier = (int)min((size_t)((double)n / (double)k * P.sampleScale), (size_t)n);
    int allOk = 1;
    for (int i = tid; i < n; i += ALL) {
        idxA[i] = i; vis[i] = -1; win[i] = 0;
        if (sc && sc[i] < 0) allOk = 0;
    }
    if (tid < 49) cur[tid] = (tid % 8 == 0) ? 0.00001 : 0.0;
    allOk = __syncthreads_and(allOk);
    if (!allOk) return;                      // k_select owns this problem
    if (deletion && tid < 49) {              // curMat += every block, in pool order (:2370-2380)
        double s = cur[tid];
        for (int i = 0; i < n; i++) s = s + blk[(size_t)i * 49 + tid];
        cur[tid] = s;
    }
    const int rounds = deletion ? n - k : k;
    double* pk = P.pack + (size_t)b * P.stride * 28;
    for (int c = tid; c < n; c += ALL) pack_block(blk + (size_t)c * 49, pk + (size_t)c * 28);
    __syncthreads();
    if (warp == NW) {
        // ================= sampler warp: RANDOM_ACCESS_LAZIER_GREEDY draws (:2895-2921) of round i, column space n - i
        WarpRand rng;
        rng.rs = s_rs; rng.raw = s_raw; rng.n = 34;
        rng.seed(P.seeds ? P.seeds[b] : 1u, lane);
        int rp = 0;                                   // rand() calls consumed
        int epoch = 1;                                // sampling pass counter
        const unsigned lt = (1u << lane) - 1u;
        for (int i = 0; i < rounds; i++) {
            const int p = i & 1;
            int* colp = p ? col1 : col0;
            const int m = n - i;
            const int szActual = min(szLazier, m);
            if (i >= 2) bar_sync2<B_EMPTY, ALL>(p);   // the evaluators are done with round i - 2's list
            int numHit = 0, rej = 0;
            bool stop = szActual <= 0;
            while (!stop) {
                // one pass = the next 128 draws, 4 per lane (draw number d = h * 32 + lane).  A draw is "new" if its
                // column was not visited in this round: stamp in vis[] from earlier passes, first-in-pass by atomicMax
                // on win[] (key grows with the pass number and shrinks with d, so the earliest draw of a pass wins).
                while (rng.n - 344 < rp + SA_WIN) rng.gen30(lane);
                int j[SA_DPL]; int key[SA_DPL]; bool seen[SA_DPL];
#pragma unroll
                for (int h = 0; h < SA_DPL; h++) {
                    const int d = h * 32 + lane;
                    j[h] = (int)(s_raw[(rp + d) & (RAW_RING - 1)] % (uint32_t)m);
                    seen[h] = vis[j[h]] >= i;
                    key[h] = (epoch << 7) | (SA_WIN - 1 - d);
                    if (!seen[h]) atomicMax(&win[j[h]], key[h]);
                }
                __syncwarp();
                unsigned nm[SA_DPL]; bool isNew[SA_DPL];
#pragma unroll
                for (int h = 0; h < SA_DPL; h++) {
                    isNew[h] = !seen[h] && win[j[h]] == key[h];
                    nm[h] = __ballot_sync(full, isNew[h]);
                }
                int cut = INT_MAX, before = numHit, lastNew = -1;      // lastNew: last new draw before group h (uniform)
                int hbv[SA_DPL];
#pragma unroll
                for (int h = 0; h < SA_DPL; h++) {
                    const int d = h * 32 + lane;
                    const int hb = before + __popc(nm[h] & lt);       // hits before this draw
                    hbv[h] = hb;
                    const bool hitA = isNew[h] && hb + 1 == szActual;  // the draw that completes the subset
                    int run;                                            // consecutive rejections ending at this draw
                    if (nm[h] & lt) run = lane - (31 - __clz(nm[h] & lt));
                    else run = (lastNew >= 0) ? d - lastNew : rej + d + 1;
                    const bool hitB = !isNew[h] && run >= MAX_RANDOM_QUERY_TIME;
                    const unsigned cm = __ballot_sync(full, hitA || hitB);
                    if (cm && cut == INT_MAX) cut = h * 32 + __ffs(cm) - 1;
                    before += __popc(nm[h]);
                    if (nm[h]) lastNew = h * 32 + 31 - __clz(nm[h]);
                }
                stop = cut != INT_MAX;
                if (!stop) cut = SA_WIN - 1;                            // last draw the sequential loop consumes
                int V = 0, lastTaken = -1;
#pragma unroll
                for (int h = 0; h < SA_DPL; h++) {
                    const int d = h * 32 + lane, lo = h * 32;
                    if (d <= cut && isNew[h]) { vis[j[h]] = i; colp[hbv[h]] = j[h]; }
                    if (lo <= cut) {
                        const unsigned take = (cut - lo >= 31) ? full : ((2u << (cut - lo)) - 1u);
                        const unsigned nw = nm[h] & take;
                        V += __popc(nw);
                        if (nw) lastTaken = lo + 31 - __clz(nw);
                    }
                }
                numHit += V;
                rej = (lastTaken >= 0) ? cut - lastTaken : rej + cut + 1;
                rp += cut + 1;
                epoch++;
                __syncwarp();
            }
            if (lane == 0) s_cnt2[p] = numHit;
            __threadfence_block();
            bar_arrive2<B_FULL, ALL>(p);
        }
        return;
    }
    // ================= evaluator threads
    int m = n;
    int* idx = idxA; int* idx2 = idxB;
    int nOut = 0;
    bool dead = false;                                // "if (maxLmk == -1) break": keep the barrier protocol, do nothing
    long long tprev = 0;
    const bool prof = P.prof && b == 0 && tid == 0;
    if (prof) tprev = clock64();
    for (int i = 0; i < rounds; i++) {
        const int p = i & 1;
        const int* colp = p ? col1 : col0;
        if (i == 0) bar_sync2<B_FULL, ALL>(p);            // later lists are waited for at the end of the previous round
        SEL_MARK(0)
        const int cnt = dead ? 0 : s_cnt2[p];
        // ---- logDet(cur +- block) of every drawn column's pool entry; arg-max, first drawn wins ties (strict >)
        double bestV = -DBL_MAX; int bestT = INT_MAX;
        // A round usually holds more candidates than evaluator threads (230 vs 128 in config C4): the packed blocks of this
        // thread's later passes are pulled into L1 while the first pass computes (224 B = at most three 128-B lines).
        for (int t = tid + T; t < cnt; t += T) {
            const char* pf = reinterpret_cast<const char*>(pk + (size_t)idx[colp[t]] * 28);
            asm volatile("prefetch.global.L1 [%0];" ::"l"(pf));
            asm volatile("prefetch.global.L1 [%0];" ::"l"(pf + 112));
            asm volatile("prefetch.global.L1 [%0];" ::"l"(pf + 223));
        }
        for (int t = tid; t < cnt; t += T) {
            const int q = idx[colp[t]];
            const double v = logdet_sum_packed(cur, pk + (size_t)q * 28, blk + (size_t)q * 49, sgn);
            if (v > bestV) { bestV = v; bestT = t; }
        }
        {   // warp arg-max (largest value, then smallest draw number) with three redux ops on an order-preserving integer
            // image of the double instead of five rounds of 64-bit shuffles
            const long long bits = __double_as_longlong(bestV);
            const unsigned long long key = bits < 0 ? ~(unsigned long long)bits : ((unsigned long long)bits | 0x8000000000000000ull);
            const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
            const unsigned mhi = __reduce_max_sync(full, hi);
            const unsigned mlo = __reduce_max_sync(full, hi == mhi ? lo : 0u);
            const bool top = hi == mhi && lo == mlo;
            const int mt = __reduce_min_sync(full, top ? bestT : INT_MAX);
            const int src = __ffs(__ballot_sync(full, top && bestT == mt)) - 1;
            bestV = __shfl_sync(full, bestV, src);
            bestT = mt;
        }
        if (lane == 0) { redV[i & 1][warp] = bestV; redT[i & 1][warp] = bestT; }
        bar_sync_i<B_EVAL, T>();
        double winV = redV[i & 1][0]; int winT = redT[i & 1][0];   // every thread scans the per-warp winners
#pragma unroll
        for (int w = 1; w < NW; w++)
            if (redV[i & 1][w] > winV || (redV[i & 1][w] == winV && redT[i & 1][w] < winT)) { winV = redV[i & 1][w]; winT = redT[i & 1][w]; }
        const int pick = winT == INT_MAX ? -1 : idx[colp[winT]];
        SEL_MARK(1)
        if (i + 2 < rounds) bar_arrive2<B_EMPTY, ALL>(p);     // this thread's last read of list p is behind it
        if (pick < 0) dead = true;
        if (i + 1 < rounds) {
            // The next round's list is there already (the sampler runs ahead) and its columns map to pool entries through
            // this round's index array and pick, before the erase below: pull this thread's first block of the next round
            // into L1 now, so that the L2 round trip overlaps the update, the erase and the barrier.
            bar_sync2<B_FULL, ALL>(p ^ 1);
            if (!dead && tid < s_cnt2[p ^ 1]) {
                const int c = (p ? col0 : col1)[tid];
                const int e0 = idx[c];
                const int qn = (e0 >= pick) ? idx[c + 1] : e0;
                const char* pf = reinterpret_cast<const char*>(pk + (size_t)qn * 28);
                asm volatile("prefetch.global.L1 [%0];" ::"l"(pf));
                asm volatile("prefetch.global.L1 [%0];" ::"l"(pf + 112));
                asm volatile("prefetch.global.L1 [%0];" ::"l"(pf + 223));
            }
        }
        if (!dead) {
            // ---- update the running matrix, record the pick, erase its column (order preserved)
            // the winner's packed triangle was read by its evaluator a moment ago (L1 / L2 hot); blocks are exactly
            // symmetric (H^T H with commutative products), so entry (i, j) is the packed entry (max, min)
            if (tid < 49) {
                const int ui = tid / 7, uj = tid - ui * 7;
                cur[tid] = cur[tid] + sgn * pk[(size_t)pick * 28 + LT(max(ui, uj), min(ui, uj))];
            }
            if (tid == 0 && !deletion && i < P.outStride) { outIdx[i] = pick; if (outLd) outLd[i] = winV; }
            if (!deletion) nOut = i + 1;
            // lmkIdx stays ascending (identity minus erased columns): every entry >= pick moves one column left
            for (int j = tid; j < m - 1; j += T) { const int a = idx[j]; idx2[j] = (a >= pick) ? idx[j + 1] : a; }
            m -= 1;
            { int* t = idx; idx = idx2; idx2 = t; }
        }
        bar_sync_i<B_EVAL, T>();
        SEL_MARK(2)
    }
    if (deletion) {
        // survivors in ascending pool order (:2530-2540) == the columns that are left
        for (int j = tid; j < m && j < P.outStride; j += T) { outIdx[j] = idx[j]; if (outLd) outLd[j] = 0.0; }
        if (tid == 0) P.outN[b] = min(m, P.outStride);
    } else if (tid == 0) P.outN[b] = min(nOut, P.outStride);
}

}  // namespace gforb

using namespace gforb;

// ================================================================================================ host side
struct KineSeg {
    double dt, dt_inSeg;
    double Xv[13], F_Q[16], F_Omg[12], F[169], F_Q_inSeg[16], F_Omg_inSeg[12], F_inSeg[169];
    float Tcw[16];
};

struct GforbObs {
    GforbObsConfig cfg;
    int device = 0, lastLaunches = 0;
    int selThreads = 0, numSMs = 148, selAhead = 1;
    int selResidency = 0;    // gforb_obs_set_select_residency: 0 = by waves, 2 / 3 = register budget of k_select_ahead<128, .>
    int aheadPad = 0;        // same for k_select_ahead (GFORB_SELECT_AHEAD_PAD)
    int selPad = 0;          // extra dynamic smem of k_select (GFORB_SELECT_SMEM_PAD): caps its CTAs per SM when it shares the GPU
    double Xv[13];
    std::vector<KineSeg> kine;
    Arena arena;
    int* d_err = nullptr;
    std::vector<OutCopy> outs;
};

namespace {

// ---- 13-state kinematics (include/Util.hpp:440-494, 541-605, 683-697, 787-1005), all on the host
double nrm3(const double* v) { return std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); }

void dcm_to_quat_f(const float* a, double* q) {      // DCM2QUAT_float
    const float tr = a[0] + a[4] + a[8];
    if (tr > 0) {
        const float s = sqrtf(tr + 1.0f) * 2;
        q[0] = 0.25f * s; q[1] = (a[7] - a[5]) / s; q[2] = (a[2] - a[6]) / s; q[3] = (a[3] - a[1]) / s;
    } else if (a[0] > a[4] && a[0] > a[8]) {
        const float s = 2.0f * sqrtf(1.0f + a[0] - a[4] - a[8]);
        q[0] = (a[7] - a[5]) / s; q[1] = 0.25f * s; q[2] = (a[1] + a[3]) / s; q[3] = (a[2] + a[6]) / s;
    } else if (a[4] > a[8]) {
        const float s = 2.0f * sqrtf(1.0f + a[4] - a[0] - a[8]);
        q[0] = (a[2] - a[6]) / s; q[1] = (a[1] + a[3]) / s; q[2] = 0.25f * s; q[3] = (a[5] + a[7]) / s;
    } else {
        const float s = 2.0f * sqrtf(1.0f + a[8] - a[0] - a[4]);
        q[0] = (a[3] - a[1]) / s; q[1] = (a[2] + a[6]) / s; q[2] = (a[5] + a[7]) / s; q[3] = 0.25f * s;
    }
    if (q[0] < 0) for (int i = 0; i < 4; i++) q[i] = -1.0 * q[i];
}

void inv4_f(const float* Af, float* out) {            // cv::Mat::inv() (LU, partial pivoting) on 4x4 CV_32F
    double A[16], I[16];
    for (int i = 0; i < 16; i++) { A[i] = Af[i]; I[i] = (i % 5 == 0) ? 1.0 : 0.0; }
    for (int k = 0; k < 4; k++) {
        int p = k;
        for (int i = k + 1; i < 4; i++) if (std::fabs(A[i * 4 + k]) > std::fabs(A[p * 4 + k])) p = i;
        if (p != k) for (int c = 0; c < 4; c++) { std::swap(A[k * 4 + c], A[p * 4 + c]); std::swap(I[k * 4 + c], I[p * 4 + c]); }
        const double d = 1.0 / A[k * 4 + k];
        for (int c = 0; c < 4; c++) { A[k * 4 + c] *= d; I[k * 4 + c] *= d; }
        for (int i = 0; i < 4; i++) {
            if (i == k) continue;
            const double f = A[i * 4 + k];
            for (int c = 0; c < 4; c++) { A[i * 4 + c] -= f * A[k * 4 + c]; I[i * 4 + c] -= f * I[k * 4 + c]; }
        }
    }
    for (int i = 0; i < 16; i++) out[i] = (float)I[i];
}

void dq_by_domega(const double* w, double dt, double* RES /*4x3*/) {     // dqomegadt_by_domega
    const double m = nrm3(w);
    if (std::fabs(m) < 1e-8) {
        for (int i = 0; i < 12; i++) RES[i] = 0;
        RES[3] = RES[7] = RES[11] = dt / 2;
        return;
    }
    const double sh = std::sin(m * dt / 2.0), ch = std::cos(m * dt / 2.0);
    for (int c = 0; c < 3; c++) RES[c] = (-dt / 2.0) * (w[c] / m) * sh;
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) {
            if (r == c)
                RES[(r + 1) * 3 + c] = (dt / 2.0) * w[r] * w[r] / (m * m) * ch + (1.0 / m) * (1.0 - w[r] * w[r] / (m * m)) * sh;
            else
                RES[(r + 1) * 3 + c] = (w[r] * w[c] / (m * m)) * ((dt / 2.0) * ch - (1.0 / m) * sh);
        }
}

void f_subblock(const double* Xv, double dt, double* F_Q, double* F_Omg) {   // compute_F_subblock
    const double* w = Xv + 10;
    const double* qo = Xv + 3;
    const double v[3] = {w[0] * dt, w[1] * dt, w[2] * dt};
    const double theta = nrm3(v);
    double q[4] = {1, 0, 0, 0};
    if (!(theta < 1e-6)) {
        const double vn[3] = {v[0] / theta, v[1] / theta, v[2] / theta};
        const double nn = nrm3(vn), s = std::sin(theta / 2.0);
        q[0] = std::cos(theta / 2.0);
        for (int i = 0; i < 3; i++) q[i + 1] = s * (vn[i] / nn);
    }
    {
        const double R = q[0], X = q[1], Y = q[2], Z = q[3];
        const double fq[16] = {R, -X, -Y, -Z, X, R, Z, -Y, Y, -Z, R, X, Z, Y, -X, R};
        memcpy(F_Q, fq, sizeof(fq));
    }
    const double R = qo[0], X = qo[1], Y = qo[2], Z = qo[3];
    const double d31[16] = {R, -X, -Y, -Z, X, R, -Z, Y, Y, Z, R, -X, Z, -Y, X, R};
    double dq[12];
    dq_by_domega(w, dt, dq);
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 3; c++) {
            double s = 0;
            for (int k = 0; k < 4; k++) s += d31[r * 4 + k] * dq[k * 3 + c];
            F_Omg[r * 3 + c] = s;
        }
}

void assemble_F(double dt, const double* F_Q, const double* F_Omg, double* F) {
    for (int i = 0; i < 169; i++) F[i] = (i % 14 == 0) ? 1.0 : 0.0;
    F[7] = dt; F[13 + 8] = dt; F[26 + 9] = dt;
    for (int r = 0; r < 4; r++) {
        for (int c = 0; c < 4; c++) F[(3 + r) * 13 + 3 + c] = F_Q[r * 4 + c];
        for (int c = 0; c < 3; c++) F[(3 + r) * 13 + 10 + c] = F_Omg[r * 3 + c];
    }
}

void pwls_to_homo(const double* p, float* Tcw) {      // convert_PWLS_Vec_To_Homo (+ QUAT2DCM_float)
    const double x = p[4], y = p[5], z = p[6], r = p[3];
    float Rwc[9];
    Rwc[0] = r * r + x * x - y * y - z * z; Rwc[1] = 2.0 * (x * y - r * z);         Rwc[2] = 2.0 * (z * x + r * y);
    Rwc[3] = 2.0 * (x * y + r * z);         Rwc[4] = r * r - x * x + y * y - z * z; Rwc[5] = 2.0 * (y * z - r * x);
    Rwc[6] = 2.0 * (z * x - r * y);         Rwc[7] = 2.0 * (y * z + r * x);         Rwc[8] = r * r - x * x - y * y + z * z;
    const float twc[3] = {(float)p[0], (float)p[1], (float)p[2]};
    for (int i = 0; i < 16; i++) Tcw[i] = (i % 5 == 0) ? 1.f : 0.f;
    for (int rr = 0; rr < 3; rr++) {
        double s = 0;
        for (int c = 0; c < 3; c++) { Tcw[rr * 4 + c] = Rwc[c * 3 + rr]; s += (double)(-Rwc[c * 3 + rr]) * twc[c]; }
        Tcw[rr * 4 + 3] = (float)s;
    }
}

void propagate(const KineSeg& cur, double* next) {   // propagate_PWLS
    for (int i = 0; i < 13; i++) next[i] = cur.Xv[i];
    for (int i = 0; i < 3; i++) next[i] = cur.Xv[i] + cur.Xv[7 + i] * cur.dt;
    const double v[3] = {cur.Xv[10] * cur.dt, cur.Xv[11] * cur.dt, cur.Xv[12] * cur.dt};
    double a = nrm3(v), u[3] = {0, 0, 0};
    if (a > 0.0000001) { u[0] = v[0] / a; u[1] = v[1] / a; u[2] = v[2] / a; } else a = 0;
    const double q2[4] = {std::cos(a / 2), u[0] * std::sin(a / 2), u[1] * std::sin(a / 2), u[2] * std::sin(a / 2)};
    const double* q1 = cur.Xv + 3;
    const double A = q1[0], B = q1[1], C = q1[2], D = q1[3], W = q2[0], X = q2[1], Y = q2[2], Z = q2[3];
    const double qr[4] = {A * W - B * X - C * Y - D * Z, A * X + B * W + C * Z - D * Y, A * Y - B * Z + C * W + D * X,
                          A * Z + B * Y - C * X + D * W};
    const double n = std::sqrt(qr[0] * qr[0] + qr[1] * qr[1] + qr[2] * qr[2] + qr[3] * qr[3]);
    for (int i = 0; i < 4; i++) next[3 + i] = qr[i] / n;
}

struct OCtx {
    GforbObs* O;
    cudaStream_t st;
    bool inDev, outDev;
    int rc = GFORB_OK;
    template <class T>
    const T* in(const T* p, size_t n) {
        if (!p || rc) return p;
        if (inDev) return p;
        T* d = (T*)O->arena.take(n * sizeof(T));
        if (!d) { rc = GFORB_ERR_CAPACITY; return nullptr; }
        if (cudaMemcpyAsync(d, p, n * sizeof(T), cudaMemcpyHostToDevice, st) != cudaSuccess) rc = GFORB_ERR_CUDA;
        return d;
    }
    template <class T>
    T* out(T* p, size_t n) {
        if (!p || rc) return p;
        if (outDev) return p;
        T* d = (T*)O->arena.take(n * sizeof(T));
        if (!d) { rc = GFORB_ERR_CAPACITY; return nullptr; }
        O->outs.push_back(OutCopy{p, d, n * sizeof(T)});
        return d;
    }
};

int obs_begin(GforbObs* O, int batch, int stride, int flags, void* stream, OCtx& c) {
    if (!O || batch < 0 || stride < 0) return GFORB_ERR_ARG;
    if (batch > O->cfg.max_batch || stride > O->cfg.max_points) return GFORB_ERR_CAPACITY;
    GF_CUDA_OK(cudaSetDevice(O->device));
    c.O = O; c.st = (cudaStream_t)stream;
    c.inDev = (flags & GFORB_IN_DEVICE) != 0; c.outDev = (flags & GFORB_OUT_DEVICE) != 0;
    O->arena.reset();
    O->outs.clear();
    O->lastLaunches = 0;
    return GFORB_OK;
}

int obs_end(GforbObs* O, OCtx& c, int flags) {
    if (c.rc) return c.rc;
    GF_CUDA_OK(cudaGetLastError());
    for (auto& o : O->outs) GF_CUDA_OK(cudaMemcpyAsync(o.host, o.dev, o.bytes, cudaMemcpyDeviceToHost, c.st));
    if (!(flags & GFORB_ASYNC) || !O->outs.empty()) GF_CUDA_OK(cudaStreamSynchronize(c.st));
    return GFORB_OK;
}

size_t ahead_smem(int maxPoints) { return (size_t)maxPoints * 6 * sizeof(int) + 16; }
size_t select_smem(int maxPoints) { return (size_t)maxPoints * (5 * sizeof(int) + 2) + 16; }

}  // namespace

extern "C" {

int gforb_obs_create(const GforbObsConfig* cfg, GforbObs** out) {
    if (!cfg || !out) return GFORB_ERR_ARG;
    *out = nullptr;
    if (cfg->max_batch <= 0 || cfg->max_points <= 0 || cfg->max_points > 8192 || !(cfg->fu != 0) || !(cfg->fv != 0)) return GFORB_ERR_ARG;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || cfg->device < 0 || cfg->device >= ndev) {
        fprintf(stderr, "[gforb] no usable CUDA device (requested %d of %d): the B200 path has no CPU fallback\n", cfg->device, ndev);
        return GFORB_ERR_CUDA;
    }
    GF_CUDA_OK(cudaSetDevice(cfg->device));
    GforbObs* O = new GforbObs();
    const auto init = [&]() -> int {
    O->cfg = *cfg;
    O->device = cfg->device;
    for (int i = 0; i < 13; i++) O->Xv[i] = 0;
    const size_t B = cfg->max_batch, N = cfg->max_points;
    // staging for host-pointer calls: world/octave/score/gate in, blocks + H + scores + uv + selection out
    GF_CUDA_OK(O->arena.init(B * N * (12 + 4 + 4 + 1 + 392 + 112 + 8 + 8 + 4 + 8 + 64 + 224) + B * 256 + (1u << 20)));
    GF_CUDA_OK(cudaMalloc((void**)&O->d_err, sizeof(int)));
    if (const char* e = getenv("GFORB_SELECT_SMEM_PAD")) O->selPad = atoi(e);
    cudaDeviceGetAttribute(&O->numSMs, cudaDevAttrMultiProcessorCount, cfg->device);
    const int selSmem = (int)select_smem(cfg->max_points) + O->selPad;
    GF_CUDA_OK(raise_dyn_smem(k_select<64>, selSmem, cfg->device));
    GF_CUDA_OK(raise_dyn_smem(k_select<128>, selSmem, cfg->device));
    GF_CUDA_OK(raise_dyn_smem(k_select<256>, selSmem, cfg->device));
    if (const char* e = getenv("GFORB_SELECT_AHEAD_PAD")) O->aheadPad = atoi(e);
    const int aheadSmem = (int)ahead_smem(cfg->max_points) + O->aheadPad;
    GF_CUDA_OK(raise_dyn_smem(k_select_ahead<64, 5>, aheadSmem, cfg->device));
    GF_CUDA_OK(raise_dyn_smem(k_select_ahead<128, 2>, aheadSmem, cfg->device));
    GF_CUDA_OK(raise_dyn_smem(k_select_ahead<128, 3>, aheadSmem, cfg->device));
    GF_CUDA_OK(raise_dyn_smem(k_select_ahead<256, 1>, aheadSmem, cfg->device));
    if (const char* e = getenv("GFORB_SELECT_THREADS")) O->selThreads = atoi(e);
    if (getenv("GFORB_SELECT_NO_AHEAD")) O->selAhead = 0;
        return GFORB_OK;
    };
    if (const int rc = init()) { gforb_obs_destroy(O); return rc; }   // a failure part-way frees what was allocated
    *out = O;
    return GFORB_OK;
}

int gforb_obs_destroy(GforbObs* O) {
    if (!O) return GFORB_ERR_ARG;
    cudaSetDevice(O->device);
    O->arena.release();
    if (O->d_err) cudaFree(O->d_err);
    delete O;
    return GFORB_OK;
}

int gforb_obs_poll_error(GforbObs* O, void* stream) {
    if (!O) return GFORB_ERR_ARG;
    GF_CUDA_OK(cudaSetDevice(O->device));
    GF_CUDA_OK(cudaStreamSynchronize((cudaStream_t)stream));
    int err = 0;
    GF_CUDA_OK(cudaMemcpy(&err, O->d_err, sizeof(int), cudaMemcpyDeviceToHost));
    if (err) { cudaMemset(O->d_err, 0, sizeof(int)); return err; }
    return GFORB_OK;
}

int gforb_obs_last_launches(const GforbObs* O) { return O ? O->lastLaunches : GFORB_ERR_ARG; }

int gforb_obs_set_select_residency(GforbObs* O, int ctas_per_sm) {
    if (!O || (ctas_per_sm != 0 && ctas_per_sm != 2 && ctas_per_sm != 3)) return GFORB_ERR_ARG;
    O->selResidency = ctas_per_sm;
    return GFORB_OK;
}

int gforb_obs_set_state(GforbObs* O, const double* Xv13) {
    if (!O || !Xv13) return GFORB_ERR_ARG;
    memcpy(O->Xv, Xv13, sizeof(O->Xv));
    return GFORB_OK;
}
int gforb_obs_get_state(const GforbObs* O, double* Xv13) {
    if (!O || !Xv13) return GFORB_ERR_ARG;
    memcpy(Xv13, O->Xv, sizeof(O->Xv));
    return GFORB_OK;
}

int gforb_obs_update_pwls(GforbObs* O, double t_prev, const float* Tcw_prev, double t_cur, const float* Twc_cur) {
    if (!O || !Tcw_prev || !Twc_cur) return GFORB_ERR_ARG;
    double* Xv = O->Xv;
    for (int i = 0; i < 13; i++) Xv[i] = 0;
    Xv[0] = Twc_cur[3]; Xv[1] = Twc_cur[7]; Xv[2] = Twc_cur[11];
    const float R[9] = {Twc_cur[0], Twc_cur[1], Twc_cur[2], Twc_cur[4], Twc_cur[5], Twc_cur[6], Twc_cur[8], Twc_cur[9], Twc_cur[10]};
    dcm_to_quat_f(R, Xv + 3);
    float P[16], Trel[16];
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++) {
            double s = 0;
            for (int k = 0; k < 4; k++) s += (double)Tcw_prev[r * 4 + k] * Twc_cur[k * 4 + c];
            P[r * 4 + c] = (float)s;
        }
    inv4_f(P, Trel);
    const float dtf = float(t_cur - t_prev);
    Xv[7] = Trel[3] / dtf; Xv[8] = Trel[7] / dtf; Xv[9] = Trel[11] / dtf;
    const float o21 = Trel[9] - Trel[6], o02 = Trel[2] - Trel[8], o10 = Trel[4] - Trel[1];
    Xv[10] = 0.5 * o21 / dtf; Xv[11] = 0.5 * o02 / dtf; Xv[12] = 0.5 * o10 / dtf;
    return GFORB_OK;
}

int gforb_obs_predict_pwls(GforbObs* O, double dt, int num_seg) {
    if (!O || num_seg < 0) return GFORB_ERR_ARG;
    double cur[13];
    memcpy(cur, O->Xv, sizeof(cur));
    O->kine.clear();
    for (int i = 0; i < num_seg; i++) {
        KineSeg k;
        k.dt = float(dt);
        k.dt_inSeg = k.dt / float(13);
        memcpy(k.Xv, cur, sizeof(cur));
        pwls_to_homo(k.Xv, k.Tcw);
        f_subblock(k.Xv, k.dt_inSeg, k.F_Q_inSeg, k.F_Omg_inSeg);
        assemble_F(k.dt_inSeg, k.F_Q_inSeg, k.F_Omg_inSeg, k.F_inSeg);
        f_subblock(k.Xv, k.dt, k.F_Q, k.F_Omg);
        assemble_F(k.dt, k.F_Q, k.F_Omg, k.F);
        propagate(k, cur);
        O->kine.push_back(k);
    }
    return GFORB_OK;
}

int gforb_obs_num_segments(const GforbObs* O) { return O ? (int)O->kine.size() : GFORB_ERR_ARG; }

int gforb_obs_get_segment(const GforbObs* O, int seg, double* Xv13, double* F169, double* F_Q16, double* F_Omg12,
                          double* F_Q_inSeg16, double* F_Omg_inSeg12, float* Tcw16) {
    if (!O || seg < 0 || seg >= (int)O->kine.size()) return GFORB_ERR_ARG;
    const KineSeg& k = O->kine[seg];
    if (Xv13) memcpy(Xv13, k.Xv, sizeof(k.Xv));
    if (F169) memcpy(F169, k.F, sizeof(k.F));
    if (F_Q16) memcpy(F_Q16, k.F_Q, sizeof(k.F_Q));
    if (F_Omg12) memcpy(F_Omg12, k.F_Omg, sizeof(k.F_Omg));
    if (F_Q_inSeg16) memcpy(F_Q_inSeg16, k.F_Q_inSeg, sizeof(k.F_Q_inSeg));
    if (F_Omg_inSeg12) memcpy(F_Omg_inSeg12, k.F_Omg_inSeg, sizeof(k.F_Omg_inSeg));
    if (Tcw16) memcpy(Tcw16, k.Tcw, sizeof(k.Tcw));
    return GFORB_OK;
}

int gforb_obs_build_info_frame(GforbObs* O, const double* Xv, const float* world, const int32_t* octave,
                               const int32_t* match_score, const uint8_t* valid, const int32_t* counts, int stride,
                               int batch, const float* level_sigma2, int nlevels, double* blocks, double* obs_score,
                               double* H_meas, int flags, void* stream) {
    OCtx c;
    int rc = obs_begin(O, batch, stride, flags, stream, c);
    if (rc) return rc;
    if (!Xv || !world || !octave || !match_score || !valid || !counts || !level_sigma2 || nlevels <= 0 || !blocks || !obs_score)
        return GFORB_ERR_ARG;
    if (batch == 0 || stride == 0) return GFORB_OK;
    const size_t n = (size_t)batch * stride;
    const double* dXv = c.in(Xv, (size_t)batch * 13);
    const float* dW = c.in(world, n * 3);
    const int32_t* dOct = c.in(octave, n);
    const int32_t* dSc = c.in(match_score, n);
    const uint8_t* dV = c.in(valid, n);
    const int32_t* dCnt = c.in(counts, (size_t)batch);
    const float* dLs = c.in(level_sigma2, (size_t)nlevels);
    double* dB = c.out(blocks, n * 49);
    double* dS = c.out(obs_score, n);
    double* dH = c.out(H_meas, n * 14);
    if (c.rc) return c.rc;
    CamDev cam{O->cfg.fu, O->cfg.fv, O->cfg.Cx, O->cfg.Cy};
    k_build_info<<<dim3((stride + BI_THREADS - 1) / BI_THREADS, batch), BI_THREADS, 0, c.st>>>(0, dXv, cam, dW, dOct, dSc, dV, dCnt, stride, dLs, nlevels, 0,
                                                                       0, 0, 0, 0, 0, 0, 0, dB, dS, dH, nullptr);
    O->lastLaunches = 1;
    return obs_end(O, c, flags);
}

int gforb_obs_build_info_map(GforbObs* O, const double* Xv, const float* world, const uint8_t* track_in_view,
                             const int32_t* counts, int stride, int batch, int check_viz, int min_x, int max_x,
                             int min_y, int max_y, double* blocks, double* obs_score, float* uv, double* H_meas,
                             int flags, void* stream) {
    OCtx c;
    int rc = obs_begin(O, batch, stride, flags, stream, c);
    if (rc) return rc;
    if (!Xv || !world || !track_in_view || !counts || !blocks || !obs_score) return GFORB_ERR_ARG;
    if (batch == 0 || stride == 0) return GFORB_OK;
    const size_t n = (size_t)batch * stride;
    const double* dXv = c.in(Xv, (size_t)batch * 13);
    const float* dW = c.in(world, n * 3);
    const uint8_t* dV = c.in(track_in_view, n);
    const int32_t* dCnt = c.in(counts, (size_t)batch);
    double* dB = c.out(blocks, n * 49);
    double* dS = c.out(obs_score, n);
    float* dUv = c.out(uv, n * 2);
    double* dH = c.out(H_meas, n * 14);
    if (c.rc) return c.rc;
    CamDev cam{O->cfg.fu, O->cfg.fv, O->cfg.Cx, O->cfg.Cy};
    // mBoundXInFrame = mBoundYInFrame = 20, mBoundDepth = 0 (include/Observability.h:129-131)
    k_build_info<<<dim3((stride + BI_THREADS - 1) / BI_THREADS, batch), BI_THREADS, 0, c.st>>>(1, dXv, cam, dW, nullptr, nullptr, dV, dCnt, stride, nullptr, 1,
                                                                       check_viz, min_x, max_x, min_y, max_y, 20.0, 20.0, 0.0, dB,
                                                                       dS, dH, dUv);
    O->lastLaunches = 1;
    return obs_end(O, c, flags);
}

int gforb_obs_logdet(GforbObs* O, const double* mats, int n, double* out, int flags, void* stream) {
    OCtx c;
    int rc = obs_begin(O, 0, 0, flags, stream, c);
    if (rc) return rc;
    if (!mats || !out || n < 0) return GFORB_ERR_ARG;
    if (n == 0) return GFORB_OK;
    const double* dM = c.in(mats, (size_t)n * 49);
    double* dO = c.out(out, (size_t)n);
    if (c.rc) return c.rc;
    k_logdet<<<(n + 127) / 128, 128, 0, c.st>>>(dM, n, dO);
    O->lastLaunches = 1;
    return obs_end(O, c, flags);
}

int gforb_obs_rand_stream(GforbObs* O, uint32_t seed, int n, int32_t* out_host) {
    if (!O || !out_host || n <= 0) return GFORB_ERR_ARG;
    GF_CUDA_OK(cudaSetDevice(O->device));
    int32_t* d = nullptr;
    GF_CUDA_OK(cudaMalloc((void**)&d, sizeof(int32_t) * n));
    k_rand_stream<<<1, 32>>>(seed, n, d);
    cudaError_t e = cudaMemcpy(out_host, d, sizeof(int32_t) * n, cudaMemcpyDeviceToHost);
    cudaFree(d);
    GF_CUDA_OK(e);
    return GFORB_OK;
}

int gforb_obs_select(GforbObs* O, const double* blocks, const double* obs_score, const int32_t* counts, int stride,
                     int batch, int k, int mtd, double sample_scale, const uint32_t* seeds, int32_t* out_idx,
                     double* out_logdet, int32_t* out_n, int out_stride, int flags, void* stream) {
    OCtx c;
    int rc = obs_begin(O, batch, stride, flags, stream, c);
    if (rc) return rc;
    if (!blocks || !counts || !out_idx || !out_n || k <= 0 || out_stride <= 0) return GFORB_ERR_ARG;
    if (mtd != MTD_GREEDY && mtd != MTD_LAZIER_SELECT && mtd != MTD_LAZIER_AUTO && mtd != MTD_LAZIER_DELETE) return GFORB_ERR_ARG;
    if (batch == 0) return GFORB_OK;
    const size_t n = (size_t)batch * stride;
    SelParams P;
    P.blocks = c.in(blocks, n * 49);
    P.obsScore = c.in(obs_score, n);
    P.counts = c.in(counts, (size_t)batch);
    P.seeds = c.in(seeds, (size_t)batch);
    P.stride = stride; P.k = k; P.mtd = mtd; P.maxPoints = O->cfg.max_points; P.sampleScale = sample_scale;
    P.outIdx = c.out(out_idx, (size_t)batch * out_stride);
    P.outLogdet = c.out(out_logdet, (size_t)batch * out_stride);
    P.outN = c.out(out_n, (size_t)batch);
    P.outStride = out_stride;
    P.prof = nullptr;
    P.pack = (double*)O->arena.take(n * 28 * sizeof(double));
    if (!P.pack) return GFORB_ERR_CAPACITY;
    if (c.rc) return c.rc;
    static const bool wantProf = getenv("GFORB_SELECT_PROF") != nullptr;
    if (wantProf) { GF_CUDA_OK(cudaMalloc((void**)&P.prof, 8 * sizeof(long long))); GF_CUDA_OK(cudaMemsetAsync(P.prof, 0, 8 * sizeof(long long), c.st)); }
    // threads per problem: a few problems -> shortest chain; a GPU-filling batch -> smallest register footprint
    int T = O->selThreads;
    if (T != 64 && T != 128 && T != 256) T = batch <= O->numSMs / 2 ? 256 : 128;   // B200, 256 problems: 1.62 / 1.56 / 2.28 ms for 256 / 128 / 64
    const size_t smem = select_smem(O->cfg.max_points) + O->selPad;
    // lazier modes: problems whose pool is all valid go to the sampler-ahead kernel, the others (and the trivial ones)
    // to the in-order kernel; each CTA decides for its own problem, so both kernels are launched over the batch
    P.split = (O->selAhead && mtd != MTD_GREEDY) ? 1 : 0;
    if (T == 256) k_select<256><<<batch, 256, smem, c.st>>>(P);
    else if (T == 128) k_select<128><<<batch, 128, smem, c.st>>>(P);
    else k_select<64><<<batch, 64, smem, c.st>>>(P);
    O->lastLaunches = 1;
    if (P.split) {
        const size_t asm_ = ahead_smem(O->cfg.max_points) + O->aheadPad;
        const char* mb = getenv("GFORB_SELECT_MINB");
        // waves of CTAs at 2 / 3 resident per SM; a 3-per-SM wave takes ~1.35x a 2-per-SM wave (measured, tools/bench_select.py)
        const int w2 = (batch + 2 * O->numSMs - 1) / (2 * O->numSMs), w3 = (batch + 3 * O->numSMs - 1) / (3 * O->numSMs);
        const bool three = mb ? atoi(mb) >= 3 : (O->selResidency ? O->selResidency == 3 : (w3 * 1.35 < w2));
        if (T == 256) k_select_ahead<256, 1><<<batch, 256 + 32, asm_, c.st>>>(P);
        else if (T == 128 && !three) k_select_ahead<128, 2><<<batch, 128 + 32, asm_, c.st>>>(P);
        else if (T == 128) k_select_ahead<128, 3><<<batch, 128 + 32, asm_, c.st>>>(P);
        else k_select_ahead<64, 5><<<batch, 64 + 32, asm_, c.st>>>(P);
        O->lastLaunches = 2;
    }
    if (wantProf) {
        long long h[8];
        GF_CUDA_OK(cudaMemcpyAsync(h, P.prof, sizeof(h), cudaMemcpyDeviceToHost, c.st));
        GF_CUDA_OK(cudaStreamSynchronize(c.st));
        fprintf(stderr, "[gforb] k_select problem 0 cycles: sample (ahead kernel: wait for the sampler) %lld  eval+argmax %lld  update+erase %lld\n", h[0], h[1], h[2]);
        cudaFree(P.prof);
    }
    return obs_end(O, c, flags);
}

}  // extern "C"
