// Optimizer::PoseOptimization (src/Optimizer.cc:279-413) for a batch of frames: motion-only bundle adjustment of one SE3
// vertex per frame over the frame's matched map points -- the step Tracking runs right after the matchers of this path
// (TrackWithMotionModel :1596, TrackLocalMap :2776).  The reference drives g2o: Levenberg-Marquardt
// (g2o/core/optimization_algorithm_levenberg.cpp:61-177) on EdgeSE3ProjectXYZ edges with a Huber kernel (delta = sqrt(5.991)),
// dense 6x6 LDLT solve, SE3Quat exponential update, 4 rounds of {10,10,7,5} iterations with the chi2 outlier re-test
// {9.210, 7.378, 5.991, 5.991} (outliers stay in the graph with information 1e-10).
//
// One CTA per frame.  Per LM step the edges are evaluated in parallel (error, Jacobian, Huber weight: one thread per edge,
// 128 at a time into a shared-memory chunk) and the 21 + 6 + 1 sums of the normal equations are taken by 28 threads, each
// adding its row IN EDGE ORDER -- the summation order of g2o's sequential loop -- so H, b and the robust chi2 are
// bit-identical to the scalar sequence and every accept / reject decision of the LM loop and every chi2 test falls the
// same way (this TU is built with --fmad=false; the operation order is that of oracle/pose_optimizer.cpp).  Thread 0 runs
// the scalar part (lambda schedule, pivoted LDLT, exp map).  The kernel is a latency chain (<= 32 x ~2 sequential sums of N
// terms); throughput comes from the batch: hundreds of frames in flight, two CTAs per SM.
#include <cfloat>
#include <cmath>
#include <cstring>
#include <vector>
#include "common.cuh"
#include "host_util.h"

namespace gforb {

struct PQuat { double x, y, z, w; };
struct PSE3 { PQuat r; double t[3]; };

__device__ __forceinline__ PQuat pq_from_R(const double R[9]) {       // Eigen::Quaterniond(Matrix3d)
    PQuat q;
    double t = R[0] + R[4] + R[8];
    if (t > 0.0) {
        t = sqrt(t + 1.0);
        q.w = 0.5 * t;
        t = 0.5 / t;
        q.x = (R[7] - R[5]) * t; q.y = (R[2] - R[6]) * t; q.z = (R[3] - R[1]) * t;
    } else {
        int i = 0;
        if (R[4] > R[0]) i = 1;
        if (R[8] > R[i * 4]) i = 2;
        const int j = (i + 1) % 3, k = (j + 1) % 3;
        t = sqrt(R[i * 4] - R[j * 4] - R[k * 4] + 1.0);
        double v[3];
        v[i] = 0.5 * t;
        t = 0.5 / t;
        q.w = (R[k * 3 + j] - R[j * 3 + k]) * t;
        v[j] = (R[j * 3 + i] + R[i * 3 + j]) * t;
        v[k] = (R[k * 3 + i] + R[i * 3 + k]) * t;
        q.x = v[0]; q.y = v[1]; q.z = v[2];
    }
    return q;
}
__device__ __forceinline__ void pq_normalize(PQuat& q) {              // SE3Quat::normalizeRotation
    if (q.w < 0) { q.x *= -1; q.y *= -1; q.z *= -1; q.w *= -1; }
    const double n = sqrt(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
    q.x /= n; q.y /= n; q.z /= n; q.w /= n;
}
__device__ __forceinline__ PQuat pq_mul(const PQuat& a, const PQuat& b) {
    PQuat r;
    r.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
    r.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
    r.y = a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z;
    r.z = a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x;
    return r;
}
__device__ __forceinline__ void pq_rot(const PQuat& q, const double v[3], double o[3]) {   // Eigen _transformVector
    double uv[3] = {q.y * v[2] - q.z * v[1], q.z * v[0] - q.x * v[2], q.x * v[1] - q.y * v[0]};
    uv[0] += uv[0]; uv[1] += uv[1]; uv[2] += uv[2];
    o[0] = v[0] + q.w * uv[0] + (q.y * uv[2] - q.z * uv[1]);
    o[1] = v[1] + q.w * uv[1] + (q.z * uv[0] - q.x * uv[2]);
    o[2] = v[2] + q.w * uv[2] + (q.x * uv[1] - q.y * uv[0]);
}
__device__ __forceinline__ void pq_to_R(const PQuat& q, double R[9]) {
    const double tx = 2 * q.x, ty = 2 * q.y, tz = 2 * q.z;
    const double twx = tx * q.w, twy = ty * q.w, twz = tz * q.w;
    const double txx = tx * q.x, txy = ty * q.x, txz = tz * q.x;
    const double tyy = ty * q.y, tyz = tz * q.y, tzz = tz * q.z;
    R[0] = 1 - (tyy + tzz); R[1] = txy - twz; R[2] = txz + twy;
    R[3] = txy + twz; R[4] = 1 - (txx + tzz); R[5] = tyz - twx;
    R[6] = txz - twy; R[7] = tyz + twx; R[8] = 1 - (txx + tyy);
}
__device__ PSE3 pse3_exp(const double u[6]) {                          // SE3Quat::exp (se3quat.h:223-257)
    const double om[3] = {u[0], u[1], u[2]}, up[3] = {u[3], u[4], u[5]};
    const double theta = sqrt(om[0] * om[0] + om[1] * om[1] + om[2] * om[2]);
    const double O[9] = {0, -om[2], om[1], om[2], 0, -om[0], -om[1], om[0], 0};
    double O2[9];
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) O2[r * 3 + c] = O[r * 3] * O[c] + O[r * 3 + 1] * O[3 + c] + O[r * 3 + 2] * O[6 + c];
    double R[9], V[9];
    if (theta < 0.00001) {
        for (int i = 0; i < 9; i++) { R[i] = ((i % 4 == 0) ? 1.0 : 0.0) + O[i] + O2[i]; V[i] = R[i]; }
    } else {
        const double a = sin(theta) / theta, b = (1 - cos(theta)) / (theta * theta);
        const double c = (theta - sin(theta)) / pow(theta, 3.0);
        for (int i = 0; i < 9; i++) {
            const double I = (i % 4 == 0) ? 1.0 : 0.0;
            R[i] = I + a * O[i] + b * O2[i];
            V[i] = I + b * O[i] + c * O2[i];
        }
    }
    PSE3 T;
    T.r = pq_from_R(R);
    for (int r = 0; r < 3; r++) T.t[r] = V[r * 3] * up[0] + V[r * 3 + 1] * up[1] + V[r * 3 + 2] * up[2];
    pq_normalize(T.r);
    return T;
}
__device__ PSE3 pse3_mul(const PSE3& a, const PSE3& b) {               // SE3Quat::operator*
    PSE3 r = a;
    double rt[3];
    pq_rot(a.r, b.t, rt);
    r.t[0] += rt[0]; r.t[1] += rt[1]; r.t[2] += rt[2];
    r.r = pq_mul(a.r, b.r);
    pq_normalize(r.r);
    return r;
}
// Eigen::LDLT<MatrixXd>, 6x6, lower storage, pivoting on the largest |diagonal|; Hl = full row-major with a valid lower triangle
__device__ bool p_ldlt_solve6(const double* Hl, const double* bvec, double* x) {
    const int n = 6;
    double A[36];
    for (int i = 0; i < 36; i++) A[i] = Hl[i];
    int tr[6];
    int sign = 0;
    for (int k = 0; k < n; k++) {
        int p = k;
        double big = fabs(A[k * n + k]);
        for (int i = k + 1; i < n; i++) if (fabs(A[i * n + i]) > big) { big = fabs(A[i * n + i]); p = i; }
        tr[k] = p;
        if (p != k) {
            for (int c = 0; c < k; c++) { const double t = A[k * n + c]; A[k * n + c] = A[p * n + c]; A[p * n + c] = t; }
            for (int r = p + 1; r < n; r++) { const double t = A[r * n + k]; A[r * n + k] = A[r * n + p]; A[r * n + p] = t; }
            { const double t = A[k * n + k]; A[k * n + k] = A[p * n + p]; A[p * n + p] = t; }
            for (int i = k + 1; i < p; i++) { const double t = A[i * n + k]; A[i * n + k] = A[p * n + i]; A[p * n + i] = t; }
        }
        double temp[6];
        for (int c = 0; c < k; c++) temp[c] = A[c * n + c] * A[k * n + c];
        if (k > 0) {
            double s = 0;
            for (int c = 0; c < k; c++) s += A[k * n + c] * temp[c];
            A[k * n + k] -= s;
            for (int r = k + 1; r < n; r++) {
                double s2 = 0;
                for (int c = 0; c < k; c++) s2 += A[r * n + c] * temp[c];
                A[r * n + k] -= s2;
            }
        }
        const double d = A[k * n + k];
        if (fabs(d) > 0) for (int r = k + 1; r < n; r++) A[r * n + k] /= d;
        if (sign == 0) sign = d > 0 ? 1 : (d < 0 ? -1 : 0);
        else if (sign == 1 && d < 0) sign = 2;
        else if (sign == -1 && d > 0) sign = 2;
    }
    double y[6];
    for (int i = 0; i < 6; i++) y[i] = bvec[i];
    for (int k = 0; k < n; k++) { const double t = y[k]; y[k] = y[tr[k]]; y[tr[k]] = t; }
    for (int r = 0; r < n; r++) for (int c = 0; c < r; c++) y[r] -= A[r * n + c] * y[c];
    for (int r = 0; r < n; r++) {
        const double d = A[r * n + r];
        if (fabs(d) > (1.0 / DBL_MAX)) y[r] /= d; else y[r] = 0;
    }
    for (int r = n - 1; r >= 0; r--) for (int c = r + 1; c < n; c++) y[r] -= A[c * n + r] * y[c];
    for (int k = n - 1; k >= 0; k--) { const double t = y[k]; y[k] = y[tr[k]]; y[tr[k]] = t; }
    for (int i = 0; i < 6; i++) x[i] = y[i];
    return sign == 1 || sign == 0;
}

struct PoseParams {
    float* Tcw;                  // batch x 16 in/out
    const GforbKeyPoint* kps;    // batch x stride (mvKeysUn: pt, octave)
    const int32_t* mapPoint;     // batch x stride (>= 0: matched)
    const float* world;          // batch x stride x 3
    const int32_t* counts;       // batch
    int stride;
    const float* invSigma2;      // nlevels
    int nlevels;
    float fx, fy, cx, cy;
    uint8_t* outlier;            // batch x stride in/out (only matched entries are written)
    int32_t* nInliers;           // batch
    // scratch, batch x stride each
    int32_t* eIdx; double* eErr; double* eInfo;
};

static const int PO_T = 128, PO_ROWS = 28;
struct PoseShared {
    PSE3 T, backup;
    double H[36], b[6], x[6];
    double lambda, ni, currentChi, iniChi, sumChi;
    int E, ok, again, qmax, nBadLM, nBad;
};

__global__ void __launch_bounds__(PO_T) k_pose_opt(const PoseParams P) {
    __shared__ PoseShared S;
    __shared__ double contrib[PO_ROWS][PO_T + 1];
    __shared__ int s_scan[PO_T];
    const int tid = threadIdx.x, f = blockIdx.x;
    const size_t base = (size_t)f * P.stride;
    const int N = min(P.counts[f], P.stride);
    const double delta = (double)(float)sqrt(5.991), dsqr = delta * delta;
    const double fx = P.fx, fy = P.fy, cx = P.cx, cy = P.cy;
    // ---- edge list: matched keypoints in index order (block scan over chunks of 128)
    if (tid == 0) S.E = 0;
    __syncthreads();
    for (int c0 = 0; c0 < N; c0 += PO_T) {
        const int i = c0 + tid;
        const int m = (i < N && P.mapPoint[base + i] >= 0) ? 1 : 0;
        s_scan[tid] = m;
        __syncthreads();
        for (int d = 1; d < PO_T; d <<= 1) {
            const int v = tid >= d ? s_scan[tid - d] : 0;
            __syncthreads();
            s_scan[tid] += v;
            __syncthreads();
        }
        if (m) {
            const int e = S.E + s_scan[tid] - 1;
            P.eIdx[base + e] = i;
            P.eInfo[base + e] = (double)P.invSigma2[min(max(P.kps[base + i].octave, 0), P.nlevels - 1)];
            P.outlier[base + i] = 0;                                   // pFrame->mvbOutlier[i] = false (:327)
        }
        __syncthreads();
        if (tid == 0) S.E += s_scan[PO_T - 1];
        __syncthreads();
    }
    const int E = S.E;
    if (tid == 0) {
        const float* Tm = P.Tcw + (size_t)f * 16;
        double R[9];
        for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) R[r * 3 + c] = (double)Tm[r * 4 + c]; S.T.t[r] = (double)Tm[r * 4 + 3]; }
        S.T.r = pq_from_R(R);
        pq_normalize(S.T.r);
        S.nBad = 0;
        for (int i = 0; i < 36; i++) S.H[i] = 0.0;
    }
    __syncthreads();
    if (E == 0) { if (tid == 0) P.nInliers[f] = 0; return; }

    auto load_edge = [&](int e, double X[3], double obs[2]) {
        const int i = P.eIdx[base + e];
        X[0] = (double)P.world[(base + i) * 3]; X[1] = (double)P.world[(base + i) * 3 + 1]; X[2] = (double)P.world[(base + i) * 3 + 2];
        obs[0] = (double)P.kps[base + i].x; obs[1] = (double)P.kps[base + i].y;
    };
    auto edge_error = [&](const PSE3& T, const double X[3], const double obs[2], double p[3], double er[2]) {
        pq_rot(T.r, X, p);
        p[0] = T.t[0] + p[0]; p[1] = T.t[1] + p[1]; p[2] = T.t[2] + p[2];
        const double px = p[0] / p[2], py = p[1] / p[2];
        er[0] = obs[0] - (px * fx + cx);
        er[1] = obs[1] - (py * fy + cy);
    };
    auto huber = [&](double e, double rho[3]) {
        if (e <= dsqr) { rho[0] = e; rho[1] = 1.; rho[2] = 0.; }
        else { const double sq = sqrt(e); rho[0] = 2 * sq * delta - dsqr; rho[1] = delta / sq; rho[2] = -0.5 * rho[1] / e; }
    };
    // computeActiveErrors + (FULL: buildSystem) + activeRobustChi2: rows 0..20 = lower triangle of H (row-major j >= k),
    // 21..26 = b, 27 = robust chi2; every row summed in edge order by one thread
    auto pass = [&](bool full) {
        double acc = 0.0;
        for (int c0 = 0; c0 < E; c0 += PO_T) {
            const int e = c0 + tid;
            if (e < E) {
                double X[3], obs[2], p[3], er[2];
                load_edge(e, X, obs);
                const PSE3 T = S.T;
                edge_error(T, X, obs, p, er);
                P.eErr[(base + e) * 2] = er[0]; P.eErr[(base + e) * 2 + 1] = er[1];
                const double om = P.eInfo[base + e];
                const double chi2 = er[0] * (om * er[0]) + er[1] * (om * er[1]);
                double rho[3];
                huber(chi2, rho);
                contrib[27][tid] = rho[0];
                if (full) {
                    const double x_ = p[0], y_ = p[1], z_ = p[2], z_2 = z_ * z_;
                    double J[12];
                    J[0] = x_ * y_ / z_2 * fx; J[1] = -(1 + (x_ * x_ / z_2)) * fx; J[2] = y_ / z_ * fx; J[3] = -1. / z_ * fx; J[4] = 0; J[5] = x_ / z_2 * fx;
                    J[6] = (1 + y_ * y_ / z_2) * fy; J[7] = -x_ * y_ / z_2 * fy; J[8] = -x_ / z_ * fy; J[9] = 0; J[10] = -1. / z_ * fy; J[11] = y_ / z_2 * fy;
                    double or0 = -(om * er[0]), or1 = -(om * er[1]);
                    const double w = rho[1] * om;
                    or0 *= rho[1]; or1 *= rho[1];
#pragma unroll
                    for (int j = 0; j < 6; j++) contrib[21 + j][tid] = J[j] * or0 + J[6 + j] * or1;
                    int row = 0;
#pragma unroll
                    for (int j = 0; j < 6; j++) {
                        const double t0 = J[j] * w, t1 = J[6 + j] * w;
#pragma unroll
                        for (int k = 0; k <= j; k++) contrib[row++][tid] = t0 * J[k] + t1 * J[6 + k];
                    }
                }
            }
            __syncthreads();
            const int cnt = min(PO_T, E - c0);
            if (tid < PO_ROWS && (full || tid == 27)) {
                for (int i = 0; i < cnt; i++) acc += contrib[tid][i];
            }
            __syncthreads();
        }
        if (tid < PO_ROWS && (full || tid == 27)) {
            if (tid == 27) S.sumChi = acc;
            else if (tid >= 21) S.b[tid - 21] = acc;
            else {
                int j = 0, r = tid;
                while (r > j) { r -= j + 1; j++; }          // tid = j (j + 1) / 2 + k
                S.H[j * 6 + r] = acc;
            }
        }
        __syncthreads();
    };

    const float chi2Th[4] = {9.210f, 7.378f, 5.991f, 5.991f};
    const int its[4] = {10, 10, 7, 5};
    for (int it = 0; it < 4; it++) {
        if (tid == 0) { S.ok = 1; S.lambda = -1.; S.ni = 2.; S.nBadLM = 0; }
        __syncthreads();
        for (int iter = 0; iter < its[it]; iter++) {
            if (!S.ok) break;
            pass(true);
            if (tid == 0) {
                S.currentChi = S.sumChi; S.iniChi = S.sumChi;
                if (iter == 0) {
                    double maxDiagonal = 0.;
                    for (int j = 0; j < 6; j++) maxDiagonal = fmax(fabs(S.H[j * 6 + j]), maxDiagonal);
                    S.lambda = 1e-5 * maxDiagonal; S.ni = 2; S.nBadLM = 0;
                }
                S.qmax = 0;
            }
            __syncthreads();
            double rhoLast = 0;
            do {
                if (tid == 0) {
                    S.backup = S.T;
                    double Hl[36];
                    for (int i = 0; i < 36; i++) Hl[i] = S.H[i];
                    for (int j = 0; j < 6; j++) Hl[j * 6 + j] += S.lambda;
                    S.again = p_ldlt_solve6(Hl, S.b, S.x) ? 1 : 0;        // ok2 (consumed below)
                    S.T = pse3_mul(pse3_exp(S.x), S.T);
                }
                __syncthreads();
                pass(false);
                if (tid == 0) {
                    double tempChi = S.sumChi;
                    if (!S.again) tempChi = DBL_MAX;
                    double rho = S.currentChi - tempChi;
                    double scale = 0.;
                    for (int j = 0; j < 6; j++) scale += S.x[j] * (S.lambda * S.x[j] + S.b[j]);
                    scale += 1e-3;
                    rho /= scale;
                    if (rho > 0 && isfinite(tempChi)) {
                        double alpha = 1. - pow((2 * rho - 1), 3.0);
                        alpha = fmin(alpha, 2. / 3.);
                        const double scaleFactor = fmax(1. / 3., alpha);
                        S.lambda *= scaleFactor;
                        S.ni = 2;
                        S.currentChi = tempChi;
                    } else {
                        S.lambda *= S.ni;
                        S.ni *= 2;
                        S.T = S.backup;
                    }
                    S.qmax++;
                    S.again = (rho < 0 && S.qmax < 10) ? 1 : 0;
                    // Terminate: qmax == maxTrials || rho == 0; then the "nBad" stop criterion
                    if (!S.again) {
                        if (S.qmax == 10 || rho == 0) S.ok = 0;
                        else {
                            if ((S.iniChi - S.currentChi) * 1e3 < S.iniChi) S.nBadLM++; else S.nBadLM = 0;
                            if (S.nBadLM >= 3) S.ok = 0;
                        }
                    }
                    rhoLast = rho;
                }
                __syncthreads();
            } while (S.again);
            (void)rhoLast;
        }
        __syncthreads();
        // ---- outlier re-test of this round (:370-398); nBad is an integer count, order independent
        if (tid == 0) S.nBad = 0;
        __syncthreads();
        for (int e = tid; e < E; e += PO_T) {
            const int i = P.eIdx[base + e];
            double er[2] = {P.eErr[(base + e) * 2], P.eErr[(base + e) * 2 + 1]};
            double om = P.eInfo[base + e];
            if (P.outlier[base + i]) {
                om = (double)P.invSigma2[min(max(P.kps[base + i].octave, 0), P.nlevels - 1)];
                double X[3], obs[2], p[3];
                load_edge(e, X, obs);
                const PSE3 T = S.T;
                edge_error(T, X, obs, p, er);
                P.eErr[(base + e) * 2] = er[0]; P.eErr[(base + e) * 2 + 1] = er[1];
            }
            const double c2 = er[0] * (om * er[0]) + er[1] * (om * er[1]);
            if (c2 > (double)chi2Th[it]) { P.outlier[base + i] = 1; om = 1e-10; atomicAdd(&S.nBad, 1); }
            else if (c2 <= (double)chi2Th[it]) P.outlier[base + i] = 0;
            P.eInfo[base + e] = om;
        }
        __syncthreads();
        if (E < 10) break;
    }
    if (tid == 0) {
        double Ro[9];
        pq_to_R(S.T.r, Ro);
        float* Tm = P.Tcw + (size_t)f * 16;
        for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) Tm[r * 4 + c] = (float)Ro[r * 3 + c]; Tm[r * 4 + 3] = (float)S.T.t[r]; }
        Tm[12] = 0; Tm[13] = 0; Tm[14] = 0; Tm[15] = 1;
        P.nInliers[f] = E - S.nBad;
    }
}

}  // namespace gforb

using namespace gforb;

struct GforbPoseOpt {
    int device = 0, maxBatch = 0, maxPoints = 0;
    int32_t* d_eIdx = nullptr; double* d_eErr = nullptr; double* d_eInfo = nullptr;
    // staging for host-pointer calls
    float* d_T = nullptr; GforbKeyPoint* d_kps = nullptr; int32_t* d_mp = nullptr; float* d_world = nullptr; int32_t* d_counts = nullptr;
    float* d_sig = nullptr; uint8_t* d_out = nullptr; int32_t* d_ninl = nullptr;
};

extern "C" {

int gforb_poseopt_destroy(GforbPoseOpt* h) {
    if (!h) return GFORB_ERR_ARG;
    cudaSetDevice(h->device);
    void* ptrs[] = {h->d_eIdx, h->d_eErr, h->d_eInfo, h->d_T, h->d_kps, h->d_mp, h->d_world, h->d_counts, h->d_sig, h->d_out, h->d_ninl};
    for (void* p : ptrs) if (p) cudaFree(p);
    delete h;
    return GFORB_OK;
}

int gforb_poseopt_create(int max_batch, int max_points, int device, GforbPoseOpt** out) {
    if (!out || max_batch <= 0 || max_points <= 0) return GFORB_ERR_ARG;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || device < 0 || device >= ndev) return GFORB_ERR_CUDA;    // no CPU fallback
    GF_CUDA_OK(cudaSetDevice(device));
    GforbPoseOpt* h = new GforbPoseOpt();
    h->device = device; h->maxBatch = max_batch; h->maxPoints = max_points;
    const size_t n = (size_t)max_batch * max_points;
    bool ok = cudaMalloc((void**)&h->d_eIdx, n * 4) == cudaSuccess && cudaMalloc((void**)&h->d_eErr, n * 16) == cudaSuccess &&
              cudaMalloc((void**)&h->d_eInfo, n * 8) == cudaSuccess && cudaMalloc((void**)&h->d_T, (size_t)max_batch * 64) == cudaSuccess &&
              cudaMalloc((void**)&h->d_kps, n * sizeof(GforbKeyPoint)) == cudaSuccess && cudaMalloc((void**)&h->d_mp, n * 4) == cudaSuccess &&
              cudaMalloc((void**)&h->d_world, n * 12) == cudaSuccess && cudaMalloc((void**)&h->d_counts, (size_t)max_batch * 4) == cudaSuccess &&
              cudaMalloc((void**)&h->d_sig, MAX_LEVELS * 4) == cudaSuccess && cudaMalloc((void**)&h->d_out, n) == cudaSuccess &&
              cudaMalloc((void**)&h->d_ninl, (size_t)max_batch * 4) == cudaSuccess;
    if (!ok) { gforb_poseopt_destroy(h); return GFORB_ERR_CUDA; }
    *out = h;
    return GFORB_OK;
}

int gforb_pose_optimization(GforbPoseOpt* h, float* Tcw, const GforbKeyPoint* keys_un, const int32_t* map_points, const float* world,
                            const int32_t* counts, int stride, int batch, const float* inv_level_sigma2, int nlevels, float fx, float fy,
                            float cx, float cy, uint8_t* outlier, int32_t* n_inliers, int flags, void* stream) {
    if (!h || !Tcw || !keys_un || !map_points || !world || !counts || !inv_level_sigma2 || !outlier || !n_inliers) return GFORB_ERR_ARG;
    if (batch < 0 || stride <= 0 || nlevels <= 0 || nlevels > MAX_LEVELS) return GFORB_ERR_ARG;
    if (batch > h->maxBatch || stride > h->maxPoints) return GFORB_ERR_CAPACITY;
    if (batch == 0) return GFORB_OK;
    GF_CUDA_OK(cudaSetDevice(h->device));
    cudaStream_t st = (cudaStream_t)stream;
    const bool inDev = flags & GFORB_IN_DEVICE, outDev = flags & GFORB_OUT_DEVICE;
    const size_t n = (size_t)batch * stride;
    PoseParams P;
    memset(&P, 0, sizeof(P));
    // Tcw and outlier are in/out: they follow GFORB_OUT_DEVICE; the pure inputs follow GFORB_IN_DEVICE; inv_level_sigma2 is a host array
    GF_CUDA_OK(cudaMemcpyAsync(h->d_sig, inv_level_sigma2, (size_t)nlevels * 4, cudaMemcpyHostToDevice, st));
    if (inDev) { P.kps = keys_un; P.mapPoint = map_points; P.world = world; P.counts = counts; }
    else {
        GF_CUDA_OK(cudaMemcpyAsync(h->d_kps, keys_un, n * sizeof(GforbKeyPoint), cudaMemcpyHostToDevice, st));
        GF_CUDA_OK(cudaMemcpyAsync(h->d_mp, map_points, n * 4, cudaMemcpyHostToDevice, st));
        GF_CUDA_OK(cudaMemcpyAsync(h->d_world, world, n * 12, cudaMemcpyHostToDevice, st));
        GF_CUDA_OK(cudaMemcpyAsync(h->d_counts, counts, (size_t)batch * 4, cudaMemcpyHostToDevice, st));
        P.kps = h->d_kps; P.mapPoint = h->d_mp; P.world = h->d_world; P.counts = h->d_counts;
    }
    if (outDev) { P.Tcw = Tcw; P.outlier = outlier; P.nInliers = n_inliers; }
    else {
        GF_CUDA_OK(cudaMemcpyAsync(h->d_T, Tcw, (size_t)batch * 64, cudaMemcpyHostToDevice, st));
        GF_CUDA_OK(cudaMemcpyAsync(h->d_out, outlier, n, cudaMemcpyHostToDevice, st));
        P.Tcw = h->d_T; P.outlier = h->d_out; P.nInliers = h->d_ninl;
    }
    P.stride = stride; P.invSigma2 = h->d_sig; P.nlevels = nlevels; P.fx = fx; P.fy = fy; P.cx = cx; P.cy = cy;
    P.eIdx = h->d_eIdx; P.eErr = h->d_eErr; P.eInfo = h->d_eInfo;
    k_pose_opt<<<batch, PO_T, 0, st>>>(P);
    GF_CUDA_OK(cudaGetLastError());
    if (!outDev) {
        GF_CUDA_OK(cudaMemcpyAsync(Tcw, h->d_T, (size_t)batch * 64, cudaMemcpyDeviceToHost, st));
        GF_CUDA_OK(cudaMemcpyAsync(outlier, h->d_out, n, cudaMemcpyDeviceToHost, st));
        GF_CUDA_OK(cudaMemcpyAsync(n_inliers, h->d_ninl, (size_t)batch * 4, cudaMemcpyDeviceToHost, st));
    }
    if (!(flags & GFORB_ASYNC) || !outDev) GF_CUDA_OK(cudaStreamSynchronize(st));
    return GFORB_OK;
}

}  // extern "C"
