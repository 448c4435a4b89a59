// TEST INFRASTRUCTURE ONLY.  CPU restatement of Optimizer::PoseOptimization (src/Optimizer.cc:279-413): motion-only bundle
// adjustment of one SE3 vertex over the frame's matched map points.  The reference drives g2o (vendored, Thirdparty/g2o):
//   OptimizationAlgorithmLevenberg::solve          g2o/core/optimization_algorithm_levenberg.cpp:61-177
//   SparseOptimizer::optimize / update / push/pop  g2o/core/sparse_optimizer.cpp:354-440
//   BaseBinaryEdge::constructQuadraticForm         g2o/core/base_binary_edge.hpp:55-119   (robust branch, point vertex fixed)
//   RobustKernelHuber::robustify                   g2o/core/robust_kernel_impl.cpp:78-91
//   EdgeSE3ProjectXYZ::computeError/linearizeOplus g2o/types/sba/types_six_dof_expmap.{h:172-177,cpp:384-428}
//   VertexSE3Expmap::oplusImpl, SE3Quat::exp/map/* g2o/types/sba/types_six_dof_expmap.h:104-107, g2o/types/slam3d/se3quat.h
//   LinearSolverDense (Eigen LDLT)                 g2o/solvers/dense/linear_solver_dense.h:64-113
// Eigen itself is not vendored (un-pinned system Eigen3): its small fixed-size products, Quaterniond and LDLT are restated
// from the published algorithms; north_star asks 1e-5 relative on the pose, so association order inside a 2x6 product is
// immaterial for reference parity -- but the CUDA path (csrc/poseopt.cu) follows THIS file operation for operation so that
// the accept / reject decisions of the LM loop and the chi2 outlier tests agree exactly.
#include <cmath>
#include <cstring>
#include <limits>
#include <vector>
#include "oracle.hpp"

namespace oracle {
namespace {

struct Quat { double x, y, z, w; };
struct SE3 { Quat r; double t[3]; };

// Eigen::Quaterniond(Matrix3d) (Eigen/src/Geometry/Quaternion.h, quaternionbase_assign_impl)
Quat quat_from_R(const double R[9]) {
    Quat q;
    double t = R[0] + R[4] + R[8];
    if (t > 0.0) {
        t = std::sqrt(t + 1.0);
        q.w = 0.5 * t;
        t = 0.5 / t;
        q.x = (R[7] - R[5]) * t;
        q.y = (R[2] - R[6]) * t;
        q.z = (R[3] - R[1]) * t;
    } else {
        int i = 0;
        if (R[4] > R[0]) i = 1;
        if (R[8] > R[i * 4]) i = 2;
        const int j = (i + 1) % 3, k = (j + 1) % 3;
        t = std::sqrt(R[i * 4] - R[j * 4] - R[k * 4] + 1.0);
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
// SE3Quat::normalizeRotation (se3quat.h:280-285)
void normalize_rotation(Quat& q) {
    if (q.w < 0) { q.x *= -1; q.y *= -1; q.z *= -1; q.w *= -1; }
    const double n = std::sqrt(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
    q.x /= n; q.y /= n; q.z /= n; q.w /= n;
}
// Eigen quaternion product a * b
Quat quat_mul(const Quat& a, const Quat& b) {
    Quat r;
    r.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
    r.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
    r.y = a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z;
    r.z = a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x;
    return r;
}
// Eigen QuaternionBase::_transformVector: v + w * uv + q.vec x uv, uv = 2 * (q.vec x v)
void quat_rot(const Quat& q, const double v[3], double o[3]) {
    double uv[3] = {q.y * v[2] - q.z * v[1], q.z * v[0] - q.x * v[2], q.x * v[1] - q.y * v[0]};
    uv[0] += uv[0]; uv[1] += uv[1]; uv[2] += uv[2];
    o[0] = v[0] + q.w * uv[0] + (q.y * uv[2] - q.z * uv[1]);
    o[1] = v[1] + q.w * uv[1] + (q.z * uv[0] - q.x * uv[2]);
    o[2] = v[2] + q.w * uv[2] + (q.x * uv[1] - q.y * uv[0]);
}
// Eigen QuaternionBase::toRotationMatrix (row-major out)
void quat_to_R(const Quat& q, double R[9]) {
    const double tx = 2 * q.x, ty = 2 * q.y, tz = 2 * q.z;
    const double twx = tx * q.w, twy = ty * q.w, twz = tz * q.w;
    const double txx = tx * q.x, txy = ty * q.x, txz = tz * q.x;
    const double tyy = ty * q.y, tyz = tz * q.y, tzz = tz * q.z;
    R[0] = 1 - (tyy + tzz); R[1] = txy - twz; R[2] = txz + twy;
    R[3] = txy + twz; R[4] = 1 - (txx + tzz); R[5] = tyz - twx;
    R[6] = txz - twy; R[7] = tyz + twx; R[8] = 1 - (txx + tyy);
}
// SE3Quat::exp (se3quat.h:223-257) then SE3Quat(Quaterniond(R), V * upsilon)
SE3 se3_exp(const double u[6]) {
    const double om[3] = {u[0], u[1], u[2]}, up[3] = {u[3], u[4], u[5]};
    const double theta = std::sqrt(om[0] * om[0] + om[1] * om[1] + om[2] * om[2]);
    const double O[9] = {0, -om[2], om[1], om[2], 0, -om[0], -om[1], om[0], 0};
    double O2[9];
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) O2[r * 3 + c] = O[r * 3] * O[c] + O[r * 3 + 1] * O[3 + c] + O[r * 3 + 2] * O[6 + c];
    double R[9], V[9];
    if (theta < 0.00001) {
        for (int i = 0; i < 9; i++) { R[i] = ((i % 4 == 0) ? 1.0 : 0.0) + O[i] + O2[i]; V[i] = R[i]; }
    } else {
        const double a = std::sin(theta) / theta, b = (1 - std::cos(theta)) / (theta * theta);
        const double c = (theta - std::sin(theta)) / std::pow(theta, 3);
        for (int i = 0; i < 9; i++) {
            const double I = (i % 4 == 0) ? 1.0 : 0.0;
            R[i] = I + a * O[i] + b * O2[i];
            V[i] = I + b * O[i] + c * O2[i];
        }
    }
    SE3 T;
    T.r = quat_from_R(R);
    for (int r = 0; r < 3; r++) T.t[r] = V[r * 3] * up[0] + V[r * 3 + 1] * up[1] + V[r * 3 + 2] * up[2];
    normalize_rotation(T.r);
    return T;
}
// SE3Quat::operator* (se3quat.h:104-110)
SE3 se3_mul(const SE3& a, const SE3& b) {
    SE3 r = a;
    double rt[3];
    quat_rot(a.r, b.t, rt);
    r.t[0] += rt[0]; r.t[1] += rt[1]; r.t[2] += rt[2];
    r.r = quat_mul(a.r, b.r);
    normalize_rotation(r.r);
    return r;
}

// Eigen::LDLT<MatrixXd> of a 6x6 SPD matrix: in-place unblocked, diagonal pivoting on the largest |diagonal| (lower part),
// then solve P^T L D L^T P x = b.  Returns isPositive().
bool ldlt_solve6(const double H[36], const double b[6], double x[6]) {
    const int n = 6;
    double A[36];
    memcpy(A, H, sizeof(A));
    int tr[6];
    int sign = 0;     // 0 zero, 1 positive, -1 negative, 2 indefinite
    for (int k = 0; k < n; k++) {
        int p = k;
        double big = std::fabs(A[k * n + k]);
        for (int i = k + 1; i < n; i++) if (std::fabs(A[i * n + i]) > big) { big = std::fabs(A[i * n + i]); p = i; }
        tr[k] = p;
        if (p != k) {          // symmetric swap of rows / columns k and p, lower triangle storage
            for (int c = 0; c < k; c++) std::swap(A[k * n + c], A[p * n + c]);
            for (int r = p + 1; r < n; r++) std::swap(A[r * n + k], A[r * n + p]);
            std::swap(A[k * n + k], A[p * n + p]);
            for (int i = k + 1; i < p; i++) std::swap(A[i * n + k], A[p * n + i]);
        }
        // A21 -= A20 * (D0 .* A10^T); A11 -= A10 * temp
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
        if (std::fabs(d) > 0) for (int r = k + 1; r < n; r++) A[r * n + k] /= d;     // (Eigen: if abs(pivot) > tolerance 1/highest)
        if (sign == 0) sign = d > 0 ? 1 : (d < 0 ? -1 : 0);
        else if (sign == 1 && d < 0) sign = 2;
        else if (sign == -1 && d > 0) sign = 2;
    }
    // solve: y = P b; L z = y; w = D^-1 z; L^T v = w; x = P^T v
    double y[6];
    memcpy(y, b, sizeof(y));
    for (int k = 0; k < n; k++) std::swap(y[k], y[tr[k]]);
    for (int r = 0; r < n; r++) for (int c = 0; c < r; c++) y[r] -= A[r * n + c] * y[c];
    for (int r = 0; r < n; r++) {
        const double d = A[r * n + r];
        // Eigen: tolerance = 1 / max double; pseudo-inverse of D
        if (std::fabs(d) > (1.0 / std::numeric_limits<double>::max())) y[r] /= d; else y[r] = 0;
    }
    for (int r = n - 1; r >= 0; r--) for (int c = r + 1; c < n; c++) y[r] -= A[c * n + r] * y[c];
    for (int k = n - 1; k >= 0; k--) std::swap(y[k], y[tr[k]]);
    memcpy(x, y, sizeof(y));
    return sign == 1 || sign == 0;
}

struct Edge { double X[3], obs[2], info, invSigma2; int idx; };

struct Problem {
    SE3 T;
    std::vector<Edge> edges;
    std::vector<double> err;       // 2 per edge: _error of the last computeError
    double fx, fy, cx, cy;
    double H[36], b[6], x[6];

    void map(const double X[3], double o[3]) const { quat_rot(T.r, X, o); o[0] = T.t[0] + o[0]; o[1] = T.t[1] + o[1]; o[2] = T.t[2] + o[2]; }
    void computeError(size_t i) {      // EdgeSE3ProjectXYZ::computeError + cam_project + project2d
        double p[3];
        map(edges[i].X, p);
        const double px = p[0] / p[2], py = p[1] / p[2];
        err[2 * i] = edges[i].obs[0] - (px * fx + cx);
        err[2 * i + 1] = edges[i].obs[1] - (py * fy + cy);
    }
    double chi2(size_t i) const { return err[2 * i] * (edges[i].info * err[2 * i]) + err[2 * i + 1] * (edges[i].info * err[2 * i + 1]); }
    void computeActiveErrors() { for (size_t i = 0; i < edges.size(); i++) computeError(i); }
    static void huber(double e, double delta, double dsqr, double rho[3]) {
        if (e <= dsqr) { rho[0] = e; rho[1] = 1.; rho[2] = 0.; }
        else { const double sqrte = std::sqrt(e); rho[0] = 2 * sqrte * delta - dsqr; rho[1] = delta / sqrte; rho[2] = -0.5 * rho[1] / e; }
    }
    double activeRobustChi2(double delta, double dsqr) const {
        double chi = 0.0;
        for (size_t i = 0; i < edges.size(); i++) { double rho[3]; huber(chi2(i), delta, dsqr, rho); chi += rho[0]; }
        return chi;
    }
    void buildSystem(double delta, double dsqr) {
        memset(H, 0, sizeof(H)); memset(b, 0, sizeof(b));
        for (size_t i = 0; i < edges.size(); i++) {
            double p[3];
            map(edges[i].X, p);
            const double x_ = p[0], y_ = p[1], z_ = p[2], z_2 = z_ * z_;
            double J[12];      // _jacobianOplusXj (2 x 6)
            J[0] = x_ * y_ / z_2 * fx; J[1] = -(1 + (x_ * x_ / z_2)) * fx; J[2] = y_ / z_ * fx; J[3] = -1. / z_ * fx; J[4] = 0; J[5] = x_ / z_2 * fx;
            J[6] = (1 + y_ * y_ / z_2) * fy; J[7] = -x_ * y_ / z_2 * fy; J[8] = -x_ / z_ * fy; J[9] = 0; J[10] = -1. / z_ * fy; J[11] = y_ / z_2 * fy;
            const double om = edges[i].info;
            double omega_r[2] = {-(om * err[2 * i]), -(om * err[2 * i + 1])};
            double rho[3];
            huber(chi2(i), delta, dsqr, rho);
            const double w = rho[1] * om;                       // robustInformation
            omega_r[0] *= rho[1]; omega_r[1] *= rho[1];
            for (int j = 0; j < 6; j++) b[j] += J[j] * omega_r[0] + J[6 + j] * omega_r[1];
            for (int j = 0; j < 6; j++) {
                const double t0 = J[j] * w, t1 = J[6 + j] * w;   // (B^T * weightedOmega) row j
                for (int k = 0; k < 6; k++) H[j * 6 + k] += t0 * J[k] + t1 * J[6 + k];
            }
        }
    }
};

}  // namespace

// Returns nInitialCorrespondences - nBad; Tcw (row-major 4x4 float) is updated in place, outlier[i] <- mvbOutlier[i] for the
// matched keypoints (entries with mapPoint[i] < 0 are left alone).
int PoseOptimization(float* Tcw, int N, const float* kpXY, const int* octave, const int* mapPoint, const float* world,
                     const float* invLevelSigma2, float fx, float fy, float cx, float cy, uint8_t* outlier) {
    Problem P;
    double R[9], t[3];
    for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) R[r * 3 + c] = Tcw[r * 4 + c]; t[r] = Tcw[r * 4 + 3]; }
    P.T.r = quat_from_R(R);
    normalize_rotation(P.T.r);
    P.T.t[0] = t[0]; P.T.t[1] = t[1]; P.T.t[2] = t[2];
    P.fx = fx; P.fy = fy; P.cx = cx; P.cy = cy;
    int nInitialCorrespondences = 0;
    const float deltaF = std::sqrt(5.991);                     // const float delta = sqrt(5.991) (:316)
    const double delta = deltaF, dsqr = delta * delta;          // rk->setDelta(delta): dsqr = delta * delta
    for (int i = 0; i < N; i++)
        if (mapPoint[i] >= 0) {
            Edge e;
            e.X[0] = world[3 * i]; e.X[1] = world[3 * i + 1]; e.X[2] = world[3 * i + 2];
            e.obs[0] = kpXY[2 * i]; e.obs[1] = kpXY[2 * i + 1];
            const float invSigma2 = invLevelSigma2[octave[i]];
            e.info = invSigma2; e.invSigma2 = invSigma2; e.idx = i;
            P.edges.push_back(e);
            nInitialCorrespondences++;
            outlier[i] = 0;
        }
    P.err.assign(2 * P.edges.size(), 0.0);
    if (P.edges.empty()) return 0;      // g2o: "0 vertices to optimize" -> nothing changes (pose is written back unchanged below in the reference)
    const float chi2Th[4] = {9.210f, 7.378f, 5.991f, 5.991f};
    const int its[4] = {10, 10, 7, 5};
    const double tau = 1e-5, goodStepUpperScale = 2. / 3., goodStepLowerScale = 1. / 3.;
    const int maxTrialsAfterFailure = 10;
    int nBad = 0;
    for (int it = 0; it < 4; it++) {
        // optimizer.initializeOptimization(); optimizer.optimize(its[it])
        double currentLambda = -1., ni = 2.;
        int nBadLM = 0;
        bool ok = true;
        for (int iter = 0; iter < its[it] && ok; iter++) {
            P.computeActiveErrors();
            double currentChi = P.activeRobustChi2(delta, dsqr);
            double tempChi = currentChi;
            const double iniChi = currentChi;
            P.buildSystem(delta, dsqr);
            if (iter == 0) {
                double maxDiagonal = 0.;
                for (int j = 0; j < 6; j++) maxDiagonal = std::max(std::fabs(P.H[j * 6 + j]), maxDiagonal);
                currentLambda = tau * maxDiagonal;
                ni = 2; nBadLM = 0;
            }
            double rho = 0;
            int qmax = 0;
            do {
                const SE3 backup = P.T;                                 // _optimizer->push()
                double Hl[36];
                memcpy(Hl, P.H, sizeof(Hl));
                for (int j = 0; j < 6; j++) Hl[j * 6 + j] += currentLambda;     // setLambda(_currentLambda, true)
                const bool ok2 = ldlt_solve6(Hl, P.b, P.x);
                P.T = se3_mul(se3_exp(P.x), P.T);                       // oplusImpl: exp(update) * estimate
                P.computeActiveErrors();
                tempChi = P.activeRobustChi2(delta, dsqr);
                if (!ok2) tempChi = std::numeric_limits<double>::max();
                rho = (currentChi - tempChi);
                double scale = 0.;
                for (int j = 0; j < 6; j++) scale += P.x[j] * (currentLambda * P.x[j] + P.b[j]);
                scale += 1e-3;
                rho /= scale;
                if (rho > 0 && std::isfinite(tempChi)) {
                    double alpha = 1. - std::pow((2 * rho - 1), 3);
                    alpha = (std::min)(alpha, goodStepUpperScale);
                    const double scaleFactor = (std::max)(goodStepLowerScale, alpha);
                    currentLambda *= scaleFactor;
                    ni = 2;
                    currentChi = tempChi;
                } else {
                    currentLambda *= ni;
                    ni *= 2;
                    P.T = backup;                                        // pop(): errors stay those of the rejected trial
                }
                qmax++;
            } while (rho < 0 && qmax < maxTrialsAfterFailure);
            if (qmax == maxTrialsAfterFailure || rho == 0) { ok = false; break; }      // Terminate
            if ((iniChi - currentChi) * 1e3 < iniChi) nBadLM++; else nBadLM = 0;
            if (nBadLM >= 3) { ok = false; break; }
        }
        nBad = 0;
        for (size_t i = 0; i < P.edges.size(); i++) {
            const int idx = P.edges[i].idx;
            if (outlier[idx]) { P.edges[i].info = P.edges[i].invSigma2; P.computeError(i); }
            const double c2 = P.chi2(i);
            if (c2 > chi2Th[it]) { outlier[idx] = 1; P.edges[i].info = 1e-10; nBad++; }
            else if (c2 <= chi2Th[it]) outlier[idx] = 0;
        }
        if (P.edges.size() < 10) break;
    }
    double Ro[9];
    quat_to_R(P.T.r, Ro);
    for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) Tcw[r * 4 + c] = (float)Ro[r * 3 + c]; Tcw[r * 4 + 3] = (float)P.T.t[r]; }
    Tcw[12] = 0; Tcw[13] = 0; Tcw[14] = 0; Tcw[15] = 1;
    return nInitialCorrespondences - nBad;
}

}  // namespace oracle
