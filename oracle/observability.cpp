// TEST INFRASTRUCTURE ONLY (see oracle.hpp).  CPU restatement of ORB_SLAM::Observability's good-feature path:
// include/Observability.h:117-295 (ctors, PWLS update/predict), :376-596 (H sub-blocks, re-weighting);
// include/Util.hpp:131-167 (PinHoleCamera), :295-438 (quaternion helpers), :714-731 (logDet), :787-1005 (kinematics);
// src/Observability.cc:386-644 (batchInfoMat_Frame/_Map), :2339-2542, :2815-3155 (greedy selection).
// Armadillo/LAPACK (absent from /root/reference; pinned Armadillo 9.200.7 + OpenBLAS 0.3.5 by build_dep.sh:17-33)
// are restated as plain loops: closed-form 3x3 inverse, unblocked left-looking Cholesky (OpenBLAS potf2_L order),
// partial-pivot LU for the log_det fallback.
#include "oracle.hpp"
#include <queue>
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <cstring>

namespace oracle {

// ------------------------------------------------------------------------------------------------ small helpers
static void q2r(const double* q, double* R) {   // include/Util.hpp:302-317
    double x = q[1], y = q[2], z = q[3], r = q[0];
    R[0] = r * r + x * x - y * y - z * z; R[1] = 2.0 * (x * y - r * z);         R[2] = 2.0 * (z * x + r * y);
    R[3] = 2.0 * (x * y + r * z);         R[4] = r * r - x * x + y * y - z * z; R[5] = 2.0 * (y * z - r * x);
    R[6] = 2.0 * (z * x - r * y);         R[7] = 2.0 * (y * z + r * x);         R[8] = r * r - x * x - y * y + z * z;
}

static void inv3(const double* A, double* B) {   // arma::inv tiny-matrix path (cofactors / det)
    double c00 = A[4] * A[8] - A[5] * A[7], c01 = A[5] * A[6] - A[3] * A[8], c02 = A[3] * A[7] - A[4] * A[6];
    double det = A[0] * c00 + A[1] * c01 + A[2] * c02;
    double id = 1.0 / det;
    B[0] = c00 * id; B[1] = (A[2] * A[7] - A[1] * A[8]) * id; B[2] = (A[1] * A[5] - A[2] * A[4]) * id;
    B[3] = c01 * id; B[4] = (A[0] * A[8] - A[2] * A[6]) * id; B[5] = (A[2] * A[3] - A[0] * A[5]) * id;
    B[6] = c02 * id; B[7] = (A[1] * A[6] - A[0] * A[7]) * id; B[8] = (A[0] * A[4] - A[1] * A[3]) * id;
}

// include/Util.hpp:391-438: columns = dR/dq{0,x,y,z} * a   (3x4, row-major)
static void dRq_times_a_by_dq(const double* q, const double* a, double* RES) {
    double q0 = q[0], qx = q[1], qy = q[2], qz = q[3];
    const double d0[9] = {2 * q0, -2 * qz, 2 * qy, 2 * qz, 2 * q0, -2 * qx, -2 * qy, 2 * qx, 2 * q0};
    const double dx[9] = {2 * qx, 2 * qy, 2 * qz, 2 * qy, -2 * qx, -2 * q0, 2 * qz, 2 * q0, -2 * qx};
    const double dy[9] = {-2 * qy, 2 * qx, 2 * q0, 2 * qx, 2 * qy, 2 * qz, -2 * q0, 2 * qz, -2 * qy};
    const double dz[9] = {-2 * qz, -2 * q0, 2 * qx, 2 * q0, -2 * qz, 2 * qy, 2 * qx, 2 * qy, 2 * qz};
    const double* D[4] = {d0, dx, dy, dz};
    for (int c = 0; c < 4; c++)
        for (int r = 0; r < 3; r++) RES[r * 4 + c] = D[c][r * 3] * a[0] + D[c][r * 3 + 1] * a[1] + D[c][r * 3 + 2] * a[2];
}

// ------------------------------------------------------------------------------------------------ logDet
// include/Util.hpp:714-731: chol(M,'lower') ok -> 2*log(prod(diag)); else real(log_det(M)) via LU.
double logDet(const double* M, int n) {
    double L[13 * 13];
    bool ok = true;
    for (int i = 0; i < n * n; i++) L[i] = M[i];
    for (int j = 0; j < n && ok; j++) {
        double ajj = L[j * n + j];
        for (int k = 0; k < j; k++) ajj -= L[j * n + k] * L[j * n + k];
        if (!(ajj > 0.0)) { ok = false; break; }
        ajj = std::sqrt(ajj);
        L[j * n + j] = ajj;
        double rinv = 1.0 / ajj;
        for (int i = j + 1; i < n; i++) {
            double s = L[i * n + j];
            for (int k = 0; k < j; k++) s -= L[i * n + k] * L[j * n + k];
            L[i * n + j] = s * rinv;
        }
    }
    if (ok) {
        double p = 1.0;
        for (int j = 0; j < n; j++) p *= L[j * n + j];
        return 2 * std::log(p);
    }
    // arma::log_det: LU with partial pivoting, sum of log|u_ii| (real part)
    double A[13 * 13];
    for (int i = 0; i < n * n; i++) A[i] = M[i];
    double ld = 0;
    for (int k = 0; k < n; k++) {
        int p = k;
        for (int i = k + 1; i < n; i++) if (std::fabs(A[i * n + k]) > std::fabs(A[p * n + k])) p = i;
        if (p != k) for (int c = 0; c < n; c++) std::swap(A[k * n + c], A[p * n + c]);
        double piv = A[k * n + k];
        ld += std::log(std::fabs(piv));
        if (piv == 0) continue;
        for (int i = k + 1; i < n; i++) {
            double f = A[i * n + k] / piv;
            for (int c = k + 1; c < n; c++) A[i * n + c] -= f * A[k * n + c];
        }
    }
    return ld;
}

// ------------------------------------------------------------------------------------------------ ctors
static void initCommon(Observability* o) {
    for (int i = 0; i < 13; i++) o->Xv[i] = 0;
}
// include/Util.hpp:134-145 + include/Observability.h:117
Observability::Observability(double f, int nRows, int nCols, double Cx, double Cy, double k1, double k2, double dx, double dy) {
    camera = PinHoleCamera{f / dx, f / dy, nRows, nCols, Cx, Cy, k1, k2, dx, dy};
    initCommon(this);
}
// include/Util.hpp:148-158 + include/Observability.h:165
Observability::Observability(double fu, double fv, int nRows, int nCols, double Cx, double Cy, double k1, double k2) {
    double dx = 0.0112;
    camera = PinHoleCamera{fu, fv, nRows, nCols, Cx, Cy, k1, k2, dx, fu * dx / fv};
    initCommon(this);
}

// ------------------------------------------------------------------------------------------------ kinematics
static void DCM2QUAT_float(const float* a /*3x3 row-major*/, double* q) {   // include/Util.hpp:541-583
    float trace = a[0] + a[4] + a[8];
    if (trace > 0) {
        float s = sqrtf(trace + 1.0f) * 2;
        q[0] = 0.25f * s; q[1] = (a[7] - a[5]) / s; q[2] = (a[2] - a[6]) / s; q[3] = (a[3] - a[1]) / s;
    } else if (a[0] > a[4] && a[0] > a[8]) {
        float s = 2.0f * sqrtf(1.0f + a[0] - a[4] - a[8]);
        q[0] = (a[7] - a[5]) / s; q[1] = 0.25f * s; q[2] = (a[1] + a[3]) / s; q[3] = (a[2] + a[6]) / s;
    } else if (a[4] > a[8]) {
        float s = 2.0f * sqrtf(1.0f + a[4] - a[0] - a[8]);
        q[0] = (a[2] - a[6]) / s; q[1] = (a[1] + a[3]) / s; q[2] = 0.25f * s; q[3] = (a[5] + a[7]) / s;
    } else {
        float s = 2.0f * sqrtf(1.0f + a[8] - a[0] - a[4]);
        q[0] = (a[3] - a[1]) / s; q[1] = (a[2] + a[6]) / s; q[2] = (a[5] + a[7]) / s; q[3] = 0.25f * s;
    }
    if (q[0] < 0) for (int i = 0; i < 4; i++) q[i] = -1.0 * q[i];
}

static void inv4f(const float* Af, float* out) {   // cv::Mat::inv() (LU) on a 4x4 CV_32F
    double A[16], I[16];
    for (int i = 0; i < 16; i++) { A[i] = Af[i]; I[i] = (i % 5 == 0); }
    for (int k = 0; k < 4; k++) {
        int p = k;
        for (int i = k + 1; i < 4; i++) if (std::fabs(A[i * 4 + k]) > std::fabs(A[p * 4 + k])) p = i;
        if (p != k) for (int c = 0; c < 4; c++) { std::swap(A[k * 4 + c], A[p * 4 + c]); std::swap(I[k * 4 + c], I[p * 4 + c]); }
        double d = 1.0 / A[k * 4 + k];
        for (int c = 0; c < 4; c++) { A[k * 4 + c] *= d; I[k * 4 + c] *= d; }
        for (int i = 0; i < 4; i++) if (i != k) {
            double f = A[i * 4 + k];
            for (int c = 0; c < 4; c++) { A[i * 4 + c] -= f * A[k * 4 + c]; I[i * 4 + c] -= f * I[k * 4 + c]; }
        }
    }
    for (int i = 0; i < 16; i++) out[i] = (float)I[i];
}

// include/Util.hpp:907-954 + include/Observability.h:222-259
void Observability::updatePWLSVec(double t_0, const float* Tcw_0, double t_1, const float* Twc_1) {
    for (int i = 0; i < 13; i++) Xv[i] = 0;
    Xv[0] = Twc_1[3]; Xv[1] = Twc_1[7]; Xv[2] = Twc_1[11];
    float R[9] = {Twc_1[0], Twc_1[1], Twc_1[2], Twc_1[4], Twc_1[5], Twc_1[6], Twc_1[8], Twc_1[9], Twc_1[10]};
    DCM2QUAT_float(R, Xv + 3);
    float P[16], T_rel[16];
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++) {
            double s = 0;
            for (int k = 0; k < 4; k++) s += (double)Tcw_0[r * 4 + k] * Twc_1[k * 4 + c];
            P[r * 4 + c] = (float)s;
        }
    inv4f(P, T_rel);
    float dtf = float(t_1 - t_0);
    Xv[7] = T_rel[3] / dtf; Xv[8] = T_rel[7] / dtf; Xv[9] = T_rel[11] / dtf;
    // get_angular_velocity_with_H (include/Util.hpp:683-697): Omega = R - R^T (float), omega = 0.5 * {O(2,1), O(0,2), O(1,0)}
    float o21 = T_rel[9] - T_rel[6], o02 = T_rel[2] - T_rel[8], o10 = T_rel[4] - T_rel[1];
    Xv[10] = 0.5 * o21 / dtf; Xv[11] = 0.5 * o02 / dtf; Xv[12] = 0.5 * o10 / dtf;
}

static double norm3(const double* v) { return std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); }

static double dq0_by_domegaA(double omegaA, double omega, double dt) { return (-dt / 2.0) * (omegaA / omega) * std::sin(omega * dt / 2.0); }
static double dqA_by_domegaA(double omegaA, double omega, double dt) {
    return (dt / 2.0) * omegaA * omegaA / (omega * omega) * std::cos(omega * dt / 2.0) +
           (1.0 / omega) * (1.0 - omegaA * omegaA / (omega * omega)) * std::sin(omega * dt / 2.0);
}
static double dqA_by_domegaB(double omegaA, double omegaB, double omega, double dt) {
    return (omegaA * omegaB / (omega * omega)) * ((dt / 2.0) * std::cos(omega * dt / 2.0) - (1.0 / omega) * std::sin(omega * dt / 2.0));
}
// include/Util.hpp:440-494  (4x3 row-major)
static void dqomegadt_by_domega(const double* w, double dt, double* RES) {
    double m = norm3(w);
    if (std::fabs(m) < 1e-8) {
        const double z[12] = {0, 0, 0, dt / 2, 0, 0, 0, dt / 2, 0, 0, 0, dt / 2};
        memcpy(RES, z, sizeof(z));
        return;
    }
    for (int c = 0; c < 3; c++) RES[c] = dq0_by_domegaA(w[c], m, dt);
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++)
            RES[(r + 1) * 3 + c] = (r == c) ? dqA_by_domegaA(w[r], m, dt) : dqA_by_domegaB(w[r], w[c], m, dt);
}

// include/Util.hpp:787-846
static void compute_F_subblock(const double* Xv, double dt, double* F_Q, double* F_Omg) {
    const double* omegaOld = Xv + 10;
    const double* qOld = Xv + 3;
    double v[3] = {omegaOld[0] * dt, omegaOld[1] * dt, omegaOld[2] * dt};
    double theta = norm3(v);
    double q[4];
    if (theta < 1e-6) { q[0] = 1; q[1] = q[2] = q[3] = 0; }
    else {
        double vn[3] = {v[0] / theta, v[1] / theta, v[2] / theta};
        double nn = norm3(vn), s = std::sin(theta / 2.0);
        q[0] = std::cos(theta / 2.0);
        for (int i = 0; i < 3; i++) q[i + 1] = s * (vn[i] / nn);
    }
    double R = q[0], X = q[1], Y = q[2], Z = q[3];
    const double fq[16] = {R, -X, -Y, -Z, X, R, Z, -Y, Y, -Z, R, X, Z, Y, -X, R};
    memcpy(F_Q, fq, sizeof(fq));
    R = qOld[0]; X = qOld[1]; Y = qOld[2]; Z = qOld[3];
    const double d31[16] = {R, -X, -Y, -Z, X, R, -Z, Y, Y, Z, R, -X, Z, -Y, X, R};
    double dq[12];
    dqomegadt_by_domega(omegaOld, dt, dq);
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 3; c++) {
            double s = 0;
            for (int k = 0; k < 4; k++) s += d31[r * 4 + k] * dq[k * 3 + c];
            F_Omg[r * 3 + c] = s;
        }
}

// include/Util.hpp:956-966
static void assemble_F_matrix(double dt, const double* F_Q, const double* F_Omg, double* F) {
    for (int i = 0; i < 169; i++) F[i] = (i % 14 == 0);
    F[0 * 13 + 7] = dt; F[1 * 13 + 8] = dt; F[2 * 13 + 9] = dt;
    for (int r = 0; r < 4; r++) {
        for (int c = 0; c < 4; c++) F[(3 + r) * 13 + 3 + c] = F_Q[r * 4 + c];
        for (int c = 0; c < 3; c++) F[(3 + r) * 13 + 10 + c] = F_Omg[r * 3 + c];
    }
}

// include/Util.hpp:883-905 (+ QUAT2DCM_float :585-605)
static void convert_PWLS_Vec_To_Homo(const double* p, float* Tcw) {
    double x = p[4], y = p[5], z = p[6], r = p[3];
    float Rwc[9];
    Rwc[0] = r * r + x * x - y * y - z * z; Rwc[1] = 2.0 * (x * y - r * z);         Rwc[2] = 2.0 * (z * x + r * y);
    Rwc[3] = 2.0 * (x * y + r * z);         Rwc[4] = r * r - x * x + y * y - z * z; Rwc[5] = 2.0 * (y * z - r * x);
    Rwc[6] = 2.0 * (z * x - r * y);         Rwc[7] = 2.0 * (y * z + r * x);         Rwc[8] = r * r - x * x - y * y + z * z;
    float twc[3] = {(float)p[0], (float)p[1], (float)p[2]};
    for (int i = 0; i < 16; i++) Tcw[i] = (i % 5 == 0) ? 1.f : 0.f;
    for (int rr = 0; rr < 3; rr++) {
        double s = 0;
        for (int c = 0; c < 3; c++) { Tcw[rr * 4 + c] = Rwc[c * 3 + rr]; s += (double)(-Rwc[c * 3 + rr]) * twc[c]; }
        Tcw[rr * 4 + 3] = (float)s;
    }
}

// include/Util.hpp:968-1005 (v2q :319-337, qProd :339-360, qNormalize :362-367)
static void propagate_PWLS(const KineStruct& cur, double* next) {
    for (int i = 0; i < 13; i++) next[i] = cur.Xv[i];
    for (int i = 0; i < 3; i++) next[i] = cur.Xv[i] + cur.Xv[7 + i] * cur.dt;
    double v[3] = {cur.Xv[10] * cur.dt, cur.Xv[11] * cur.dt, cur.Xv[12] * cur.dt};
    double a = norm3(v), u[3];
    if (a > 0.0000001) { u[0] = v[0] / a; u[1] = v[1] / a; u[2] = v[2] / a; }
    else { a = 0; u[0] = u[1] = u[2] = 0; }
    double q2[4] = {std::cos(a / 2), u[0] * std::sin(a / 2), u[1] * std::sin(a / 2), u[2] * std::sin(a / 2)};
    const double* q1 = cur.Xv + 3;
    double A = q1[0], B = q1[1], C = q1[2], D = q1[3], W = q2[0], X = q2[1], Y = q2[2], Z = q2[3];
    double qr[4] = {A * W - B * X - C * Y - D * Z, A * X + B * W + C * Z - D * Y, A * Y - B * Z + C * W + D * X,
                    A * Z + B * Y - C * X + D * W};
    double n = std::sqrt(qr[0] * qr[0] + qr[1] * qr[1] + qr[2] * qr[2] + qr[3] * qr[3]);
    for (int i = 0; i < 4; i++) next[3 + i] = qr[i] / n;
}

// include/Observability.h:261-295
void Observability::predictPWLSVec(double dt, size_t num_seg_pred) {
    double curXv[13];
    memcpy(curXv, Xv, sizeof(curXv));
    kinematic.clear();
    for (size_t i = 0; i < num_seg_pred; ++i) {
        KineStruct k;
        k.dt = float(dt);
        k.dt_inSeg = k.dt / float(13);
        memcpy(k.Xv, curXv, sizeof(curXv));
        convert_PWLS_Vec_To_Homo(k.Xv, k.Tcw);
        compute_F_subblock(k.Xv, k.dt_inSeg, k.F_Q_inSeg, k.F_Omg_inSeg);
        assemble_F_matrix(k.dt_inSeg, k.F_Q_inSeg, k.F_Omg_inSeg, k.F_inSeg);
        compute_F_subblock(k.Xv, k.dt, k.F_Q, k.F_Omg);
        assemble_F_matrix(k.dt, k.F_Q, k.F_Omg, k.F);
        propagate_PWLS(k, curXv);
        kinematic.push_back(k);
    }
}

// ------------------------------------------------------------------------------------------------ Jacobians
static void pinholeJac(const PinHoleCamera& cam, const double* hrl, double* J) {
    if (std::fabs(hrl[2]) < 1e-6) { for (int i = 0; i < 6; i++) J[i] = 0; return; }
    J[0] = cam.fu / hrl[2]; J[1] = 0.0; J[2] = -hrl[0] * cam.fu / std::pow(hrl[2], 2.0);
    J[3] = 0.0; J[4] = cam.fv / hrl[2]; J[5] = -hrl[1] * cam.fv / std::pow(hrl[2], 2.0);
}

static void assembleH(const double* dh /*2x3*/, const double* R_rw, const double* q_wr, const double* t_rw,
                      double* H13, double* H47) {
    for (int r = 0; r < 2; r++)
        for (int c = 0; c < 3; c++) {
            double s = 0;
            for (int k = 0; k < 3; k++) s += dh[r * 3 + k] * R_rw[k * 3 + c];
            H13[r * 3 + c] = -1.0 * s;
        }
    double qc[4] = {q_wr[0], -q_wr[1], -q_wr[2], -q_wr[3]};   // qconj (include/Util.hpp:295-300)
    double D[12];
    dRq_times_a_by_dq(qc, t_rw, D);
    const double dqbar[4] = {1.0, -1.0, -1.0, -1.0};          // dqbar_by_dq = diag(1,-1,-1,-1)
    for (int r = 0; r < 3; r++) for (int c = 0; c < 4; c++) D[r * 4 + c] *= dqbar[c];
    for (int r = 0; r < 2; r++)
        for (int c = 0; c < 4; c++) {
            double s = 0;
            for (int k = 0; k < 3; k++) s += dh[r * 3 + k] * D[k * 4 + c];
            H47[r * 4 + c] = s;
        }
}

// include/Observability.h:460-515
bool Observability::compute_H_subblock_simplied(const double* Xv_, const double* yi, double* H13, double* H47,
                                                double* dhu_dhrl, bool check_viz, float& u, float& v) const {
    const double* q_wr = Xv_ + 3;
    double R[9], R_rw[9];
    q2r(q_wr, R);
    inv3(R, R_rw);
    double t_rw[3] = {yi[0] - Xv_[0], yi[1] - Xv_[1], yi[2] - Xv_[2]};
    double hrl[3];
    for (int r = 0; r < 3; r++) hrl[r] = R_rw[r * 3] * t_rw[0] + R_rw[r * 3 + 1] * t_rw[1] + R_rw[r * 3 + 2] * t_rw[2];
    if (hrl[2] > 0) {
        u = float(camera.fu) * hrl[0] / hrl[2] + float(camera.Cx);
        v = float(camera.fv) * hrl[1] / hrl[2] + float(camera.Cy);
    } else { u = FLT_MAX; v = FLT_MAX; }
    if (check_viz) {
        if (hrl[2] < 0.0 + mBoundDepth) return false;
        if (u < vizMinX - mBoundXInFrame || u > vizMaxX + mBoundXInFrame) return false;
        if (v < vizMinY - mBoundYInFrame || v > vizMaxY + mBoundYInFrame) return false;
    }
    pinholeJac(camera, hrl, dhu_dhrl);
    assembleH(dhu_dhrl, R_rw, q_wr, t_rw, H13, H47);
    return true;
}

// include/Observability.h:376-456
void Observability::compute_H_subblock_complete(const double* Xv_, const double* yi, const double* zi, double* H13,
                                                double* H47, double* dhu_dhrl, float& residual_u, float& residual_v) const {
    const double* q_wr = Xv_ + 3;
    double R[9], R_rw[9];
    q2r(q_wr, R);
    inv3(R, R_rw);
    double t_rw[3] = {yi[0] - Xv_[0], yi[1] - Xv_[1], yi[2] - Xv_[2]};
    const PinHoleCamera& c = camera;
    double ud = zi[0], vd = zi[1];
    double xd = (zi[0] - c.Cx) * c.dx, yd = (zi[1] - c.Cy) * c.dy;
    double rd2 = xd * xd + yd * yd, rd4 = rd2 * rd2;
    double uu_ud = (1 + c.k1 * rd2 + c.k2 * rd4) + (ud - c.Cx) * (c.k1 + 2 * c.k2 * rd2) * (2 * (ud - c.Cx) * c.dx * c.dx);
    double vu_vd = (1 + c.k1 * rd2 + c.k2 * rd4) + (vd - c.Cy) * (c.k1 + 2 * c.k2 * rd2) * (2 * (vd - c.Cy) * c.dy * c.dy);
    double uu_vd = (ud - c.Cx) * (c.k1 + 2 * c.k2 * rd2) * (2 * (vd - c.Cy) * c.dy * c.dy);
    double vu_ud = (vd - c.Cy) * (c.k1 + 2 * c.k2 * rd2) * (2 * (ud - c.Cx) * c.dx * c.dx);
    double det = uu_ud * vu_vd - uu_vd * vu_ud;
    double Jinv[4] = {vu_vd / det, -uu_vd / det, -vu_ud / det, uu_ud / det};
    double hrl[3];
    for (int r = 0; r < 3; r++) hrl[r] = R_rw[r * 3] * t_rw[0] + R_rw[r * 3 + 1] * t_rw[1] + R_rw[r * 3 + 2] * t_rw[2];
    pinholeJac(camera, hrl, dhu_dhrl);
    const float invz = 1.0 / hrl[2];
    const float u = float(c.fu) * hrl[0] * invz + float(c.Cx);
    const float v = float(c.fv) * hrl[1] * invz + float(c.Cy);
    residual_u = std::fabs(zi[0] - u);
    residual_v = std::fabs(zi[1] - v);
    double dh[6];
    for (int r = 0; r < 2; r++)
        for (int cc = 0; cc < 3; cc++) dh[r * 3 + cc] = Jinv[r * 2] * dhu_dhrl[cc] + Jinv[r * 2 + 1] * dhu_dhrl[3 + cc];
    assembleH(dh, R_rw, q_wr, t_rw, H13, H47);
}

// reWeightInfoMat (include/Observability.h:517-596) with WITH_OCT_LEVELED_NOISE: Sigma_r = sigma2*I,
// W = chol -> sqrt(sigma2) I, H_rw = inv(W) * H_meas; then infMat = H_rw^T H_rw (src/Observability.cc:467-470)
static void infoBlock(const double* H13, const double* H47, double sigma2, double* block, double* H_meas_out) {
    double H[14];
    for (int r = 0; r < 2; r++) {
        for (int c = 0; c < 3; c++) H[r * 7 + c] = H13[r * 3 + c];
        for (int c = 0; c < 4; c++) H[r * 7 + 3 + c] = H47[r * 4 + c];
    }
    if (H_meas_out) memcpy(H_meas_out, H, sizeof(H));
    double w = std::sqrt(sigma2), wi = 1.0 / w;
    double Hr[14];
    for (int i = 0; i < 14; i++) Hr[i] = wi * H[i];
    for (int r = 0; r < 7; r++)
        for (int c = 0; c < 7; c++) block[r * 7 + c] = Hr[r] * Hr[c] + Hr[7 + r] * Hr[7 + c];
}

// src/Observability.cc:386-554 (FRAME_INFO_MATRIX; time cap = infinity; valid[i] = pMP && !outlier && candidate)
void Observability::batchInfoMat_Frame(const float* worldPos, const int* octave, const int* matchScore,
                                       const uint8_t* valid, int n, const float* levelSigma2, double* blocks,
                                       double* obsScore, double* H_meas) {
    if (mKineIdx >= kinematic.size()) return;
    for (int i = 0; i < n; i++) {
        if (!valid[i]) { obsScore[i] = -1.0; continue; }
        double Y[3] = {worldPos[3 * i], worldPos[3 * i + 1], worldPos[3 * i + 2]};
        double H13[6], H47[8], Hp[6];
        float u, v;
        compute_H_subblock_simplied(kinematic[mKineIdx].Xv, Y, H13, H47, Hp, false, u, v);
        float Sigma2 = levelSigma2[octave[i]];
        infoBlock(H13, H47, (double)Sigma2, blocks + (size_t)49 * i, H_meas ? H_meas + (size_t)14 * i : nullptr);
        obsScore[i] = matchScore[i];
    }
}

// src/Observability.cc:556-644 (MAP_INFO_MATRIX; identity noise because F == NULL)
void Observability::batchInfoMat_Map(const float* worldPos, const uint8_t* trackInView, int n, bool check_viz,
                                     int mnMinX, int mnMaxX, int mnMinY, int mnMaxY, double* blocks, double* obsScore,
                                     float* uv, double* H_meas) {
    if (mKineIdx >= kinematic.size()) return;
    vizMinX = mnMinX; vizMaxX = mnMaxX; vizMinY = mnMinY; vizMaxY = mnMaxY;
    for (int i = 0; i < n; i++) {
        obsScore[i] = -2.0;   // untouched
        if (check_viz == false && !trackInView[i]) continue;
        double Y[3] = {worldPos[3 * i], worldPos[3 * i + 1], worldPos[3 * i + 2]};
        double H13[6], H47[8], Hp[6];
        float u, v;
        bool flag = compute_H_subblock_simplied(kinematic[mKineIdx].Xv, Y, H13, H47, Hp, check_viz, u, v);
        if (uv) { uv[2 * i] = u; uv[2 * i + 1] = v; }
        if (!flag) { obsScore[i] = -1.0; continue; }
        infoBlock(H13, H47, 1.0, blocks + (size_t)49 * i, H_meas ? H_meas + (size_t)14 * i : nullptr);
        obsScore[i] = 1.0;
    }
}

// ------------------------------------------------------------------------------------------------ selection
static void addBlock(double* cur, const double* b, double sgn) { for (int i = 0; i < 49; i++) cur[i] = cur[i] + sgn * b[i]; }

// src/Observability.cc:2815-3029 (RANDOM_ACCESS_LAZIER_GREEDY, MAX_RANDOM_QUERY_TIME 2000; time cap ignored)
bool Observability::maxVolSelection_LazierGreedy(size_t stIdx, size_t edIdx, std::vector<GoodPoint>* subVec,
                                                 size_t mpLimit, double sampleScale, double max_time_for_select) {
    if (lmkSelectPool.size() == 0 || subVec == nullptr) return false;
    if (stIdx >= edIdx || edIdx > lmkSelectPool.size() || max_time_for_select < 0) return false;
    subVec->clear();
    if (mpLimit >= edIdx - stIdx) {
        for (size_t i = stIdx; i < edIdx; ++i) { lmkSelectPool[i].selected = true; subVec->push_back(lmkSelectPool[i]); }
        return true;
    }
    size_t szLazierSubset = static_cast<size_t>(double(edIdx - stIdx) / double(mpLimit) * sampleScale);
    double curMat[49];
    for (int i = 0; i < 49; i++) curMat[i] = (i % 8 == 0) ? 0.00001 : 0.0;
    std::vector<double> lmkIdx(edIdx - stIdx), lmkVisited(edIdx - stIdx);
    for (size_t i = stIdx, l = 0; i < edIdx; ++i, ++l) { lmkIdx[l] = i; lmkVisited[l] = -1; }
    auto rnd = rand_fn ? rand_fn : &std::rand;
    for (size_t i = 0; i < mpLimit; ++i) {
        int maxLmk = -1;
        double maxDet = -DBL_MAX;
        size_t numHit = 0, numRndQue = 0, szActualSubset = szLazierSubset;
        if (lmkIdx.size() < szActualSubset) szActualSubset = lmkIdx.size();
        while (numHit < szActualSubset) {
            size_t j = 0;
            numRndQue = 0;
            while (numRndQue < 2000) {
                j = (rnd() % lmkIdx.size());
                if (lmkVisited[j] < i) { lmkVisited[j] = i; break; }
                ++numRndQue;
            }
            if (numRndQue >= 2000) break;
            int queIdx = lmkIdx[j];
            ++numHit;
            if (lmkSelectPool[queIdx].selected == true || lmkSelectPool[queIdx].obs_score < 0) { --numHit; continue; }
            double M[49];
            for (int t = 0; t < 49; t++) M[t] = curMat[t] + lmkSelectPool[queIdx].obs_block[t];
            double curDet = logDet(M, 7);
            if (curDet > maxDet) { maxDet = curDet; maxLmk = queIdx; }
        }
        if (maxLmk == -1) break;
        addBlock(curMat, lmkSelectPool[maxLmk].obs_block, 1.0);
        lmkSelectPool[maxLmk].selected = true;
        subVec->push_back(lmkSelectPool[maxLmk]);
        subVec->back().upper_bound = maxDet;    // (oracle extra) running log-det after this pick
        size_t k = 0;
        for (size_t j = 0; j < lmkIdx.size(); ++j)
            if (lmkIdx[j] != maxLmk) { lmkIdx[k] = lmkIdx[j]; lmkVisited[k] = lmkVisited[j]; k++; }
        lmkIdx.resize(k); lmkVisited.resize(k);
    }
    return true;
}

// src/Observability.cc:2339-2542
bool Observability::maxVolDeletion_LazierGreedy(size_t stIdx, size_t edIdx, std::vector<GoodPoint>* subVec,
                                                size_t mpLimit, double sampleScale, double max_time_for_select) {
    if (lmkSelectPool.size() == 0 || subVec == nullptr) return false;
    if (stIdx >= edIdx || edIdx > lmkSelectPool.size()) return false;
    subVec->clear();
    if (mpLimit >= edIdx - stIdx) {
        for (size_t i = stIdx; i < edIdx; ++i) { lmkSelectPool[i].selected = true; subVec->push_back(lmkSelectPool[i]); }
        return true;
    }
    (void)max_time_for_select;
    size_t szLazierSubset = static_cast<size_t>(double(edIdx - stIdx) / double(mpLimit) * sampleScale);
    double curMat[49];
    for (int i = 0; i < 49; i++) curMat[i] = (i % 8 == 0) ? 0.00001 : 0.0;
    std::vector<double> lmkIdx(edIdx - stIdx), lmkVisited(edIdx - stIdx);
    for (size_t i = stIdx, l = 0; i < edIdx; ++i, ++l) {
        addBlock(curMat, lmkSelectPool[i].obs_block, 1.0);
        lmkSelectPool[i].selected = true;
        lmkIdx[l] = i; lmkVisited[l] = -1;
    }
    auto rnd = rand_fn ? rand_fn : &std::rand;
    size_t mDelLim = edIdx - stIdx - mpLimit;
    for (size_t i = 0; i < mDelLim; ++i) {
        int maxLmk = -1;
        double maxDet = -DBL_MAX;
        size_t numHit = 0, numRndQue = 0, szActualSubset = szLazierSubset;
        if (lmkIdx.size() < szActualSubset) szActualSubset = lmkIdx.size();
        while (numHit < szActualSubset) {
            size_t j = 0;
            numRndQue = 0;
            while (numRndQue < 2000) {
                j = (rnd() % lmkIdx.size());
                if (lmkVisited[j] < i) { lmkVisited[j] = i; break; }
                ++numRndQue;
            }
            if (numRndQue >= 2000) break;
            int queIdx = lmkIdx[j];
            ++numHit;
            if (lmkSelectPool[queIdx].selected == false || lmkSelectPool[queIdx].obs_score < 0) { --numHit; continue; }
            double M[49];
            for (int t = 0; t < 49; t++) M[t] = curMat[t] - lmkSelectPool[queIdx].obs_block[t];
            double curDet = logDet(M, 7);
            if (curDet > maxDet) { maxDet = curDet; maxLmk = queIdx; }
        }
        if (maxLmk == -1) break;
        addBlock(curMat, lmkSelectPool[maxLmk].obs_block, -1.0);
        lmkSelectPool[maxLmk].selected = false;
        size_t k = 0;
        for (size_t j = 0; j < lmkIdx.size(); ++j)
            if (lmkIdx[j] != maxLmk) { lmkIdx[k] = lmkIdx[j]; lmkVisited[k] = lmkVisited[j]; k++; }
        lmkIdx.resize(k); lmkVisited.resize(k);
    }
    for (size_t i = stIdx; i < edIdx; ++i)
        if (lmkSelectPool[i].selected == true) subVec->push_back(lmkSelectPool[i]);
    return true;
}

// src/Observability.cc:3141-3155
bool Observability::maxVolAutomatic_LazierGreedy(size_t stIdx, size_t edIdx, std::vector<GoodPoint>* subVec,
                                                 size_t mpLimit, double sampleScale, double max_time_for_select) {
    if (mpLimit * 2 > edIdx - stIdx) return maxVolDeletion_LazierGreedy(stIdx, edIdx, subVec, mpLimit, sampleScale, max_time_for_select);
    return maxVolSelection_LazierGreedy(stIdx, edIdx, subVec, mpLimit, sampleScale, max_time_for_select);
}

// src/Observability.cc:3031-3139 (lazy greedy with sum-log-diag upper bound + std::sort)
void Observability::maxVolSelection_BaselineGreedy(std::vector<GoodPoint>* subVec, size_t mpLimit) {
    if (lmkSelectPool.size() == 0 || subVec == nullptr) return;
    subVec->clear();
    if (mpLimit >= lmkSelectPool.size()) {
        for (auto& g : lmkSelectPool) { g.selected = true; subVec->push_back(g); }
        return;
    }
    double curMat[49];
    for (int i = 0; i < 49; i++) curMat[i] = (i % 8 == 0) ? 0.00001 : 0.0;
    std::vector<std::vector<double>> sum_mat(lmkSelectPool.size(), std::vector<double>(49));
    struct Item { GoodPoint g; };
    for (size_t i = 0; i < mpLimit; ++i) {
        int maxLmk = -1;
        double maxDet = -DBL_MAX;
        for (auto& g : lmkSelectPool) {
            if (g.selected == true || g.obs_score < 0) { g.upper_bound = -DBL_MAX; continue; }
            double ub = 0;
            for (int d = 0; d < 7; d++) ub += std::log(curMat[d * 8] + g.obs_block[d * 8]);
            g.upper_bound = ub;
        }
        std::sort(lmkSelectPool.begin(), lmkSelectPool.end(),
                  [](const GoodPoint& a, const GoodPoint& b) { return a.upper_bound > b.upper_bound; });
        for (int j = 0; j < (int)lmkSelectPool.size(); ++j) {
            GoodPoint& g = lmkSelectPool[j];
            if (g.selected == true || g.obs_score < 0) continue;
            if (g.upper_bound < maxDet) break;
            double M[49];
            for (int t = 0; t < 49; t++) M[t] = curMat[t] + g.obs_block[t];
            double curDet = logDet(M, 7);
            if (curDet > maxDet) { maxDet = curDet; maxLmk = j; }
        }
        if (maxLmk == -1) break;
        addBlock(curMat, lmkSelectPool[maxLmk].obs_block, 1.0);
        lmkSelectPool[maxLmk].selected = true;
        subVec->push_back(lmkSelectPool[maxLmk]);
    }
}


// ------------------------------------------------------------------------------------------------ active matching
// include/Util.hpp:47-62: the heap entry of runActiveMapMatching (std::priority_queue<SimplePoint>, operator< on score)
namespace {
struct SimplePoint {
    int idx; double score;
    SimplePoint(int i, double s) : idx(i), score(s) {}
    bool operator<(const SimplePoint& r) const { return score < r.score; }
};
}  // namespace

// src/Observability.cc:1249-1524
int runActiveMapMatching(Frame& F, const std::vector<MapPointView>& mps, const uint8_t* inPool, const double* obsMat,
                         const double* H_meas, double* curMat, float th, ORBmatcher& matcher, int num_to_match,
                         std::vector<int>& left, int (*rand_fn)()) {
    const size_t M = mps.size();
    if (M == 0 || num_to_match <= 0) {                        // :1259-1272
        for (size_t i = 0; i < M; ++i) {
            if (mps[i].isBad) continue;
            if (!mps[i].mbTrackInView) continue;
            left.push_back((int)i);
        }
        return 0;
    }
    auto rnd = rand_fn ? rand_fn : &std::rand;
    int nMatched = 0;
    std::vector<double> lmkIdx, lmkVisited;                  // arma::mat rows of doubles (:1286-1310)
    for (size_t i = 0; i < M; ++i) {
        if (mps[i].isBad) continue;
        if (!mps[i].mbTrackInView) continue;
        if (!inPool[i]) continue;
        lmkIdx.push_back((double)i); lmkVisited.push_back(-1);
    }
    const size_t N = lmkIdx.size();
    const size_t szLazierSubset = static_cast<size_t>(float(N) / float(num_to_match) * 1.0);   // :1320
    for (size_t i = 0; i < (size_t)num_to_match; ++i) {
        std::priority_queue<SimplePoint> heapSubset;
        std::vector<size_t> removeIdx;
        size_t numHit = 0, numRndQue = 0, szActualSubset = szLazierSubset;
        if (lmkIdx.size() < szActualSubset) szActualSubset = lmkIdx.size();
        while (numHit < szActualSubset) {
            size_t j = 0;
            numRndQue = 0;
            while (numRndQue < 2000) {
                j = (rnd() % lmkIdx.size());
                if (lmkVisited[j] < i) { lmkVisited[j] = i; break; }
                ++numRndQue;
            }
            if (numRndQue >= 2000) break;
            const int queIdx = (int)lmkIdx[j];
            ++numHit;
            double Mx[49];
            for (int t = 0; t < 49; t++) Mx[t] = curMat[t] + obsMat[(size_t)49 * queIdx + t];
            heapSubset.push(SimplePoint(queIdx, logDet(Mx, 7)));
            if (numHit >= szActualSubset) {
                const SimplePoint heapTop = heapSubset.top();
                const int bestIdx = matcher.SearchByProjection_OnePoint(F, mps[heapTop.idx], heapTop.idx, th);
                if (bestIdx >= 0) {
                    // reWeightInfoMat (include/Observability.h:517-596) with the matched keypoint's octave, then
                    // curMat += H_rw^T H_rw (:1394-1417)
                    const float Sigma2 = F.mvLevelSigma2[F.mvKeysUn[bestIdx].octave];
                    const double w = std::sqrt((double)Sigma2), wi = 1.0 / w;
                    double Hr[14];
                    for (int t = 0; t < 14; t++) Hr[t] = wi * H_meas[(size_t)14 * heapTop.idx + t];
                    for (int r = 0; r < 7; r++)
                        for (int c = 0; c < 7; c++) curMat[r * 7 + c] = curMat[r * 7 + c] + (Hr[r] * Hr[c] + Hr[7 + r] * Hr[7 + c]);
                    removeIdx.push_back(heapTop.idx);
                    ++nMatched;
                    break;
                } else {
                    removeIdx.push_back(heapTop.idx);
                    heapSubset.pop();
                    --numHit;
                }
            }
        }
        if (numRndQue >= 2000 || heapSubset.size() == 0 || removeIdx.size() == 0) break;     // "early termination"
        if (lmkIdx.size() == removeIdx.size()) break;                                        // "went through all map points"
        std::sort(removeIdx.begin(), removeIdx.end());
        std::vector<double> ni, nv;
        size_t j = 0, k = 0;
        while (j < lmkIdx.size()) {                           // :1459-1480
            if (k >= removeIdx.size()) { ni.push_back(lmkIdx[j]); nv.push_back(lmkVisited[j]); ++j; }
            else if (lmkIdx[j] < removeIdx[k]) { ni.push_back(lmkIdx[j]); nv.push_back(lmkVisited[j]); ++j; }
            else { if (lmkIdx[j] == removeIdx[k]) ++j; ++k; }
        }
        lmkIdx.swap(ni); lmkVisited.swap(nv);
    }
    for (size_t i = 0; i < lmkIdx.size(); ++i) left.push_back((int)lmkIdx[i]);
    return nMatched;
}

}  // namespace oracle
