// CPU ORACLE (test infrastructure only) — residual functors of the reference, restated.
//   rotate_point        : ceres::AngleAxisRotatePoint semantics (Ceres 1.14 rotation.h; call sites Sim3BA.h:61,77,177,216)
//   reproj_residual<T>  : ReprojCost::operator() (Sim3BA.h:34-88) and ReprojCostShape::operator() (Sim3BA.h:126-227)
//   reproj_analytic     : same residual with a hand-derived Jacobian (what the CUDA kernels use)
#pragma once
#include "oracle.hpp"

namespace orc {

struct Skeleton {
    int J = NJ, S = NS;
    int parent[NJ];
    double offset[3 * NJ];
    const double* JSR = nullptr;   // (3J x S) row-major or null (functor built without jointShapeReg)
};

// full per-frame parameter layout used by the functors: [s, rootAA(3), rootT(3), jointAA[1..J-1](3 each), w(S)]
constexpr int IDX_S = 0, IDX_RAA = 1, IDX_T = 4, IDX_JAA = 7, IDX_W = 7 + 3 * (NJ - 1);   // 76
constexpr int NFULL = IDX_W + NS;                                                        // 86
inline int idx_jaa(int j) { return IDX_JAA + 3 * (j - 1); }

template <class T>
inline void rotate_point(const T aa[3], const T pt[3], T out[3]) {
    const T theta2 = aa[0] * aa[0] + aa[1] * aa[1] + aa[2] * aa[2];
    if (theta2 > T(std::numeric_limits<double>::epsilon())) {
        const T theta = sqrt(theta2);
        const T costheta = cos(theta);
        const T sintheta = sin(theta);
        const T theta_inverse = T(1.0) / theta;
        const T w[3] = {aa[0] * theta_inverse, aa[1] * theta_inverse, aa[2] * theta_inverse};
        const T wxp[3] = {w[1] * pt[2] - w[2] * pt[1], w[2] * pt[0] - w[0] * pt[2], w[0] * pt[1] - w[1] * pt[0]};
        const T tmp = (w[0] * pt[0] + w[1] * pt[1] + w[2] * pt[2]) * (T(1.0) - costheta);
        const T r0 = pt[0] * costheta + wxp[0] * sintheta + w[0] * tmp;
        const T r1 = pt[1] * costheta + wxp[1] * sintheta + w[1] * tmp;
        const T r2 = pt[2] * costheta + wxp[2] * sintheta + w[2] * tmp;
        out[0] = r0; out[1] = r1; out[2] = r2;
    } else {
        const T wxp[3] = {aa[1] * pt[2] - aa[2] * pt[1], aa[2] * pt[0] - aa[0] * pt[2], aa[0] * pt[1] - aa[1] * pt[0]};
        const T r0 = pt[0] + wxp[0], r1 = pt[1] + wxp[1], r2 = pt[2] + wxp[2];
        out[0] = r0; out[1] = r1; out[2] = r2;
    }
}

// params: pointer to the 86-layout above (w ignored unless sk.JSR).
template <class T>
inline void reproj_residual(const Skeleton& sk, int jid, double u_obs, double v_obs, const double cam[4] /*fx fy cx cy*/,
                            const double R0[9] /*col-major*/, const T* x, T res[2]) {
    const T* w = x + IDX_W;
    auto bone = [&](int j, T b[3]) {   // offset[j] + (delta_j - delta_parent(j)) when shape is on
        for (int c = 0; c < 3; ++c) b[c] = T(sk.offset[3 * j + c]);
        if (sk.JSR) {
            const int pj = sk.parent[j];
            for (int c = 0; c < 3; ++c) {
                T dj = T(0.0), dp = T(0.0);
                for (int s = 0; s < sk.S; ++s) {
                    dj += T(sk.JSR[(size_t)(3 * j + c) * sk.S + s]) * w[s];
                    if (pj >= 0) dp += T(sk.JSR[(size_t)(3 * pj + c) * sk.S + s]) * w[s];
                }
                b[c] += (dj - dp);
            }
        }
    };
    T pos[3];
    bone(jid, pos);
    int cur = jid;
    while (sk.parent[cur] != -1 && sk.parent[cur] != 0) {
        const int p = sk.parent[cur];
        rotate_point(x + idx_jaa(p), pos, pos);
        T b[3];
        bone(p, b);
        for (int c = 0; c < 3; ++c) pos[c] += b[c];
        cur = p;
    }
    T pc[3];
    for (int r = 0; r < 3; ++r) pc[r] = T(R0[r]) * pos[0] + T(R0[3 + r]) * pos[1] + T(R0[6 + r]) * pos[2];
    rotate_point(x + IDX_RAA, pc, pc);
    for (int c = 0; c < 3; ++c) pc[c] = x[IDX_S] * pc[c] + x[IDX_T + c];
    const T u = T(cam[0]) * pc[0] / pc[2] + T(cam[2]);
    const T v = T(cam[1]) * pc[1] / pc[2] + T(cam[3]);
    res[0] = u - T(u_obs);
    res[1] = v - T(v_obs);
}

// rotate with derivatives: dpt, daa are row-major 3x3 (d out_r / d pt_c, d out_r / d aa_c)
inline void rotate_point_jac(const double aa[3], const double p[3], double out[3], double dpt[9], double daa[9]) {
    const double th2 = aa[0] * aa[0] + aa[1] * aa[1] + aa[2] * aa[2];
    if (th2 > std::numeric_limits<double>::epsilon()) {
        const double th = std::sqrt(th2), c = std::cos(th), s = std::sin(th), ti = 1.0 / th;
        const double w[3] = {aa[0] * ti, aa[1] * ti, aa[2] * ti};
        const double wxp[3] = {w[1] * p[2] - w[2] * p[1], w[2] * p[0] - w[0] * p[2], w[0] * p[1] - w[1] * p[0]};
        const double wp = w[0] * p[0] + w[1] * p[1] + w[2] * p[2];
        const double tmp = wp * (1.0 - c);
        for (int r = 0; r < 3; ++r) out[r] = p[r] * c + wxp[r] * s + w[r] * tmp;
        // d/dp = c I + s [w]x + (1-c) w w^T
        const double W[9] = {0, -w[2], w[1], w[2], 0, -w[0], -w[1], w[0], 0};
        for (int r = 0; r < 3; ++r)
            for (int k = 0; k < 3; ++k) dpt[3 * r + k] = (r == k ? c : 0.0) + s * W[3 * r + k] + (1.0 - c) * w[r] * w[k];
        // dw/daa = (I - w w^T)/th
        double dw[9];
        for (int r = 0; r < 3; ++r)
            for (int k = 0; k < 3; ++k) dw[3 * r + k] = ((r == k ? 1.0 : 0.0) - w[r] * w[k]) * ti;
        // -[p]x dw
        const double Pm[9] = {0, p[2], -p[1], -p[2], 0, p[0], p[1], -p[0], 0};   // = -[p]x
        double pdw[3];   // p^T dw (row)
        for (int k = 0; k < 3; ++k) pdw[k] = p[0] * dw[k] + p[1] * dw[3 + k] + p[2] * dw[6 + k];
        for (int r = 0; r < 3; ++r)
            for (int k = 0; k < 3; ++k) {
                double cross = Pm[3 * r] * dw[k] + Pm[3 * r + 1] * dw[3 + k] + Pm[3 * r + 2] * dw[6 + k];
                daa[3 * r + k] = (-s * p[r] + c * wxp[r]) * w[k] + s * cross + tmp * dw[3 * r + k] +
                                 w[r] * ((1.0 - c) * pdw[k] + wp * s * w[k]);
            }
    } else {
        out[0] = p[0] + (aa[1] * p[2] - aa[2] * p[1]);
        out[1] = p[1] + (aa[2] * p[0] - aa[0] * p[2]);
        out[2] = p[2] + (aa[0] * p[1] - aa[1] * p[0]);
        const double A[9] = {1, -aa[2], aa[1], aa[2], 1, -aa[0], -aa[1], aa[0], 1};
        const double Pm[9] = {0, p[2], -p[1], -p[2], 0, p[0], p[1], -p[0], 0};
        for (int k = 0; k < 9; ++k) { dpt[k] = A[k]; daa[k] = Pm[k]; }
    }
}

// residual + dense Jacobian rows Jout[2][NFULL] (zeros outside the active set)
inline void reproj_analytic(const Skeleton& sk, int jid, double u_obs, double v_obs, const double cam[4], const double R0[9],
                            const double* x, double res[2], double* Jout /*2*NFULL*/) {
    double D[3][NFULL];
    for (int r = 0; r < 3; ++r)
        for (int k = 0; k < NFULL; ++k) D[r][k] = 0.0;
    const double* w = x + IDX_W;
    auto bone = [&](int j, double b[3], bool add_deriv) {
        for (int c = 0; c < 3; ++c) b[c] = sk.offset[3 * j + c];
        if (sk.JSR) {
            const int pj = sk.parent[j];
            for (int c = 0; c < 3; ++c) {
                double dj = 0.0, dp = 0.0;
                for (int s = 0; s < sk.S; ++s) {
                    const double a = sk.JSR[(size_t)(3 * j + c) * sk.S + s];
                    const double b2 = pj >= 0 ? sk.JSR[(size_t)(3 * pj + c) * sk.S + s] : 0.0;
                    dj += a * w[s];
                    dp += b2 * w[s];
                    if (add_deriv) D[c][IDX_W + s] += a - b2;
                }
                b[c] += (dj - dp);
            }
        }
    };
    auto left_mul = [&](const double M[9]) {   // D <- M D
        for (int k = 0; k < NFULL; ++k) {
            const double a = D[0][k], b = D[1][k], c = D[2][k];
            if (a == 0.0 && b == 0.0 && c == 0.0) continue;
            D[0][k] = M[0] * a + M[1] * b + M[2] * c;
            D[1][k] = M[3] * a + M[4] * b + M[5] * c;
            D[2][k] = M[6] * a + M[7] * b + M[8] * c;
        }
    };
    double pos[3];
    bone(jid, pos, true);
    int cur = jid;
    while (sk.parent[cur] != -1 && sk.parent[cur] != 0) {
        const int p = sk.parent[cur];
        double dpt[9], daa[9], q[3];
        rotate_point_jac(x + idx_jaa(p), pos, q, dpt, daa);
        left_mul(dpt);
        for (int r = 0; r < 3; ++r)
            for (int k = 0; k < 3; ++k) D[r][idx_jaa(p) + k] += daa[3 * r + k];
        double b[3];
        bone(p, b, true);
        for (int c = 0; c < 3; ++c) pos[c] = q[c] + b[c];
        cur = p;
    }
    const double R0rm[9] = {R0[0], R0[3], R0[6], R0[1], R0[4], R0[7], R0[2], R0[5], R0[8]};
    double pc[3];
    for (int r = 0; r < 3; ++r) pc[r] = R0rm[3 * r] * pos[0] + R0rm[3 * r + 1] * pos[1] + R0rm[3 * r + 2] * pos[2];
    left_mul(R0rm);
    double dpt[9], daa[9], q[3];
    rotate_point_jac(x + IDX_RAA, pc, q, dpt, daa);
    left_mul(dpt);
    for (int r = 0; r < 3; ++r)
        for (int k = 0; k < 3; ++k) D[r][IDX_RAA + k] = daa[3 * r + k];
    const double s = x[IDX_S];
    double X[3];
    for (int c = 0; c < 3; ++c) X[c] = s * q[c] + x[IDX_T + c];
    for (int r = 0; r < 3; ++r) {
        for (int k = 0; k < NFULL; ++k) D[r][k] *= s;
        D[r][IDX_S] = q[r];
        D[r][IDX_T + r] = 1.0;
    }
    const double iz = 1.0 / X[2];
    res[0] = cam[0] * X[0] * iz + cam[2] - u_obs;
    res[1] = cam[1] * X[1] * iz + cam[3] - v_obs;
    for (int k = 0; k < NFULL; ++k) {
        Jout[k] = cam[0] * (D[0][k] - X[0] * iz * D[2][k]) * iz;
        Jout[NFULL + k] = cam[1] * (D[1][k] - X[1] * iz * D[2][k]) * iz;
    }
}

}  // namespace orc
