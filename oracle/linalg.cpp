// CPU ORACLE (test infrastructure only) — small dense linear algebra used by the oracle.
// Stands in for the Eigen calls on the reference path: LLT (GaussianMixture.cpp:45-59),
// HouseholderQR::solve (Ceres 1.14 DenseQRSolver), FullPivLU (Ceres polynomial.cc),
// AngleAxisd::toRotationMatrix (Sim3BA.h:481-507).
#include "oracle.hpp"

namespace orc {

bool cholesky_lower(double* A, int n) {
    for (int j = 0; j < n; ++j) {
        double d = A[(size_t)j * n + j];
        for (int k = 0; k < j; ++k) d -= A[(size_t)j * n + k] * A[(size_t)j * n + k];
        if (!(d > 0.0)) return false;
        d = std::sqrt(d);
        A[(size_t)j * n + j] = d;
        for (int i = j + 1; i < n; ++i) {
            double s = A[(size_t)i * n + j];
            for (int k = 0; k < j; ++k) s -= A[(size_t)i * n + k] * A[(size_t)j * n + k];
            A[(size_t)i * n + j] = s / d;
        }
        for (int k = j + 1; k < n; ++k) A[(size_t)j * n + k] = 0.0;
    }
    return true;
}

void chol_solve(const double* L, int n, double* b) {
    for (int i = 0; i < n; ++i) {
        double s = b[i];
        for (int k = 0; k < i; ++k) s -= L[(size_t)i * n + k] * b[k];
        b[i] = s / L[(size_t)i * n + i];
    }
    for (int i = n - 1; i >= 0; --i) {
        double s = b[i];
        for (int k = i + 1; k < n; ++k) s -= L[(size_t)k * n + i] * b[k];
        b[i] = s / L[(size_t)i * n + i];
    }
}

// Unpivoted Householder QR least squares (what Eigen's householderQr().solve does for full-rank A).
bool householder_qr_solve(double* A, int m, int n, double* b, double* x) {
    if (m < n) return false;
    std::vector<double> vbuf(m);
    for (int k = 0; k < n; ++k) {
        double norm2 = 0.0;
        for (int i = k; i < m; ++i) norm2 += A[(size_t)i * n + k] * A[(size_t)i * n + k];
        const double norm = std::sqrt(norm2);
        if (norm == 0.0) continue;
        const double akk = A[(size_t)k * n + k];
        const double alpha = akk > 0 ? -norm : norm;
        // v = x - alpha e1
        for (int i = k; i < m; ++i) vbuf[i] = A[(size_t)i * n + k];
        vbuf[k] -= alpha;
        double vnorm2 = 0.0;
        for (int i = k; i < m; ++i) vnorm2 += vbuf[i] * vbuf[i];
        if (vnorm2 == 0.0) continue;
        const double beta = 2.0 / vnorm2;
        for (int j = k; j < n; ++j) {
            double dot = 0.0;
            for (int i = k; i < m; ++i) dot += vbuf[i] * A[(size_t)i * n + j];
            dot *= beta;
            for (int i = k; i < m; ++i) A[(size_t)i * n + j] -= dot * vbuf[i];
        }
        double dot = 0.0;
        for (int i = k; i < m; ++i) dot += vbuf[i] * b[i];
        dot *= beta;
        for (int i = k; i < m; ++i) b[i] -= dot * vbuf[i];
    }
    for (int i = n - 1; i >= 0; --i) {
        double s = b[i];
        for (int k = i + 1; k < n; ++k) s -= A[(size_t)i * n + k] * x[k];
        const double d = A[(size_t)i * n + i];
        if (d == 0.0) return false;
        x[i] = s / d;
    }
    return true;
}

bool lu_fullpiv_solve(double* A, int n, double* b, double* x) {
    std::vector<int> cp(n);
    for (int i = 0; i < n; ++i) cp[i] = i;
    for (int k = 0; k < n; ++k) {
        int pr = k, pc = k;
        double best = -1.0;
        for (int i = k; i < n; ++i)
            for (int j = k; j < n; ++j)
                if (std::fabs(A[i * n + j]) > best) { best = std::fabs(A[i * n + j]); pr = i; pc = j; }
        if (best == 0.0) return false;
        if (pr != k) { for (int j = 0; j < n; ++j) std::swap(A[k * n + j], A[pr * n + j]); std::swap(b[k], b[pr]); }
        if (pc != k) { for (int i = 0; i < n; ++i) std::swap(A[i * n + k], A[i * n + pc]); std::swap(cp[k], cp[pc]); }
        for (int i = k + 1; i < n; ++i) {
            const double f = A[i * n + k] / A[k * n + k];
            for (int j = k; j < n; ++j) A[i * n + j] -= f * A[k * n + j];
            b[i] -= f * b[k];
        }
    }
    std::vector<double> y(n);
    for (int i = n - 1; i >= 0; --i) {
        double s = b[i];
        for (int j = i + 1; j < n; ++j) s -= A[i * n + j] * y[j];
        y[i] = s / A[i * n + i];
    }
    for (int i = 0; i < n; ++i) x[cp[i]] = y[i];
    return true;
}

// theta > 1e-12 ? AngleAxisd(theta, aa/theta).toRotationMatrix() : I   (Sim3BA.h:483-487 etc.)
// Eigen's AngleAxis::toRotationMatrix: R = c I + (1-c) a a^T + s [a]x
void aa_to_matrix(const double aa[3], double R[9]) {
    const double th = std::sqrt(aa[0] * aa[0] + aa[1] * aa[1] + aa[2] * aa[2]);
    for (int k = 0; k < 9; ++k) R[k] = 0.0;
    R[0] = R[4] = R[8] = 1.0;
    if (!(th > 1e-12)) return;
    const double ax = aa[0] / th, ay = aa[1] / th, az = aa[2] / th;
    const double s = std::sin(th), c = std::cos(th);
    const double cx = (1.0 - c) * ax, cy = (1.0 - c) * ay, cz = (1.0 - c) * az;
    double t;
    // col-major: R[col*3+row]
    t = cx * ay; R[3 * 0 + 1] = t + s * az; R[3 * 1 + 0] = t - s * az;   // (1,0), (0,1)
    t = cx * az; R[3 * 0 + 2] = t - s * ay; R[3 * 2 + 0] = t + s * ay;   // (2,0), (0,2)
    t = cy * az; R[3 * 1 + 2] = t + s * ax; R[3 * 2 + 1] = t - s * ax;   // (2,1), (1,2)
    R[0] = cx * ax + c; R[4] = cy * ay + c; R[8] = cz * az + c;
}

void matmul3_cm(const double A[9], const double B[9], double C[9]) {
    double T[9];
    for (int col = 0; col < 3; ++col)
        for (int row = 0; row < 3; ++row)
            T[3 * col + row] = A[row] * B[3 * col] + A[3 + row] * B[3 * col + 1] + A[6 + row] * B[3 * col + 2];
    for (int k = 0; k < 9; ++k) C[k] = T[k];
}

}  // namespace orc
