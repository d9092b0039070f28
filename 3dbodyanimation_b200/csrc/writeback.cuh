// The AA -> rotation write-back after a solve (include/Sim3BA.h:481-507, 649-679; include/MultiFrameBA.h:154-172) as device code, shared by
// writeback_kernel (util.cu) and the window kernel's chain mode (fit_window.cu), so that both produce the same bits.
#pragma once

namespace b3d {

// Eigen::AngleAxisd(theta, aa/theta).toRotationMatrix() for theta > 1e-12, identity otherwise (Sim3BA.h:483-489). Col-major out.
__device__ __forceinline__ void aa_to_matrix_cm(const double* aa, double* R) {
    const double th = sqrt(aa[0] * aa[0] + aa[1] * aa[1] + aa[2] * aa[2]);
    if (th > 1e-12) {
        const double x = aa[0] / th, y = aa[1] / th, z = aa[2] / th;
        double s, c;
        sincos(th, &s, &c);
        const double t = 1.0 - c;
        R[0] = t * x * x + c;     R[3] = t * x * y - s * z; R[6] = t * x * z + s * y;
        R[1] = t * x * y + s * z; R[4] = t * y * y + c;     R[7] = t * y * z - s * x;
        R[2] = t * x * z - s * y; R[5] = t * y * z + s * x; R[8] = t * z * z + c;
    } else {
        R[0] = 1; R[1] = 0; R[2] = 0; R[3] = 0; R[4] = 1; R[5] = 0; R[6] = 0; R[7] = 0; R[8] = 1;
    }
}

// joint j of frame f: R0 <- R(rootAA) R0, r[0] = R0, r[j] = R(jointAA_j), p = rootT
__device__ __forceinline__ void writeback_entry(int f, int j, const double* poses, double* R0, double* r, double* p) {
    const double* pose = poses + (size_t)f * NPF;
    double R[9];
    if (j == 0) {
        double Rr[9], A[9];
        aa_to_matrix_cm(pose + IDX_RAA, Rr);
#pragma unroll
        for (int k = 0; k < 9; ++k) A[k] = R0[(size_t)f * 9 + k];
        // r[0] <- R(rootAA) * r[0]  (col-major product)
#pragma unroll
        for (int col = 0; col < 3; ++col)
#pragma unroll
            for (int row = 0; row < 3; ++row)
                R[3 * col + row] = Rr[row] * A[3 * col] + Rr[3 + row] * A[3 * col + 1] + Rr[6 + row] * A[3 * col + 2];
#pragma unroll
        for (int k = 0; k < 9; ++k) R0[(size_t)f * 9 + k] = R[k];
        if (p) { p[(size_t)f * 3] = pose[IDX_T]; p[(size_t)f * 3 + 1] = pose[IDX_T + 1]; p[(size_t)f * 3 + 2] = pose[IDX_T + 2]; }
    } else {
        aa_to_matrix_cm(pose + IDX_JAA + 3 * (j - 1), R);
    }
    if (r) {
#pragma unroll
        for (int k = 0; k < 9; ++k) r[((size_t)f * NJ + j) * 9 + k] = R[k];
    }
}

}  // namespace b3d
