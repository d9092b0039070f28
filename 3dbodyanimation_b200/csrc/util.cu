// Small per-frame kernels around the solves: the AA -> rotation write-back (include/Sim3BA.h:481-507, 649-679;
// include/MultiFrameBA.h:154-172) and mean_pixel_error (include/Utils.h:102-115).
#include <algorithm>

#include "internal.hpp"
#include "writeback.cuh"

namespace b3d {

// one thread per (frame, joint)
__global__ void writeback_kernel(int F, const double* __restrict__ poses, double* __restrict__ R0, double* __restrict__ r,
                                 double* __restrict__ p) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= F * NJ) return;
    writeback_entry(t / NJ, t % NJ, poses, R0, r, p);
}

__global__ void pixel_error_kernel(int F, const b3d_kp* __restrict__ kps, const int32_t* __restrict__ nkp,
                                   const double* __restrict__ jointPos, double fx, double fy, double cx, double cy,
                                   double* __restrict__ err) {
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= F) return;
    const int n = min(max(nkp[f], 0), MAXKP);
    double sum = 0.0;
    for (int k = 0; k < n; ++k) {
        const b3d_kp kp = kps[(size_t)f * MAXKP + k];
        const double* J = jointPos + ((size_t)f * NJ + clamp_jid(kp.jid)) * 3;
        const double u = fx * J[0] / J[2] + cx, v = fy * J[1] / J[2] + cy;
        sum += hypot(u - kp.u, v - kp.v);
    }
    err[f] = n > 0 ? sum / n : 0.0;
}

__global__ void convert_kernel(const double* __restrict__ src, float* __restrict__ dst, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = (float)src[i];
}

int validate_keypoints(const b3d_kp* kps, const int32_t* nkp, size_t frames, const char* who) {
    for (size_t f = 0; f < frames; ++f) {
        const int n = std::min(std::max(nkp[f], 0), MAXKP);
        for (int k = 0; k < n; ++k) {
            const int jid = kps[f * MAXKP + k].jid;
            if (jid < 0 || jid >= NJ) return set_error(B3D_EINVAL, "%s: frame %zu keypoint %d has jid %d (must be in [0, %d))", who, f, k, jid, NJ);
        }
    }
    return B3D_OK;
}

int b3d_convert_f64_f32_dev(const double* src, float* dst, size_t n, cudaStream_t st) {
    if (n == 0) return B3D_OK;
    convert_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(src, dst, n);
    B3D_LAUNCH_CHECK();
    return B3D_OK;
}

}  // namespace b3d

using namespace b3d;

extern "C" {

int b3d_writeback_dev(int32_t F, const double* poses, double* R0, double* r, double* p, void* stream) {
    if (!poses || !R0) return set_error(B3D_EINVAL, "b3d_writeback: null argument");
    if (F <= 0) return F < 0 ? set_error(B3D_EINVAL, "b3d_writeback: F < 0") : B3D_OK;
    const int n = F * NJ;
    writeback_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(F, poses, R0, r, p);
    B3D_LAUNCH_CHECK();
    return B3D_OK;
}

int b3d_writeback(int32_t F, const double* poses, double* R0, double* r, double* p) {
    if (!poses || !R0) return set_error(B3D_EINVAL, "b3d_writeback: null argument");
    if (F <= 0) return F < 0 ? set_error(B3D_EINVAL, "b3d_writeback: F < 0") : B3D_OK;
    const size_t nf = (size_t)F;
    double *dp = nullptr;
    B3D_CUDA(cudaMalloc(&dp, 8 * nf * (NPF + 9 + NJ * 9 + 3)));
    double *dR0 = dp + nf * NPF, *dr = dR0 + nf * 9, *dpp = dr + nf * NJ * 9;
    cudaError_t e = cudaMemcpy(dp, poses, 8 * nf * NPF, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(dR0, R0, 8 * nf * 9, cudaMemcpyHostToDevice);
    int rc = B3D_OK;
    if (e == cudaSuccess) rc = b3d_writeback_dev(F, dp, dR0, dr, dpp, nullptr);
    if (e == cudaSuccess && !rc) {
        e = cudaStreamSynchronize(nullptr);
        if (e == cudaSuccess) e = cudaMemcpy(R0, dR0, 8 * nf * 9, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess && r) e = cudaMemcpy(r, dr, 8 * nf * NJ * 9, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess && p) e = cudaMemcpy(p, dpp, 8 * nf * 3, cudaMemcpyDeviceToHost);
    }
    cudaFree(dp);
    if (e != cudaSuccess) return set_error(B3D_ECUDA, "b3d_writeback: %s", cudaGetErrorString(e));
    return rc;
}

int b3d_mean_pixel_error_dev(int32_t F, const b3d_kp* kps, const int32_t* nkp, const double* jointPos, double fx, double fy, double cx,
                             double cy, double* err, void* stream) {
    if (!kps || !nkp || !jointPos || !err) return set_error(B3D_EINVAL, "b3d_mean_pixel_error: null argument");
    if (F <= 0) return F < 0 ? set_error(B3D_EINVAL, "b3d_mean_pixel_error: F < 0") : B3D_OK;
    pixel_error_kernel<<<(F + 127) / 128, 128, 0, (cudaStream_t)stream>>>(F, kps, nkp, jointPos, fx, fy, cx, cy, err);
    B3D_LAUNCH_CHECK();
    return B3D_OK;
}

int b3d_mean_pixel_error(int32_t F, const b3d_kp* kps, const int32_t* nkp, const double* jointPos, double fx, double fy, double cx,
                         double cy, double* err) {
    if (!kps || !nkp || !jointPos || !err) return set_error(B3D_EINVAL, "b3d_mean_pixel_error: null argument");
    if (F <= 0) return F < 0 ? set_error(B3D_EINVAL, "b3d_mean_pixel_error: F < 0") : B3D_OK;
    const size_t nf = (size_t)F;
    if (int rc = validate_keypoints(kps, nkp, nf, "b3d_mean_pixel_error")) return rc;
    char* base = nullptr;
    const size_t s_kp = sizeof(b3d_kp) * nf * MAXKP, s_jp = 8 * nf * NJ * 3, s_nk = (4 * nf + 7) / 8 * 8;
    B3D_CUDA(cudaMalloc(&base, s_kp + s_jp + 8 * nf + s_nk));
    b3d_kp* dk = (b3d_kp*)base; double* dj = (double*)(base + s_kp); double* de = dj + nf * NJ * 3; int32_t* dn = (int32_t*)(de + nf);
    cudaError_t e = cudaMemcpy(dk, kps, s_kp, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(dj, jointPos, s_jp, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(dn, nkp, 4 * nf, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        pixel_error_kernel<<<(F + 127) / 128, 128>>>(F, dk, dn, dj, fx, fy, cx, cy, de);
        count_launch();
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy(err, de, 8 * nf, cudaMemcpyDeviceToHost);
    cudaFree(base);
    if (e != cudaSuccess) return set_error(B3D_ECUDA, "b3d_mean_pixel_error: %s", cudaGetErrorString(e));
    return B3D_OK;
}

}  // extern "C"
