// GMM pose prior handle: replaces ark::GaussianMixture::load (external/avatar/GaussianMixture.cpp:12-75).
// Derivation runs once on the host in float64 (as the reference does at model load); the fitting kernels read the
// device copies.  b3d_gmm_residual evaluates GaussianMixture::residual (:94-113) on the GPU for a batch of poses.
#include <cmath>
#include <cfloat>
#include <fstream>
#include <limits>

#include "internal.hpp"

namespace b3d {

// in-place lower Cholesky of a dense row-major n x n matrix; false if a pivot is not positive
static bool llt(std::vector<double>& A, int n) {
    for (int c = 0; c < n; ++c) {
        double piv = A[(size_t)c * n + c];
        for (int k = 0; k < c; ++k) piv -= A[(size_t)c * n + k] * A[(size_t)c * n + k];
        if (!(piv > 0.0)) return false;
        piv = std::sqrt(piv);
        A[(size_t)c * n + c] = piv;
        for (int r = c + 1; r < n; ++r) {
            double acc = A[(size_t)r * n + c];
            for (int k = 0; k < c; ++k) acc -= A[(size_t)r * n + k] * A[(size_t)c * n + k];
            A[(size_t)r * n + c] = acc / piv;
        }
        for (int k = c + 1; k < n; ++k) A[(size_t)c * n + k] = 0.0;
    }
    return true;
}

// inverse of an SPD matrix from its Cholesky factor (stands in for Eigen's .inverse(), GaussianMixture.cpp:57)
static void spd_inverse(const std::vector<double>& L, int n, std::vector<double>& inv) {
    inv.assign((size_t)n * n, 0.0);
    std::vector<double> col(n);
    for (int c = 0; c < n; ++c) {
        for (int i = 0; i < n; ++i) {          // L z = e_c
            double acc = (i == c) ? 1.0 : 0.0;
            for (int k = 0; k < i; ++k) acc -= L[(size_t)i * n + k] * col[k];
            col[i] = acc / L[(size_t)i * n + i];
        }
        for (int i = n - 1; i >= 0; --i) {     // L^T x = z
            double acc = col[i];
            for (int k = i + 1; k < n; ++k) acc -= L[(size_t)k * n + i] * col[k];
            col[i] = acc / L[(size_t)i * n + i];
        }
        for (int i = 0; i < n; ++i) inv[(size_t)i * n + c] = col[i];
    }
    for (int a = 0; a < n; ++a)
        for (int b = 0; b < a; ++b) {
            const double s = 0.5 * (inv[(size_t)a * n + b] + inv[(size_t)b * n + a]);
            inv[(size_t)a * n + b] = inv[(size_t)b * n + a] = s;
        }
}

template <int T>
__global__ void __launch_bounds__(T) gmm_residual_kernel(int nc, const double* __restrict__ mean, const double* __restrict__ L,
                                                         const double* __restrict__ consts, int B, const double* __restrict__ x,
                                                         double* __restrict__ out, int32_t* __restrict__ comp) {
    __shared__ double s_x[NPRI];
    __shared__ double s_h[16][NPRI];
    __shared__ double s_p[16];
    const int b = blockIdx.x, tid = threadIdx.x;
    if (b >= B) return;
    for (int k = tid; k < NPRI; k += T) s_x[k] = x[(size_t)b * NPRI + k];
    __syncthreads();
    const double rh = sqrt(0.5);
    for (int t = tid; t < nc * NPRI; t += T) {
        const int i = t / NPRI, k = t % NPRI;
        const double* Li = L + (size_t)i * NPRI * NPRI;
        double acc = 0.0;
        for (int m = k; m < NPRI; ++m) acc += Li[m * NPRI + k] * (s_x[m] - mean[i * NPRI + m]);
        s_h[i][k] = acc * rh;
    }
    __syncthreads();
    if (tid < nc) {
        double sq = 0.0;
        for (int k = 0; k < NPRI; ++k) sq += s_h[tid][k] * s_h[tid][k];
        s_p[tid] = sq - consts[tid];
    }
    __syncthreads();
    int c = 0;
    double best = DBL_MAX;
    for (int i = 0; i < nc; ++i) if (s_p[i] < best) { best = s_p[i]; c = i; }
    for (int k = tid; k < NPRI; k += T) out[(size_t)b * (NPRI + 1) + k] = s_h[c][k];
    if (tid == 0) { out[(size_t)b * (NPRI + 1) + NPRI] = sqrt(-consts[c]); comp[b] = c; }
}

}  // namespace b3d

using namespace b3d;

extern "C" {

int b3d_gmm_create(int32_t nc, int32_t nd, const double* weights, const double* means, const double* covs, b3d_gmm** out) {
    if (!out) return set_error(B3D_EINVAL, "b3d_gmm_create: null out");
    *out = nullptr;
    if (!weights || !means || !covs || nc <= 0) return set_error(B3D_EINVAL, "b3d_gmm_create: null / empty argument");
    if (nd != NPRI) return set_error(B3D_EINVAL, "b3d_gmm_create: nDims must be %d (23 joints x 3), got %d", NPRI, nd);
    if (nc > 16) return set_error(B3D_EINVAL, "b3d_gmm_create: at most 16 components");
    if (int rc = require_device()) return rc;
    b3d_gmm* g = new b3d_gmm();
    cudaGetDevice(&g->device);
    g->nComps = nc; g->nDims = nd;
    const size_t nn = (size_t)nd * nd;
    g->mean.assign(means, means + (size_t)nc * nd);
    g->prec_cho.assign(nc * nn, 0.0);
    g->consts_log.resize(nc);
    std::vector<double> prec(nc * nn), precT(nc * nn);
    const double half_log_2pi_n = nd * 0.5 * std::log(2.0 * M_PI);
    double min_det = std::numeric_limits<double>::max();
    for (int i = 0; i < nc; ++i) {
        std::vector<double> cho(covs + i * nn, covs + (i + 1) * nn), inv;
        if (!llt(cho, nd)) { delete g; return set_error(B3D_ENOTPD, "GMM component %d: covariance is not positive definite (reference throws \"Decomposition failed!\")", i); }
        spd_inverse(cho, nd, inv);
        std::vector<double> pc(inv);
        if (!llt(pc, nd)) { delete g; return set_error(B3D_ENOTPD, "GMM component %d: precision is not positive definite", i); }
        double det = 1.0;
        for (int k = 0; k < nd; ++k) det *= cho[(size_t)k * nd + k];
        if (det < min_det) min_det = det;
        g->consts_log[i] = std::log(weights[i]) - half_log_2pi_n - std::log(det);
        for (size_t e = 0; e < nn; ++e) g->prec_cho[i * nn + e] = pc[e];
        for (int a = 0; a < nd; ++a)
            for (int b = 0; b < nd; ++b) {
                precT[i * nn + (size_t)a * nd + b] = pc[(size_t)b * nd + a];
                double acc = 0.0;                     // P = L L^T
                for (int k = 0; k <= (a < b ? a : b); ++k) acc += pc[(size_t)a * nd + k] * pc[(size_t)b * nd + k];
                prec[i * nn + (size_t)a * nd + b] = acc;
            }
    }
    for (int i = 0; i < nc; ++i) g->consts_log[i] += std::log(min_det);
    // An upper bound of the spectral norm of every L_i (= sqrt(lambda_max(L_i L_i^T))): power iteration on P_i plus a margin, capped
    // by the Frobenius norm (always valid).  The single-frame kernel uses it to bound |L_i^T (x - mu_i)| away from a point where
    // it is known, so that components that cannot win the argmin of GaussianMixture::residual are not evaluated (fit_single.cu).
    std::vector<double> sigma(nc);
    for (int i = 0; i < nc; ++i) {
        const double* P = &prec[i * nn];
        std::vector<double> v(nd, 1.0 / std::sqrt((double)nd)), u(nd);
        double lam = 0.0, fro = 0.0;
        for (size_t e = 0; e < nn; ++e) fro += g->prec_cho[i * nn + e] * g->prec_cho[i * nn + e];
        for (int it = 0; it < 300; ++it) {
            double nrm = 0.0;
            for (int a = 0; a < nd; ++a) { double acc = 0.0; for (int b = 0; b < nd; ++b) acc += P[(size_t)a * nd + b] * v[b]; u[a] = acc; nrm += acc * acc; }
            nrm = std::sqrt(nrm);
            lam = nrm;
            if (!(nrm > 0.0)) break;
            for (int a = 0; a < nd; ++a) v[a] = u[a] / nrm;
        }
        sigma[i] = std::min(std::sqrt(fro), 1.05 * std::sqrt(lam));     // power iteration approaches lambda_max from below: 5 % margin on the norm
    }
    cudaError_t e = cudaMalloc(&g->d_mean, sizeof(double) * nc * nd);
    if (e == cudaSuccess) e = cudaMalloc(&g->d_prec_cho, sizeof(double) * nc * nn);
    if (e == cudaSuccess) e = cudaMalloc(&g->d_prec_cho70, sizeof(double) * nc * nd * (nd + 1));
    if (e == cudaSuccess) e = cudaMalloc(&g->d_prec_choT, sizeof(double) * nc * nn);
    if (e == cudaSuccess) e = cudaMalloc(&g->d_prec, sizeof(double) * nc * nn);
    if (e == cudaSuccess) e = cudaMalloc(&g->d_consts, sizeof(double) * nc);
    if (e == cudaSuccess) e = cudaMalloc(&g->d_sigma, sizeof(double) * nc);
    if (e == cudaSuccess) e = cudaMemcpy(g->d_sigma, sigma.data(), sizeof(double) * nc, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(g->d_mean, g->mean.data(), sizeof(double) * nc * nd, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(g->d_prec_cho, g->prec_cho.data(), sizeof(double) * nc * nn, cudaMemcpyHostToDevice);
    {
        std::vector<double> pad((size_t)nc * nd * (nd + 1), 0.0);
        for (int i = 0; i < nc; ++i)
            for (int a = 0; a < nd; ++a)
                for (int b = 0; b < nd; ++b) pad[((size_t)i * nd + a) * (nd + 1) + b] = g->prec_cho[(size_t)i * nn + (size_t)a * nd + b];
        if (e == cudaSuccess) e = cudaMemcpy(g->d_prec_cho70, pad.data(), sizeof(double) * pad.size(), cudaMemcpyHostToDevice);
    }
    if (e == cudaSuccess) e = cudaMemcpy(g->d_prec_choT, precT.data(), sizeof(double) * nc * nn, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(g->d_prec, prec.data(), sizeof(double) * nc * nn, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(g->d_consts, g->consts_log.data(), sizeof(double) * nc, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { b3d_gmm_destroy(g); return set_error(B3D_ECUDA, "b3d_gmm_create: %s", cudaGetErrorString(e)); }
    *out = g;
    return B3D_OK;
}

int b3d_gmm_load(const char* path, b3d_gmm** out) {
    if (!out) return set_error(B3D_EINVAL, "b3d_gmm_load: null out");
    *out = nullptr;
    if (!path) return set_error(B3D_EINVAL, "b3d_gmm_load: null path");
    std::ifstream ifs(path);
    if (!ifs) return set_error(B3D_ENOFILE, "pose prior file '%s' not found (the reference warns and falls back to the L2 prior)", path);
    int nc = 0, nd = 0;
    ifs >> nc >> nd;
    if (!ifs || nc <= 0 || nd <= 0 || nc > 4096 || nd > 4096) return set_error(B3D_EFORMAT, "pose prior '%s': bad header", path);
    std::vector<double> wts(nc), means((size_t)nc * nd), covs((size_t)nc * nd * nd);
    for (auto& v : wts) ifs >> v;
    for (auto& v : means) ifs >> v;
    for (auto& v : covs) ifs >> v;
    if (!ifs) return set_error(B3D_EFORMAT, "pose prior '%s': truncated", path);
    return b3d_gmm_create(nc, nd, wts.data(), means.data(), covs.data(), out);
}

void b3d_gmm_destroy(b3d_gmm* g) {
    if (!g) return;
    cudaFree(g->d_mean); cudaFree(g->d_prec_cho); cudaFree(g->d_prec_cho70); cudaFree(g->d_prec_choT); cudaFree(g->d_prec); cudaFree(g->d_consts); cudaFree(g->d_sigma);
    delete g;
}

int b3d_gmm_dims(const b3d_gmm* g, int32_t* nc, int32_t* nd) {
    if (!g) return set_error(B3D_EINVAL, "b3d_gmm_dims: null gmm");
    if (nc) *nc = g->nComps;
    if (nd) *nd = g->nDims;
    return B3D_OK;
}

int b3d_gmm_get(const b3d_gmm* g, double* prec_cho, double* consts_log) {
    if (!g) return set_error(B3D_EINVAL, "b3d_gmm_get: null gmm");
    if (prec_cho) std::memcpy(prec_cho, g->prec_cho.data(), sizeof(double) * g->prec_cho.size());
    if (consts_log) std::memcpy(consts_log, g->consts_log.data(), sizeof(double) * g->consts_log.size());
    return B3D_OK;
}

int b3d_gmm_residual(const b3d_gmm* g, int32_t B, const double* x, double* out, int32_t* comp) {
    if (!g || !x || !out || !comp) return set_error(B3D_EINVAL, "b3d_gmm_residual: null argument");
    if (B <= 0) return B < 0 ? set_error(B3D_EINVAL, "b3d_gmm_residual: B < 0") : B3D_OK;
    double *dx = nullptr, *dout = nullptr;
    int32_t* dc = nullptr;
    B3D_CUDA(cudaMalloc(&dx, sizeof(double) * (size_t)B * NPRI));
    cudaError_t e = cudaMalloc(&dout, sizeof(double) * (size_t)B * (NPRI + 1));
    if (e == cudaSuccess) e = cudaMalloc(&dc, sizeof(int32_t) * (size_t)B);
    if (e == cudaSuccess) e = cudaMemcpy(dx, x, sizeof(double) * (size_t)B * NPRI, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        gmm_residual_kernel<128><<<B, 128>>>(g->nComps, g->d_mean, g->d_prec_cho, g->d_consts, B, dx, dout, dc);
        count_launch();
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy(out, dout, sizeof(double) * (size_t)B * (NPRI + 1), cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(comp, dc, sizeof(int32_t) * (size_t)B, cudaMemcpyDeviceToHost);
    cudaFree(dx); cudaFree(dout); cudaFree(dc);
    if (e != cudaSuccess) return set_error(B3D_ECUDA, "b3d_gmm_residual: %s", cudaGetErrorString(e));
    return B3D_OK;
}

}  // extern "C"
