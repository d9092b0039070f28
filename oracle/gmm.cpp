// CPU ORACLE (test infrastructure only) — Gaussian mixture pose prior.
// Follows external/avatar/GaussianMixture.cpp:12-75 (load) and :94-113 (residual).
#include "oracle.hpp"

#include <fstream>

namespace orc {

int GMM::build(int nc, int nd, const double* weights, const double* means, const double* covs) {
    nComps = nc; nDims = nd;
    const double log_sqrt_2_pi_n = nd * 0.5 * std::log(2 * M_PI);   // :24
    weight.assign(weights, weights + nc);
    consts_log.resize(nc);
    for (int i = 0; i < nc; ++i) consts_log[i] = std::log(weight[i]) - log_sqrt_2_pi_n;   // :31
    mean.assign(means, means + (size_t)nc * nd);
    cov.resize(nc); cov_cho.resize(nc); prec_cho.resize(nc);
    double minDet = std::numeric_limits<double>::max();
    const size_t nn = (size_t)nd * nd;
    for (int i = 0; i < nc; ++i) {
        cov[i].assign(covs + i * nn, covs + (i + 1) * nn);
        cov_cho[i] = cov[i];
        if (!cholesky_lower(cov_cho[i].data(), nd)) return -2;   // reference: throw "Decomposition failed!" (:55)
        // cov^-1 (reference: Eigen .inverse(), :57) — here via the Cholesky factor: inv = L^-T L^-1
        std::vector<double> inv(nn, 0.0), col(nd);
        for (int c = 0; c < nd; ++c) {
            for (int k = 0; k < nd; ++k) col[k] = (k == c) ? 1.0 : 0.0;
            chol_solve(cov_cho[i].data(), nd, col.data());
            for (int k = 0; k < nd; ++k) inv[(size_t)k * nd + c] = col[k];
        }
        // symmetrise (LLT reads the lower triangle only)
        for (int a = 0; a < nd; ++a)
            for (int b = 0; b < a; ++b) {
                const double s = 0.5 * (inv[(size_t)a * nd + b] + inv[(size_t)b * nd + a]);
                inv[(size_t)a * nd + b] = inv[(size_t)b * nd + a] = s;
            }
        prec_cho[i] = inv;
        if (!cholesky_lower(prec_cho[i].data(), nd)) return -2;
        double det = 1.0;                                       // det(cov_cho) = prod diag (:59)
        for (int k = 0; k < nd; ++k) det *= cov_cho[i][(size_t)k * nd + k];
        minDet = std::min(det, minDet);
        consts_log[i] -= std::log(det);                         // :64
    }
    for (int i = 0; i < nc; ++i) consts_log[i] += std::log(minDet);   // :70
    return 0;
}

int GMM::load_text(const char* path) {
    std::ifstream ifs(path);
    if (!ifs) { nComps = -1; return -1; }                       // :15-19
    int nc = 0, nd = 0;
    ifs >> nc >> nd;
    std::vector<double> wts(nc), means((size_t)nc * nd), covs((size_t)nc * nd * nd);
    for (auto& x : wts) ifs >> x;
    for (auto& x : means) ifs >> x;
    for (auto& x : covs) ifs >> x;
    return build(nc, nd, wts.data(), means.data(), covs.data());
}

// residual(): for each comp h = sqrt(.5) L^T (x - mu); p = |h|^2 - consts_log; strict '<' argmin, first wins.
int GMM::residual(const double* x, double* out) const {
    double best = std::numeric_limits<double>::max();
    int bi = 0;
    std::vector<double> d(nDims), h(nDims);
    const double rh = std::sqrt(0.5);
    for (int i = 0; i < nComps; ++i) {
        const double* L = prec_cho[i].data();
        for (int k = 0; k < nDims; ++k) d[k] = x[k] - mean[(size_t)i * nDims + k];
        double sq = 0.0;
        for (int k = 0; k < nDims; ++k) {
            double acc = 0.0;
            for (int m = k; m < nDims; ++m) acc += L[(size_t)m * nDims + k] * d[m];
            h[k] = acc * rh;
            sq += h[k] * h[k];
        }
        const double p = sq - consts_log[i];
        if (p < best) {
            best = p; bi = i;
            for (int k = 0; k < nDims; ++k) out[k] = h[k];
            out[nDims] = std::sqrt(-consts_log[i]);
        }
    }
    return bi;
}

}  // namespace orc
