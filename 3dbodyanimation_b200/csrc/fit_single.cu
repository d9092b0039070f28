// Entry (2), single-frame: one CTA runs the whole Levenberg-Marquardt solve of one frame.
//
// Replaces the ceres::Problem built and solved by OptimizePoseReprojection (include/Sim3BA.h:515-683) and
// OptimizePoseShapeReprojection (include/Sim3BA.h:348-511).  The loop below restates Ceres Solver 1.14's
// TrustRegionMinimizer + LevenbergMarquardtStrategy (defaults, as the reference sets only DENSE_QR / max iterations /
// threads, Sim3BA.h:473-477): Jacobi scaling fixed at iteration 0, D^2 = clamp(diag(J^T J), 1e-6, 1e32) refreshed after
// successful steps only, step from (J^T J + D^2/radius) y = J^T r, rho-based acceptance and radius update, projected
// Armijo line search because the scale is bounded (Sim3BA.h:450-451), tolerance checks in Ceres' order.
// The linear solve is a per-frame Cholesky of the (<= 86 x 86) normal equations held in shared memory; DENSE_QR of the
// stacked Jacobian is the same step in exact arithmetic (DESIGN.md "QR vs Cholesky").
//
// Residual blocks: <= 17 reprojection blocks + Huber (lm_device.cuh), pose prior PosePriorAAAnalytic
// (Sim3BA.h:245-320: GMM min-component with the reference's Jacobian that omits sqrt(1/2), or L2), shape prior
// ShapePriorL2Analytic (Sim3BA.h:323-346).
#include "lm_device.cuh"

namespace b3d {

constexpr int FS_THREADS = 128;
constexpr int NAUG = NFULL + 1;                       // augmented with the right-hand side row
constexpr int LTRI = NAUG * (NAUG + 1) / 2;           // 3828 packed doubles
constexpr int HTRI = NFULL * (NFULL + 1) / 2;         // 3741
constexpr int GMM_MAXC = 16;

constexpr double kInitialRadius = 1e4, kMaxRadius = 1e16, kMinRadius = 1e-32;
constexpr double kMinDiagonal = 1e-6, kMaxDiagonal = 1e32;
constexpr double kMinRelativeDecrease = 1e-3;
constexpr double kFunctionTolerance = 1e-6, kGradientTolerance = 1e-10, kParameterTolerance = 1e-8;
constexpr int kMaxConsecutiveInvalidSteps = 5;
constexpr double kSufficientDecrease = 1e-4, kMinLineSearchStepSize = 1e-9;
constexpr int kMaxLineSearchIters = 20;
constexpr double kLsContractLo = 1e-3, kLsContractHi = 0.6;   // max_/min_line_search_step_contraction

struct GmmTmp {
    double h[GMM_MAXC][NPRI];     // L_i^T (x - mu_i)
    double d[GMM_MAXC][NPRI];     // x - mu_i
    double p[GMM_MAXC];           // |h_i|^2 / 2 - consts_log_i
    double sq[GMM_MAXC];          // |sqrt(1/2) h_i|^2
};

union SingleUnion {
    EvalTmp ev;
    GmmTmp gm;
    double L[LTRI];
};

struct SingleSmem {
    EvalOut out;
    SingleUnion u;
    double xfull[NFULL], cfull[NFULL];     // current / candidate parameters, 86-layout
    double g[NFULL], gs[NFULL], gc[NFULL]; // gradient (free layout): unscaled, Jacobi-scaled, at a line-search point
    double scale[NFULL], diag[NFULL], hdiag[NFULL], y[NFULL], delta[NFULL];
    double gpri[NPRI];                     // pose-prior gradient, pose-dim layout
    double R0[9];
    double red[40];
    double ls_step;
    int32_t comp, flag;
    int16_t fcol[MAXKP][NNZPAD];          // free-layout column of each Jacobian slot (-1 = held constant)
};

// ---- polynomial helpers for the (rare) Armijo interpolation: thread 0 only --------------------------------------
struct LsSample { double x, value, gradient; bool valid; };

__device__ double poly_eval(const double* p, int nc, double x) { double v = 0; for (int i = 0; i < nc; ++i) v = v * x + p[i]; return v; }

// Gaussian elimination with full pivoting (Eigen FullPivLU as used by ceres polynomial.cc); n <= 6
__device__ bool lu_solve(double* A, int n, double* b, double* x) {
    int perm[6];
    for (int i = 0; i < n; ++i) perm[i] = i;
    for (int k = 0; k < n; ++k) {
        int pr = k, pc = k; double best = 0.0;
        for (int i = k; i < n; ++i) for (int j = k; j < n; ++j) if (fabs(A[i * n + j]) > best) { best = fabs(A[i * n + j]); pr = i; pc = j; }
        if (best == 0.0) return false;
        if (pr != k) { for (int j = 0; j < n; ++j) { double t = A[k * n + j]; A[k * n + j] = A[pr * n + j]; A[pr * n + j] = t; } double t = b[k]; b[k] = b[pr]; b[pr] = t; }
        if (pc != k) { for (int i = 0; i < n; ++i) { double t = A[i * n + k]; A[i * n + k] = A[i * n + pc]; A[i * n + pc] = t; } int t = perm[k]; perm[k] = perm[pc]; perm[pc] = t; }
        for (int i = k + 1; i < n; ++i) {
            const double f = A[i * n + k] / A[k * n + k];
            for (int j = k; j < n; ++j) A[i * n + j] -= f * A[k * n + j];
            b[i] -= f * b[k];
        }
    }
    double z[6];
    for (int i = n - 1; i >= 0; --i) { double s = b[i]; for (int j = i + 1; j < n; ++j) s -= A[i * n + j] * z[j]; z[i] = s / A[i * n + i]; }
    for (int i = 0; i < n; ++i) x[perm[i]] = z[i];
    return true;
}

// real parts of the roots of a polynomial of degree <= 4 (coefficients highest first): Aberth-Ehrlich iteration
__device__ int poly_root_real_parts(const double* pin, int nc, double* out) {
    int lead = 0;
    while (lead + 1 < nc && pin[lead] == 0.0) ++lead;
    const double* p = pin + lead;
    const int deg = nc - lead - 1;
    if (deg <= 0) return 0;
    if (deg == 1) { out[0] = -p[1] / p[0]; return 1; }
    if (deg == 2) {
        const double a = p[0], b = p[1], c = p[2];
        const double D = b * b - 4 * a * c, sD = sqrt(fabs(D));
        if (D >= 0) {
            if (b >= 0) { out[0] = (-b - sD) / (2.0 * a); out[1] = (2.0 * c) / (-b - sD); }
            else { out[0] = (2.0 * c) / (-b + sD); out[1] = (-b + sD) / (2.0 * a); }
        } else { out[0] = out[1] = -b / (2.0 * a); }
        return 2;
    }
    double ar[6], zr[5], zi[5];
    for (int i = 0; i <= deg; ++i) ar[i] = p[i] / p[0];
    double rad = 0;
    for (int i = 1; i <= deg; ++i) rad = fmax(rad, fabs(ar[i]));
    rad = 1.0 + rad;
    for (int i = 0; i < deg; ++i) { double s, c; sincos(2 * 3.14159265358979323846 * i / deg + 0.4, &s, &c); zr[i] = rad * 0.5 * c; zi[i] = rad * 0.5 * s; }
    for (int it = 0; it < 500; ++it) {
        double change = 0;
        for (int i = 0; i < deg; ++i) {
            double fr = ar[0], fi = 0, dr = 0, di = 0;
            for (int k = 1; k <= deg; ++k) {
                double t = dr * zr[i] - di * zi[i] + fr; di = dr * zi[i] + di * zr[i] + fi; dr = t;
                t = fr * zr[i] - fi * zi[i] + ar[k]; fi = fr * zi[i] + fi * zr[i]; fr = t;
            }
            if (fr == 0.0 && fi == 0.0) continue;
            const double dd = dr * dr + di * di;
            const double rr = (fr * dr + fi * di) / dd, ri = (fi * dr - fr * di) / dd;   // ratio = f / df
            double sr = 0, si = 0;
            for (int j = 0; j < deg; ++j) if (j != i) { const double ur = zr[i] - zr[j], ui = zi[i] - zi[j], uu = ur * ur + ui * ui; sr += ur / uu; si -= ui / uu; }
            const double er = 1.0 - (rr * sr - ri * si), ei = -(rr * si + ri * sr);
            const double ee = er * er + ei * ei;
            const double zr_d = (rr * er + ri * ei) / ee, zi_d = (ri * er - rr * ei) / ee;
            zr[i] -= zr_d; zi[i] -= zi_d;
            change = fmax(change, sqrt(zr_d * zr_d + zi_d * zi_d));
        }
        if (change < 1e-15 * rad) break;
    }
    for (int i = 0; i < deg; ++i) out[i] = zr[i];
    return deg;
}

// ceres MinimizeInterpolatingPolynomial: every sample carries value and gradient
__device__ double minimize_interpolating_polynomial(const LsSample* s, int ns, double x_min, double x_max) {
    const int nc = 2 * ns, degree = nc - 1;
    double lhs[36], rhs[6], coef[6];
    for (int i = 0; i < nc * nc; ++i) lhs[i] = 0.0;
    int row = 0;
    for (int q = 0; q < ns; ++q) {
        for (int j = 0; j <= degree; ++j) lhs[row * nc + j] = pow(s[q].x, (double)(degree - j));
        rhs[row++] = s[q].value;
        for (int j = 0; j < degree; ++j) lhs[row * nc + j] = (degree - j) * pow(s[q].x, (double)(degree - j - 1));
        rhs[row++] = s[q].gradient;
    }
    double best_x = (x_min + x_max) / 2.0;
    if (!lu_solve(lhs, nc, rhs, coef)) return best_x;
    double best_v = poly_eval(coef, nc, best_x);
    const double vmin = poly_eval(coef, nc, x_min);
    if (vmin < best_v) { best_v = vmin; best_x = x_min; }
    const double vmax = poly_eval(coef, nc, x_max);
    if (vmax < best_v) { best_v = vmax; best_x = x_max; }
    double dcoef[5], roots[5];
    for (int i = 0; i < degree; ++i) dcoef[i] = (degree - i) * coef[i];
    const int nr = poly_root_real_parts(dcoef, degree, roots);
    for (int i = 0; i < nr; ++i) {
        if (roots[i] < x_min || roots[i] > x_max) continue;
        const double v = poly_eval(coef, nc, roots[i]);
        if (v < best_v) { best_v = v; best_x = roots[i]; }
    }
    return best_x;
}

// ---- GMM pose prior (GaussianMixture::residual, external/avatar/GaussianMixture.cpp:94-113) --------------------
// Returns 0.5 |beta r|^2 of the prior block and sets S.comp; with JAC also S.gpri = J^T r (pose-dim layout), where
// J = beta L_c^T WITHOUT the sqrt(1/2) the residual carries (reference quirk, Sim3BA.h:298-299).
template <int T, bool JAC>
__device__ double gmm_eval(const FitConst* __restrict__ C, const double* xj /*69*/, SingleSmem& S) {
    const int tid = threadIdx.x, nc = C->gmm_ncomp;
    GmmTmp& G = S.u.gm;
    for (int t = tid; t < nc * NPRI; t += T) G.d[t / NPRI][t % NPRI] = xj[t % NPRI] - C->gmm_mean[t];
    __syncthreads();
    for (int t = tid; t < nc * NPRI; t += T) {
        const int i = t / NPRI, k = t % NPRI;
        const double* __restrict__ L = C->gmm_L + (size_t)i * NPRI * NPRI + k;    // column k: consecutive threads, consecutive addresses
        const double* dd = G.d[i];
        double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
        int m = k;
        for (; m + 4 <= NPRI; m += 4) {
            a0 += __ldg(L + m * NPRI) * dd[m];           a1 += __ldg(L + (m + 1) * NPRI) * dd[m + 1];
            a2 += __ldg(L + (m + 2) * NPRI) * dd[m + 2]; a3 += __ldg(L + (m + 3) * NPRI) * dd[m + 3];
        }
        for (; m < NPRI; ++m) a0 += __ldg(L + m * NPRI) * dd[m];
        G.h[i][k] = (a0 + a1) + (a2 + a3);
    }
    __syncthreads();
    if (tid < nc) {
        const double rh = sqrt(0.5);
        double sq = 0.0;
        for (int k = 0; k < NPRI; ++k) { const double h = G.h[tid][k] * rh; sq += h * h; }
        G.sq[tid] = sq;
        G.p[tid] = sq - C->gmm_consts[tid];
    }
    __syncthreads();
    int c = 0;
    double best = DBL_MAX;
    for (int i = 0; i < nc; ++i) { const double p = G.p[i]; if (p < best) { best = p; c = i; } }
    if (tid == 0) S.comp = c;
    // residual = beta [h_c ; sqrt(-consts_log_c)]  =>  cost = 0.5 beta^2 (|h_c|^2 - consts_log_c)
    const double cost = 0.5 * C->beta_pose * C->beta_pose * (G.sq[c] + (-C->gmm_consts[c]));
    if (JAC) {
        const double f = C->beta_pose * C->beta_pose * sqrt(0.5);
        const double* LT = C->gmm_LT + (size_t)c * NPRI * NPRI;
        for (int m = tid; m < NPRI; m += T) {
            double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
            int k = 0;
            for (; k + 4 <= m + 1; k += 4) {
                a0 += __ldg(LT + k * NPRI + m) * G.h[c][k];           a1 += __ldg(LT + (k + 1) * NPRI + m) * G.h[c][k + 1];
                a2 += __ldg(LT + (k + 2) * NPRI + m) * G.h[c][k + 2]; a3 += __ldg(LT + (k + 3) * NPRI + m) * G.h[c][k + 3];
            }
            for (; k <= m; ++k) a0 += __ldg(LT + k * NPRI + m) * G.h[c][k];
            S.gpri[m] = f * ((a0 + a1) + (a2 + a3));
        }
    }
    __syncthreads();
    return cost;
}

// ---- cost (and with JAC the normal equations) at parameters xf --------------------------------------------------
// JAC: S.u.L holds the packed UNSCALED J^T J (free layout) on return and S.g the unscaled gradient.
template <int T, bool JAC>
__device__ double evaluate(const FitConst* __restrict__ C, const double* xf, const b3d_kp* __restrict__ kps, int nkp, SingleSmem& S) {
    const int tid = threadIdx.x, n = C->n;
    eval_reproj<T, JAC>(C, xf, kps, nkp, S.R0, S.u.ev, S.out);
    double cost = 0.0;
    for (int k = 0; k < nkp; ++k) cost += S.out.kcost[k];
    double pcost = 0.0;
    const bool pose_prior = C->beta_pose > 0.0;
    if (pose_prior && C->use_gmm) pcost = gmm_eval<T, JAC>(C, xf + IDX_JAA, S);   // aliases S.u.ev: eval temporaries are dead
    // the remaining prior costs are sums of squares over <= 79 terms: fixed-order block sum
    double v = 0.0;
    if (pose_prior && !C->use_gmm)
        for (int k = tid; k < NPRI; k += T) { const double r = xf[IDX_JAA + k] * C->beta_pose; v += 0.5 * r * r; }
    if (C->has_w)
        for (int k = tid; k < NS; k += T) { const double r = xf[IDX_W + k] * C->beta_shape; v += 0.5 * r * r; }
    cost += pcost + block_sum<T>(v, S.red);
    if (!JAC) return cost;

    // ---- J^T J and J^T r ----
    __syncthreads();
    double* H = S.u.L;
    for (int e = tid; e < n * (n + 1) / 2; e += T) H[e] = 0.0;
    for (int k = tid; k < n; k += T) S.g[k] = 0.0;
    for (int t = tid; t < nkp * NNZPAD; t += T) {
        const int k = t / NNZPAD, sl = t % NNZPAD;
        S.fcol[k][sl] = sl < S.out.jnnz[k] ? (int16_t)C->col_of[S.out.jcol[k][sl]] : (int16_t)-1;
    }
    __syncthreads();
    for (int k = 0; k < nkp; ++k) {
        const int nnz = S.out.jnnz[k];
        const double r0 = S.out.res[k][0], r1 = S.out.res[k][1];
        const double* J0 = S.out.J[k][0];
        const double* J1 = S.out.J[k][1];
        for (int i = tid >> 5; i < nnz; i += T / 32) {          // row of the outer product: one warp; columns: lanes
            const int ci = S.fcol[k][i];
            if (ci < 0) continue;
            const double a0 = J0[i], a1 = J1[i];
            for (int j = tid & 31; j <= i; j += 32) {
                const int cj = S.fcol[k][j];
                if (cj < 0) continue;
                H[ci >= cj ? tri(ci, cj) : tri(cj, ci)] += a0 * J0[j] + a1 * J1[j];
                if (i == j) S.g[ci] += a0 * r0 + a1 * r1;
            }
        }
        __syncthreads();
    }
    // priors
    if (pose_prior) {
        const double b2 = C->beta_pose * C->beta_pose;
        if (C->use_gmm) {
            const double* P = C->gmm_P + (size_t)S.comp * NPRI * NPRI;
            for (int i = tid >> 5; i < NPRI; i += T / 32) {
                const int ci = C->col_of[IDX_JAA + i];
                if (ci < 0) continue;
                for (int j = tid & 31; j <= i; j += 32) {
                    const int cj = C->col_of[IDX_JAA + j];
                    if (cj >= 0) H[tri(ci, cj)] += b2 * P[i * NPRI + j];     // free order preserves pose-dim order, so ci >= cj
                }
            }
            for (int k = tid; k < NPRI; k += T) { const int ck = C->col_of[IDX_JAA + k]; if (ck >= 0) S.g[ck] += S.gpri[k]; }
        } else {
            for (int k = tid; k < NPRI; k += T) {
                const int ck = C->col_of[IDX_JAA + k];
                if (ck >= 0) { H[tri(ck, ck)] += b2; S.g[ck] += b2 * xf[IDX_JAA + k]; }
            }
        }
    }
    if (C->has_w) {
        const double b2 = C->beta_shape * C->beta_shape;
        for (int k = tid; k < NS; k += T) { const int ck = C->col_of[IDX_W + k]; H[tri(ck, ck)] += b2; S.g[ck] += b2 * xf[IDX_W + k]; }
    }
    __syncthreads();
    return cost;
}

// Cholesky (Crout / left-looking) of the packed lower matrix A (np x np; the last row is the right-hand side): on return rows
// 0..np-2 hold L and row np-1 holds z = L^-1 rhs.  Every thread recomputes the pivot of column j (a length-j dot product over
// row j, broadcast reads), so one barrier per column suffices; thread i owns row i.  Returns false when a pivot is not
// positive (Eigen LLT / oracle criterion).
template <int T>
__device__ bool chol_factor_aug(double* A, int np) {
    const int tid = threadIdx.x, n = np - 1;
    bool ok = true;
    for (int j = 0; j < n; ++j) {
        const double* rj = A + tri(j, 0);
        double d0 = rj[j], d1 = 0.0, d2 = 0.0, d3 = 0.0;
        int k = 0;
        for (; k + 4 <= j; k += 4) { d0 -= rj[k] * rj[k]; d1 -= rj[k + 1] * rj[k + 1]; d2 -= rj[k + 2] * rj[k + 2]; d3 -= rj[k + 3] * rj[k + 3]; }
        for (; k < j; ++k) d0 -= rj[k] * rj[k];
        const double d = (d0 + d1) + (d2 + d3);
        if (!(d > 0.0)) { ok = false; break; }              // uniform: every thread computed the same d from the same data
        const double ljj = sqrt(d), inv = 1.0 / ljj;
        for (int i = j + 1 + tid; i < np; i += T) {
            double* ri = A + tri(i, 0);
            double s0 = ri[j], s1 = 0.0, s2 = 0.0, s3 = 0.0;
            int q = 0;
            for (; q + 4 <= j; q += 4) { s0 -= ri[q] * rj[q]; s1 -= ri[q + 1] * rj[q + 1]; s2 -= ri[q + 2] * rj[q + 2]; s3 -= ri[q + 3] * rj[q + 3]; }
            for (; q < j; ++q) s0 -= ri[q] * rj[q];
            ri[j] = ((s0 + s1) + (s2 + s3)) * inv;
        }
        __syncthreads();
        if (tid == 0) A[tri(j, j)] = ljj;                   // nobody reads the raw diagonal of column j any more
    }
    __syncthreads();
    return ok;
}

// back substitution L^T y = z with one warp (z = row n of the augmented factor)
__device__ void chol_backsolve_warp(const double* A, int n, double* y) {
    const int lane = threadIdx.x & 31;
    double z[3];
#pragma unroll
    for (int q = 0; q < 3; ++q) { const int i = lane + 32 * q; z[q] = i < n ? A[tri(n, i)] : 0.0; }
    for (int j = n - 1; j >= 0; --j) {
        const int owner = j & 31, slot = j >> 5;
        double zj = slot == 0 ? z[0] : (slot == 1 ? z[1] : z[2]);
        zj = __shfl_sync(0xffffffffu, zj, owner);
        const double yj = zj / A[tri(j, j)];
        if (lane == owner) y[j] = yj;
#pragma unroll
        for (int q = 0; q < 3; ++q) { const int i = lane + 32 * q; if (i < j) z[q] -= A[tri(j, i)] * yj; }
    }
}

template <int T>
__global__ void __launch_bounds__(T, 4) fit_single_kernel(const FitConst* __restrict__ C, int B, const b3d_kp* __restrict__ kps_all,
                                                       const int32_t* __restrict__ nkp_all, const double* __restrict__ R0_all,
                                                       double* __restrict__ sim3, double* __restrict__ w, double* __restrict__ jointAA,
                                                       uint8_t* __restrict__ usable, b3d_iter_record* __restrict__ trace, int trace_stride,
                                                       int32_t* __restrict__ niters, int32_t* __restrict__ term, double* __restrict__ Hs_all) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SingleSmem& S = *reinterpret_cast<SingleSmem*>(smem_raw);
    const int tid = threadIdx.x, b = blockIdx.x;
    if (b >= B) return;
    const int n = C->n, ntri = n * (n + 1) / 2;
    const b3d_kp* kps = kps_all + (size_t)b * MAXKP;
    const int nkp = min(max(nkp_all[b], 0), MAXKP);
    double* Hs = Hs_all + (size_t)b * HTRI;
    b3d_iter_record* tr = trace ? trace + (size_t)b * trace_stride : nullptr;

    // ---- problem set-up (Sim3BA.h:395-404 / 558-569): sim3 from the caller, jointAA = 0, w from the caller ----
    for (int k = tid; k < NFULL; k += T) {
        double v = 0.0;
        if (k < 7) v = sim3[(size_t)b * 7 + k];
        else if (k >= IDX_W) v = w[(size_t)b * NS + (k - IDX_W)];
        S.xfull[k] = v;
    }
    if (tid < 9) S.R0[tid] = R0_all[(size_t)b * 9 + tid];
    __syncthreads();
    // IterationZero: project onto the bounds (only the scale is bounded)
    if (tid == 0) S.xfull[IDX_S] = fmin(fmax(S.xfull[IDX_S], C->scale_lo), C->scale_hi);
    __syncthreads();

    double radius = kInitialRadius, decrease_factor = 2.0, x_cost = 0.0, cand_cost = 0.0, mcc = 0.0;
    bool reuse_diagonal = false, step_successful = true;
    int num_invalid = 0, iteration = 0, nrec = 0, termination = 0;
    b3d_iter_record it;
    it.cost = it.cost_change = it.gradient_max_norm = it.step_norm = it.relative_decrease = it.radius = 0.0;
    it.flags = 0; it.aux = -1;

    // evaluate at x with Jacobian, apply Jacobi scaling, publish H_s to global memory.  Returns false on non-finite cost.
    auto eval_grad_jac = [&](bool first) -> bool {
        x_cost = evaluate<T, true>(C, S.xfull, kps, nkp, S);
        if (!isfinite(x_cost)) return false;
        it.cost = x_cost;
        double* H = S.u.L;
        if (first)
            for (int k = tid; k < n; k += T) S.scale[k] = 1.0 / (1.0 + sqrt(H[tri(k, k)]));
        __syncthreads();
        for (int i = tid >> 5; i < n; i += T / 32) {
            const double si = S.scale[i];
            for (int j = tid & 31; j <= i; j += 32) {
                const double hv = H[tri(i, j)] * si * S.scale[j];
                Hs[tri(i, j)] = hv;
                if (i == j) S.hdiag[i] = hv;
            }
        }
        double gm = 0.0;
        for (int k = tid; k < n; k += T) {
            const double gk = S.g[k];
            S.gs[k] = gk * S.scale[k];
            const int fk = C->free_idx[k];
            double xn = S.xfull[fk] - gk;
            if (fk == IDX_S) xn = fmin(fmax(xn, C->scale_lo), C->scale_hi);
            gm = fmax(gm, fabs(S.xfull[fk] - xn));
        }
        // max-reduce through the sum helper's scratch (order independent)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) gm = fmax(gm, __shfl_down_sync(0xffffffffu, gm, o));
        __syncthreads();
        if ((tid & 31) == 0) S.red[tid >> 5] = gm;
        __syncthreads();
        gm = 0.0;
        for (int i = 0; i < T / 32; ++i) gm = fmax(gm, S.red[i]);
        it.gradient_max_norm = gm;
        __syncthreads();
        return true;
    };
    auto push = [&]() {
        it.radius = radius;
        it.aux = C->use_gmm && C->beta_pose > 0.0 ? S.comp : -1;
        if (tr && nrec < trace_stride && tid == 0) tr[nrec] = it;
        ++nrec;
    };
    auto x_norm_of = [&](const double* xf) -> double {
        double v = 0.0;
        for (int k = tid; k < n; k += T) { const double a = xf[C->free_idx[k]]; v += a * a; }
        return sqrt(block_sum<T>(v, S.red));
    };
    // cfull = Plus(xfull, a * delta): add, then clamp the bounded coordinate
    auto make_candidate = [&](double a) {
        for (int k = tid; k < NFULL; k += T) S.cfull[k] = S.xfull[k];
        __syncthreads();
        for (int k = tid; k < n; k += T) {
            const int fk = C->free_idx[k];
            double v = S.xfull[fk] + a * S.delta[k];
            if (fk == IDX_S) v = fmin(fmax(v, C->scale_lo), C->scale_hi);
            S.cfull[fk] = v;
        }
        __syncthreads();
    };

    double x_norm = x_norm_of(S.xfull);
    if (!eval_grad_jac(true)) {
        termination = B3D_TERM_EVAL_FAILURE;
    } else {
        it.flags = 3;
        for (;;) {
            // ---- FinalizeIterationAndCheckIfMinimizerCanContinue ----
            push();
            if (iteration >= C->max_iters) { termination = B3D_TERM_MAX_ITERS; break; }
            if (step_successful && it.gradient_max_norm <= kGradientTolerance) { termination = B3D_TERM_GRAD_TOL; break; }
            if (!(radius > kMinRadius)) { termination = B3D_TERM_MIN_RADIUS; break; }
            const double prev_gmax = it.gradient_max_norm;
            it.cost = it.cost_change = it.gradient_max_norm = it.step_norm = it.relative_decrease = 0.0;
            it.flags = 0;
            ++iteration;
            step_successful = false;

            // ---- LevenbergMarquardtStrategy::ComputeStep ----
            if (!reuse_diagonal)
                for (int k = tid; k < n; k += T) S.diag[k] = fmin(fmax(S.hdiag[k], kMinDiagonal), kMaxDiagonal);
            __syncthreads();
            double* A = S.u.L;
            for (int e = tid; e < ntri; e += T) A[e] = Hs[e];
            for (int k = tid; k < n; k += T) A[tri(n, k)] = S.gs[k];
            __syncthreads();
            for (int k = tid; k < n; k += T) { const double l = sqrt(S.diag[k] / radius); A[tri(k, k)] += l * l; }
            __syncthreads();
            reuse_diagonal = true;
            bool step_valid = false;
            if (chol_factor_aug<T>(A, n + 1)) {
                if (tid < 32) chol_backsolve_warp(A, n, S.y);
                __syncthreads();
                // model_cost_change = -(J step)^T (r + J step / 2), step = -y:  y^T gs - y^T H_s y / 2
                double v = 0.0, fin = 0.0;
                for (int i = tid >> 5; i < n; i += T / 32) {
                    const double yi = S.y[i];
                    for (int j = tid & 31; j <= i; j += 32) v -= (i == j ? 0.5 : 1.0) * Hs[tri(i, j)] * yi * S.y[j];
                }
                for (int k = tid; k < n; k += T) { v += S.y[k] * S.gs[k]; fin += isfinite(S.y[k]) ? 0.0 : 1.0; }
                mcc = block_sum<T>(v, S.red);
                const double bad = block_sum<T>(fin, S.red);
                step_valid = (bad == 0.0) && (mcc > 0.0);
            }
            if (!step_valid) {
                // ---- HandleInvalidStep ----
                if (++num_invalid >= kMaxConsecutiveInvalidSteps) { termination = B3D_TERM_INVALID_STEPS; break; }
                radius = radius / decrease_factor; decrease_factor *= 2.0;
                it.cost = x_cost; it.gradient_max_norm = prev_gmax;
                continue;
            }
            num_invalid = 0;
            for (int k = tid; k < n; k += T) S.delta[k] = -S.y[k] * S.scale[k];
            __syncthreads();
            it.flags = 1;

            // ---- DoLineSearch (the problem is bounded: projected Armijo along delta) ----
            make_candidate(1.0);
            cand_cost = evaluate<T, false>(C, S.cfull, kps, nkp, S);
            {
                double v = 0.0;
                for (int k = tid; k < n; k += T) v += S.g[k] * S.delta[k];
                const double gd = block_sum<T>(v, S.red);
                const bool valid1 = isfinite(cand_cost);
                if (!valid1 || cand_cost > x_cost + kSufficientDecrease * gd * 1.0) {
                    // rare path: interpolate.  Needs the directional derivative at every trial point.
                    double dmax = 0.0;
                    for (int k = 0; k < n; ++k) dmax = fmax(dmax, fabs(S.delta[k]));
                    LsSample init{0.0, x_cost, gd, true}, prev{0, 0, 0, false}, cur{1.0, 0, 0, false};
                    // S.g must survive: gradient at trial points goes to S.gc
                    auto ls_eval = [&](double a, LsSample& s) {
                        s.x = a; s.valid = false;
                        make_candidate(a);
                        for (int k = tid; k < n; k += T) S.gc[k] = S.g[k];
                        __syncthreads();
                        const double c = evaluate<T, true>(C, S.cfull, kps, nkp, S);   // overwrites S.g
                        double vv = 0.0;
                        for (int k = tid; k < n; k += T) { const double t = S.g[k]; S.g[k] = S.gc[k]; vv += S.delta[k] * t; }
                        const double gg = block_sum<T>(vv, S.red);
                        if (!isfinite(c) || !isfinite(gg)) return;
                        s.value = c; s.gradient = gg; s.valid = true;
                    };
                    ls_eval(1.0, cur);
                    bool success = true;
                    int ls_iters = 0;
                    while (!cur.valid || cur.value > x_cost + kSufficientDecrease * gd * cur.x) {
                        if (++ls_iters >= kMaxLineSearchIters) { success = false; break; }
                        const double lo = kLsContractLo * cur.x, hi = kLsContractHi * cur.x;
                        if (tid == 0) {
                            double ss;
                            if (!cur.valid) ss = fmin(fmax(cur.x * 0.5, lo), hi);
                            else {
                                LsSample smp[3] = {init, cur, prev};
                                ss = minimize_interpolating_polynomial(smp, prev.valid ? 3 : 2, lo, hi);
                            }
                            S.ls_step = ss;
                        }
                        __syncthreads();
                        const double step_size = S.ls_step;
                        __syncthreads();
                        if (step_size * dmax < kMinLineSearchStepSize) { success = false; break; }
                        prev = cur;
                        ls_eval(step_size, cur);
                    }
                    if (success && cur.x != 1.0) {
                        for (int k = tid; k < n; k += T) S.delta[k] *= cur.x;
                        it.flags |= 4;
                    }
                    __syncthreads();
                    // ---- ComputeCandidatePointAndEvaluateCost ----
                    make_candidate(1.0);
                    cand_cost = evaluate<T, false>(C, S.cfull, kps, nkp, S);
                }
            }
            if (!isfinite(cand_cost)) cand_cost = DBL_MAX;

            // ---- ParameterToleranceReached ----
            {
                double v = 0.0;
                for (int k = tid; k < n; k += T) { const int fk = C->free_idx[k]; const double d = S.xfull[fk] - S.cfull[fk]; v += d * d; }
                it.step_norm = sqrt(block_sum<T>(v, S.red));
            }
            if (it.step_norm <= kParameterTolerance * (x_norm + kParameterTolerance)) { termination = B3D_TERM_PARAM_TOL; break; }
            // ---- FunctionToleranceReached ----
            it.cost_change = x_cost - cand_cost;
            if (fabs(it.cost_change) <= kFunctionTolerance * x_cost) { termination = B3D_TERM_FUNC_TOL; break; }
            // ---- IsStepSuccessful ----
            it.relative_decrease = it.cost_change / mcc;
            if (it.relative_decrease > kMinRelativeDecrease) {
                __syncthreads();
                for (int k = tid; k < NFULL; k += T) S.xfull[k] = S.cfull[k];
                __syncthreads();
                x_norm = x_norm_of(S.xfull);
                if (!eval_grad_jac(false)) { termination = B3D_TERM_EVAL_FAILURE; break; }
                step_successful = true;
                it.flags |= 2;
                const double q = 2.0 * it.relative_decrease - 1.0;
                radius = radius / fmax(1.0 / 3.0, 1.0 - q * q * q);
                radius = fmin(kMaxRadius, radius);
                decrease_factor = 2.0;
                reuse_diagonal = false;
            } else {
                radius = radius / decrease_factor; decrease_factor *= 2.0;
                it.cost = cand_cost;
            }
        }
    }
    __syncthreads();
    // ---- results: optimised blocks back to the caller's arrays (Sim3BA.h:481-507 reads them from the blocks) ----
    if (tid < 7) sim3[(size_t)b * 7 + tid] = S.xfull[tid];
    for (int k = tid; k < NJ * 3; k += T) jointAA[(size_t)b * NJ * 3 + k] = k < 3 ? 0.0 : S.xfull[IDX_JAA + k - 3];
    if (C->has_w && tid < NS) w[(size_t)b * NS + tid] = S.xfull[IDX_W + tid];
    if (tid == 0) {
        if (usable) usable[b] = (termination >= B3D_TERM_PARAM_TOL && termination <= B3D_TERM_MIN_RADIUS) ? 1 : 0;
        if (niters) niters[b] = nrec;
        if (term) term[b] = termination;
    }
}

// Host side: launch-constant block shared by the single-frame and window drivers.
int build_fit_const(const b3d_model* m, const b3d_gmm* gmm, const b3d_fit_opts* o, bool multi, FitConst* fc) {
    std::memset(fc, 0, sizeof(FitConst));
    for (int j = 0; j < NJ; ++j) {
        fc->parent[j] = m->sk.parent[j];
        int len = 0, cur = j;
        fc->chain[j][len++] = j;
        while (m->sk.parent[cur] != -1 && m->sk.parent[cur] != 0) {
            cur = m->sk.parent[cur];
            if (len >= MAXCHAIN) return set_error(B3D_EINVAL, "kinematic chain of joint %d is deeper than %d", j, MAXCHAIN - 1);
            fc->chain[j][len++] = cur;
        }
        fc->chain_len[j] = len;
        for (int c = 0; c < 3; ++c) fc->offset[j][c] = m->sk.offset[3 * j + c];
    }
    // single: jointShapeReg is handed to the functor iff mode == pose+shape && betaShape > 0 (Sim3BA.h:417); pose-only never
    // (Sim3BA.h:581-590).  multi: iff betaShape > 0 (MultiFrameBA.h:87).
    fc->with_shape = multi ? (o->beta_shape > 0.0) : (o->mode == B3D_FIT_POSE_SHAPE && o->beta_shape > 0.0);
    if (fc->with_shape)
        for (int j = 0; j < NJ; ++j)
            for (int c = 0; c < 3; ++c)
                for (int s = 0; s < NS; ++s) {
                    const int pa = m->sk.parent[j];
                    fc->DB[j][c][s] = m->sk.JSR[(3 * j + c) * NS + s] - (pa >= 0 ? m->sk.JSR[(3 * pa + c) * NS + s] : 0.0);
                }
    fc->cam[0] = o->fx; fc->cam[1] = o->fy; fc->cam[2] = o->cx; fc->cam[3] = o->cy;
    fc->huber = o->huber; fc->beta_pose = o->beta_pose; fc->beta_shape = o->beta_shape; fc->lambda_temp = o->lambda_temp;
    fc->scale_lo = o->scale_lo; fc->scale_hi = o->scale_hi;
    fc->mode = o->mode; fc->max_iters = o->max_iters;
    fc->use_gmm = (!multi && gmm && gmm->nComps > 0) ? 1 : 0;      // silent L2 degrade otherwise (Sim3BA.h:266); multi never (MultiFrameBA.h:109)
    fc->has_w = multi ? 1 : (o->beta_shape > 0.0 ? 1 : 0);         // shape-prior block (Sim3BA.h:466-470 / 630-638)
    int nfree = 0;
    for (int k = 0; k < NFULL; ++k) fc->col_of[k] = -1;
    auto add = [&](int fi) { fc->col_of[fi] = nfree; fc->free_idx[nfree++] = fi; };
    for (int k = 0; k < 7; ++k) add(k);
    for (int j = 1; j < NJ; ++j) {
        // pose-only mode holds joints 10, 11, 22, 23 constant (Sim3BA.h:608-611)
        if (!multi && o->mode == B3D_FIT_POSE && (j == 10 || j == 11 || j == 22 || j == 23)) continue;
        for (int c = 0; c < 3; ++c) add(IDX_JAA + 3 * (j - 1) + c);
    }
    if (fc->has_w) for (int s = 0; s < NS; ++s) add(IDX_W + s);
    fc->n = nfree;
    if (fc->use_gmm) {
        if (gmm->nComps > GMM_MAXC) return set_error(B3D_EINVAL, "GMM with %d components (max %d)", gmm->nComps, GMM_MAXC);
        fc->gmm_ncomp = gmm->nComps;
        fc->gmm_mean = gmm->d_mean; fc->gmm_L = gmm->d_prec_cho; fc->gmm_LT = gmm->d_prec_choT; fc->gmm_P = gmm->d_prec; fc->gmm_consts = gmm->d_consts;
    }
    return B3D_OK;
}

}  // namespace b3d

using namespace b3d;

extern "C" {

int b3d_fit_single_dev(const b3d_model* m, const b3d_gmm* gmm, const b3d_fit_opts* opts, int32_t B, const b3d_kp* kps,
                       const int32_t* nkp, const double* R0, double* sim3, double* w, double* jointAA, uint8_t* usable,
                       b3d_iter_record* trace, int32_t trace_stride, int32_t* niters, int32_t* term, void* stream) {
    if (!m || !opts || !kps || !nkp || !R0 || !sim3 || !w || !jointAA) return set_error(B3D_EINVAL, "b3d_fit_single: null argument");
    if (opts->mode != B3D_FIT_POSE && opts->mode != B3D_FIT_POSE_SHAPE) return set_error(B3D_EINVAL, "b3d_fit_single: bad mode %d", opts->mode);
    if (B < 0 || opts->max_iters < 0) return set_error(B3D_EINVAL, "b3d_fit_single: negative size");
    if (trace && trace_stride <= 0) return set_error(B3D_EINVAL, "b3d_fit_single: trace without trace_stride");
    if (B == 0) return B3D_OK;
    cudaStream_t st = (cudaStream_t)stream;
    FitConst fc;
    if (int rc = build_fit_const(m, gmm, opts, false, &fc)) return rc;
    FitConst* dfc = nullptr;
    double* Hs = nullptr;
    B3D_CUDA(cudaMallocAsync(&dfc, sizeof(FitConst), st));
    B3D_CUDA(cudaMallocAsync(&Hs, sizeof(double) * (size_t)B * HTRI, st));
    // pageable source: the copy is staged before the call returns, so the stack object may go out of scope
    B3D_CUDA(cudaMemcpyAsync(dfc, &fc, sizeof(FitConst), cudaMemcpyHostToDevice, st));
    static bool attr_set = false;
    if (!attr_set) {
        B3D_CUDA(cudaFuncSetAttribute(fit_single_kernel<FS_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SingleSmem)));
        attr_set = true;
    }
    fit_single_kernel<FS_THREADS><<<B, FS_THREADS, sizeof(SingleSmem), st>>>(dfc, B, kps, nkp, R0, sim3, w, jointAA, usable, trace, trace_stride,
                                                                            niters, term, Hs);
    B3D_LAUNCH_CHECK();
    B3D_CUDA(cudaFreeAsync(Hs, st));
    B3D_CUDA(cudaFreeAsync(dfc, st));
    return B3D_OK;
}

int b3d_fit_single(const b3d_model* m, const b3d_gmm* gmm, const b3d_fit_opts* opts, int32_t B, const b3d_kp* kps, const int32_t* nkp,
                   const double* R0, double* sim3, double* w, double* jointAA, uint8_t* usable, b3d_iter_record* trace,
                   int32_t trace_stride, int32_t* niters, int32_t* term) {
    if (!m || !opts || !kps || !nkp || !R0 || !sim3 || !w || !jointAA) return set_error(B3D_EINVAL, "b3d_fit_single: null argument");
    if (B < 0) return set_error(B3D_EINVAL, "b3d_fit_single: B < 0");
    if (B == 0) return B3D_OK;
    const size_t nb = (size_t)B;
    const size_t s_kp = sizeof(b3d_kp) * nb * MAXKP, s_nk = 4 * nb, s_r0 = 8 * nb * 9, s_s3 = 8 * nb * 7, s_w = 8 * nb * NS,
                 s_ja = 8 * nb * NJ * 3, s_tr = trace ? sizeof(b3d_iter_record) * nb * (size_t)trace_stride : 0;
    auto al = [](size_t x) { return (x + 255) / 256 * 256; };
    const size_t total = al(s_kp) + al(s_nk) + al(s_r0) + al(s_s3) + al(s_w) + al(s_ja) + al(s_tr) + al(nb) + 2 * al(4 * nb);
    char* base = nullptr;
    B3D_CUDA(cudaMalloc(&base, total));
    char* p = base;
    auto take = [&](size_t s) { char* q = p; p += al(s); return q; };
    b3d_kp* d_kp = (b3d_kp*)take(s_kp); int32_t* d_nk = (int32_t*)take(s_nk); double* d_r0 = (double*)take(s_r0);
    double* d_s3 = (double*)take(s_s3); double* d_w = (double*)take(s_w); double* d_ja = (double*)take(s_ja);
    b3d_iter_record* d_tr = trace ? (b3d_iter_record*)take(s_tr) : nullptr;
    uint8_t* d_us = (uint8_t*)take(nb); int32_t* d_ni = (int32_t*)take(4 * nb); int32_t* d_te = (int32_t*)take(4 * nb);
    cudaError_t e = cudaMemcpy(d_kp, kps, s_kp, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_nk, nkp, s_nk, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_r0, R0, s_r0, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_s3, sim3, s_s3, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_w, w, s_w, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && d_tr) e = cudaMemset(d_tr, 0, s_tr);
    int rc = B3D_OK;
    if (e == cudaSuccess) {
        rc = b3d_fit_single_dev(m, gmm, opts, B, d_kp, d_nk, d_r0, d_s3, d_w, d_ja, d_us, d_tr, trace_stride, d_ni, d_te, nullptr);
        if (!rc) {
            e = cudaStreamSynchronize(nullptr);
            if (e == cudaSuccess) e = cudaMemcpy(sim3, d_s3, s_s3, cudaMemcpyDeviceToHost);
            if (e == cudaSuccess) e = cudaMemcpy(w, d_w, s_w, cudaMemcpyDeviceToHost);
            if (e == cudaSuccess) e = cudaMemcpy(jointAA, d_ja, s_ja, cudaMemcpyDeviceToHost);
            if (e == cudaSuccess && usable) e = cudaMemcpy(usable, d_us, nb, cudaMemcpyDeviceToHost);
            if (e == cudaSuccess && trace) e = cudaMemcpy(trace, d_tr, s_tr, cudaMemcpyDeviceToHost);
            if (e == cudaSuccess && niters) e = cudaMemcpy(niters, d_ni, 4 * nb, cudaMemcpyDeviceToHost);
            if (e == cudaSuccess && term) e = cudaMemcpy(term, d_te, 4 * nb, cudaMemcpyDeviceToHost);
        }
    }
    cudaFree(base);
    if (e != cudaSuccess) return set_error(B3D_ECUDA, "b3d_fit_single: %s", cudaGetErrorString(e));
    return rc;
}

}  // extern "C"
