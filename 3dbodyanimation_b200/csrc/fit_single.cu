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
//
// The kernel is latency-bound (one frame's iterations are a serial chain), so the layout aims at few barriers and short
// dependent chains: J^T J is assembled by element owners walking a per-column keypoint bitmask (no per-keypoint phases),
// the Cholesky keeps running diagonals in registers and splits each row's dot product over a thread pair (one barrier per
// column), and the scaled normal equations stay in shared memory across rejected steps.
#include "lm_device.cuh"
#include "ls_poly.cuh"

namespace b3d {

constexpr int FS_THREADS = 256;
constexpr int NAUG = NFULL + 1;                       // augmented with the right-hand side row
constexpr int LTRI = NAUG * (NAUG + 1) / 2;           // 3828 packed doubles
constexpr int HTRI = NFULL * (NFULL + 1) / 2;         // 3741
constexpr int GMM_MAXC = 16;

constexpr double kInitialRadius = 1e4, kMaxRadius = 1e16, kMinRadius = 1e-32;
constexpr double kMinDiagonal = 1e-6, kMaxDiagonal = 1e32;
constexpr double kMinRelativeDecrease = 1e-3;
constexpr double kFunctionTolerance = 1e-6, kGradientTolerance = 1e-10, kParameterTolerance = 1e-8;
constexpr int kMaxConsecutiveInvalidSteps = 5;

// Optional phase timing (debug hook, off unless b3d_debug_phase_timing(1) was called): cycles of block 0 per phase.
__device__ long long g_phase_cycles[32];
__device__ long long g_block_cycles[1024];   // total cycles and line-search evaluations per block (debug)
__device__ int g_block_ls[1024];
__device__ int g_phase_on;
struct PhaseTimer {
    long long t;
    bool on;
    __device__ void start() { on = g_phase_on != 0 && blockIdx.x == 0 && threadIdx.x == 0; if (on) t = clock64(); }
    __device__ void mark(int i) { if (on) { const long long u = clock64(); g_phase_cycles[i] += u - t; t = u; } }
};

struct GmmTmp {
    double h[GMM_MAXC][NPRI];     // L_i^T (x - mu_i)
    double d[GMM_MAXC][NPRI];     // x - mu_i
    double p[GMM_MAXC];           // |h_i|^2 / 2 - consts_log_i
    double sq[GMM_MAXC];          // |sqrt(1/2) h_i|^2
    double part[3][NPRI];         // partial sums of L_c h
};

union SingleUnion {
    EvalTmp ev;
    GmmTmp gm;
    double L[LTRI];               // damped system being factored, last row = right-hand side
};

struct SingleSmem {
    EvalOut out;
    SingleUnion u;
    double Hs[HTRI];                       // Jacobi-scaled J^T J (free layout, packed lower), kept across rejected steps
    double xfull[NFULL], cfull[NFULL];     // current / candidate parameters, 86-layout
    double g[NFULL], gs[NFULL], gc[NFULL]; // gradient (free layout): unscaled, Jacobi-scaled, saved copy during a line search
    double scale[NFULL], diag[NFULL], y[NFULL], delta[NFULL], invd[NFULL];
    double gpri[NPRI];                     // pose-prior gradient, pose-dim layout
    double R0[9];
    double red[40];
    double piv[2];                         // pivot of the next Cholesky column: value, 1/sqrt
    double ls_step;
    int32_t comp;
    uint32_t kmask[NFULL];                 // per full-layout column: keypoints whose Jacobian touches it
    uint8_t slot[MAXKP][NFULL + 2];        // per keypoint: local slot of each full-layout column
    int8_t fcol[MAXKP][NNZPAD];            // per keypoint slot: free-layout column, -1 = held constant
};

// ---- GMM pose prior (GaussianMixture::residual, external/avatar/GaussianMixture.cpp:94-113) --------------------
// Returns 0.5 |beta r|^2 of the prior block and sets S.comp; with JAC also S.gpri = J^T r (pose-dim layout), where
// J = beta L_c^T WITHOUT the sqrt(1/2) the residual carries (reference quirk, Sim3BA.h:298-299).
template <int T, bool JAC>
__device__ double gmm_eval(const FitConst* __restrict__ C, const double* xj /*69*/, SingleSmem& S) {
    const int tid = threadIdx.x, nc = C->gmm_ncomp;
    GmmTmp& G = S.u.gm;
    PhaseTimer pg;
    pg.start();
    for (int t = tid; t < nc * NPRI; t += T) G.d[t / NPRI][t % NPRI] = xj[t % NPRI] - C->gmm_mean[t];
    __syncthreads();
    pg.mark(16);
    // h_i[k] = sum_{m >= k} L_i[m][k] d_i[m].  Warp i takes component i; lane q < 17 owns the adjacent column pairs (2q, 2q+1) and
    // (66-2q, 67-2q): 72 sixteen-byte loads whatever q (rows are padded to 70 doubles; entries above the diagonal are stored
    // zeros).  All lanes walk the SAME row in the same instruction, so a warp load touches 3 lines instead of one line per lane
    // (ncu: the per-lane row order made every load a 17-line gather); rows above a lane's columns are predicated off.
    constexpr int QT = 17, LD = NPRI + 1, RS = LD / 2;           // row stride in double2
    static_assert(T / 32 >= 8, "one warp per GMM component");
    for (int i = tid >> 5; i < nc; i += T / 32) {
        const int q = tid & 31;
        if (q < QT) {
            const int ca = 2 * q, cb = 66 - 2 * q;
            const double* __restrict__ L = C->gmm_L70 + (size_t)i * NPRI * LD;
            const double* dd = G.d[i];
            const double2* __restrict__ La = reinterpret_cast<const double2*>(L + ca);
            const double2* __restrict__ Lb = reinterpret_cast<const double2*>(L + cb);
            double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0, b0 = 0.0, b1 = 0.0, b2 = 0.0, b3 = 0.0;
            const double2 zero2 = make_double2(0.0, 0.0);
            // rows 0 .. 31 can only touch pair A (cb >= 34); rows 32 .. 68 may touch both
            for (int m = 0; m < 32; m += 8) {
                double2 v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) v[u] = (m + u >= ca) ? __ldg(La + (m + u) * RS) : zero2;
#pragma unroll
                for (int u = 0; u < 8; u += 2) {
                    a0 += v[u].x * dd[m + u]; a1 += v[u].y * dd[m + u];
                    a2 += v[u + 1].x * dd[m + u + 1]; a3 += v[u + 1].y * dd[m + u + 1];
                }
            }
            for (int m = 32; m < NPRI; m += 4) {
                double2 v[4], w4[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int r = m + u;
                    v[u] = (r < NPRI && r >= ca) ? __ldg(La + r * RS) : zero2;
                    w4[u] = (r < NPRI && r >= cb) ? __ldg(Lb + r * RS) : zero2;
                }
#pragma unroll
                for (int u = 0; u < 4; u += 2) {
                    const double d0 = m + u < NPRI ? dd[m + u] : 0.0, d1 = m + u + 1 < NPRI ? dd[m + u + 1] : 0.0;
                    a0 += v[u].x * d0; a1 += v[u].y * d0; a2 += v[u + 1].x * d1; a3 += v[u + 1].y * d1;
                    b0 += w4[u].x * d0; b1 += w4[u].y * d0; b2 += w4[u + 1].x * d1; b3 += w4[u + 1].y * d1;
                }
            }
            G.h[i][ca] = a0 + a2; G.h[i][ca + 1] = a1 + a3;
            G.h[i][cb] = b0 + b2; G.h[i][cb + 1] = b1 + b3;
            if (q == 0) G.h[i][NPRI - 1] = __ldg(L + (size_t)(NPRI - 1) * LD + (NPRI - 1)) * dd[NPRI - 1];   // column 68: one entry
        }
    }
    __syncthreads();
    pg.mark(17);
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
    for (int i = 0; i < nc; ++i) { const double p = G.p[i]; if (p < best) { best = p; c = i; } }   // strict '<': first wins ties
    if (tid == 0) S.comp = c;
    pg.mark(18);
    // residual = beta [h_c ; sqrt(-consts_log_c)]  =>  cost = 0.5 beta^2 (|h_c|^2 - consts_log_c)
    const double cost = 0.5 * C->beta_pose * C->beta_pose * (G.sq[c] + (-C->gmm_consts[c]));
    if (JAC) {
        // J^T r = beta^2 sqrt(1/2) L_c h: row m split over three threads
        const double* __restrict__ LT = C->gmm_LT + (size_t)c * NPRI * NPRI;
        if (tid < 3 * NPRI) {
            const int part = tid / NPRI, m = tid % NPRI;
            double a0 = 0.0, a1 = 0.0;
            int k = part;
            for (; k + 3 <= m; k += 6) { a0 += __ldg(LT + k * NPRI + m) * G.h[c][k]; a1 += __ldg(LT + (k + 3) * NPRI + m) * G.h[c][k + 3]; }
            for (; k <= m; k += 3) a0 += __ldg(LT + k * NPRI + m) * G.h[c][k];
            G.part[part][m] = a0 + a1;
        }
        __syncthreads();
        const double f = C->beta_pose * C->beta_pose * sqrt(0.5);
        for (int m = tid; m < NPRI; m += T) S.gpri[m] = f * ((G.part[0][m] + G.part[1][m]) + G.part[2][m]);
    }
    __syncthreads();
    pg.mark(19);
    return cost;
}

// ---- cost (and with JAC the normal equations) at parameters xf --------------------------------------------------
// JAC: on return S.Hs holds the packed J^T J in the free layout — scaled by S.scale unless `first`, in which case it is
// unscaled and the caller derives the Jacobi scaling from its diagonal — and S.g the unscaled gradient J^T r.
// MODE 0: cost only.  MODE 1: cost + normal equations (S.Hs, S.g).  MODE 2: cost + gradient only, into S.gc — what the
// Armijo line search needs at its trial points; S.Hs / S.g / S.gs of the current iterate stay untouched.  The gradient entries
// are accumulated in the same order as in MODE 1, so both give the same bits.
template <int T, int MODE>
__device__ __noinline__ double evaluate(const FitConst* __restrict__ C, const double* xf, const b3d_kp* __restrict__ kps, int nkp, SingleSmem& S,
                           bool first) {
    constexpr bool JAC = MODE != 0;
    const int tid = threadIdx.x, n = C->n;
    PhaseTimer pe;
    pe.start();
    eval_reproj<T, JAC>(C, xf, kps, nkp, S.R0, S.u.ev, S.out);
    pe.mark(JAC ? 9 : 12);
    double cost = 0.0;
    for (int k = 0; k < nkp; ++k) cost += S.out.kcost[k];
    double pcost = 0.0;
    const bool pose_prior = C->beta_pose > 0.0;
    const bool gmm = pose_prior && C->use_gmm;
    if (gmm) pcost = gmm_eval<T, JAC>(C, xf + IDX_JAA, S);   // aliases S.u.ev: the evaluation temporaries are dead
    // the remaining prior costs are sums of squares over <= 79 terms: fixed-order block sum
    double v = 0.0;
    if (pose_prior && !gmm)
        for (int k = tid; k < NPRI; k += T) { const double r = xf[IDX_JAA + k] * C->beta_pose; v += 0.5 * r * r; }
    if (C->has_w)
        for (int k = tid; k < NS; k += T) { const double r = xf[IDX_W + k] * C->beta_shape; v += 0.5 * r * r; }
    if (JAC) {   // column -> keypoint tables for the assembly below (published by the barriers inside block_sum)
        for (int k = tid; k < NFULL; k += T) S.kmask[k] = 0u;
    }
    cost += pcost + block_sum<T>(v, S.red);
    pe.mark(JAC ? 10 : 13);
    if (!JAC) return cost;

    for (int t = tid; t < nkp * NNZPAD; t += T) {
        const int k = t / NNZPAD, sl = t % NNZPAD;
        if (sl < S.out.jnnz[k]) {
            const int col = S.out.jcol[k][sl];
            S.slot[k][col] = (uint8_t)sl;
            S.fcol[k][sl] = (int8_t)C->col_of[col];
            atomicOr(&S.kmask[col], 1u << k);
        }
    }
    __syncthreads();
    const double bp2 = C->beta_pose * C->beta_pose, bs2 = C->beta_shape * C->beta_shape;
    // ---- gradient J^T r: thread i sums the keypoints that touch column i in keypoint order, then the priors ----
    {
        double* gdst = MODE == 2 ? S.gc : S.g;
        for (int i = tid; i < n; i += T) {
            const int a = C->free_idx[i];
            uint32_t m = S.kmask[a];
            double gacc = 0.0;
            while (m) {
                const int k = __ffs(m) - 1;
                m &= m - 1;
                const int sa = S.slot[k][a];
                gacc += S.out.J[k][0][sa] * S.out.res[k][0] + S.out.J[k][1][sa] * S.out.res[k][1];
            }
            if (pose_prior && a >= IDX_JAA && a < IDX_W) gacc += gmm ? S.gpri[a - IDX_JAA] : bp2 * xf[a];
            if (C->has_w && a >= IDX_W) gacc += bs2 * xf[a];
            gdst[i] = gacc;
        }
    }
    if (MODE == 2) {
        __syncthreads();
        return cost;
    }
    // ---- J^T J: a thread owns element (i, j) of the packed lower triangle and sums, in keypoint order and in a register, the
    // keypoints whose Jacobian touches both columns; then the prior term and the Jacobi scaling, one store per element.  Four
    // elements per pass, so that the four loads of the GMM precision matrix (L2) are in flight while the keypoints are walked. ----
    {
        const double* __restrict__ P = gmm ? C->gmm_P + (size_t)S.comp * NPRI * NPRI : nullptr;
        const int ne = n * (n + 1) / 2;
        for (int e0 = tid; e0 < ne; e0 += 4 * T) {
            int ei[4], ej[4];
            double pri[4], acc[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int e = e0 + u * T;
                int i = (int)((sqrtf(8.0f * (float)e + 1.0f) - 1.0f) * 0.5f);
                while ((i + 1) * (i + 2) / 2 <= e) ++i;
                while (i * (i + 1) / 2 > e) --i;
                ei[u] = i; ej[u] = e - i * (i + 1) / 2;
                pri[u] = 0.0;
                if (e < ne) {
                    const int a = C->free_idx[i], b = C->free_idx[ej[u]];
                    if (pose_prior && a >= IDX_JAA && a < IDX_W && b >= IDX_JAA) {
                        if (gmm) pri[u] = bp2 * __ldg(P + (a - IDX_JAA) * NPRI + (b - IDX_JAA));
                        else if (a == b) pri[u] = bp2;
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                acc[u] = 0.0;
                if (e0 + u * T < ne) {
                    const int a = C->free_idx[ei[u]], b = C->free_idx[ej[u]];
                    uint32_t m = S.kmask[a] & S.kmask[b];
                    while (m) {
                        const int k = __ffs(m) - 1;
                        m &= m - 1;
                        const int sa = S.slot[k][a], sb = S.slot[k][b];
                        acc[u] += S.out.J[k][0][sa] * S.out.J[k][0][sb] + S.out.J[k][1][sa] * S.out.J[k][1][sb];
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int e = e0 + u * T;
                if (e < ne) {
                    const int i = ei[u], j = ej[u];
                    const int a = C->free_idx[i], b = C->free_idx[j];
                    double v2 = acc[u];
                    if (pose_prior && a >= IDX_JAA && a < IDX_W && b >= IDX_JAA && (gmm || a == b)) v2 += pri[u];
                    if (C->has_w && a >= IDX_W && a == b) v2 += bs2;
                    S.Hs[e] = first ? v2 : v2 * S.scale[i] * S.scale[j];
                }
            }
        }
    }
    __syncthreads();
    pe.mark(11);
    return cost;
}

// Cholesky of the damped system in S.u.L, packed lower, (n + 1) rows: the last row is the right-hand side, so on return it
// holds z = L^-1 rhs.  Blocked left-looking, 4 columns per panel, 2 barriers per panel (44 instead of one per column):
//   update   the thread pair of row p subtracts sum_{q < j0} L[p][q] L[c][q] from its entries of the panel's columns c — the
//            four dot products share the loads of row p and the two halves of the pair split q;
//   factor   every row thread factors the (updated) 4 x 4 diagonal block redundantly in registers — the same arithmetic in
//            every thread, hence the same bits — and solves its own row against it.
// Returns false when a pivot is not positive (Eigen LLT / oracle criterion).  1 / L_jj goes to S.invd.
template <int T>
__device__ bool chol_factor_aug(SingleSmem& S, int n) {
    static_assert(T >= 2 * NAUG, "one thread pair per row");
    constexpr int PW = 4;
    double* A = S.u.L;
    const int tid = threadIdx.x, p = tid >> 1, half = tid & 1, np = n + 1;
    bool ok = true;
    for (int j0 = 0; j0 < n; j0 += PW) {
        const int w = min(PW, n - j0);
        const bool active = p >= j0 && p < np;
        double* ri = A + tri(active ? p : 0, 0);
        // ---- update ----
        if (j0 > 0) {
            double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
            if (active) {
                const double* r0 = A + tri(j0, 0);
                const double* r1 = A + tri(min(j0 + 1, n - 1), 0);
                const double* r2 = A + tri(min(j0 + 2, n - 1), 0);
                const double* r3 = A + tri(min(j0 + 3, n - 1), 0);
                int q = half;
                for (; q + 2 < j0; q += 4) {
                    const double a0 = ri[q], a1 = ri[q + 2];
                    s0 -= a0 * r0[q]; s1 -= a0 * r1[q]; s2 -= a0 * r2[q]; s3 -= a0 * r3[q];
                    s0 -= a1 * r0[q + 2]; s1 -= a1 * r1[q + 2]; s2 -= a1 * r2[q + 2]; s3 -= a1 * r3[q + 2];
                }
                for (; q < j0; q += 2) {
                    const double a0 = ri[q];
                    s0 -= a0 * r0[q]; s1 -= a0 * r1[q]; s2 -= a0 * r2[q]; s3 -= a0 * r3[q];
                }
            }
            s0 += __shfl_xor_sync(0xffffffffu, s0, 1);
            s1 += __shfl_xor_sync(0xffffffffu, s1, 1);
            s2 += __shfl_xor_sync(0xffffffffu, s2, 1);
            s3 += __shfl_xor_sync(0xffffffffu, s3, 1);
            if (active && half == 0) {
                const int last = min(w - 1, p - j0);       // columns j0 .. j0 + last exist in row p
                ri[j0] += s0;
                if (last >= 1) ri[j0 + 1] += s1;
                if (last >= 2) ri[j0 + 2] += s2;
                if (last >= 3) ri[j0 + 3] += s3;
            }
            __syncthreads();
        }
        // ---- factor the diagonal block (redundantly) and solve the own row ----
        {
            const double d00 = A[tri(j0, j0)];
            if (!(d00 > 0.0)) { ok = false; break; }
            const double i0 = rsqrt(d00);
            double l10 = 0.0, l20 = 0.0, l21 = 0.0, l30 = 0.0, l31 = 0.0, l32 = 0.0, i1 = 0.0, i2 = 0.0, i3 = 0.0;
            if (w > 1) {
                l10 = A[tri(j0 + 1, j0)] * i0;
                const double d11 = A[tri(j0 + 1, j0 + 1)] - l10 * l10;
                if (!(d11 > 0.0)) { ok = false; break; }
                i1 = rsqrt(d11);
            }
            if (w > 2) {
                l20 = A[tri(j0 + 2, j0)] * i0;
                l21 = (A[tri(j0 + 2, j0 + 1)] - l20 * l10) * i1;
                const double d22 = A[tri(j0 + 2, j0 + 2)] - l20 * l20 - l21 * l21;
                if (!(d22 > 0.0)) { ok = false; break; }
                i2 = rsqrt(d22);
            }
            if (w > 3) {
                l30 = A[tri(j0 + 3, j0)] * i0;
                l31 = (A[tri(j0 + 3, j0 + 1)] - l30 * l10) * i1;
                l32 = (A[tri(j0 + 3, j0 + 2)] - l30 * l20 - l31 * l21) * i2;
                const double d33 = A[tri(j0 + 3, j0 + 3)] - l30 * l30 - l31 * l31 - l32 * l32;
                if (!(d33 > 0.0)) { ok = false; break; }
                i3 = rsqrt(d33);
            }
            __syncthreads();    // every thread has read the block before its owners overwrite it
            if (active && half == 0) {
                const int last = min(w - 1, p - j0);
                const double v0 = ri[j0] * i0;
                double v1 = 0.0, v2 = 0.0, v3 = 0.0;
                if (last >= 1) v1 = (ri[j0 + 1] - v0 * l10) * i1;
                if (last >= 2) v2 = (ri[j0 + 2] - v0 * l20 - v1 * l21) * i2;
                if (last >= 3) v3 = (ri[j0 + 3] - v0 * l30 - v1 * l31 - v2 * l32) * i3;
                // a row inside the panel stops before its diagonal entry (kept unfactored; 1 / L_jj lives in S.invd)
                const int stop = p - j0;   // first column index (relative) that is the diagonal or beyond
                if (stop > 0) ri[j0] = v0;
                if (last >= 1 && stop > 1) ri[j0 + 1] = v1;
                if (last >= 2 && stop > 2) ri[j0 + 2] = v2;
                if (last >= 3 && stop > 3) ri[j0 + 3] = v3;
                if (p < n && stop < w) S.invd[p] = stop == 0 ? i0 : (stop == 1 ? i1 : (stop == 2 ? i2 : i3));
            }
            __syncthreads();
        }
    }
    __syncthreads();
    return ok;
}

// back substitution L^T y = z with one warp (z = row n of the augmented factor; 1 / L_jj in S.invd)
__device__ void chol_backsolve_warp(const SingleSmem& S, int n, double* y) {
    const double* A = S.u.L;
    const int lane = threadIdx.x & 31;
    double z[3];
#pragma unroll
    for (int q = 0; q < 3; ++q) { const int i = lane + 32 * q; z[q] = i < n ? A[tri(n, i)] : 0.0; }
    for (int j = n - 1; j >= 0; --j) {
        const int owner = j & 31, sl = j >> 5;
        double zj = sl == 0 ? z[0] : (sl == 1 ? z[1] : z[2]);
        zj = __shfl_sync(0xffffffffu, zj, owner);
        const double yj = zj * S.invd[j];
        if (lane == owner) y[j] = yj;
        const double* rj = A + tri(j, 0);
#pragma unroll
        for (int q = 0; q < 3; ++q) { const int i = lane + 32 * q; if (i < j) z[q] -= rj[i] * yj; }
    }
}

template <int T>
__global__ void __launch_bounds__(T, 2) fit_single_kernel(const FitConst* __restrict__ C, int B, const b3d_kp* __restrict__ kps_all,
                                                          const int32_t* __restrict__ nkp_all, const double* __restrict__ R0_all,
                                                          double* __restrict__ sim3, double* __restrict__ w, double* __restrict__ jointAA,
                                                          uint8_t* __restrict__ usable, b3d_iter_record* __restrict__ trace, int trace_stride,
                                                          int32_t* __restrict__ niters, int32_t* __restrict__ term) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SingleSmem& S = *reinterpret_cast<SingleSmem*>(smem_raw);
    const int tid = threadIdx.x, b = blockIdx.x;
    if (b >= B) return;
    const int n = C->n;
    const b3d_kp* kps = kps_all + (size_t)b * MAXKP;
    const int nkp = min(max(nkp_all[b], 0), MAXKP);
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

    // evaluate at x with Jacobian and apply the Jacobi scaling.  Returns false on non-finite cost.
    auto eval_grad_jac = [&](bool first) -> bool {
        x_cost = evaluate<T, 1>(C, S.xfull, kps, nkp, S, first);
        if (!isfinite(x_cost)) return false;
        it.cost = x_cost;
        if (first) {
            for (int k = tid; k < n; k += T) S.scale[k] = 1.0 / (1.0 + sqrt(S.Hs[tri(k, k)]));
            __syncthreads();
            for (int i = tid >> 5; i < n; i += T / 32) {
                const double si = S.scale[i];
                for (int j = tid & 31; j <= i; j += 32) S.Hs[tri(i, j)] *= si * S.scale[j];
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

    PhaseTimer pt;
    pt.start();
    const long long t_begin = clock64();
    int n_ls_evals = 0;
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
            pt.mark(0);
            if (!reuse_diagonal)
                for (int k = tid; k < n; k += T) S.diag[k] = fmin(fmax(S.Hs[tri(k, k)], kMinDiagonal), kMaxDiagonal);
            __syncthreads();
            double* A = S.u.L;
            for (int i = tid >> 5; i < n; i += T / 32)
                for (int j = tid & 31; j <= i; j += 32) {
                    double v = S.Hs[tri(i, j)];
                    if (i == j) { const double l = sqrt(S.diag[i] / radius); v += l * l; }
                    A[tri(i, j)] = v;
                }
            for (int k = tid; k < n; k += T) A[tri(n, k)] = S.gs[k];
            __syncthreads();
            reuse_diagonal = true;
            bool step_valid = false;
            pt.mark(1);
            const bool chol_ok = chol_factor_aug<T>(S, n);
            pt.mark(2);
            if (chol_ok) {
                if (tid < 32) chol_backsolve_warp(S, n, S.y);
                __syncthreads();
                pt.mark(3);
                // model_cost_change = -(J step)^T (r + J step / 2), step = -y:  y^T gs - y^T H_s y / 2
                double v = 0.0, fin = 0.0;
                for (int i = tid >> 5; i < n; i += T / 32) {
                    const double yi = S.y[i];
                    for (int j = tid & 31; j <= i; j += 32) v -= (i == j ? 0.5 : 1.0) * S.Hs[tri(i, j)] * yi * S.y[j];
                }
                for (int k = tid; k < n; k += T) { v += S.y[k] * S.gs[k]; fin += isfinite(S.y[k]) ? 0.0 : 1.0; }
                mcc = block_sum<T>(v, S.red);
                const double bad = block_sum<T>(fin, S.red);
                step_valid = (bad == 0.0) && (mcc > 0.0);
                pt.mark(4);
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
            pt.mark(5);
            cand_cost = evaluate<T, 0>(C, S.cfull, kps, nkp, S, false);
            pt.mark(6);
            {
                double v = 0.0;
                for (int k = tid; k < n; k += T) v += S.g[k] * S.delta[k];
                const double gd = block_sum<T>(v, S.red);
                const bool valid1 = isfinite(cand_cost);
                if (!valid1 || cand_cost > x_cost + kSufficientDecrease * gd * 1.0) {
                    // rare path: interpolate.  Needs the directional derivative at every trial point: gradient-only evaluations
                    // (MODE 2) that leave the normal equations of x in place.
                    ++n_ls_evals;
                    double dmax = 0.0;
                    for (int k = 0; k < n; ++k) dmax = fmax(dmax, fabs(S.delta[k]));
                    LsSample init{0.0, x_cost, gd, true}, prev{0, 0, 0, false}, cur{1.0, 0, 0, false};
                    auto ls_eval = [&](double a, LsSample& s) {
                        s.x = a; s.valid = false;
                        make_candidate(a);
                        const double c = evaluate<T, 2>(C, S.cfull, kps, nkp, S, false);
                        double vv = 0.0;
                        for (int k = tid; k < n; k += T) vv += S.delta[k] * S.gc[k];
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
                        if (tid < 32) {   // one warp: the quartic's roots are found with a lane per root
                            double ss;
                            if (!cur.valid) ss = fmin(fmax(cur.x * 0.5, lo), hi);
                            else {
                                LsSample smp[3] = {init, cur, prev};
                                ss = minimize_interpolating_polynomial(smp, prev.valid ? 3 : 2, lo, hi);
                            }
                            if (tid == 0) S.ls_step = ss;
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
                    cand_cost = evaluate<T, 0>(C, S.cfull, kps, nkp, S, false);
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
                pt.mark(7);
                if (!eval_grad_jac(false)) { termination = B3D_TERM_EVAL_FAILURE; break; }
                pt.mark(8);
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
    if (tid == 0 && g_phase_on && b < 1024) { g_block_cycles[b] = clock64() - t_begin; g_block_ls[b] = n_ls_evals; }
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
        fc->gmm_mean = gmm->d_mean; fc->gmm_L = gmm->d_prec_cho; fc->gmm_L70 = gmm->d_prec_cho70; fc->gmm_LT = gmm->d_prec_choT; fc->gmm_P = gmm->d_prec; fc->gmm_consts = gmm->d_consts;
    }
    return B3D_OK;
}

}  // namespace b3d

using namespace b3d;

extern "C" {

// debug hooks (not part of include/b3d.h): per-phase cycle counters of block 0
int b3d_debug_phase_timing(int on) {
    long long z[32] = {0};
    cudaMemcpyToSymbol(g_phase_cycles, z, sizeof(z));
    return cudaMemcpyToSymbol(g_phase_on, &on, sizeof(int)) == cudaSuccess ? 0 : 2;
}
int b3d_debug_block_read(long long* cyc1024, int* ls1024) {
    cudaMemcpyFromSymbol(ls1024, g_block_ls, 1024 * sizeof(int));
    return cudaMemcpyFromSymbol(cyc1024, g_block_cycles, 1024 * sizeof(long long)) == cudaSuccess ? 0 : 2;
}
int b3d_debug_phase_read(long long* out32) { return cudaMemcpyFromSymbol(out32, g_phase_cycles, 32 * sizeof(long long)) == cudaSuccess ? 0 : 2; }

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
    B3D_CUDA(cudaMallocAsync(&dfc, sizeof(FitConst), st));
    // pageable source: the copy is staged before the call returns, so the stack object may go out of scope
    B3D_CUDA(cudaMemcpyAsync(dfc, &fc, sizeof(FitConst), cudaMemcpyHostToDevice, st));
    // per device and context, cheap: set on every call (a process-wide flag would miss a second GPU)
    B3D_CUDA(cudaFuncSetAttribute(fit_single_kernel<FS_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SingleSmem)));
    fit_single_kernel<FS_THREADS><<<B, FS_THREADS, sizeof(SingleSmem), st>>>(dfc, B, kps, nkp, R0, sim3, w, jointAA, usable, trace, trace_stride,
                                                                            niters, term);
    B3D_LAUNCH_CHECK();
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
    if (int rc = validate_keypoints(kps, nkp, nb, "b3d_fit_single")) return rc;
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
