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
// The kernel is latency-bound (one frame's iterations are a serial chain), so the layout aims at few barriers, short dependent
// chains and little shared-memory traffic.  A thread owns one 4 x 4 tile (ti, tj) of the lower triangle of the (<= 88 x 88)
// system for the whole solve: the reprojection Jacobian is kept dense ([34][88], lm_device.cuh eval_reproj_dense), so J^T J is a
// SYRK over register tiles; the Cholesky is right-looking with the matrix held IN REGISTERS (one tile per thread): per 4-column
// panel the owner of the diagonal tile factors it, the tile column solves against it and publishes its rows ([4][88] doubles), and
// every trailing tile subtracts the rank-4 product of two 4 x 4 row groups — two barriers per panel, no read-modify-write of the
// matrix in shared memory.  The scaled normal equations stay in shared memory (packed) across rejected steps.
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
__device__ long long g_block_lscyc[1024];     // cycles inside line-search fall-backs, trial evaluations (debug)
__device__ int g_block_lstrials[1024];
__device__ long long g_block_lsparts[1024][4];  // set-up, trial evaluations, polynomial minimisation, final evaluation
__device__ long long g_block_t0[1024], g_block_t1[1024];   // globaltimer (ns) at the first start and the last end of a frame (debug)
__device__ __forceinline__ long long global_ns() { long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
__device__ int g_phase_on;
// Cycles accumulate in shared memory (a global read-modify-write per mark would stall the timing thread for an L2 round trip and,
// through the next barrier, the whole block); the kernel adds them to g_phase_cycles once at its end.
__device__ __forceinline__ long long* phase_acc() {
    __shared__ long long s_phase[32];
    return s_phase;
}
struct PhaseTimer {
    long long t;
    bool on;
    __device__ void start() { on = g_phase_on != 0 && blockIdx.x == 0 && threadIdx.x == 0; if (on) t = clock64(); }
    __device__ void mark(int i) { if (on) { const long long u = clock64(); phase_acc()[i] += u - t; t = u; } }
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
    double L[LTRI];               // Cholesky factor (strict lower part) + the forward-substituted right-hand side as last row
};

constexpr int NTILE = 22;                  // 4-wide tile rows / columns of the augmented system (87 rows -> 22, 86 columns -> 22)
static_assert(NFULL + 1 <= 4 * NTILE && NTILE * (NTILE + 1) / 2 <= FS_THREADS, "one 4 x 4 tile of the lower triangle per thread");

struct SingleSmem {
    EvalDense out;
    SingleUnion u;
    alignas(16) double2 pan[4][2][NTILE];  // the current panel's solved rows (Cholesky): [panel column][rows 0-1 | rows 2-3 of a tile][tile row]
    double dk[2][4][4];                    // diagonal tiles (by panel parity): strict lower part = L, diagonal = 1 / L_jj
    double damp[NFULL];                    // D^2 / radius of the current step
    int32_t chol_fail;
    double Hs[HTRI];                       // Jacobi-scaled J^T J (free layout, packed lower), kept across rejected steps
    double xfull[NFULL], cfull[NFULL];     // current / candidate parameters, 86-layout
    double g[NFULL], gs[NFULL], gc[NFULL]; // gradient (free layout): unscaled, Jacobi-scaled, saved copy during a line search
    double scale[NFULL], diag[NFULL], y[NFULL], delta[NFULL], invd[NFULL];
    double gpri[NPRI];                     // pose-prior gradient, pose-dim layout
    double R0[9];
    double red[40];
    double ls_step;
    double dfull[NFULL];                   // line search: the step in the 86-layout (zero for held-constant parameters)
    double lsq[GMM_MAXC][3];               // line search: |h_i|^2, h_i . u_i, |u_i|^2 with h_i = L_i^T (x - mu_i), u_i = L_i^T step
    double lsdphi[MAXKP];
    double hc[NPRI], hc_sq;                // h of the component chosen at the last cost-only evaluation (reused when that point is accepted)
    double xref[GMM_MAXC][NPRI];           // per component: the pose at which |h_i| was last computed exactly, and that norm
    double href[GMM_MAXC];
    int32_t gskip[GMM_MAXC];               // component skipped at this evaluation (it cannot be the argmin)
    int32_t ref_valid;
    int32_t comp;
    int32_t fidx[NFULL + 2];               // free -> full column (copy of FitConst::free_idx: read on every pass over the system)
};

// ---- GMM pose prior (GaussianMixture::residual, external/avatar/GaussianMixture.cpp:94-113) --------------------
// Returns 0.5 |beta r|^2 of the prior block and sets S.comp; with JAC also S.gpri = J^T r (pose-dim layout), where
// J = beta L_c^T WITHOUT the sqrt(1/2) the residual carries (reference quirk, Sim3BA.h:298-299).
__device__ int g_gmm_screen = 1;      // debug hook b3d_debug_gmm_screen: 0 evaluates every component at every point

// G.h[i] = L_i^T (v - mu_i) for every component (sub_mean) or L_i^T v (the direction of a line search); ends with a barrier.
// With `screen` (residual evaluations): component i is skipped (S.gskip[i] = 1) when it provably cannot be the argmin of
// p_i = |h_i|^2 / 2 - consts_i.  With h_i known exactly at an earlier pose xref_i, |h_i(v)| >= |h_i(xref_i)| - sigma_i |v - xref_i|
// (sigma_i >= ||L_i||_2) bounds p_i(v) from below, and the same inequality the other way bounds p_c(v) of the previous argmin c from
// above: if the lower bound of i exceeds the upper bound of c (with slack for rounding), i loses whatever its exact value is.  The
// chosen component is always evaluated exactly by the same code, so results are bit-identical to evaluating all of them.  Late
// iterations of a long solve take tiny steps: there seven of the eight 38 KB matrices of the SMPLify prior stay untouched.
template <int T>
__device__ void gmm_LTv(const FitConst* __restrict__ C, const double* v /*69*/, bool sub_mean, SingleSmem& S, bool screen = false) {
    const int tid = threadIdx.x, nc = C->gmm_ncomp;
    GmmTmp& G = S.u.gm;
    PhaseTimer pg;
    pg.start();
    for (int t = tid; t < nc * NPRI; t += T) G.d[t / NPRI][t % NPRI] = v[t % NPRI] - (sub_mean ? C->gmm_mean[t] : 0.0);
    __syncthreads();
    pg.mark(16);
    // h_i[k] = sum_{m >= k} L_i[m][k] d_i[m].  Warp i takes component i; lane q < 17 owns the adjacent column pairs (2q, 2q+1) and
    // (66-2q, 67-2q): 72 sixteen-byte loads whatever q (rows are padded to 70 doubles; entries above the diagonal are stored
    // zeros).  All lanes walk the SAME row in the same instruction, so a warp load touches 3 lines instead of one line per lane
    // (ncu: the per-lane row order made every load a 17-line gather); rows above a lane's columns are predicated off.
    constexpr int QT = 17, LD = NPRI + 1, RS = LD / 2;           // row stride in double2
    static_assert(T / 32 >= 8, "one warp per GMM component");
    const bool do_screen = screen && g_gmm_screen != 0;
    const int cprev = S.comp;                 // read before anything of this evaluation writes it
    const bool have_ref = do_screen && S.ref_valid != 0 && cprev >= 0 && cprev < nc;
    for (int i = tid >> 5; i < nc; i += T / 32) {
        const int q = tid & 31;
        if (do_screen) {
            bool skip = false;
            if (have_ref && i != cprev) {
                double di = 0.0, dc = 0.0;
                for (int k = q; k < NPRI; k += 32) { const double a = v[k] - S.xref[i][k], b2 = v[k] - S.xref[cprev][k]; di += a * a; dc += b2 * b2; }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) { di += __shfl_xor_sync(0xffffffffu, di, o); dc += __shfl_xor_sync(0xffffffffu, dc, o); }
                const double hu = S.href[cprev] + C->gmm_sigma[cprev] * sqrt(dc), hl = S.href[i] - C->gmm_sigma[i] * sqrt(di);
                const double ub = 0.5 * hu * hu - C->gmm_consts[cprev], lb = (hl > 0.0 ? 0.5 * hl * hl : 0.0) - C->gmm_consts[i];
                skip = lb > ub + 1e-9 * (fabs(ub) + fabs(lb)) + 1e-12;
            }
            if (q == 0) S.gskip[i] = skip ? 1 : 0;
            if (skip) continue;                                   // warp-uniform
            for (int k = q; k < NPRI; k += 32) S.xref[i][k] = v[k];   // this component is evaluated exactly at v
        }
        if (q < QT) {
            const int ca = 2 * q, cb = 66 - 2 * q;
            const double* __restrict__ L = C->gmm_L70 + (size_t)i * NPRI * LD;
            const double* dd = G.d[i];
            const double2* __restrict__ La = reinterpret_cast<const double2*>(L + ca);
            const double2* __restrict__ Lb = reinterpret_cast<const double2*>(L + cb);
            double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0, b0 = 0.0, b1 = 0.0, b2 = 0.0, b3 = 0.0;
            const double2 zero2 = make_double2(0.0, 0.0);
            // rows 0 .. 31 can only touch pair A (cb >= 34); rows 32 .. 68 may touch both.  Sixteen 16-byte loads are requested
            // before the first is used: the matrices stream from L2 (~800 cycles away), and the bytes in flight per warp are what
            // paces this loop (8 loads per batch: 13 k cycles per evaluation of the 8 components; the accumulation order is kept).
            for (int m = 0; m < 32; m += 16) {
                double2 v[16];
#pragma unroll
                for (int u = 0; u < 16; ++u) v[u] = (m + u >= ca) ? __ldg(La + (m + u) * RS) : zero2;
#pragma unroll
                for (int u = 0; u < 16; u += 2) {
                    a0 += v[u].x * dd[m + u]; a1 += v[u].y * dd[m + u];
                    a2 += v[u + 1].x * dd[m + u + 1]; a3 += v[u + 1].y * dd[m + u + 1];
                }
            }
            for (int m = 32; m < NPRI; m += 8) {
                double2 v[8], w4[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int r = m + u;
                    v[u] = (r < NPRI && r >= ca) ? __ldg(La + r * RS) : zero2;
                    w4[u] = (r < NPRI && r >= cb) ? __ldg(Lb + r * RS) : zero2;
                }
#pragma unroll
                for (int u = 0; u < 8; u += 2) {
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
}

// `reuse`: xj is the point of the previous (cost-only) evaluation, whose component and h_c are still in S.comp / S.hc.
template <int T, bool JAC>
__device__ double gmm_eval(const FitConst* __restrict__ C, const double* xj /*69*/, SingleSmem& S, bool reuse) {
    const int tid = threadIdx.x, nc = C->gmm_ncomp;
    GmmTmp& G = S.u.gm;
    PhaseTimer pg;
    if (!reuse) {
        gmm_LTv<T>(C, xj, true, S, true);
        pg.start();
        if (tid < nc) {
            if (g_gmm_screen != 0 && S.gskip[tid]) {
                G.sq[tid] = DBL_MAX; G.p[tid] = DBL_MAX;
            } else {
                const double rh = sqrt(0.5);
                double sq = 0.0, nn = 0.0;
                for (int k = 0; k < NPRI; ++k) { const double h0 = G.h[tid][k], h = h0 * rh; sq += h * h; nn += h0 * h0; }
                G.sq[tid] = sq;
                G.p[tid] = sq - C->gmm_consts[tid];
                S.href[tid] = sqrt(nn);
            }
        }
        __syncthreads();
        int cc = 0;
        double best = DBL_MAX;
        for (int i = 0; i < nc; ++i) { const double p = G.p[i]; if (p < best) { best = p; cc = i; } }   // strict '<': first wins ties
        for (int k = tid; k < NPRI; k += T) S.hc[k] = G.h[cc][k];
        if (tid == 0) { S.comp = cc; S.hc_sq = G.sq[cc]; S.ref_valid = 1; }
        __syncthreads();
    } else {
        pg.start();
    }
    const int c = S.comp;
    pg.mark(18);
    // residual = beta [h_c ; sqrt(-consts_log_c)]  =>  cost = 0.5 beta^2 (|h_c|^2 - consts_log_c)
    const double cost = 0.5 * C->beta_pose * C->beta_pose * (S.hc_sq + (-C->gmm_consts[c]));
    if (JAC) {
        // J^T r = beta^2 sqrt(1/2) L_c h: row m split over three threads
        const double* __restrict__ LT = C->gmm_LT + (size_t)c * NPRI * NPRI;
        if (tid < 3 * NPRI) {
            const int part = tid / NPRI, m = tid % NPRI;
            double a0 = 0.0, a1 = 0.0;
            int k = part;
            for (; k + 3 <= m; k += 6) { a0 += __ldg(LT + k * NPRI + m) * S.hc[k]; a1 += __ldg(LT + (k + 3) * NPRI + m) * S.hc[k + 3]; }
            for (; k <= m; k += 3) a0 += __ldg(LT + k * NPRI + m) * S.hc[k];
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
// (ti, tj): this thread's 4 x 4 tile of the lower triangle (tile_ok: it exists for this problem size).
template <int T, int MODE>
__device__ __noinline__ double evaluate(const FitConst* __restrict__ C, const double* xf, const b3d_kp* __restrict__ kps, int nkp, SingleSmem& S,
                           bool first, int ti, int tj, bool tile_ok, bool reuse_gmm = false) {
    constexpr bool JAC = MODE != 0;
    const int tid = threadIdx.x, n = C->n;
    PhaseTimer pe;
    pe.start();
    eval_reproj_dense<T, JAC>(C, xf, kps, nkp, S.R0, S.u.ev, S.out);
    pe.mark(JAC ? 9 : 12);
    double cost = 0.0;
    for (int k = 0; k < nkp; ++k) cost += S.out.kcost[k];
    double pcost = 0.0;
    const bool pose_prior = C->beta_pose > 0.0;
    const bool gmm = pose_prior && C->use_gmm;
    if (gmm) pcost = gmm_eval<T, JAC>(C, xf + IDX_JAA, S, reuse_gmm);   // aliases S.u.ev: the evaluation temporaries are dead
    // the remaining prior costs are sums of squares over <= 79 terms: fixed-order block sum
    double v = 0.0;
    if (pose_prior && !gmm)
        for (int k = tid; k < NPRI; k += T) { const double r = xf[IDX_JAA + k] * C->beta_pose; v += 0.5 * r * r; }
    if (C->has_w)
        for (int k = tid; k < NS; k += T) { const double r = xf[IDX_W + k] * C->beta_shape; v += 0.5 * r * r; }
    cost += pcost + block_sum<T>(v, S.red);
    pe.mark(JAC ? 10 : 13);
    if (!JAC) return cost;

    const double bp2 = C->beta_pose * C->beta_pose, bs2 = C->beta_shape * C->beta_shape;
    const int nrow = 2 * nkp;
    // ---- gradient J^T r: thread i sums column i of the dense Jacobian in row (= keypoint) order, then the priors ----
    {
        double* gdst = MODE == 2 ? S.gc : S.g;
        for (int i = tid; i < n; i += T) {
            const int a = S.fidx[i];
            double gacc = 0.0;
            for (int r = 0; r < nrow; r += 2)
                gacc += S.out.J[r][jpos(i)] * S.out.res[r >> 1][0] + S.out.J[r + 1][jpos(i)] * S.out.res[r >> 1][1];
            if (pose_prior && a >= IDX_JAA && a < IDX_W) gacc += gmm ? S.gpri[a - IDX_JAA] : bp2 * xf[a];
            if (C->has_w && a >= IDX_W) gacc += bs2 * xf[a];
            gdst[i] = gacc;
        }
    }
    if (MODE == 2) {
        __syncthreads();
        return cost;
    }
    // ---- J^T J: SYRK over the thread's 4 x 4 register tile, then the prior term and the Jacobi scaling, one store per element.
    // The 16 entries of the GMM precision matrix (L2) are requested first, so that their latency hides behind the SYRK. ----
    if (tile_ok) {
        double pri[4][4];
        int fa[4], fb[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) { fa[a] = S.fidx[4 * ti + a]; fb[a] = S.fidx[4 * tj + a]; }   // NFULL + 1 beyond the last column
        {
            const double* __restrict__ P = gmm ? C->gmm_P + (size_t)S.comp * NPRI * NPRI : nullptr;
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    const bool in_pose = fa[a] >= IDX_JAA && fa[a] < IDX_W && fb[b] >= IDX_JAA && fb[b] < IDX_W;
                    double v2 = 0.0;
                    if (pose_prior && in_pose) {
                        if (gmm) v2 = bp2 * __ldg(P + (fa[a] - IDX_JAA) * NPRI + (fb[b] - IDX_JAA));
                        else if (fa[a] == fb[b]) v2 = bp2;
                    }
                    if (C->has_w && fa[a] >= IDX_W && fa[a] <= NFULL - 1 && fa[a] == fb[b]) v2 += bs2;
                    pri[a][b] = v2;
                }
        }
        double acc[4][4];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
        const double2* Ja = reinterpret_cast<const double2*>(&S.out.J[0][0]) + ti;     // row layout: lm_device.cuh jpos()
        const double2* Jb = reinterpret_cast<const double2*>(&S.out.J[0][0]) + tj;
#pragma unroll 2
        for (int r = 0; r < nrow; ++r) {
            const double2 x01 = Ja[r * (LDJ / 2)], x23 = Ja[r * (LDJ / 2) + LDJ / 4];
            const double2 y01 = Jb[r * (LDJ / 2)], y23 = Jb[r * (LDJ / 2) + LDJ / 4];
            const double xv[4] = {x01.x, x01.y, x23.x, x23.y}, yv[4] = {y01.x, y01.y, y23.x, y23.y};
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) acc[a][b] = fma(xv[a], yv[b], acc[a][b]);
        }
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            const int i = 4 * ti + a;
            const double si = (first || i >= n) ? 1.0 : S.scale[i];
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const int j = 4 * tj + b;
                if (i < n && j <= i) {
                    const double v2 = acc[a][b] + pri[a][b];
                    S.Hs[tri(i, j)] = first ? v2 : v2 * si * S.scale[j];
                }
            }
        }
    }
    __syncthreads();
    pe.mark(11);
    return cost;
}

// Cholesky of the damped system (H_s + D^2 / radius) with the right-hand side gs as an extra row, the matrix held in registers:
// thread (ti, tj) owns rows 4 ti .. 4 ti + 3, columns 4 tj .. 4 tj + 3 of the lower triangle (row n = right-hand side, rows > n
// and columns >= n are zero padding).  Right-looking over 4-column panels; per panel
//   factor   the owner of the diagonal tile factors it in registers and publishes it (S.dk: L below the diagonal, 1 / L_jj on it);
//   solve    the tiles below it solve their rows against it and publish them as S.pan[column][row];
//   update   every tile to the right subtracts (its rows of the panel) x (its columns' rows of the panel)^T.
// The factorisation of diagonal tile k + 1 is folded into the update phase of panel k: two barriers per panel.
// On return (true) S.u.L holds the strict lower part of L and, as row n, z = L^-1 gs; 1 / L_jj is in S.invd.  Returns false when a
// pivot is not positive (Eigen LLT / oracle criterion).
template <int T>
__device__ bool chol_tiles(SingleSmem& S, int n, double radius, int ti, int tj, bool tile_ok) {
    const int ntc = (n + 3) >> 2;                         // panels
    PhaseTimer pc;
    pc.start();
    double a[4][4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int i = 4 * ti + r;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int j = 4 * tj + c;
            double v = 0.0;
            if (tile_ok && j < n) {
                if (i < n) {
                    if (j <= i) {
                        v = S.Hs[tri(i, j)];
                        if (i == j) v += S.damp[i];
                    }
                } else if (i == n) {
                    v = S.gs[j];
                }
            }
            a[r][c] = v;
        }
    }
    if (threadIdx.x == 0) S.chol_fail = 0;
    __syncthreads();
    pc.mark(20);
    // factor the thread's own (diagonal) tile in registers and publish it; padding columns (>= n) have no pivot
    auto factor_own = [&](int k) {
        // branch-free: the pivot chain is a serial stretch of ONE lane, paced by its instruction count.  A non-positive (or NaN)
        // pivot only raises the flag; what the tile then holds is never used.  Padding columns (>= n) get 1 / L_jj = 0.
        bool bad = false;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int j = 4 * k + c;
            const bool exists = j < n;
            const double piv = a[c][c];
            bad = bad || (exists && !(piv > 0.0));
            const double inv = exists ? rsqrt(piv) : 0.0;
            a[c][c] = inv;
#pragma unroll
            for (int r = c + 1; r < 4; ++r) a[r][c] *= inv;
#pragma unroll
            for (int c2 = c + 1; c2 < 4; ++c2)
#pragma unroll
                for (int r = c2; r < 4; ++r) a[r][c2] -= a[r][c] * a[c2][c];
            if (exists) S.invd[j] = inv;
        }
        if (bad) S.chol_fail = 1;
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int c = 0; c <= r; ++c) S.dk[k & 1][r][c] = a[r][c];
    };
    if (tile_ok && ti == 0 && tj == 0) factor_own(0);
    __syncthreads();
    pc.mark(21);
    for (int k = 0; k < ntc; ++k) {
        if (S.chol_fail) break;                           // block-uniform (written before the barrier every thread just passed)
        // ---- solve the tile column against the diagonal tile ----
        if (tile_ok && tj == k && ti > k) {
            const double (*D)[4] = S.dk[k & 1];
            const double d00 = D[0][0], d10 = D[1][0], d11 = D[1][1], d20 = D[2][0], d21 = D[2][1], d22 = D[2][2],
                         d30 = D[3][0], d31 = D[3][1], d32 = D[3][2], d33 = D[3][3];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const double x0 = a[r][0] * d00;
                const double x1 = (a[r][1] - x0 * d10) * d11;
                const double x2 = (a[r][2] - x0 * d20 - x1 * d21) * d22;
                const double x3 = (a[r][3] - x0 * d30 - x1 * d31 - x2 * d32) * d33;
                a[r][0] = x0; a[r][1] = x1; a[r][2] = x2; a[r][3] = x3;
            }
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                S.pan[c][0][ti] = make_double2(a[0][c], a[1][c]);
                S.pan[c][1][ti] = make_double2(a[2][c], a[3][c]);
            }
        }
        // Look-ahead: tile (k + 1, k) and the next diagonal tile (k + 1, k + 1) are neighbouring lanes of one warp (their linear
        // indices differ by one and never straddle a multiple of 32 for <= 22 tile rows), so a warp barrier is enough for the
        // diagonal owner to see the rows just solved: it applies panel k to its tile and factors it while the rest of the column
        // is still being solved, and the pivot chain of panel k + 1 leaves the block-wide critical path.
        __syncwarp();
        const bool next_diag = tile_ok && ti == k + 1 && tj == k + 1 && k + 1 < ntc;
        if (next_diag) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const double2 r01 = S.pan[q][0][ti], r23 = S.pan[q][1][ti];
                const double rv[4] = {r01.x, r01.y, r23.x, r23.y};
#pragma unroll
                for (int r = 0; r < 4; ++r)
#pragma unroll
                    for (int c = 0; c <= r; ++c) a[r][c] = fma(-rv[r], rv[c], a[r][c]);
            }
            factor_own(k + 1);
        }
        __syncthreads();
        pc.mark(22);
        // ---- rank-4 update of the trailing tiles ----
        if (tile_ok && tj > k && !next_diag) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const double2 r01 = S.pan[q][0][ti], r23 = S.pan[q][1][ti], c01 = S.pan[q][0][tj], c23 = S.pan[q][1][tj];
                const double rv[4] = {r01.x, r01.y, r23.x, r23.y}, cv[4] = {c01.x, c01.y, c23.x, c23.y};
#pragma unroll
                for (int r = 0; r < 4; ++r)
#pragma unroll
                    for (int c = 0; c < 4; ++c) a[r][c] = fma(-rv[r], cv[c], a[r][c]);
            }
        }
        __syncthreads();
        pc.mark(23);
    }
    const bool ok = S.chol_fail == 0;
    if (ok && tile_ok) {       // the factor's strict lower part and the substituted right-hand side row -> S.u.L for the back substitution
        double* A = S.u.L;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int i = 4 * ti + r;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int j = 4 * tj + c;
                if (j < n && i <= n && i > j) A[tri(i, j)] = a[r][c];
            }
        }
    }
    __syncthreads();
    pc.mark(24);
    return ok;
}

// back substitution L^T y = z with one warp (z = row n of the factor; 1 / L_jj in S.invd): lm_device.cuh backsolve_warp_blocked
__device__ void chol_backsolve_warp(const SingleSmem& S, int n, double* y) {
    const double* A = S.u.L;
    const int lane = threadIdx.x & 31;
    double z[3];
#pragma unroll
    for (int q = 0; q < 3; ++q) { const int i = lane + 32 * q; z[q] = i < n ? A[tri(n, i)] : 0.0; }
    // packed rows: reads past a row's end stay inside S.u.L for every row < n (tri(n - 1, 95) < LTRI)
    static_assert((NFULL - 1) * NFULL / 2 + 95 < LTRI, "row reads of the blocked back substitution stay inside the factor");
    backsolve_warp_blocked(n, [A](int r) { return A + tri(r, 0); }, S.invd, z, y);
}

// Scheduling of frames onto CTAs.  A frame's iterations are a serial chain (up to 101 of them, against a mean of 20), so a batch
// ends with its longest frames, and with one CTA per frame such a frame starts whenever the block scheduler reaches it — up to
// several milliseconds late once the batch exceeds the 2 CTAs per SM that fit.  Instead the kernel is persistent (one CTA per slot)
// and takes frames from a device-side queue in time slices: a frame runs a capped number of iterations; if it is not finished and
// another frame is waiting for a CTA, it checkpoints itself (what Ceres' minimiser carries between iterations; the normal equations
// are rebuilt by re-evaluating at x: same inputs, same bits) and goes to the back of the queue.  Fresh frames have priority.  All
// unfinished frames thus advance at the same rate until no more of them are left than there are CTAs, and from then on the long
// ones run undisturbed (a barrier between "passes" loses more than it gains: each pass ends with its own slowest frame).
struct FitCkpt {
    double xfull[NFULL], scale[NFULL], diag[NFULL];
    double radius, decrease_factor, x_norm, gmax;
    int32_t reuse_diagonal, num_invalid, iteration, nrec, n_ls, pad;
};
struct FitQueue {
    FitCkpt* ckpt;            // B entries, or null: no queue, frame = blockIdx.x, one launch slot per frame
    int32_t* ctl;             // [0] next fresh frame, [1] next queue slot to claim, [2] queue slots handed out, [3] frames finished
    int32_t* slots;           // queue of checkpointed frames, -1 = not written yet (B + grid entries)
    int32_t stop_at;          // a fresh frame checkpoints itself once this many iterations are done
};
enum { Q_FRESH = 0, Q_HEAD = 1, Q_TAIL = 2, Q_DONE = 3 };

template <int T>
__device__ void fit_frame(const FitConst* __restrict__ C, int B, const b3d_kp* __restrict__ kps_all, const int32_t* __restrict__ nkp_all,
                          const double* __restrict__ R0_all, double* __restrict__ sim3, double* __restrict__ w, double* __restrict__ jointAA,
                          uint8_t* __restrict__ usable, b3d_iter_record* __restrict__ trace, int trace_stride, int32_t* __restrict__ niters,
                          int32_t* __restrict__ term, const FitQueue& P, SingleSmem& S, int b, bool resumed);

// Persistent driver of fit_frame (or one frame per CTA when there is no queue).
template <int T>
__global__ void __launch_bounds__(T, 2) fit_single_kernel(const FitConst* __restrict__ C, int B, const b3d_kp* __restrict__ kps_all,
                                                          const int32_t* __restrict__ nkp_all, const double* __restrict__ R0_all,
                                                          double* __restrict__ sim3, double* __restrict__ w, double* __restrict__ jointAA,
                                                          uint8_t* __restrict__ usable, b3d_iter_record* __restrict__ trace, int trace_stride,
                                                          int32_t* __restrict__ niters, int32_t* __restrict__ term, const FitQueue P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SingleSmem& S = *reinterpret_cast<SingleSmem*>(smem_raw);
    __shared__ int s_item[2];
    for (;;) {
        if (!P.ckpt) {
            if (threadIdx.x == 0) { s_item[0] = (int)blockIdx.x < B ? (int)blockIdx.x : -1; s_item[1] = 0; }
        } else if (threadIdx.x == 0) {
            int item = -1, res = 0;
            if (*(volatile int32_t*)&P.ctl[Q_FRESH] < B) {
                const int f = atomicAdd(&P.ctl[Q_FRESH], 1);
                if (f < B) item = f;
            }
            if (item < 0) {   // no fresh frame left: claim the next queue slot and wait until it is filled or every frame is finished
                const int sl = atomicAdd(&P.ctl[Q_HEAD], 1);
                for (;;) {
                    const int v = *(volatile int32_t*)&P.slots[sl];
                    if (v >= 0) { __threadfence(); item = v; res = 1; break; }
                    if (*(volatile int32_t*)&P.ctl[Q_DONE] >= B) break;
                    __nanosleep(200);
                }
            }
            s_item[0] = item; s_item[1] = res;
        }
        __syncthreads();
        const int b = s_item[0];
        const bool resumed = s_item[1] != 0;
        __syncthreads();
        if (b < 0) return;
        fit_frame<T>(C, B, kps_all, nkp_all, R0_all, sim3, w, jointAA, usable, trace, trace_stride, niters, term, P, S, b, resumed);
        if (!P.ckpt) return;
        __syncthreads();
    }
}

template <int T>
__device__ void fit_frame(const FitConst* __restrict__ C, int B, const b3d_kp* __restrict__ kps_all, const int32_t* __restrict__ nkp_all,
                          const double* __restrict__ R0_all, double* __restrict__ sim3, double* __restrict__ w, double* __restrict__ jointAA,
                          uint8_t* __restrict__ usable, b3d_iter_record* __restrict__ trace, int trace_stride, int32_t* __restrict__ niters,
                          int32_t* __restrict__ term, const FitQueue& P, SingleSmem& S, int b, bool resumed) {
    const int tid = threadIdx.x;
    const int n = C->n;
    const b3d_kp* kps = kps_all + (size_t)b * MAXKP;
    const int nkp = min(max(nkp_all[b], 0), MAXKP);
    b3d_iter_record* tr = trace ? trace + (size_t)b * trace_stride : nullptr;
    // this thread's 4 x 4 tile of the lower triangle of the augmented system (rows 0 .. n, columns 0 .. n - 1)
    int ti = (int)((sqrtf(8.0f * (float)tid + 1.0f) - 1.0f) * 0.5f);
    while ((ti + 1) * (ti + 2) / 2 <= tid) ++ti;
    while (ti * (ti + 1) / 2 > tid) --ti;
    const int tj = tid - ti * (ti + 1) / 2;
    const bool tile_ok = ti < ((n + 4) >> 2) && tj < ((n + 3) >> 2);

    // ---- problem set-up (Sim3BA.h:395-404 / 558-569): sim3 from the caller, jointAA = 0, w from the caller ----
    for (int k = tid; k < NFULL + 2; k += T) S.fidx[k] = k < n ? C->free_idx[k] : NFULL + 1;
    for (int k = tid; k < NFULL; k += T) {
        double v = 0.0;
        if (k < 7) v = sim3[(size_t)b * 7 + k];
        else if (k >= IDX_W) v = w[(size_t)b * NS + (k - IDX_W)];
        S.xfull[k] = v;
    }
    if (tid < 9) S.R0[tid] = R0_all[(size_t)b * 9 + tid];
    if (tid == 0) { S.ref_valid = 0; S.comp = -1; }
    FitCkpt* ck = P.ckpt ? P.ckpt + b : nullptr;
    if (resumed) {
        for (int k = tid; k < NFULL; k += T) { S.xfull[k] = ck->xfull[k]; S.scale[k] = ck->scale[k]; S.diag[k] = ck->diag[k]; }
    }
    __syncthreads();
    // IterationZero: project onto the bounds (only the scale is bounded)
    if (tid == 0 && !resumed) S.xfull[IDX_S] = fmin(fmax(S.xfull[IDX_S], C->scale_lo), C->scale_hi);
    __syncthreads();

    double radius = kInitialRadius, decrease_factor = 2.0, x_cost = 0.0, cand_cost = 0.0, mcc = 0.0;
    bool reuse_diagonal = false, step_successful = true;
    int num_invalid = 0, iteration = 0, nrec = 0, termination = 0;
    b3d_iter_record it;
    it.cost = it.cost_change = it.gradient_max_norm = it.step_norm = it.relative_decrease = it.radius = 0.0;
    it.flags = 0; it.aux = -1;

    // evaluate at x with Jacobian and apply the Jacobi scaling.  Returns false on non-finite cost.
    auto eval_grad_jac = [&](bool first, bool reuse_gmm = false) -> bool {
        x_cost = evaluate<T, 1>(C, S.xfull, kps, nkp, S, first, ti, tj, tile_ok, reuse_gmm);
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

    if (tid < 32) phase_acc()[tid] = 0;
    __syncthreads();
    PhaseTimer pt;
    pt.start();
    const long long t_begin = clock64();
    if (tid == 0 && g_phase_on && b < 1024 && !resumed) g_block_t0[b] = global_ns();
    int n_ls_evals = 0, slice_end = P.stop_at;
    long long ls_cyc = 0, ls_part[4] = {0, 0, 0, 0};
    int ls_trials = 0;
    double x_norm;
    bool unfinished = false, skip_head = false, started;
    if (resumed) {
        radius = ck->radius; decrease_factor = ck->decrease_factor; x_norm = ck->x_norm;
        reuse_diagonal = ck->reuse_diagonal != 0; num_invalid = ck->num_invalid; iteration = ck->iteration; nrec = ck->nrec; n_ls_evals = ck->n_ls;
        slice_end = iteration + P.stop_at;
        started = eval_grad_jac(false);     // the normal equations, gradient and cost at x, bit for bit what the previous slice held
        it.gradient_max_norm = ck->gmax;    // (zero after a rejected step: what the next record of an invalid step would repeat)
        skip_head = true;                   // the checks at the head of this iteration were passed before the checkpoint
    } else {
        x_norm = x_norm_of(S.xfull);
        started = eval_grad_jac(true);
        it.flags = 3;
    }
    if (!started) {
        termination = B3D_TERM_EVAL_FAILURE;
    } else {
        for (;;) {
            if (!skip_head) {
                // ---- FinalizeIterationAndCheckIfMinimizerCanContinue ----
                push();
                if (iteration >= C->max_iters) { termination = B3D_TERM_MAX_ITERS; break; }
                if (step_successful && it.gradient_max_norm <= kGradientTolerance) { termination = B3D_TERM_GRAD_TOL; break; }
                if (!(radius > kMinRadius)) { termination = B3D_TERM_MIN_RADIUS; break; }
                if (ck && iteration >= slice_end) {
                    // end of this frame's time slice: yield (checkpoint, go to the back of the queue) if any other frame is waiting
                    // for a CTA, otherwise carry on.  One thread looks, so that the decision is the same for the whole block.
                    if (tid == 0)
                        S.chol_fail = (*(volatile int32_t*)&P.ctl[Q_FRESH] < B || *(volatile int32_t*)&P.ctl[Q_TAIL] > *(volatile int32_t*)&P.ctl[Q_HEAD]) ? 1 : 0;
                    __syncthreads();
                    const bool yield = S.chol_fail != 0;
                    __syncthreads();
                    if (yield) { unfinished = true; break; }
                    slice_end += P.stop_at;
                }
            }
            skip_head = false;
            const double prev_gmax = it.gradient_max_norm;
            it.cost = it.cost_change = it.gradient_max_norm = it.step_norm = it.relative_decrease = 0.0;
            it.flags = 0;
            ++iteration;
            step_successful = false;

            // ---- LevenbergMarquardtStrategy::ComputeStep ----
            pt.mark(0);
            for (int k = tid; k < n; k += T) {
                if (!reuse_diagonal) S.diag[k] = fmin(fmax(S.Hs[tri(k, k)], kMinDiagonal), kMaxDiagonal);
                const double l = sqrt(S.diag[k] / radius);
                S.damp[k] = l * l;
            }
            __syncthreads();
            reuse_diagonal = true;
            bool step_valid = false;
            pt.mark(1);
            const bool chol_ok = chol_tiles<T>(S, n, radius, ti, tj, tile_ok);
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
            cand_cost = evaluate<T, 0>(C, S.cfull, kps, nkp, S, false, ti, tj, tile_ok);
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
                    const long long ls_t0 = clock64();
                    double dmax = 0.0;
                    for (int k = 0; k < n; ++k) dmax = fmax(dmax, fabs(S.delta[k]));
                    LsSample init{0.0, x_cost, gd, true}, prev{0, 0, 0, false}, cur{1.0, 0, 0, false};
                    // Along the search line every prior is a quadratic in the step size a: the GMM components through
                    // h_i(a) = h_i + a u_i (so the argmin component and its cost need no matrix at a trial point), the L2 priors
                    // through three sums.  Only the reprojection blocks are evaluated per trial (eval_reproj_dir).
                    const bool pose_prior = C->beta_pose > 0.0, gmm = pose_prior && C->use_gmm;
                    for (int k = tid; k < NFULL; k += T) S.dfull[k] = 0.0;
                    __syncthreads();
                    for (int k = tid; k < n; k += T) S.dfull[S.fidx[k]] = S.delta[k];
                    __syncthreads();
                    double sxx = 0.0, sxd = 0.0, sdd = 0.0, pxx = 0.0, pxd = 0.0, pdd = 0.0;     // shape / L2 pose prior sums (every thread)
                    if (C->has_w)
                        for (int k = 0; k < NS; ++k) { const double xv = S.xfull[IDX_W + k], dv = S.dfull[IDX_W + k]; sxx += xv * xv; sxd += xv * dv; sdd += dv * dv; }
                    if (pose_prior && !gmm)
                        for (int k = 0; k < NPRI; ++k) { const double xv = S.xfull[IDX_JAA + k], dv = S.dfull[IDX_JAA + k]; pxx += xv * xv; pxd += xv * dv; pdd += dv * dv; }
                    if (gmm) {
                        const int nc = C->gmm_ncomp;
                        double* stash = &S.out.J[0][0];                       // the dense Jacobian is dead during a line search
                        gmm_LTv<T>(C, S.xfull + IDX_JAA, true, S);
                        for (int t = tid; t < nc * NPRI; t += T) stash[t] = S.u.gm.h[t / NPRI][t % NPRI];
                        __syncthreads();
                        gmm_LTv<T>(C, S.dfull + IDX_JAA, false, S);
                        for (int i = tid >> 5; i < nc; i += T / 32) {
                            double hh = 0.0, hu = 0.0, uu = 0.0;
                            for (int k = tid & 31; k < NPRI; k += 32) { const double h = stash[i * NPRI + k], u = S.u.gm.h[i][k]; hh += h * h; hu += h * u; uu += u * u; }
#pragma unroll
                            for (int o = 16; o > 0; o >>= 1) {
                                hh += __shfl_xor_sync(0xffffffffu, hh, o); hu += __shfl_xor_sync(0xffffffffu, hu, o); uu += __shfl_xor_sync(0xffffffffu, uu, o);
                            }
                            if ((tid & 31) == 0) { S.lsq[i][0] = hh; S.lsq[i][1] = hu; S.lsq[i][2] = uu; }
                        }
                        __syncthreads();
                    }
                    const double bp2 = C->beta_pose * C->beta_pose, bs2 = C->beta_shape * C->beta_shape;
                    auto ls_eval = [&](double a, LsSample& s) {
                        s.x = a; s.valid = false;
                        ++ls_trials;
                        make_candidate(a);
                        eval_reproj_dir<T>(C, S.cfull, S.dfull, kps, nkp, S.R0, S.u.ev, S.out.kcost, S.lsdphi);
                        double c = 0.0, gg = 0.0;
                        for (int k = 0; k < nkp; ++k) { c += S.out.kcost[k]; gg += S.lsdphi[k]; }
                        if (gmm) {
                            int best_c = 0;
                            double best = DBL_MAX, best_sq = 0.0;
                            for (int i = 0; i < C->gmm_ncomp; ++i) {
                                const double sq = 0.5 * (S.lsq[i][0] + a * (2.0 * S.lsq[i][1] + a * S.lsq[i][2]));
                                const double p = sq - C->gmm_consts[i];
                                if (p < best) { best = p; best_c = i; best_sq = sq; }
                            }
                            c += 0.5 * bp2 * (best_sq + (-C->gmm_consts[best_c]));
                            gg += bp2 * sqrt(0.5) * (S.lsq[best_c][1] + a * S.lsq[best_c][2]);      // the reference's gradient (no sqrt(1/2) in J)
                        } else if (pose_prior) {
                            c += 0.5 * bp2 * (pxx + a * (2.0 * pxd + a * pdd));
                            gg += bp2 * (pxd + a * pdd);
                        }
                        if (C->has_w) {
                            c += 0.5 * bs2 * (sxx + a * (2.0 * sxd + a * sdd));
                            gg += bs2 * (sxd + a * sdd);
                        }
                        if (!isfinite(c) || !isfinite(gg)) return;
                        s.value = c; s.gradient = gg; s.valid = true;
                    };
                    long long ls_t1 = clock64();
                    ls_part[0] += ls_t1 - ls_t0;
                    ls_eval(1.0, cur);
                    { const long long u = clock64(); ls_part[1] += u - ls_t1; ls_t1 = u; }
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
                        { const long long u = clock64(); ls_part[2] += u - ls_t1; ls_t1 = u; }
                        if (step_size * dmax < kMinLineSearchStepSize) { success = false; break; }
                        prev = cur;
                        ls_eval(step_size, cur);
                        { const long long u = clock64(); ls_part[1] += u - ls_t1; ls_t1 = u; }
                    }
                    if (success && cur.x != 1.0) {
                        for (int k = tid; k < n; k += T) S.delta[k] *= cur.x;
                        it.flags |= 4;
                    }
                    __syncthreads();
                    // ---- ComputeCandidatePointAndEvaluateCost ----
                    make_candidate(1.0);
                    cand_cost = evaluate<T, 0>(C, S.cfull, kps, nkp, S, false, ti, tj, tile_ok);
                    { const long long u = clock64(); ls_part[3] += u - ls_t1; ls_cyc += u - ls_t0; }
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
                if (!eval_grad_jac(false, true)) { termination = B3D_TERM_EVAL_FAILURE; break; }   // x is the point just evaluated: reuse its GMM component
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
    if (tid == 0 && g_phase_on && b < 1024) { g_block_cycles[b] = (resumed ? g_block_cycles[b] : 0) + (clock64() - t_begin); g_block_ls[b] = n_ls_evals; g_block_lscyc[b] = (resumed ? g_block_lscyc[b] : 0) + ls_cyc; g_block_lstrials[b] = (resumed ? g_block_lstrials[b] : 0) + ls_trials;
        for (int q = 0; q < 4; ++q) g_block_lsparts[b][q] = (resumed ? g_block_lsparts[b][q] : 0) + ls_part[q]; }
    if (b == 0 && tid < 32 && g_phase_on) g_phase_cycles[tid] += phase_acc()[tid];
    if (tid == 0 && g_phase_on && b < 1024) g_block_t1[b] = global_ns();
    if (unfinished) {
        for (int k = tid; k < NFULL; k += T) { ck->xfull[k] = S.xfull[k]; ck->scale[k] = S.scale[k]; ck->diag[k] = S.diag[k]; }
        if (tid == 0) {
            ck->radius = radius; ck->decrease_factor = decrease_factor; ck->x_norm = x_norm; ck->gmax = it.gradient_max_norm;
            ck->reuse_diagonal = reuse_diagonal ? 1 : 0; ck->num_invalid = num_invalid; ck->iteration = iteration; ck->nrec = nrec; ck->n_ls = n_ls_evals;
        }
        __syncthreads();
        if (tid == 0) {
            __threadfence();                                       // the checkpoint is visible before the slot names the frame
            const int sl = atomicAdd(&P.ctl[Q_TAIL], 1);
            *(volatile int32_t*)&P.slots[sl] = b;
        }
        return;
    }
    // ---- results: optimised blocks back to the caller's arrays (Sim3BA.h:481-507 reads them from the blocks) ----
    if (tid < 7) sim3[(size_t)b * 7 + tid] = S.xfull[tid];
    for (int k = tid; k < NJ * 3; k += T) jointAA[(size_t)b * NJ * 3 + k] = k < 3 ? 0.0 : S.xfull[IDX_JAA + k - 3];
    if (C->has_w && tid < NS) w[(size_t)b * NS + tid] = S.xfull[IDX_W + tid];
    if (tid == 0) {
        if (usable) usable[b] = (termination >= B3D_TERM_PARAM_TOL && termination <= B3D_TERM_MIN_RADIUS) ? 1 : 0;
        if (niters) niters[b] = nrec;
        if (term) term[b] = termination;
        if (P.ckpt) { __threadfence(); atomicAdd(&P.ctl[Q_DONE], 1); }
    }
}

static std::atomic<int> g_fit_stop_at{-1};

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
        for (int a = 0; a < NJ; ++a) fc->chain_pos[j][a] = -1;
        for (int i = 0; i < len; ++i) fc->chain_pos[j][fc->chain[j][i]] = (int8_t)i;
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
        fc->gmm_mean = gmm->d_mean; fc->gmm_L = gmm->d_prec_cho; fc->gmm_L70 = gmm->d_prec_cho70; fc->gmm_LT = gmm->d_prec_choT; fc->gmm_P = gmm->d_prec; fc->gmm_consts = gmm->d_consts; fc->gmm_sigma = gmm->d_sigma;
    }
    return B3D_OK;
}

}  // namespace b3d

using namespace b3d;

extern "C" {

// debug hook: iterations after which a fresh frame of the persistent kernel checkpoints itself (0: no queue, one CTA per frame;
// < 0: automatic)
int b3d_debug_fit_stop_at(int iters) { g_fit_stop_at.store(iters); return 0; }

// debug hook: 0 switches the GMM component screening off (every component evaluated at every point); results must not change
int b3d_debug_gmm_screen(int on) { return cudaMemcpyToSymbol(g_gmm_screen, &on, sizeof(int)) == cudaSuccess ? 0 : 2; }

// debug hooks (not part of include/b3d.h): per-phase cycle counters of block 0
int b3d_debug_phase_timing(int on) {
    long long z[32] = {0};
    cudaMemcpyToSymbol(g_phase_cycles, z, sizeof(z));
    return cudaMemcpyToSymbol(g_phase_on, &on, sizeof(int)) == cudaSuccess ? 0 : 2;
}
// debug / test hook: ceres' MinimizeInterpolatingPolynomial as the line search runs it (ls_poly.cuh), one warp per problem.
// in: n x 12 doubles (number of samples 2 | 3, then x, value, gradient of up to 3 samples, x_min, x_max); out: n step sizes
__global__ void poly_min_kernel(int n, const double* __restrict__ in, double* __restrict__ out) {
    const int w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (w >= n) return;
    const double* q = in + (size_t)w * 12;
    LsSample s[3];
    for (int i = 0; i < 3; ++i) { s[i].x = q[1 + 3 * i]; s[i].value = q[2 + 3 * i]; s[i].gradient = q[3 + 3 * i]; s[i].valid = true; }
    const double r = minimize_interpolating_polynomial(s, (int)q[0], q[10], q[11]);
    if ((threadIdx.x & 31) == 0) out[w] = r;
}
int b3d_debug_poly_min(int n, const double* in_host, double* out_host) {
    double *din = nullptr, *dout = nullptr;
    if (cudaMalloc(&din, sizeof(double) * 12 * (size_t)n) != cudaSuccess) return 2;
    if (cudaMalloc(&dout, sizeof(double) * (size_t)n) != cudaSuccess) { cudaFree(din); return 2; }
    cudaMemcpy(din, in_host, sizeof(double) * 12 * (size_t)n, cudaMemcpyHostToDevice);
    poly_min_kernel<<<(n + 3) / 4, 128>>>(n, din, dout);
    const cudaError_t e = cudaMemcpy(out_host, dout, sizeof(double) * (size_t)n, cudaMemcpyDeviceToHost);
    cudaFree(din); cudaFree(dout);
    return e == cudaSuccess ? 0 : 2;
}
int b3d_debug_poly_read(unsigned long long* out8) { return cudaMemcpyFromSymbol(out8, g_poly_cyc, 8 * sizeof(unsigned long long)) == cudaSuccess ? 0 : 2; }
int b3d_debug_block_lsparts_read(long long* parts4096) { return cudaMemcpyFromSymbol(parts4096, g_block_lsparts, 4096 * sizeof(long long)) == cudaSuccess ? 0 : 2; }
int b3d_debug_block_ls_read(long long* cyc1024, int* trials1024) {
    cudaMemcpyFromSymbol(trials1024, g_block_lstrials, 1024 * sizeof(int));
    return cudaMemcpyFromSymbol(cyc1024, g_block_lscyc, 1024 * sizeof(long long)) == cudaSuccess ? 0 : 2;
}
int b3d_debug_block_read(long long* cyc1024, int* ls1024) {
    cudaMemcpyFromSymbol(ls1024, g_block_ls, 1024 * sizeof(int));
    return cudaMemcpyFromSymbol(cyc1024, g_block_cycles, 1024 * sizeof(long long)) == cudaSuccess ? 0 : 2;
}
int b3d_debug_block_times(long long* t0_1024, long long* t1_1024) {
    cudaMemcpyFromSymbol(t0_1024, g_block_t0, 1024 * sizeof(long long));
    return cudaMemcpyFromSymbol(t1_1024, g_block_t1, 1024 * sizeof(long long)) == cudaSuccess ? 0 : 2;
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
    // Batches larger than the machine holds at once (2 CTAs per SM) go through the persistent kernel and its frame queue (struct
    // FitCkpt); smaller ones start every frame at once, one CTA each.
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int slots = 2 * sms;
    int stop_at = g_fit_stop_at.load();
    if (stop_at < 0) stop_at = 16;                         // automatic: a little under the mean iteration count of the bench frames (8 - 20 measure alike)
    FitQueue Q;
    std::memset(&Q, 0, sizeof(Q));
    if (B <= slots || stop_at == 0 || stop_at >= opts->max_iters) {
        fit_single_kernel<FS_THREADS><<<B, FS_THREADS, sizeof(SingleSmem), st>>>(dfc, B, kps, nkp, R0, sim3, w, jointAA, usable, trace, trace_stride,
                                                                                niters, term, Q);
        B3D_LAUNCH_CHECK();
    } else {
        char* scratch = nullptr;
        const size_t pushes = (size_t)B * (size_t)(opts->max_iters / stop_at + 2);      // a frame yields at most once per slice
        const size_t s_ck = (sizeof(FitCkpt) * (size_t)B + 255) / 256 * 256, s_q = (4 * (pushes + slots + 8) + 255) / 256 * 256;
        B3D_CUDA(cudaMallocAsync(&scratch, s_ck + s_q + 256, st));
        Q.ckpt = (FitCkpt*)scratch;
        Q.slots = (int32_t*)(scratch + s_ck);
        Q.ctl = (int32_t*)(scratch + s_ck + s_q);
        Q.stop_at = stop_at;
        B3D_CUDA(cudaMemsetAsync(Q.slots, 0xFF, s_q, st));
        B3D_CUDA(cudaMemsetAsync(Q.ctl, 0, 256, st));
        fit_single_kernel<FS_THREADS><<<slots, FS_THREADS, sizeof(SingleSmem), st>>>(dfc, B, kps, nkp, R0, sim3, w, jointAA, usable, trace,
                                                                                    trace_stride, niters, term, Q);
        B3D_LAUNCH_CHECK();
        B3D_CUDA(cudaFreeAsync(scratch, st));
    }
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
