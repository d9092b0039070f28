// Entry (2), multi-frame: one CTA runs the whole Levenberg-Marquardt solve of one window / anchor chain of F frames.
//
// Replaces the ceres::Problem built and solved by OptimizeMultiFrame (include/MultiFrameBA.h:33-177): per frame a 76-vector
// [scale, rootAA, rootT, jointAA[1..23]], one shared shape block w (10), per keypoint a ReprojCostShape block + Huber(3)
// (:71-103), per frame the L2 pose prior (:106-111; the GMM is never used here), the shape prior (:115-118) and, between
// consecutive frames, 25 Vec3DiffCost blocks lambda (a - b) on rootT, rootAA and every jointAA (:121-142; the scale is not
// smoothed).  No bounds, hence no line search.  The LM loop is the same Ceres 1.14 restatement as fit_single.cu.
//
// Where the reference hands the dense (m x n) Jacobian to DENSE_QR (m = 3,495, n = 1,530 for a 20-frame window: 1.6e10 flop
// per iteration, all but 0.2 % of it on structural zeros), the normal equations here keep their structure: block tridiagonal
// (76 x 76 blocks, off-diagonal blocks are the diagonal matrices -lambda^2 S_f E S_{f+1}) plus a 10-wide arrowhead for w.
// They are factored frame by frame by one augmented Crout elimination per frame: the rows of the next off-diagonal block, the
// 10 arrow rows and the right-hand side ride along below the 76 x 76 diagonal block, so that L_{f+1,f}, W_f and the forward
// substitution come out of the same column sweep; the 10 x 10 Schur complement of w closes the system.
#include <algorithm>
#include <atomic>
#include <vector>

#include "lm_device.cuh"
#include "writeback.cuh"

namespace b3d {

constexpr int FW_THREADS = 512;         // one 4 x 4 tile of a frame's 164 x 76 rows per thread in the sweep (446 tiles)
constexpr int NB = NPF;                 // 76: diagonal block size
constexpr int LDM = NB + 1;             // padded leading dimension of the shared-memory blocks (bank-conflict free columns)
constexpr int ROW_LO = NB;              // rows 76..151 of M: next off-diagonal block L_{f+1,f}
constexpr int ROW_W = 2 * NB;           // rows 152..161: arrow rows W_f
constexpr int ROW_Z = 2 * NB + NS;      // row 162: right-hand side
constexpr int MROWS = ROW_Z + 1;        // 163
constexpr int PROWS = NB + NS + 1;      // 87: previous [L_{f,f-1} ; W_{f-1} ; z_{f-1}]
constexpr int HFTRI = NFULL * (NFULL + 1) / 2;

constexpr double wInitialRadius = 1e4, wMaxRadius = 1e16, wMinRadius = 1e-32;
constexpr double wMinDiagonal = 1e-6, wMaxDiagonal = 1e32;
constexpr double wMinRelativeDecrease = 1e-3;
constexpr double wFunctionTolerance = 1e-6, wGradientTolerance = 1e-10, wParameterTolerance = 1e-8;
constexpr int wMaxConsecutiveInvalidSteps = 5;

// Optional phase timing (debug hook b3d_debug_win_timing): cycles of block 0 per phase, accumulated in shared memory.
__device__ long long g_win_phase[16];
__device__ int g_win_phase_on;
__device__ __forceinline__ long long* win_phase_acc() {
    __shared__ long long s_win_phase[16];
    return s_win_phase;
}
struct WinTimer {
    long long t;
    bool on;
    __device__ void start() { on = g_win_phase_on != 0 && blockIdx.x == 0 && threadIdx.x == 0; if (on) t = clock64(); }
    __device__ void mark(int i) { if (on) { const long long u = clock64(); win_phase_acc()[i] += u - t; t = u; } }
};

struct WinWs {
    double *x, *cand, *g, *gs, *scale, *diag, *y, *delta, *z, *tmp;   // n each
    double *Hd, *Ld, *Lo;     // F x 76 x 76: scaled J^T J diagonal blocks; factor diagonal blocks; factor sub-diagonal blocks
    double *Ha, *Wf;          // F x 10 x 76: scaled arrow rows; their factor
    double *Hw, *Lw;          // 10 x 10
    double* invd;             // F x 76: 1 / L_jj
    __device__ void carve(double* base, int F) {
        const size_t n = (size_t)F * NB + NS;
        double* p = base;
        x = p; p += n; cand = p; p += n; g = p; p += n; gs = p; p += n; scale = p; p += n;
        diag = p; p += n; y = p; p += n; delta = p; p += n; z = p; p += n; tmp = p; p += n;
        Hd = p; p += (size_t)F * NB * NB; Ld = p; p += (size_t)F * NB * NB; Lo = p; p += (size_t)F * NB * NB;
        Ha = p; p += (size_t)F * NS * NB; Wf = p; p += (size_t)F * NS * NB;
        Hw = p; p += NS * NS; Lw = p; p += NS * NS;
        invd = p; p += (size_t)F * NB;
        xc = p;
    }
    double* xc;               // exchange area of the cluster's two CTAs (offsets XC_*)
};
// exchange area, in doubles: the master's command and its operands, the helper's partial results
constexpr int XC_CMD = 0, XC_RADIUS = 1, XC_OK = 2, XC_COST = 4 /* 2 */, XC_HW = 16 /* 2 x 100 */, XC_GW = 216 /* 2 x 10 */, XC_SACC = 240 /* 144 */,
              XC_BP = 512;  /* the helper's last frame's panels: WT_NC x 4 x 2 x 22 double2 */
enum { CMD_EVAL_JAC = 1, CMD_EVAL_COST = 2, CMD_SOLVE = 3, CMD_EVAL_JAC0 = 4 /* first evaluation: no Jacobi scales yet */, CMD_EXIT = 9 };

// Both CTAs of a problem's cluster: global writes before, reads after (release / acquire at cluster scope).
__device__ __forceinline__ void cluster_sync_all() {
    __threadfence();
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ unsigned cluster_ctarank() { unsigned r; asm volatile("mov.u32 %0, %cluster_ctarank;" : "=r"(r)); return r; }

struct WinEval {
    EvalDense out;                        // dense robustified Jacobian [34][88] (lm_device.cuh) + residuals
    EvalTmp ev;
    double gf[NFULL];
};

struct WinChain {
    double M[MROWS][LDM];
    double P[PROWS][LDM];
    double yn[NB];                        // y of frame f + 1 during the backward sweep
    double rhs[NB];
};

// Forward sweep with the frame's rows held in registers, one 4 x 4 tile per thread (win_solve): the 164 rows of a frame are 41 tile
// rows — 19 of the diagonal block A (lower triangle), 19 of the next off-diagonal block B (row block R is zero left of column block
// R: B = diagonal matrix x L^-T), 3 of the arrow rows and the right-hand side — times 19 tile columns.
constexpr int WT_NC = NB / 4;                 // 19 tile columns / panels
constexpr int WT_RA = 0, WT_RB = WT_NC, WT_RW = 2 * WT_NC, WT_NR = 2 * WT_NC + 3;   // first tile row of A / B / W; 41 tile rows
constexpr int WT_TRI = WT_NC * (WT_NC + 1) / 2;                                       // 190 tiles in a triangle
constexpr int WT_TILES = 2 * WT_TRI + 3 * WT_NC + 9;                                  // + 57 arrow tiles + 9 tiles of the w Schur block = 446
static_assert(NB % 4 == 0 && NS + 1 <= 12, "tile grid of the multi-frame sweep");
struct WinTiles {
    double2 pan[4][2][WT_NR + 1];             // current panel's solved rows: [panel column][rows 0-1 | 2-3 of a tile][tile row]
    double2 bp[WT_NC][4][2][WT_NR - WT_RB];   // the B and W rows of EVERY panel of the frame just swept: the next frame's Schur update
    double dk[2][4][4];                       // diagonal tiles by panel parity: strict lower part = L, diagonal = 1 / L_jj
};

union WinUnion {
    WinEval e;
    WinChain c;
    WinTiles t;
};

struct WinSmem {
    WinUnion u;
    double xfull[NFULL];
    double sc[NFULL];                     // Jacobi scales of the frame being evaluated (76) and of the shape (10)
    double R0[9];
    double red[40];
    double Hw[NS * NS], gw[NS], yw[NS], Sw[NS * NS], zw[NS];
    double piv[4];                        // pivot (value, 1/sqrt) of the Crout column, double-buffered by column parity
    int32_t flag;
};

// ---- evaluate every frame of the problem at parameters xp (n = 76 F + 10) -----------------------------------------------
// Frames [f_lo, f_hi) only (the cluster's two CTAs split the frames); returns this CTA's share of the cost (rank 0 carries the shape
// prior).  JAC: fills Hd / Ha and g of those frames in the workspace and leaves this CTA's share of the shape block in the
// exchange area (XC_HW / XC_GW, slot `rank`); the master adds the two shares.  With `scaled` the Jacobi scaling of Ceres
// (H_ij s_i s_j, gs = g s) is applied while the blocks are written (W.scale is fixed after the first evaluation of a solve);
// without it (first evaluation) the blocks are left unscaled and the master scales them once the scales exist.  A frame's 86 x 86 normal equations are a SYRK of its
// dense Jacobian over 4 x 4 register tiles (thread (ti, tj), 253 tiles), scattered straight into the block structure.
template <int T, bool JAC>
__device__ double win_evaluate(const FitConst* __restrict__ C, int F, const double* xp /* written by the master during the kernel: no read-only path */, const b3d_kp* __restrict__ kps,
                               const int32_t* __restrict__ nkp_all, const double* R0_all, WinWs& W, WinSmem& S, int f_lo, int f_hi,
                               int rank, bool scaled = false) {
    const int tid = threadIdx.x;
    WinEval& E = S.u.e;
    const double bp2 = C->beta_pose * C->beta_pose, bs2 = C->beta_shape * C->beta_shape, lam = C->lambda_temp, lam2 = lam * lam;
    const bool pose_prior = C->beta_pose > 0.0, shape_prior = C->beta_shape > 0.0, temporal = lam > 0.0 && F > 1;
    double part = 0.0;                      // this thread's share of the cost
    WinTimer wt;
    wt.start();
    // this thread's tile of the 86 x 86 lower triangle (22 tile rows)
    int ti = 0;
    while ((ti + 1) * (ti + 2) / 2 <= tid) ++ti;
    const int tj = tid - ti * (ti + 1) / 2;
    const bool tile_ok = ti < (NFULL + 3) / 4;
    if (JAC) {
        for (int k = tid; k < NS * NS; k += T) S.Hw[k] = 0.0;
        if (tid < NS) S.gw[tid] = 0.0;
    }
    if (tid < NS) {
        S.xfull[IDX_W + tid] = xp[(size_t)F * NB + tid];
        if (JAC) S.sc[NB + tid] = scaled ? W.scale[(size_t)F * NB + tid] : 1.0;
    }
    for (int f = f_lo; f < f_hi; ++f) {
        const int nkp = min(max(nkp_all[f], 0), MAXKP);
        const double* xf = xp + (size_t)f * NB;
        __syncthreads();                    // previous frame's consumers are done with xfull / R0 / E
        for (int k = tid; k < NB; k += T) {
            S.xfull[k] = xf[k];
            if (JAC) S.sc[k] = scaled ? W.scale[(size_t)f * NB + k] : 1.0;
        }
        if (tid < 9) S.R0[tid] = R0_all[(size_t)f * 9 + tid];
        __syncthreads();
        wt.mark(5);
        eval_reproj_dense<T, JAC>(C, S.xfull, kps + (size_t)f * MAXKP, nkp, S.R0, E.ev, E.out);
        wt.mark(JAC ? 6 : 4);
        if (tid < nkp) part += E.out.kcost[tid];
        if (pose_prior)
            for (int k = tid; k < NPRI; k += T) { const double r = S.xfull[IDX_JAA + k] * C->beta_pose; part += 0.5 * r * r; }
        if (!JAC) continue;
        const int nrow = 2 * nkp;
        // gradient of the reprojection blocks: column sums in row order (all 86 columns are free in the multi-frame problem)
        for (int i = tid; i < NFULL; i += T) {
            double gacc = 0.0;
            for (int r = 0; r < nrow; r += 2)
                gacc += E.out.J[r][jpos(i)] * E.out.res[r >> 1][0] + E.out.J[r + 1][jpos(i)] * E.out.res[r >> 1][1];
            E.gf[i] = gacc;
        }
        double* Hd = W.Hd + (size_t)f * NB * NB;
        double* Ha = W.Ha + (size_t)f * NS * NB;
        const int nadj = temporal ? (f > 0) + (f < F - 1) : 0;
        if (tile_ok) {
            double acc[4][4];
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
            const double2* Ja = reinterpret_cast<const double2*>(&E.out.J[0][0]) + ti;
            const double2* Jb = reinterpret_cast<const double2*>(&E.out.J[0][0]) + tj;
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
            // scatter into the block structure, adding the L2 pose prior and the temporal terms on the diagonal
#pragma unroll
            for (int a = 0; a < 4; ++a) {
                const int i = 4 * ti + a;
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    const int j = 4 * tj + b;
                    if (i >= NFULL || j > i) continue;
                    double v = acc[a][b];
                    if (i < NB) {
                        if (i == j) {
                            if (pose_prior && i >= IDX_JAA) v += bp2;
                            if (i >= 1) v += nadj * lam2;                 // Vec3DiffCost on every parameter but the scale
                        }
                        if (scaled) v *= S.sc[i] * S.sc[j];
                        Hd[i * NB + j] = v;
                        Hd[j * NB + i] = v;
                    } else if (j < NB) {
                        if (scaled) v *= S.sc[i] * S.sc[j];
                        Ha[(i - NB) * NB + j] = v;
                    } else {                                              // shape x shape: summed over the frames (one owner per entry)
                        S.Hw[(i - NB) * NS + (j - NB)] += v;
                        if (i != j) S.Hw[(j - NB) * NS + (i - NB)] += v;
                    }
                }
            }
        }
        __syncthreads();
        wt.mark(7);
        if (tid < NS) S.gw[tid] += E.gf[IDX_W + tid];
        for (int k = tid; k < NB; k += T) {
            double v = E.gf[k];
            if (pose_prior && k >= IDX_JAA) v += bp2 * S.xfull[k];
            if (temporal && k >= 1) {
                if (f < F - 1) v += lam2 * (xf[k] - xf[NB + k]);
                if (f > 0) v -= lam2 * (xf[k - NB] - xf[k]);
            }
            W.g[(size_t)f * NB + k] = v;
            if (scaled) W.gs[(size_t)f * NB + k] = v * S.sc[k];
        }
        wt.mark(8);
    }
    // temporal residuals lambda (x_f - x_{f+1}) and the shape prior
    if (temporal) {      // the pair (f, f + 1) belongs to the CTA that owns f
        const int p_hi = min(f_hi, F - 1);
        for (int idx = tid; idx < (p_hi - f_lo) * (NB - 1); idx += T) {
            const int f = f_lo + idx / (NB - 1), k = 1 + idx % (NB - 1);
            const double d = lam * (xp[(size_t)f * NB + k] - xp[(size_t)(f + 1) * NB + k]);
            part += 0.5 * d * d;
        }
    }
    if (rank == 0 && shape_prior && tid < NS) { const double r = C->beta_shape * S.xfull[IDX_W + tid]; part += 0.5 * r * r; }
    __syncthreads();
    if (JAC) {
        for (int idx = tid; idx < NS * NS; idx += T) W.xc[XC_HW + rank * NS * NS + idx] = S.Hw[idx];
        if (tid < NS) W.xc[XC_GW + rank * NS + tid] = S.gw[tid];
    }
    const double total = block_sum<T>(part, S.red);
    if (tid == 0) W.xc[XC_COST + rank] = total;
    return total;
}

// master, after the cluster barrier that follows an evaluation with Jacobian: the shape block = both shares + the shape prior
template <int T>
__device__ void win_combine_shape(const FitConst* __restrict__ C, int F, const double* xp, WinWs& W) {
    const int tid = threadIdx.x;
    const double bs2 = C->beta_shape * C->beta_shape;
    const bool shape_prior = C->beta_shape > 0.0;
    for (int idx = tid; idx < NS * NS; idx += T)
        W.Hw[idx] = (W.xc[XC_HW + idx] + W.xc[XC_HW + NS * NS + idx]) + ((shape_prior && idx / NS == idx % NS) ? bs2 : 0.0);
    if (tid < NS) W.g[(size_t)F * NB + tid] = (W.xc[XC_GW + tid] + W.xc[XC_GW + NS + tid]) + (shape_prior ? bs2 * xp[(size_t)F * NB + tid] : 0.0);
}

// ---- block-tridiagonal + arrowhead solve of (H_s + D^2 / radius) y = gs ---------------------------------------------------
// H_s: scaled J^T J held as Hd (diagonal blocks), Ha (arrow rows), Hw, and the implicit off-diagonal blocks
// -lambda^2 diag(scale_f,i scale_{f+1},i), i >= 1.
//
// The chain is eliminated FROM BOTH ENDS by the two CTAs of the problem's cluster: the master sweeps frames 0, 1, ... up to the
// junction frame m = F / 2, the helper sweeps F - 1, F - 2, ... down to m + 1 (a block-tridiagonal matrix read backwards is block
// tridiagonal again: the same code with the frame order reversed).  Before the master eliminates the junction frame it waits for
// the helper and subtracts the helper's last block row as well (its panels travel through the exchange area), adds the helper's share
// of the w Schur complement, solves the 10 x 10 arrow block and the junction frame; then both halves are back-substituted at the
// same time, each away from the junction.
//
// win_forward: frames f0, f0 + dir, ... (nsteps of them).  Every frame couples to the next one visited; the last one does iff
// `last_couples` (the helper's last frame couples to the junction).  `sync_last`: a cluster barrier precedes the last step, and the
// block row in the exchange area (the helper's) is subtracted there too if `extra`.  sacc: this thread's tile of
// sum [W ; z] [W ; z]^T (group-3 threads).  Returns false when a pivot is not positive.
template <int T>
__device__ bool win_forward(const FitConst* __restrict__ C, int F, double radius, WinWs& W, WinSmem& S, int f0, int dir, int nsteps, bool last_couples,
                            bool sync_last, bool extra, double (&sacc)[4][4], int& grp_out, int& ti_out, int& tj_out) {
    const int tid = threadIdx.x;
    const double lam2 = C->lambda_temp * C->lambda_temp;
    const bool temporal = C->lambda_temp > 0.0 && F > 1;
    WinTimer wt;
    wt.start();
    WinTiles& Q = S.u.t;
    // ---- this thread's tile: group 0 = A (ti >= tj), 1 = B (row block R <= column block C), 2 = arrow rows / right-hand side,
    //      3 = one of the 3 x 3 tiles of sum_f [W_f ; z_f] [W_f ; z_f]^T (the w Schur complement), kept across the frames ----
    int grp = -1, ti = 0, tj = 0;
    if (tid < WT_TRI) {
        int r = 0;
        while ((r + 1) * (r + 2) / 2 <= tid) ++r;
        grp = 0; ti = r; tj = tid - r * (r + 1) / 2;
    } else if (tid < 2 * WT_TRI) {
        const int u = tid - WT_TRI;
        int c = 0;
        while ((c + 1) * (c + 2) / 2 <= u) ++c;
        grp = 1; tj = c; ti = WT_RB + (u - c * (c + 1) / 2);
    } else if (tid < 2 * WT_TRI + 3 * WT_NC) {
        const int u = tid - 2 * WT_TRI;
        grp = 2; ti = WT_RW + u / WT_NC; tj = u % WT_NC;
    } else if (tid < WT_TILES) {
        const int u = tid - 2 * WT_TRI - 3 * WT_NC;
        grp = 3; ti = WT_RW + u / 3; tj = WT_RW + u % 3;      // tj is a tile ROW here: the tile is rows(ti) x rows(tj)^T
    }
    static_assert(T >= WT_TILES, "one tile per thread");
    grp_out = grp; ti_out = ti; tj_out = tj;
    double a[4][4];
    bool ok = true, synced = false;
    if (tid == 0) S.flag = 1;
    __syncthreads();
    for (int t = 0; t < nsteps; ++t) {
        const int f = f0 + dir * t, fn = f + dir;             // this frame and the next one visited
        const bool last = t == nsteps - 1;
        const bool has_next = temporal && (!last || last_couples);
        const bool has_prev = temporal && t > 0;
        if (last && sync_last) { cluster_sync_all(); synced = true; }   // the helper's block row and its share of sacc are in the exchange area
        const bool use_extra = last && extra && temporal;
        // (1) this frame's rows: [Hd_f + damping ; -lambda^2 S_f E S_f+1 ; Ha_f ; gs_f], minus the products with the previous frame's
        //     block row [B ; W ; z] (its panels are still in Q.bp)
        const double* Hd = W.Hd + (size_t)f * NB * NB;
        const double* Ha = W.Ha + (size_t)f * NS * NB;
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int j = 4 * tj + c;
                double v = 0.0;
                if (grp == 0) {
                    const int i = 4 * ti + r;
                    if (j <= i) {
                        v = Hd[i * NB + j];
                        if (i == j) { const double l = sqrt(W.diag[(size_t)f * NB + i] / radius); v += l * l; }
                    }
                } else if (grp == 1) {
                    const int i = 4 * (ti - WT_RB) + r;
                    if (has_next && i == j && i >= 1) v = -lam2 * W.scale[(size_t)f * NB + i] * W.scale[(size_t)fn * NB + i];
                } else if (grp == 2) {
                    const int sidx = 4 * (ti - WT_RW) + r;
                    if (sidx < NS) v = Ha[sidx * NB + j];
                    else if (sidx == NS) v = W.gs[(size_t)f * NB + j];
                }
                a[r][c] = v;
            }
        if (has_prev && (grp == 0 || grp == 2)) {
            // A_f -= B B^T, [W ; z]_f -= [W ; z] B^T over the panels in which both row blocks are non-zero (B is upper triangular)
            const int rb = (grp == 0 ? ti : ti - WT_RB), cb = tj;      // row / column tile rows inside Q.bp (B rows 0..18, W rows 19..21)
            for (int k = (grp == 0 ? ti : tj); k < WT_NC; ++k) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const double2 r01 = Q.bp[k][q][0][rb], r23 = Q.bp[k][q][1][rb], c01 = Q.bp[k][q][0][cb], c23 = Q.bp[k][q][1][cb];
                    const double rv[4] = {r01.x, r01.y, r23.x, r23.y}, cv[4] = {c01.x, c01.y, c23.x, c23.y};
#pragma unroll
                    for (int r = 0; r < 4; ++r)
#pragma unroll
                        for (int c = 0; c < 4; ++c) a[r][c] = fma(-rv[r], cv[c], a[r][c]);
                }
            }
        }
        if (use_extra && (grp == 0 || grp == 2)) {
            // the same products with the helper's last block row, read from the exchange area (layout of Q.bp)
            const double2* xb = reinterpret_cast<const double2*>(W.xc + XC_BP);
            constexpr int XR = WT_NR - WT_RB;
            const int rb = (grp == 0 ? ti : ti - WT_RB), cb = tj;
            for (int k = (grp == 0 ? ti : tj); k < WT_NC; ++k) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const double2* base = xb + (size_t)((k * 4 + q) * 2) * XR;
                    const double2 r01 = base[rb], r23 = base[XR + rb], c01 = base[cb], c23 = base[XR + cb];
                    const double rv[4] = {r01.x, r01.y, r23.x, r23.y}, cv[4] = {c01.x, c01.y, c23.x, c23.y};
#pragma unroll
                    for (int r = 0; r < 4; ++r)
#pragma unroll
                        for (int c = 0; c < 4; ++c) a[r][c] = fma(-rv[r], cv[c], a[r][c]);
                }
            }
        }
        __syncthreads();           // every thread is done with the previous frame's Q.bp
        wt.mark(10);
        // factor the thread's own diagonal tile in registers and publish it
        auto factor_own = [&](int k) {
            // branch-free (see fit_single.cu): a non-positive or NaN pivot only raises the flag
            bool bad = false;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const double piv = a[c][c];
                bad = bad || !(piv > 0.0);
                const double inv = rsqrt(piv);
                a[c][c] = inv;
#pragma unroll
                for (int r = c + 1; r < 4; ++r) a[r][c] *= inv;
#pragma unroll
                for (int c2 = c + 1; c2 < 4; ++c2)
#pragma unroll
                    for (int r = c2; r < 4; ++r) a[r][c2] -= a[r][c] * a[c2][c];
                W.invd[(size_t)f * NB + 4 * k + c] = inv;
            }
            if (bad) S.flag = 0;
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c <= r; ++c) Q.dk[k & 1][r][c] = a[r][c];
        };
        if (grp == 0 && ti == 0 && tj == 0) factor_own(0);
        __syncthreads();
        for (int k = 0; k < WT_NC; ++k) {
            if (!S.flag) { ok = false; break; }               // block-uniform (written before the barrier every thread just passed)
            // ---- solve the tile column k against the diagonal tile: A rows below it, the B row blocks that have reached column k,
            //      the arrow rows and the right-hand side ----
            const bool in_col = tj == k && ((grp == 0 && ti > k) || (grp == 1 && has_next) || grp == 2);
            if (in_col) {
                const double (*D)[4] = Q.dk[k & 1];
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
                    const double2 lo = make_double2(a[0][c], a[1][c]), hi = make_double2(a[2][c], a[3][c]);
                    Q.pan[c][0][ti] = lo; Q.pan[c][1][ti] = hi;
                    if (ti >= WT_RB) { Q.bp[k][c][0][ti - WT_RB] = lo; Q.bp[k][c][1][ti - WT_RB] = hi; }
                }
            }
            // look-ahead: tile (k + 1, k) and the next diagonal tile (k + 1, k + 1) are neighbouring lanes of one warp
            __syncwarp();
            const bool next_diag = grp == 0 && ti == k + 1 && tj == k + 1;
            if (next_diag) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const double2 r01 = Q.pan[q][0][ti], r23 = Q.pan[q][1][ti];
                    const double rv[4] = {r01.x, r01.y, r23.x, r23.y};
#pragma unroll
                    for (int r = 0; r < 4; ++r)
#pragma unroll
                        for (int c = 0; c <= r; ++c) a[r][c] = fma(-rv[r], rv[c], a[r][c]);
                }
                factor_own(k + 1);
            }
            __syncthreads();
            // ---- rank-4 update of the trailing tiles ----
            if (grp == 3) {
                // [W ; z] [W ; z]^T accumulates over every panel of every frame
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const double2 r01 = Q.pan[q][0][ti], r23 = Q.pan[q][1][ti], c01 = Q.pan[q][0][tj], c23 = Q.pan[q][1][tj];
                    const double rv[4] = {r01.x, r01.y, r23.x, r23.y}, cv[4] = {c01.x, c01.y, c23.x, c23.y};
#pragma unroll
                    for (int r = 0; r < 4; ++r)
#pragma unroll
                        for (int c = 0; c < 4; ++c) sacc[r][c] = fma(rv[r], cv[c], sacc[r][c]);
                }
            } else if (grp >= 0 && tj > k && !next_diag && (grp != 1 || (has_next && ti - WT_RB <= k))) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const double2 r01 = Q.pan[q][0][ti], r23 = Q.pan[q][1][ti], c01 = Q.pan[q][0][tj], c23 = Q.pan[q][1][tj];
                    const double rv[4] = {r01.x, r01.y, r23.x, r23.y}, cv[4] = {c01.x, c01.y, c23.x, c23.y};
#pragma unroll
                    for (int r = 0; r < 4; ++r)
#pragma unroll
                        for (int c = 0; c < 4; ++c) a[r][c] = fma(-rv[r], cv[c], a[r][c]);
                }
            }
            __syncthreads();
        }
        wt.mark(11);
        if (!ok) break;
        // (4) results of this frame for the backward sweep: L_ff (strict lower part), L_{f+1,f} (upper triangle), W_f, z_f
        if (grp == 0) {
            double* Ld = W.Ld + (size_t)f * NB * NB;
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c < 4; ++c) { const int i = 4 * ti + r, j = 4 * tj + c; if (j < i) Ld[i * NB + j] = a[r][c]; }
        } else if (grp == 1 && has_next) {
            double* Lo = W.Lo + (size_t)f * NB * NB;
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c < 4; ++c) { const int i = 4 * (ti - WT_RB) + r, j = 4 * tj + c; Lo[i * NB + j] = j >= i ? a[r][c] : 0.0; }
        } else if (grp == 2) {
            double* Wf = W.Wf + (size_t)f * NS * NB;
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int sidx = 4 * (ti - WT_RW) + r, j = 4 * tj + c;
                    if (sidx < NS) Wf[sidx * NB + j] = a[r][c];
                    else if (sidx == NS) W.z[(size_t)f * NB + j] = a[r][c];
                }
        }
        wt.mark(12);
    }
    __syncthreads();
    if (sync_last && !synced) cluster_sync_all();        // a failed sweep still meets the helper at the barrier
    return ok;
}

// The helper hands its last block row (the panels of the last frame it swept, still in shared memory) and its share of the w Schur
// complement to the master through the exchange area.
template <int T>
__device__ void win_publish_half(WinWs& W, WinSmem& S, const double (&sacc)[4][4], int grp, int ti, int tj, bool ok) {
    const int tid = threadIdx.x;
    const double2* src = reinterpret_cast<const double2*>(&S.u.t.bp[0][0][0][0]);
    double2* dst = reinterpret_cast<double2*>(W.xc + XC_BP);
    for (int i = tid; i < WT_NC * 4 * 2 * (WT_NR - WT_RB); i += T) dst[i] = src[i];
    if (grp == 3) {
        const int u = (ti - WT_RW) * 3 + (tj - WT_RW);
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int c = 0; c < 4; ++c) W.xc[XC_SACC + u * 16 + r * 4 + c] = sacc[r][c];
    }
    if (tid == 0) W.xc[XC_OK] = ok ? 1.0 : 0.0;
}

// master: Sw = Hw + damping - sum_f W_f W_f^T, zw = gs_w - sum_f W_f z_f from the group-3 tiles of both CTAs; 10 x 10 Cholesky; y_w.
template <int T>
__device__ bool win_arrow(int F, double radius, WinWs& W, WinSmem& S, const double (&sacc)[4][4], int grp, int ti, int tj, bool helper_has) {
    const int tid = threadIdx.x;
    const size_t nw0 = (size_t)F * NB;
    if (grp == 3) {
        const int u = (ti - WT_RW) * 3 + (tj - WT_RW);
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int sr = 4 * (ti - WT_RW) + r, sc = 4 * (tj - WT_RW) + c;
                const double acc = sacc[r][c] + (helper_has ? W.xc[XC_SACC + u * 16 + r * 4 + c] : 0.0);
                if (sr < NS && sc <= sr) {
                    double v = W.Hw[sr * NS + sc] - acc;
                    if (sr == sc) { const double l = sqrt(W.diag[nw0 + sr] / radius); v += l * l; }
                    S.Sw[sr * NS + sc] = v;
                } else if (sr == NS && sc < NS) {
                    S.zw[sc] = W.gs[nw0 + sc] - acc;
                }
            }
    }
    __syncthreads();
    // 10 x 10 Cholesky by one thread
    if (tid == 0) {
        int good = 1;
        for (int j = 0; j < NS && good; ++j) {
            double d = S.Sw[j * NS + j];
            for (int k = 0; k < j; ++k) d -= S.Sw[j * NS + k] * S.Sw[j * NS + k];
            if (!(d > 0.0)) { good = 0; break; }
            d = sqrt(d);
            S.Sw[j * NS + j] = d;
            for (int i = j + 1; i < NS; ++i) {
                double s = S.Sw[i * NS + j];
                for (int k = 0; k < j; ++k) s -= S.Sw[i * NS + k] * S.Sw[j * NS + k];
                S.Sw[i * NS + j] = s / d;
            }
        }
        if (good) {
            for (int i = 0; i < NS; ++i) { double s = S.zw[i]; for (int k = 0; k < i; ++k) s -= S.Sw[i * NS + k] * S.zw[k]; S.zw[i] = s / S.Sw[i * NS + i]; }
            for (int i = NS - 1; i >= 0; --i) { double s = S.zw[i]; for (int k = i + 1; k < NS; ++k) s -= S.Sw[k * NS + i] * S.yw[k]; S.yw[i] = s / S.Sw[i * NS + i]; }
        }
        S.flag = good;
    }
    __syncthreads();
    if (!S.flag) return false;
    if (tid < NS) W.y[nw0 + tid] = S.yw[tid];
    return true;
}

// ---- backward sweep: L_ff^T y_f = z_f - W_f^T y_w - Lo_f^T y_next(f), frames f0, f0 + dir, ... (nsteps), each after its `next` =
//      f - dir (the frame it was coupled to in the forward sweep).  first_has_next: the first frame has one (the helper's first
//      frame: the junction, whose y the master has left in W.y). ----
template <int T>
__device__ void win_back(const FitConst* __restrict__ C, int F, WinWs& W, WinSmem& S, int f0, int dir, int nsteps, bool first_has_next) {
    const int tid = threadIdx.x;
    WinChain& K = S.u.c;
    const bool temporal = C->lambda_temp > 0.0 && F > 1;
    WinTimer wt;
    wt.start();
    if (tid < NS) S.yw[tid] = W.y[(size_t)F * NB + tid];
    if (nsteps > 0 && first_has_next && temporal)
        for (int i = tid; i < NB; i += T) K.yn[i] = W.y[(size_t)(f0 - dir) * NB + i];
    __syncthreads();
    for (int t = 0; t < nsteps; ++t) {
        const int f = f0 + dir * t;
        const bool has_next = temporal && (t > 0 || first_has_next);
        const double* Ld = W.Ld + (size_t)f * NB * NB;
        for (int idx = tid; idx < NB * NB; idx += T) { const int r = idx / NB, c = idx % NB; if (c < r) K.M[r][c] = Ld[idx]; }
        if (has_next) {
            const double* Lo = W.Lo + (size_t)f * NB * NB;
            for (int idx = tid; idx < NB * NB; idx += T) K.M[ROW_LO + idx / NB][idx % NB] = Lo[idx];
        }
        const double* Wf = W.Wf + (size_t)f * NS * NB;
        for (int idx = tid; idx < NS * NB; idx += T) K.M[ROW_W + idx / NB][idx % NB] = Wf[idx];
        __syncthreads();
        for (int i = tid; i < NB; i += T) {
            double v = W.z[(size_t)f * NB + i];
#pragma unroll
            for (int s = 0; s < NS; ++s) v -= K.M[ROW_W + s][i] * S.yw[s];
            if (has_next)
                for (int k = 0; k <= i; ++k) v -= K.M[ROW_LO + k][i] * K.yn[k];      // the coupling block is upper triangular
            K.rhs[i] = v;
        }
        __syncthreads();
        if (tid < 32) {      // back substitution with one warp (lm_device.cuh backsolve_warp_blocked)
            const int lane = tid;
            double zz[3];
#pragma unroll
            for (int q = 0; q < 3; ++q) { const int i = lane + 32 * q; zz[q] = i < NB ? K.rhs[i] : 0.0; }
            static_assert(LDM >= NB && MROWS * LDM >= (NB - 1) * LDM + 96, "row reads of the blocked back substitution stay inside K.M");
            backsolve_warp_blocked(NB, [&K](int r) { return (const double*)K.M[r]; }, W.invd + (size_t)f * NB, zz, K.yn);
            __syncwarp();
            for (int i = lane; i < NB; i += 32) W.y[(size_t)f * NB + i] = K.yn[i];
        }
        __syncthreads();
    }
    wt.mark(13);
}

// One problem = one cluster of two CTAs.  The master (cluster rank 0) runs the Levenberg-Marquardt loop; the helper waits for its
// commands at a cluster barrier: evaluate the upper half of the frames, or eliminate / back-substitute the far half of the chain.
template <int T>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(T, 1) fit_window_kernel(const FitConst* __restrict__ C, int nprob, const WinProblem* __restrict__ probs,
                                                          const b3d_kp* __restrict__ kps_all, const int32_t* __restrict__ nkp_all,
                                                          double* R0_all /* chain mode rotates it between solves: no read-only path */, double* poses,
                                                          double* __restrict__ Wshape, uint8_t* __restrict__ usable,
                                                          b3d_iter_record* __restrict__ trace, int trace_stride, int32_t* __restrict__ niters,
                                                          int32_t* __restrict__ term, double* __restrict__ workspace,
                                                          const int32_t* __restrict__ chain_order, int chain_len, double* r_out, double* p_out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    WinSmem& S = *reinterpret_cast<WinSmem*>(smem_raw);
    const int tid = threadIdx.x, cid = blockIdx.x >> 1;
    const int rank = (int)cluster_ctarank();
    // Chain mode (chain_order != null): this cluster solves the problems chain_order[cid * chain_len + 0, 1, ...] one after the other
    // (the windows of one shard: window k + 1 starts from what window k left in its overlap frames), each followed by the write-back
    // of its frames (R0 <- R(rootAA) R0, r, p: include/MultiFrameBA.h:154-172) — no host round trip and no wave barrier across shards.
    const int nsteps = chain_order ? chain_len : 1;
    for (int step = 0; step < nsteps; ++step) {
    const int pb = chain_order ? chain_order[(size_t)cid * chain_len + step] : cid;
    if (pb < 0 || pb >= nprob) break;        // (both CTAs of the cluster)
    const WinProblem pr = probs[pb];
    const int F = pr.count, n = F * NB + NS;
    const b3d_kp* kps = kps_all + (size_t)pr.first * MAXKP;
    const int32_t* nkp = nkp_all + pr.first;
    const double* R0 = R0_all + (size_t)pr.first * 9;
    double* xin = poses + (size_t)pr.first * NB;
    double* win = Wshape + (size_t)pr.first * NS;
    b3d_iter_record* tr = trace ? trace + (size_t)pb * trace_stride : nullptr;
    WinWs W;
    W.carve(workspace + pr.ws_offset, F);
    const double lam2 = C->lambda_temp * C->lambda_temp;
    const bool temporal = C->lambda_temp > 0.0 && F > 1;
    const int m = F / 2, nh = F - 1 - m;                     // junction frame; frames the helper eliminates (F - 1 .. m + 1)
    const int ev_split = (F + 1) / 2;                        // the master evaluates frames [0, ev_split), the helper the rest
    if (tid < 16) win_phase_acc()[tid] = 0;
    __syncthreads();

    if (rank == 1) {
        // ---------------- helper ----------------
        double sacc[4][4];
        int grp, ti, tj;
        for (;;) {
            cluster_sync_all();
            const int cmd = (int)*(volatile double*)&W.xc[XC_CMD];
            if (cmd == CMD_EXIT) break;
            if (cmd == CMD_EVAL_JAC || cmd == CMD_EVAL_JAC0) {
                win_evaluate<T, true>(C, F, W.x, kps, nkp, R0, W, S, ev_split, F, 1, cmd == CMD_EVAL_JAC);
                cluster_sync_all();
            } else if (cmd == CMD_EVAL_COST) {
                win_evaluate<T, false>(C, F, W.cand, kps, nkp, R0, W, S, ev_split, F, 1);
                cluster_sync_all();
            } else if (cmd == CMD_SOLVE) {
                const double radius_h = *(volatile double*)&W.xc[XC_RADIUS];
#pragma unroll
                for (int r = 0; r < 4; ++r)
#pragma unroll
                    for (int c = 0; c < 4; ++c) sacc[r][c] = 0.0;
                const bool okh = win_forward<T>(C, F, radius_h, W, S, F - 1, -1, nh, true, false, false, sacc, grp, ti, tj);
                win_publish_half<T>(W, S, sacc, grp, ti, tj, okh);
                cluster_sync_all();                              // (A) the master takes the junction
                cluster_sync_all();                              // (B) y_w and y of the junction frame are in W.y, or the solve failed
                if (*(volatile double*)&W.xc[XC_OK] > 0.5) win_back<T>(C, F, W, S, m + 1, +1, nh, true);
                cluster_sync_all();                              // (C)
            }
        }
    } else {
    // ---------------- master ----------------
    for (int k = tid; k < F * NB; k += T) W.x[k] = xin[k];
    if (tid < NS) W.x[(size_t)F * NB + tid] = win[tid];
    __syncthreads();
    WinTimer kt;
    kt.start();
    // evaluation of all frames by both CTAs: command, barrier, own half, barrier, sum of the two shares
    auto evaluate_all = [&](const double* xvec, bool jac, bool scaled) -> double {
        if (tid == 0) *(volatile double*)&W.xc[XC_CMD] = jac ? (double)(scaled ? CMD_EVAL_JAC : CMD_EVAL_JAC0) : (double)CMD_EVAL_COST;
        cluster_sync_all();
        const double part = jac ? win_evaluate<T, true>(C, F, xvec, kps, nkp, R0, W, S, 0, ev_split, 0, scaled)
                                : win_evaluate<T, false>(C, F, xvec, kps, nkp, R0, W, S, 0, ev_split, 0);
        cluster_sync_all();
        const double total = part + *(volatile double*)&W.xc[XC_COST + 1];
        if (jac) win_combine_shape<T>(C, F, xvec, W);
        __syncthreads();
        return total;
    };
    // the damped linear solve by both CTAs (see win_forward)
    auto solve_all = [&](double rad) -> bool {
        double sacc[4][4];
        int grp, ti, tj;
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int c = 0; c < 4; ++c) sacc[r][c] = 0.0;
        if (tid == 0) { *(volatile double*)&W.xc[XC_CMD] = (double)CMD_SOLVE; *(volatile double*)&W.xc[XC_RADIUS] = rad; }
        cluster_sync_all();
        bool ok = win_forward<T>(C, F, rad, W, S, 0, +1, m + 1, false, true, nh > 0, sacc, grp, ti, tj);      // barrier (A) inside
        if (nh > 0) ok = ok && *(volatile double*)&W.xc[XC_OK] > 0.5;
        if (ok) ok = win_arrow<T>(F, rad, W, S, sacc, grp, ti, tj, nh > 0);
        if (ok) win_back<T>(C, F, W, S, m, -1, 1, false);          // the junction frame
        __syncthreads();
        if (tid == 0) *(volatile double*)&W.xc[XC_OK] = ok ? 1.0 : 0.0;
        cluster_sync_all();                                      // (B)
        if (ok) win_back<T>(C, F, W, S, m - 1, -1, m, true);
        cluster_sync_all();                                      // (C)
        return ok;
    };

    double radius = wInitialRadius, decrease_factor = 2.0, x_cost = 0.0, cand_cost = 0.0, mcc = 0.0;
    bool reuse_diagonal = false, step_successful = true;
    int num_invalid = 0, iteration = 0, nrec = 0, termination = 0;
    b3d_iter_record it;
    it.cost = it.cost_change = it.gradient_max_norm = it.step_norm = it.relative_decrease = it.radius = 0.0;
    it.flags = 0; it.aux = -1;

    auto block_max = [&](double v) -> double {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_down_sync(0xffffffffu, v, o));
        __syncthreads();
        if ((tid & 31) == 0) S.red[tid >> 5] = v;
        __syncthreads();
        double m = 0.0;
        for (int i = 0; i < T / 32; ++i) m = fmax(m, S.red[i]);
        __syncthreads();
        return m;
    };
    auto eval_grad_jac = [&](bool first) -> bool {
        x_cost = evaluate_all(W.x, true, !first);
        if (!isfinite(x_cost)) return false;
        it.cost = x_cost;
        __syncthreads();
        if (first) {
            for (int k = tid; k < n; k += T) {
                const double d = k < F * NB ? W.Hd[(size_t)(k / NB) * NB * NB + (k % NB) * (NB + 1)] : W.Hw[(k - F * NB) * (NS + 1)];
                W.scale[k] = 1.0 / (1.0 + sqrt(d));
            }
            __syncthreads();
            // Jacobi scaling of the blocks (column scaling of J on both sides of J^T J).  Only here: every later evaluation applies
            // the (now fixed) scales while it writes the blocks (win_evaluate, `scaled`)
            for (size_t idx = tid; idx < (size_t)F * NB * NB; idx += T) {
                const int f = (int)(idx / (NB * NB)), r = (int)(idx % (NB * NB)) / NB, c = (int)(idx % NB);
                W.Hd[idx] *= W.scale[(size_t)f * NB + r] * W.scale[(size_t)f * NB + c];
            }
            for (size_t idx = tid; idx < (size_t)F * NS * NB; idx += T) {
                const int f = (int)(idx / (NS * NB)), s = (int)(idx % (NS * NB)) / NB, c = (int)(idx % NB);
                W.Ha[idx] *= W.scale[(size_t)F * NB + s] * W.scale[(size_t)f * NB + c];
            }
            for (int k = tid; k < F * NB; k += T) W.gs[k] = W.g[k] * W.scale[k];
        }
        if (tid < NS * NS) W.Hw[tid] *= W.scale[(size_t)F * NB + tid / NS] * W.scale[(size_t)F * NB + tid % NS];
        if (tid < NS) W.gs[(size_t)F * NB + tid] = W.g[(size_t)F * NB + tid] * W.scale[(size_t)F * NB + tid];
        double gm = 0.0;
        for (int k = tid; k < n; k += T) gm = fmax(gm, fabs(W.g[k]));
        it.gradient_max_norm = block_max(gm);     // unconstrained: |x - (x - g)| = |g|
        return true;
    };
    auto push = [&]() {
        it.radius = radius;
        it.aux = -1;
        if (tr && nrec < trace_stride && tid == 0) tr[nrec] = it;
        ++nrec;
    };
    auto norm_of = [&](const double* v) -> double {
        double a = 0.0;
        for (int k = tid; k < n; k += T) a += v[k] * v[k];
        return sqrt(block_sum<T>(a, S.red));
    };

    double x_norm = norm_of(W.x);
    if (!eval_grad_jac(true)) {
        termination = B3D_TERM_EVAL_FAILURE;
    } else {
        it.flags = 3;
        for (;;) {
            push();
            if (iteration >= C->max_iters) { termination = B3D_TERM_MAX_ITERS; break; }
            if (step_successful && it.gradient_max_norm <= wGradientTolerance) { termination = B3D_TERM_GRAD_TOL; break; }
            if (!(radius > wMinRadius)) { termination = B3D_TERM_MIN_RADIUS; break; }
            const double prev_gmax = it.gradient_max_norm;
            it.cost = it.cost_change = it.gradient_max_norm = it.step_norm = it.relative_decrease = 0.0;
            it.flags = 0;
            ++iteration;
            step_successful = false;
            if (!reuse_diagonal) {
                for (int k = tid; k < n; k += T) {
                    const double d = k < F * NB ? W.Hd[(size_t)(k / NB) * NB * NB + (k % NB) * (NB + 1)] : W.Hw[(k - F * NB) * (NS + 1)];
                    W.diag[k] = fmin(fmax(d, wMinDiagonal), wMaxDiagonal);
                }
            }
            __syncthreads();
            reuse_diagonal = true;
            bool step_valid = false;
            kt.mark(0);
            const bool solved = solve_all(radius);
            kt.mark(2);
            if (solved) {
                __syncthreads();
                // model_cost_change = y^T gs - y^T H_s y / 2 (step = -y)
                double v = 0.0, fin = 0.0;
                for (size_t idx = tid; idx < (size_t)F * NB * NB; idx += T) {
                    const int f = (int)(idx / (NB * NB)), r = (int)(idx % (NB * NB)) / NB, c = (int)(idx % NB);
                    v -= 0.5 * W.Hd[idx] * W.y[(size_t)f * NB + r] * W.y[(size_t)f * NB + c];
                }
                for (size_t idx = tid; idx < (size_t)F * NS * NB; idx += T) {
                    const int f = (int)(idx / (NS * NB)), s = (int)(idx % (NS * NB)) / NB, c = (int)(idx % NB);
                    v -= W.Ha[idx] * W.y[(size_t)F * NB + s] * W.y[(size_t)f * NB + c];
                }
                if (tid < NS * NS) v -= 0.5 * W.Hw[tid] * W.y[(size_t)F * NB + tid / NS] * W.y[(size_t)F * NB + tid % NS];
                if (temporal)
                    for (int idx = tid; idx < (F - 1) * (NB - 1); idx += T) {
                        const int f = idx / (NB - 1), k = 1 + idx % (NB - 1);
                        const size_t a = (size_t)f * NB + k, b = a + NB;
                        v += lam2 * W.scale[a] * W.scale[b] * W.y[a] * W.y[b];     // off-diagonal entry is -lambda^2 s_a s_b, counted twice, halved
                    }
                for (int k = tid; k < n; k += T) { v += W.y[k] * W.gs[k]; fin += isfinite(W.y[k]) ? 0.0 : 1.0; }
                mcc = block_sum<T>(v, S.red);
                const double bad = block_sum<T>(fin, S.red);
                step_valid = (bad == 0.0) && (mcc > 0.0);
            }
            if (!step_valid) {
                if (++num_invalid >= wMaxConsecutiveInvalidSteps) { termination = B3D_TERM_INVALID_STEPS; break; }
                radius = radius / decrease_factor; decrease_factor *= 2.0;
                it.cost = x_cost; it.gradient_max_norm = prev_gmax;
                continue;
            }
            num_invalid = 0;
            double sn = 0.0;
            for (int k = tid; k < n; k += T) {
                const double d = -W.y[k] * W.scale[k];
                const double c = W.x[k] + d;
                W.cand[k] = c;
                const double e = W.x[k] - c;
                sn += e * e;
            }
            it.flags = 1;
            __syncthreads();
            kt.mark(3);
            cand_cost = evaluate_all(W.cand, false, false);
            kt.mark(14);
            if (!isfinite(cand_cost)) cand_cost = DBL_MAX;
            it.step_norm = sqrt(block_sum<T>(sn, S.red));
            if (it.step_norm <= wParameterTolerance * (x_norm + wParameterTolerance)) { termination = B3D_TERM_PARAM_TOL; break; }
            it.cost_change = x_cost - cand_cost;
            if (fabs(it.cost_change) <= wFunctionTolerance * x_cost) { termination = B3D_TERM_FUNC_TOL; break; }
            it.relative_decrease = it.cost_change / mcc;
            if (it.relative_decrease > wMinRelativeDecrease) {
                __syncthreads();
                for (int k = tid; k < n; k += T) W.x[k] = W.cand[k];
                __syncthreads();
                x_norm = norm_of(W.x);
                kt.mark(0);
                if (!eval_grad_jac(false)) { termination = B3D_TERM_EVAL_FAILURE; break; }
                kt.mark(1);
                step_successful = true;
                it.flags |= 2;
                const double q = 2.0 * it.relative_decrease - 1.0;
                radius = radius / fmax(1.0 / 3.0, 1.0 - q * q * q);
                radius = fmin(wMaxRadius, radius);
                decrease_factor = 2.0;
                reuse_diagonal = false;
            } else {
                radius = radius / decrease_factor; decrease_factor *= 2.0;
                it.cost = cand_cost;
            }
        }
    }
    __syncthreads();
    if (tid == 0) *(volatile double*)&W.xc[XC_CMD] = (double)CMD_EXIT;
    cluster_sync_all();                      // releases the helper
    if (pb == 0 && tid < 16 && g_win_phase_on) g_win_phase[tid] += win_phase_acc()[tid];
    for (int k = tid; k < F * NB; k += T) xin[k] = W.x[k];
    if (tid < NS) win[tid] = W.x[(size_t)F * NB + tid];
    if (tid == 0) {
        if (usable) usable[pb] = (termination >= B3D_TERM_PARAM_TOL && termination <= B3D_TERM_MIN_RADIUS) ? 1 : 0;
        if (niters) niters[pb] = nrec;
        if (term) term[pb] = termination;
    }
    }   // master
    if (chain_order) {
        cluster_sync_all();                  // the master's poses are visible
        if (rank == 0)
            for (int t = tid; t < F * NJ; t += T)
                writeback_entry(pr.first + t / NJ, t % NJ, poses, R0_all, r_out, p_out);
        cluster_sync_all();                  // ... and so are the rotated R0 / r / p, before the chain's next window reads them
    }
    }   // chain steps
}

}  // namespace b3d

#include "fit_chain.cuh"

namespace b3d {

static std::atomic<int> g_chain_min_frames{CH_MIN_FRAMES};      // debug hook b3d_debug_chain_min_frames (tests force either kernel)

// one long chain on a cooperative grid: G CTAs, one per SM at most
static int chain_launch(const FitConst* dfc, const WinProblem& pr, int pb, const b3d_kp* kps, const int32_t* nkp, const double* R0, double* poses,
                        double* W, uint8_t* usable, b3d_iter_record* trace, int trace_stride, int32_t* niters, int32_t* term, double* ws, cudaStream_t st) {
    int dev = 0, sms = 0;
    B3D_CUDA(cudaGetDevice(&dev));
    B3D_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    B3D_CUDA(cudaFuncSetAttribute(fit_chain_kernel<FW_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ChainSmem)));
    int per_sm = 0;
    B3D_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fit_chain_kernel<FW_THREADS>, FW_THREADS, sizeof(ChainSmem)));
    if (per_sm < 1) return set_error(B3D_ECUDA, "fit_chain_kernel does not fit on an SM");
    const int G = std::max(1, std::min(std::min(sms * per_sm, (int)CH_MAX_GRID), (pr.count + 1) / 2));
    B3D_CUDA(cudaMemsetAsync(ws + pr.ws_offset + chain_bar_offset(pr.count, G), 0, 64, st));
    WinProblem prv = pr;
    void* args[] = {(void*)&dfc, (void*)&prv, (void*)&pb, (void*)&kps, (void*)&nkp, (void*)&R0, (void*)&poses, (void*)&W, (void*)&usable,
                    (void*)&trace, (void*)&trace_stride, (void*)&niters, (void*)&term, (void*)&ws};
    B3D_CUDA(cudaLaunchCooperativeKernel((const void*)fit_chain_kernel<FW_THREADS>, dim3(G), dim3(FW_THREADS), args, sizeof(ChainSmem), st));
    B3D_LAUNCH_CHECK();
    return B3D_OK;
}

int win_launch(const FitConst* dfc, const WinProblem* dprobs, int nprob, const b3d_kp* kps, const int32_t* nkp, const double* R0, double* poses,
               double* W, uint8_t* usable, b3d_iter_record* trace, int trace_stride, int32_t* niters, int32_t* term, double* ws, cudaStream_t st,
               const WinProblem* hprobs) {
    if (nprob <= 0) return B3D_OK;
    if (hprobs) {      // a launch made only of long chains (stage 1): each takes the whole machine, one after the other
        const int min_f = g_chain_min_frames.load();
        bool all_long = true;
        for (int p = 0; p < nprob; ++p) all_long = all_long && hprobs[p].count >= std::max(min_f, (int)CH_MIN_FRAMES);
        if (all_long) {
            for (int p = 0; p < nprob; ++p)
                if (int rc = chain_launch(dfc, hprobs[p], p, kps, nkp, R0, poses, W, usable, trace, trace_stride, niters, term, ws, st)) return rc;
            return B3D_OK;
        }
    }
    B3D_CUDA(cudaFuncSetAttribute(fit_window_kernel<FW_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(WinSmem)));   // per device: every call
    fit_window_kernel<FW_THREADS><<<2 * nprob, FW_THREADS, sizeof(WinSmem), st>>>(dfc, nprob, dprobs, kps, nkp, const_cast<double*>(R0), poses, W, usable, trace,
                                                                             trace_stride, niters, term, ws, nullptr, 0, nullptr, nullptr);
    B3D_LAUNCH_CHECK();
    return B3D_OK;
}

int win_launch_chains(const FitConst* dfc, const WinProblem* dprobs, int nprob, const int32_t* order_d, int nchain, int chain_len, const b3d_kp* kps,
                      const int32_t* nkp, double* R0, double* poses, double* W, double* r, double* p, uint8_t* usable, int32_t* niters, int32_t* term,
                      double* ws, cudaStream_t st) {
    if (nchain <= 0 || chain_len <= 0) return B3D_OK;
    B3D_CUDA(cudaFuncSetAttribute(fit_window_kernel<FW_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(WinSmem)));
    fit_window_kernel<FW_THREADS><<<2 * nchain, FW_THREADS, sizeof(WinSmem), st>>>(dfc, nprob, dprobs, kps, nkp, R0, poses, W, usable, nullptr, 0, niters, term, ws,
                                                                              order_d, chain_len, r, p);
    B3D_LAUNCH_CHECK();
    return B3D_OK;
}

}  // namespace b3d

using namespace b3d;

extern "C" {

int b3d_fit_window_chains_dev(const b3d_model* m, const b3d_fit_opts* opts, int32_t nchain, const int32_t* chain_ptr, const int32_t* first,
                              const int32_t* count, const b3d_kp* kps, const int32_t* nkp, double* R0, double* poses, double* W, double* r, double* p,
                              uint8_t* usable, int32_t* niters, int32_t* term, void* stream) {
    if (!m || !opts || !chain_ptr || !first || !count || !kps || !nkp || !R0 || !poses || !W) return set_error(B3D_EINVAL, "b3d_fit_window_chains: null argument");
    if (nchain < 0 || opts->max_iters < 0) return set_error(B3D_EINVAL, "b3d_fit_window_chains: negative size");
    if (nchain == 0) return B3D_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const int nprob = chain_ptr[nchain];
    if (chain_ptr[0] != 0 || nprob < 0) return set_error(B3D_EINVAL, "b3d_fit_window_chains: chain_ptr must start at 0 and ascend");
    int chain_len = 0, max_count = 0;
    for (int c = 0; c < nchain; ++c) {
        if (chain_ptr[c + 1] < chain_ptr[c]) return set_error(B3D_EINVAL, "b3d_fit_window_chains: chain_ptr must ascend");
        chain_len = std::max(chain_len, chain_ptr[c + 1] - chain_ptr[c]);
    }
    for (int q = 0; q < nprob; ++q) {
        if (count[q] <= 0 || first[q] < 0) return set_error(B3D_EINVAL, "b3d_fit_window_chains: problem %d has first=%d count=%d", q, first[q], count[q]);
        max_count = std::max(max_count, count[q]);
    }
    if (nprob == 0) return B3D_OK;
    if (max_count >= CH_MIN_FRAMES) return set_error(B3D_EINVAL, "b3d_fit_window_chains: windows of %d frames or more go through b3d_fit_windows_dev", (int)CH_MIN_FRAMES);
    const size_t slot = win_ws_doubles(max_count);
    std::vector<WinProblem> probs(nprob);
    std::vector<int32_t> order((size_t)nchain * chain_len, -1);
    for (int c = 0; c < nchain; ++c)
        for (int q = chain_ptr[c]; q < chain_ptr[c + 1]; ++q) {
            probs[q].first = first[q]; probs[q].count = count[q]; probs[q].ws_offset = (int64_t)((size_t)c * slot);
            order[(size_t)c * chain_len + (q - chain_ptr[c])] = q;
        }
    FitConst fc;
    if (int rc = build_fit_const(m, nullptr, opts, true, &fc)) return rc;
    FitConst* dfc = nullptr;
    WinProblem* dpr = nullptr;
    int32_t* dord = nullptr;
    double* ws = nullptr;
    B3D_CUDA(cudaMallocAsync(&dfc, sizeof(FitConst), st));
    B3D_CUDA(cudaMallocAsync(&dpr, sizeof(WinProblem) * nprob, st));
    B3D_CUDA(cudaMallocAsync(&dord, sizeof(int32_t) * order.size(), st));
    B3D_CUDA(cudaMallocAsync(&ws, sizeof(double) * slot * (size_t)nchain, st));
    B3D_CUDA(cudaMemcpyAsync(dfc, &fc, sizeof(FitConst), cudaMemcpyHostToDevice, st));
    B3D_CUDA(cudaMemcpyAsync(dpr, probs.data(), sizeof(WinProblem) * nprob, cudaMemcpyHostToDevice, st));
    B3D_CUDA(cudaMemcpyAsync(dord, order.data(), sizeof(int32_t) * order.size(), cudaMemcpyHostToDevice, st));
    B3D_CUDA(cudaStreamSynchronize(st));      // the staged copies must have left the host vectors
    if (int rc = win_launch_chains(dfc, dpr, nprob, dord, nchain, chain_len, kps, nkp, R0, poses, W, r, p, usable, niters, term, ws, st)) return rc;
    B3D_CUDA(cudaFreeAsync(ws, st));
    B3D_CUDA(cudaFreeAsync(dord, st));
    B3D_CUDA(cudaFreeAsync(dpr, st));
    B3D_CUDA(cudaFreeAsync(dfc, st));
    return B3D_OK;
}

// debug hooks (not part of include/b3d.h): per-phase cycle counters of block 0 of fit_window_kernel
int b3d_debug_win_timing(int on) {
    long long z[16] = {0};
    cudaMemcpyToSymbol(g_win_phase, z, sizeof(z));
    return cudaMemcpyToSymbol(g_win_phase_on, &on, sizeof(int)) == cudaSuccess ? 0 : 2;
}
// chains of at least this many frames take the grid kernel (never below CH_MIN_FRAMES: the workspace is sized by it); a huge value
// forces the cluster kernel everywhere
int b3d_debug_chain_min_frames(int frames) { g_chain_min_frames.store(frames); return 0; }
int b3d_debug_win_read(long long* out16) { return cudaMemcpyFromSymbol(out16, g_win_phase, 16 * sizeof(long long)) == cudaSuccess ? 0 : 2; }

int b3d_fit_windows_dev(const b3d_model* m, const b3d_fit_opts* opts, int32_t nprob, const int32_t* first, const int32_t* count,
                        const b3d_kp* kps, const int32_t* nkp, const double* R0, double* poses, double* W, uint8_t* usable,
                        b3d_iter_record* trace, int32_t trace_stride, int32_t* niters, int32_t* term, void* stream) {
    if (!m || !opts || !first || !count || !kps || !nkp || !R0 || !poses || !W) return set_error(B3D_EINVAL, "b3d_fit_windows: null argument");
    if (nprob < 0 || opts->max_iters < 0) return set_error(B3D_EINVAL, "b3d_fit_windows: negative size");
    if (trace && trace_stride <= 0) return set_error(B3D_EINVAL, "b3d_fit_windows: trace without trace_stride");
    if (nprob == 0) return B3D_OK;
    cudaStream_t st = (cudaStream_t)stream;
    std::vector<WinProblem> probs(nprob);
    size_t total = 0;
    for (int p = 0; p < nprob; ++p) {
        if (count[p] <= 0 || first[p] < 0) return set_error(B3D_EINVAL, "b3d_fit_windows: problem %d has first=%d count=%d", p, first[p], count[p]);
        probs[p].first = first[p]; probs[p].count = count[p]; probs[p].ws_offset = (int64_t)total;
        total += win_ws_doubles(count[p]);
    }
    FitConst fc;
    if (int rc = build_fit_const(m, nullptr, opts, true, &fc)) return rc;
    FitConst* dfc = nullptr;
    WinProblem* dpr = nullptr;
    double* ws = nullptr;
    B3D_CUDA(cudaMallocAsync(&dfc, sizeof(FitConst), st));
    B3D_CUDA(cudaMallocAsync(&dpr, sizeof(WinProblem) * nprob, st));
    B3D_CUDA(cudaMallocAsync(&ws, sizeof(double) * total, st));
    B3D_CUDA(cudaMemcpyAsync(dfc, &fc, sizeof(FitConst), cudaMemcpyHostToDevice, st));
    B3D_CUDA(cudaMemcpyAsync(dpr, probs.data(), sizeof(WinProblem) * nprob, cudaMemcpyHostToDevice, st));
    B3D_CUDA(cudaStreamSynchronize(st));      // probs / fc are stack objects: the staged copies must have left them
    if (int rc = win_launch(dfc, dpr, nprob, kps, nkp, R0, poses, W, usable, trace, trace_stride, niters, term, ws, st, probs.data())) return rc;
    B3D_CUDA(cudaFreeAsync(ws, st));
    B3D_CUDA(cudaFreeAsync(dpr, st));
    B3D_CUDA(cudaFreeAsync(dfc, st));
    return B3D_OK;
}

int b3d_fit_window(const b3d_model* m, const b3d_fit_opts* opts, int32_t F, const b3d_kp* kps, const int32_t* nkp, const double* R0,
                   double* poses, double* w, uint8_t* usable, b3d_iter_record* trace, int32_t trace_cap, int32_t* niters, int32_t* term) {
    if (!m || !opts || !kps || !nkp || !R0 || !poses || !w) return set_error(B3D_EINVAL, "b3d_fit_window: null argument");
    if (F <= 0) return F < 0 ? set_error(B3D_EINVAL, "b3d_fit_window: F < 0") : B3D_OK;
    const size_t nf = (size_t)F;
    if (int rc = validate_keypoints(kps, nkp, nf, "b3d_fit_window")) return rc;
    auto al = [](size_t x) { return (x + 255) / 256 * 256; };
    const size_t s_kp = al(sizeof(b3d_kp) * nf * MAXKP), s_nk = al(4 * nf), s_r0 = al(8 * nf * 9), s_po = al(8 * nf * NPF), s_w = al(8 * nf * NS),
                 s_tr = trace ? al(sizeof(b3d_iter_record) * (size_t)trace_cap) : 0;
    char* base = nullptr;
    B3D_CUDA(cudaMalloc(&base, s_kp + s_nk + s_r0 + s_po + s_w + s_tr + 256 * 3));
    char* q = base;
    auto take = [&](size_t s) { char* t = q; q += s; return t; };
    b3d_kp* d_kp = (b3d_kp*)take(s_kp); int32_t* d_nk = (int32_t*)take(s_nk); double* d_r0 = (double*)take(s_r0);
    double* d_po = (double*)take(s_po); double* d_w = (double*)take(s_w);
    b3d_iter_record* d_tr = trace ? (b3d_iter_record*)take(s_tr) : nullptr;
    uint8_t* d_us = (uint8_t*)take(256); int32_t* d_ni = (int32_t*)take(256); int32_t* d_te = (int32_t*)take(256);
    cudaError_t e = cudaMemcpy(d_kp, kps, sizeof(b3d_kp) * nf * MAXKP, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_nk, nkp, 4 * nf, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_r0, R0, 8 * nf * 9, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_po, poses, 8 * nf * NPF, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemset(d_w, 0, 8 * nf * NS);
    if (e == cudaSuccess) e = cudaMemcpy(d_w, w, 8 * NS, cudaMemcpyHostToDevice);     // the shared block = the window-front avatar's w
    if (e == cudaSuccess && d_tr) e = cudaMemset(d_tr, 0, s_tr);
    int rc = B3D_OK;
    if (e == cudaSuccess) {
        const int32_t first = 0, count = F;
        rc = b3d_fit_windows_dev(m, opts, 1, &first, &count, d_kp, d_nk, d_r0, d_po, d_w, d_us, d_tr, trace_cap, d_ni, d_te, nullptr);
        if (!rc) {
            e = cudaStreamSynchronize(nullptr);
            if (e == cudaSuccess) e = cudaMemcpy(poses, d_po, 8 * nf * NPF, cudaMemcpyDeviceToHost);
            if (e == cudaSuccess) e = cudaMemcpy(w, d_w, 8 * NS, cudaMemcpyDeviceToHost);
            if (e == cudaSuccess && usable) e = cudaMemcpy(usable, d_us, 1, cudaMemcpyDeviceToHost);
            if (e == cudaSuccess && trace) e = cudaMemcpy(trace, d_tr, sizeof(b3d_iter_record) * (size_t)trace_cap, cudaMemcpyDeviceToHost);
            if (e == cudaSuccess && niters) e = cudaMemcpy(niters, d_ni, 4, cudaMemcpyDeviceToHost);
            if (e == cudaSuccess && term) e = cudaMemcpy(term, d_te, 4, cudaMemcpyDeviceToHost);
        }
    }
    cudaFree(base);
    if (e != cudaSuccess) return set_error(B3D_ECUDA, "b3d_fit_window: %s", cudaGetErrorString(e));
    return rc;
}

}  // extern "C"
