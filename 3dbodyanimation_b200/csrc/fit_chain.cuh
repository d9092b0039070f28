// Entry (2), multi-frame, LONG chains (stage 1 of 3dba_multi: one OptimizeMultiFrame over all anchors, src/main_multi_frame.cpp:109-134,
// include/MultiFrameBA.h:33-177): the same Levenberg-Marquardt loop as fit_window_kernel, run by a GRID of CTAs (one per SM,
// cooperative launch) instead of one cluster, so that a chain of thousands of frames is not a serial sweep on two SMs.
//
//   * evaluation (residuals, Jacobians, per-frame normal equations): frames dealt to the CTAs in contiguous chunks (win_evaluate);
//   * the damped normal equations  [T C^T; C Hw] [y; y_w] = [gs; gs_w]  (T block tridiagonal, 76 x 76 blocks; C the 10-wide arrow):
//     BLOCK CYCLIC REDUCTION over the frames with 11 right-hand sides [gs | C^T].  Level l (stride s = 2^l) eliminates the frames
//     f = s (2 k + 1) all at once — phase 1, one CTA per eliminated frame: L = chol(A_f), U_a = L^-1 E_fa, U_c = L^-1 E_fc,
//     v = L^-1 R_f, new coupling E_ac = -U_a^T U_c; phase 2, one CTA per surviving frame: A_a -= U^T U, R_a -= U^T v for its one or
//     two eliminated neighbours (each block has ONE writer per phase: deterministic, no atomics) — then the stride doubles; log2 F
//     levels, back substitution in reverse.  The shape block closes the system: S_w = Hw - sum_f C_f X_f, y = X_0 - X_w y_w.
//     About 7x the flops of the sequential sweep per frame, but F / 2 of them run side by side.
//   * every CTA runs the scalar LM logic redundantly on the same reduced sums (sums over CTAs are taken in CTA order by every CTA:
//     identical decisions, no broadcast); grid barriers separate the phases.
// The LM loop itself (tolerances, radius updates, records) is fit_window_kernel's, statement for statement.
#pragma once

namespace b3d {

constexpr int CLD = NB + 2;          // 78: leading dimension of the shared-memory blocks (rows 16 B aligned for double2 loads)
constexpr int NRHS = NS + 1;         // right-hand sides carried through the reduction: gs and the 10 arrow columns
constexpr int RLD = CH_RLD;          // 12 (lm_device.cuh sizes the workspace with it)
static_assert(RLD == NRHS + 1, "right-hand side blocks are 76 x 12");
// CH_PART doubles of private exchange area per CTA (layout of WinWs::xc: cost / Hw / gw shares), CH_RED reduction slots per CTA,
// CH_MIN_FRAMES: lm_device.cuh

struct ChainBlocks {
    double A[NB][CLD], E1[NB][CLD], E2[NB][CLD];
    double R[NB][RLD], X1[NB][RLD], X2[NB][RLD];
    double invd[NB];
    double Sw[NS * NS], zw[NS], yw[NS];
    double red[40];
    double bc[4];                    // broadcast slots
    int flag;
};
struct ChainSmem {
    union { WinSmem w; ChainBlocks b; } u;
};

struct ChainWs {
    double *Aw, *EL, *ER;            // F x 76 x 76: working diagonal blocks (-> L), couplings to the left / right active neighbour (-> U)
    double *Rr, *Xs;                 // F x 76 x 12: right-hand sides (-> v), solutions
    double *part;                    // G x CH_PART
    double *red;                     // G x CH_RED
    unsigned* bar;                   // [0] arrivals, [1] generation, [2] failure flag of the current solve
};

// all CTAs of the (cooperative) grid: global writes before, reads after
__device__ __forceinline__ void grid_sync(unsigned* bar, unsigned nblocks) {
    __syncthreads();
    if (threadIdx.x == 0) {
        volatile unsigned* vgen = bar + 1;
        const unsigned gen = *vgen;
        __threadfence();
        if (atomicAdd(bar, 1u) == nblocks - 1) {
            atomicExch(bar, 0u);
            __threadfence();
            atomicAdd(bar + 1, 1u);
        } else {
            while (*vgen == gen) __nanosleep(32);
        }
        __threadfence();
    }
    __syncthreads();
}

// ---- dense 76 x 76 helpers on shared-memory blocks (T threads) ----
template <int T>
__device__ void blk_load(double (*dst)[CLD], const double* __restrict__ src) {     // row-major 76 x 76 from global
    for (int idx = threadIdx.x; idx < NB * NB; idx += T) dst[idx / NB][idx % NB] = src[idx];
}
template <int T>
__device__ void blk_store(double* __restrict__ dst, const double (*src)[CLD]) {
    for (int idx = threadIdx.x; idx < NB * NB; idx += T) dst[idx] = src[idx / NB][idx % NB];
}
template <int T>
__device__ void rhs_load(double (*dst)[RLD], const double* __restrict__ src) {     // 76 x 12 (12th column unused)
    for (int idx = threadIdx.x; idx < NB * RLD; idx += T) dst[idx / RLD][idx % RLD] = src[idx];
}
template <int T>
__device__ void rhs_store(double* __restrict__ dst, const double (*src)[RLD]) {
    for (int idx = threadIdx.x; idx < NB * RLD; idx += T) dst[idx] = src[idx / RLD][idx % RLD];
}

// In-place lower Cholesky of A (lower triangle read), 1 / L_jj -> invd, blocked over 4-column panels: every thread factors the
// 4 x 4 diagonal tile redundantly in registers (same shared-memory words for all: broadcast reads, no hand-over barrier), one thread
// per row solves the panel, one thread per 4 x 4 tile of the trailing lower triangle applies the rank-4 update.  Two barriers per
// panel.  Returns false on a non-positive pivot (uniform: every thread sees the same tile).
template <int T>
__device__ bool blk_chol(double (*A)[CLD], double* invd) {
    const int tid = threadIdx.x;
    constexpr int NT4 = NB / 4;                         // 19
    // this thread's tile of the lower triangle (tid < 190)
    int ti = 0;
    while ((ti + 1) * (ti + 2) / 2 <= tid) ++ti;
    const int tj = tid - ti * (ti + 1) / 2;
    const bool has_tile = tid < NT4 * (NT4 + 1) / 2;
    for (int kb = 0; kb < NT4; ++kb) {
        const int j0 = 4 * kb;
        double l[4][4], inv[4];
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int c = 0; c <= r; ++c) l[r][c] = A[j0 + r][j0 + c];
        bool bad = false;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const double piv = l[c][c];
            bad = bad || !(piv > 0.0);
            inv[c] = rsqrt(piv);
            l[c][c] = piv * inv[c];
#pragma unroll
            for (int r = c + 1; r < 4; ++r) l[r][c] *= inv[c];
#pragma unroll
            for (int c2 = c + 1; c2 < 4; ++c2)
#pragma unroll
                for (int r = c2; r < 4; ++r) l[r][c2] -= l[r][c] * l[c2][c];
        }
        if (bad) return false;
        // panel rows below the diagonal tile: x = a L_kk^-T
        for (int i = j0 + 4 + tid; i < NB; i += T) {
            const double x0 = A[i][j0] * inv[0];
            const double x1 = (A[i][j0 + 1] - x0 * l[1][0]) * inv[1];
            const double x2 = (A[i][j0 + 2] - x0 * l[2][0] - x1 * l[2][1]) * inv[2];
            const double x3 = (A[i][j0 + 3] - x0 * l[3][0] - x1 * l[3][1] - x2 * l[3][2]) * inv[3];
            A[i][j0] = x0; A[i][j0 + 1] = x1; A[i][j0 + 2] = x2; A[i][j0 + 3] = x3;
        }
        __syncthreads();                                 // every thread has read the diagonal tile; the panel is complete
        if (tid == 0) {
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                invd[j0 + r] = inv[r];
#pragma unroll
                for (int c = 0; c <= r; ++c) A[j0 + r][j0 + c] = l[r][c];
            }
        }
        if (has_tile && tj > kb) {
            double pr[4][4], pc[4][4];
#pragma unroll
            for (int a = 0; a < 4; ++a) {
                const double2 r01 = *reinterpret_cast<const double2*>(&A[4 * ti + a][j0]), r23 = *reinterpret_cast<const double2*>(&A[4 * ti + a][j0 + 2]);
                const double2 c01 = *reinterpret_cast<const double2*>(&A[4 * tj + a][j0]), c23 = *reinterpret_cast<const double2*>(&A[4 * tj + a][j0 + 2]);
                pr[a][0] = r01.x; pr[a][1] = r01.y; pr[a][2] = r23.x; pr[a][3] = r23.y;
                pc[a][0] = c01.x; pc[a][1] = c01.y; pc[a][2] = c23.x; pc[a][3] = c23.y;
            }
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    double v = A[4 * ti + a][4 * tj + b];
#pragma unroll
                    for (int r = 0; r < 4; ++r) v = fma(-pr[a][r], pc[b][r], v);
                    A[4 * ti + a][4 * tj + b] = v;
                }
        }
        __syncthreads();
    }
    return true;
}

// [E1 | E2 | R] <- L^-1 [E1 | E2 | R], blocked over 4-row panels of L: a thread per column solves the panel's 4 rows, a thread per
// 4 x 4 tile below applies the rank-4 update.  ntc tile columns: 19 of E1, then 19 of E2 (if with_e2), then 3 of R.
struct TileCol { double* base; int ld; };
__device__ __forceinline__ TileCol blk_tile_col(double (*E1)[CLD], double (*E2)[CLD], double (*R)[RLD], bool with_e2, int tc) {
    constexpr int NT4 = NB / 4;
    if (tc < NT4) return {&E1[0][4 * tc], CLD};
    if (with_e2) { if (tc < 2 * NT4) return {&E2[0][4 * (tc - NT4)], CLD}; return {&R[0][4 * (tc - 2 * NT4)], RLD}; }
    return {&R[0][4 * (tc - NT4)], RLD};
}
template <int T>
__device__ void blk_trsm(const double (*L)[CLD], const double* invd, double (*E1)[CLD], double (*E2)[CLD], double (*R)[RLD], bool with_e2) {
    const int tid = threadIdx.x;
    constexpr int NT4 = NB / 4;
    const int ntc = (with_e2 ? 2 * NT4 : NT4) + 3;
    for (int kb = 0; kb < NT4; ++kb) {
        const int j0 = 4 * kb;
        // (a) the panel's 4 rows of every column
        for (int col = tid; col < 4 * ntc; col += T) {
            const TileCol tcl = blk_tile_col(E1, E2, R, with_e2, col >> 2);
            double* b = tcl.base + (col & 3);
            const double x0 = b[(size_t)j0 * tcl.ld] * invd[j0];
            const double x1 = (b[(size_t)(j0 + 1) * tcl.ld] - L[j0 + 1][j0] * x0) * invd[j0 + 1];
            const double x2 = (b[(size_t)(j0 + 2) * tcl.ld] - L[j0 + 2][j0] * x0 - L[j0 + 2][j0 + 1] * x1) * invd[j0 + 2];
            const double x3 = (b[(size_t)(j0 + 3) * tcl.ld] - L[j0 + 3][j0] * x0 - L[j0 + 3][j0 + 1] * x1 - L[j0 + 3][j0 + 2] * x2) * invd[j0 + 3];
            b[(size_t)j0 * tcl.ld] = x0; b[(size_t)(j0 + 1) * tcl.ld] = x1; b[(size_t)(j0 + 2) * tcl.ld] = x2; b[(size_t)(j0 + 3) * tcl.ld] = x3;
        }
        __syncthreads();
        // (b) rows below: tile (tr, tc) -= L[rows of tr][panel] * B[panel rows][cols of tc]
        const int nrows = NT4 - 1 - kb;
        for (int u = tid; u < nrows * ntc; u += T) {
            const int tr = kb + 1 + u / ntc, tc = u % ntc;
            const TileCol tcl = blk_tile_col(E1, E2, R, with_e2, tc);
            double lv[4][4], bv[4][4];
#pragma unroll
            for (int a = 0; a < 4; ++a) {
                const double2 l01 = *reinterpret_cast<const double2*>(&L[4 * tr + a][j0]), l23 = *reinterpret_cast<const double2*>(&L[4 * tr + a][j0 + 2]);
                lv[a][0] = l01.x; lv[a][1] = l01.y; lv[a][2] = l23.x; lv[a][3] = l23.y;
                const double2 b01 = *reinterpret_cast<const double2*>(tcl.base + (size_t)(j0 + a) * tcl.ld), b23 = *reinterpret_cast<const double2*>(tcl.base + (size_t)(j0 + a) * tcl.ld + 2);
                bv[a][0] = b01.x; bv[a][1] = b01.y; bv[a][2] = b23.x; bv[a][3] = b23.y;
            }
#pragma unroll
            for (int a = 0; a < 4; ++a) {
                double2* row = reinterpret_cast<double2*>(tcl.base + (size_t)(4 * tr + a) * tcl.ld);
                double2 v01 = row[0], v23 = row[1];
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    v01.x = fma(-lv[a][r], bv[r][0], v01.x); v01.y = fma(-lv[a][r], bv[r][1], v01.y);
                    v23.x = fma(-lv[a][r], bv[r][2], v23.x); v23.y = fma(-lv[a][r], bv[r][3], v23.y);
                }
                row[0] = v01; row[1] = v23;
            }
        }
        __syncthreads();
    }
}

// R <- L^-T R (12 columns), blocked backwards over 4-row panels
template <int T>
__device__ void blk_trsm_t(const double (*L)[CLD], const double* invd, double (*R)[RLD]) {
    const int tid = threadIdx.x;
    constexpr int NT4 = NB / 4;
    for (int kb = NT4 - 1; kb >= 0; --kb) {
        const int j0 = 4 * kb;
        if (tid < RLD) {
            const int c = tid;
            const double x3 = R[j0 + 3][c] * invd[j0 + 3];
            const double x2 = (R[j0 + 2][c] - L[j0 + 3][j0 + 2] * x3) * invd[j0 + 2];
            const double x1 = (R[j0 + 1][c] - L[j0 + 3][j0 + 1] * x3 - L[j0 + 2][j0 + 1] * x2) * invd[j0 + 1];
            const double x0 = (R[j0][c] - L[j0 + 3][j0] * x3 - L[j0 + 2][j0] * x2 - L[j0 + 1][j0] * x1) * invd[j0];
            R[j0][c] = x0; R[j0 + 1][c] = x1; R[j0 + 2][c] = x2; R[j0 + 3][c] = x3;
        }
        __syncthreads();
        for (int idx = tid; idx < j0 * RLD; idx += T) {
            const int i = idx / RLD, c = idx % RLD;
            double v = R[i][c];
#pragma unroll
            for (int r = 0; r < 4; ++r) v = fma(-L[j0 + r][i], R[j0 + r][c], v);
            R[i][c] = v;
        }
        __syncthreads();
    }
}

// 4 x 4 register tile (ti, tj) of X^T Y for 76 x 76 blocks: acc[a][b] = sum_k X[k][4 ti + a] * Y[k][4 tj + b]
__device__ __forceinline__ void blk_tile_xty(const double (*X)[CLD], const double (*Y)[CLD], int ti, int tj, double (&acc)[4][4]) {
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
#pragma unroll 4
    for (int k = 0; k < NB; ++k) {
        const double2 x01 = *reinterpret_cast<const double2*>(&X[k][4 * ti]), x23 = *reinterpret_cast<const double2*>(&X[k][4 * ti + 2]);
        const double2 y01 = *reinterpret_cast<const double2*>(&Y[k][4 * tj]), y23 = *reinterpret_cast<const double2*>(&Y[k][4 * tj + 2]);
        const double xv[4] = {x01.x, x01.y, x23.x, x23.y}, yv[4] = {y01.x, y01.y, y23.x, y23.y};
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[a][b] = fma(xv[a], yv[b], acc[a][b]);
    }
}

// ---- the damped linear solve by block cyclic reduction: y (W.y) <- solution of (H_s + D^2 / radius) y = gs.  All CTAs. ----
template <int T>
__device__ bool chain_solve(const FitConst* __restrict__ C, int F, double radius, WinWs& W, ChainWs& CW, ChainBlocks& B, int cta, int G) {
    const int tid = threadIdx.x;
    const double lam2 = C->lambda_temp * C->lambda_temp;
    const bool temporal = C->lambda_temp > 0.0 && F > 1;
    const size_t BB = (size_t)NB * NB, RR = (size_t)NB * RLD;
    if (cta == 0 && tid == 0) CW.bar[2] = 0u;
    WinTimer ct;          // debug hook b3d_debug_win_timing: cycles of CTA 0 per part of the solve
    ct.start();
    // (0) working copies: A_f = Hd_f + D^2 / radius, couplings to the neighbours (diagonal matrices at level 0), R_f = [gs_f | Ha_f^T]
    for (int f = cta; f < F; f += G) {
        const double* Hd = W.Hd + (size_t)f * BB;
        double* Aw = CW.Aw + (size_t)f * BB;
        for (int idx = tid; idx < NB * NB; idx += T) {
            const int i = idx / NB, j = idx % NB;
            double v = Hd[idx];
            if (i == j) { const double l = sqrt(W.diag[(size_t)f * NB + i] / radius); v += l * l; }
            Aw[idx] = v;
        }
        double* EL = CW.EL + (size_t)f * BB;
        double* ER = CW.ER + (size_t)f * BB;
        for (int idx = tid; idx < NB * NB; idx += T) {
            const int i = idx / NB, j = idx % NB;
            double el = 0.0, er = 0.0;
            if (temporal && i == j && i >= 1) {
                if (f > 0) el = -lam2 * W.scale[(size_t)f * NB + i] * W.scale[(size_t)(f - 1) * NB + i];
                if (f < F - 1) er = -lam2 * W.scale[(size_t)f * NB + i] * W.scale[(size_t)(f + 1) * NB + i];
            }
            EL[idx] = el; ER[idx] = er;
        }
        const double* Ha = W.Ha + (size_t)f * NS * NB;
        double* Rr = CW.Rr + (size_t)f * RR;
        for (int idx = tid; idx < NB * RLD; idx += T) {
            const int i = idx / RLD, c = idx % RLD;
            Rr[idx] = c == 0 ? W.gs[(size_t)f * NB + i] : (c <= NS ? Ha[(c - 1) * NB + i] : 0.0);
        }
    }
    grid_sync(CW.bar, G);
    ct.mark(0);
    // (1) forward reduction
    int nlev = 0;
    for (int s = 1; s < F; s <<= 1, ++nlev) {
        // phase 1: eliminate f = s (2 t + 1)
        const int nodd = (F - s + 2 * s - 1) / (2 * s);
        for (int t = cta; t < nodd; t += G) {
            const int f = s * (2 * t + 1), a = f - s, c = f + s;
            const bool has_c = c < F;
            __syncthreads();
            blk_load<T>(B.A, CW.Aw + (size_t)f * BB);
            blk_load<T>(B.E1, CW.EL + (size_t)f * BB);
            if (has_c) blk_load<T>(B.E2, CW.ER + (size_t)f * BB);
            rhs_load<T>(B.R, CW.Rr + (size_t)f * RR);
            __syncthreads();
            ct.mark(1);
            const bool ok = blk_chol<T>(B.A, B.invd);
            if (!ok) { if (tid == 0) atomicExch(&CW.bar[2], 1u); continue; }     // uniform; the solve is void, the barriers go on
            __syncthreads();
            ct.mark(2);
            // U_a = L^-1 E_fa, U_c = L^-1 E_fc, v = L^-1 R
            blk_trsm<T>(B.A, B.invd, B.E1, B.E2, B.R, has_c);
            ct.mark(3);
            blk_store<T>(CW.Aw + (size_t)f * BB, B.A);
            for (int i = tid; i < NB; i += T) W.invd[(size_t)f * NB + i] = B.invd[i];
            blk_store<T>(CW.EL + (size_t)f * BB, B.E1);
            if (has_c) blk_store<T>(CW.ER + (size_t)f * BB, B.E2);
            rhs_store<T>(CW.Rr + (size_t)f * RR, B.R);
            if (has_c) {     // new coupling of the survivors: E_ac = -U_a^T U_c (block (a, c)) and its transpose (block (c, a))
                double acc[4][4];
                const int ti = tid / 19, tj = tid % 19;
                if (tid < 361) blk_tile_xty(B.E1, B.E2, ti, tj, acc);
                __syncthreads();                                   // the factor has left B.A: it becomes the staging tile of E_ac
                if (tid < 361) {
#pragma unroll
                    for (int x = 0; x < 4; ++x) {
                        *reinterpret_cast<double2*>(&B.A[4 * ti + x][4 * tj]) = make_double2(-acc[x][0], -acc[x][1]);
                        *reinterpret_cast<double2*>(&B.A[4 * ti + x][4 * tj + 2]) = make_double2(-acc[x][2], -acc[x][3]);
                    }
                }
                __syncthreads();
                double* Eac = CW.ER + (size_t)a * BB;
                double* Eca = CW.EL + (size_t)c * BB;
                for (int idx = tid; idx < NB * NB; idx += T) {     // both orientations as coalesced rows
                    Eac[idx] = B.A[idx / NB][idx % NB];
                    Eca[idx] = B.A[idx % NB][idx / NB];
                }
            }
            ct.mark(9);
        }
        ct.mark(9);
        grid_sync(CW.bar, G);
        ct.mark(10);
        // phase 2: the survivors a = 2 s t take the Schur updates of their eliminated neighbours
        const int nev = (F + 2 * s - 1) / (2 * s);
        for (int t = cta; t < nev; t += G) {
            const int a = 2 * s * t, fl = a - s, fr = a + s;
            double upd[4][4];
            int ti = 0;                                  // this thread's tile of the lower triangle (tid < 190): U^T U is symmetric
            while ((ti + 1) * (ti + 2) / 2 <= tid) ++ti;
            const int tj = tid - ti * (ti + 1) / 2;
            const bool has_tile = tid < 190;
#pragma unroll
            for (int x = 0; x < 4; ++x)
#pragma unroll
                for (int y = 0; y < 4; ++y) upd[x][y] = 0.0;
            double rupd[2] = {0.0, 0.0};     // this thread's entries of R_a: idx = tid and tid + T (76 x 11 = 836 <= 2 T)
            for (int side = 0; side < 2; ++side) {
                const int fo = side == 0 ? fl : fr;
                if (fo < 0 || fo >= F) continue;       // uniform
                __syncthreads();
                // U of the eliminated neighbour towards a: its right coupling if it sits on the left of a, else its left coupling
                blk_load<T>(B.E1, (side == 0 ? CW.ER : CW.EL) + (size_t)fo * BB);
                rhs_load<T>(B.R, CW.Rr + (size_t)fo * RR);
                __syncthreads();
                if (has_tile) {
                    double acc[4][4];
                    blk_tile_xty(B.E1, B.E1, ti, tj, acc);
#pragma unroll
                    for (int x = 0; x < 4; ++x)
#pragma unroll
                        for (int y = 0; y < 4; ++y) upd[x][y] += acc[x][y];
                }
                for (int q = 0; q < 2; ++q) {
                    const int idx = tid + q * T;
                    if (idx < NB * NRHS) {
                        const int i = idx / NRHS, c = idx % NRHS;
                        double sacc = 0.0;
                        for (int k = 0; k < NB; ++k) sacc += B.E1[k][i] * B.R[k][c];
                        rupd[q] += sacc;
                    }
                }
            }
            double* Aw = CW.Aw + (size_t)a * BB;
            __syncthreads();                                       // B.E1 has been read by every tile: B.A (unused in this phase) stages the update
            if (has_tile) {
#pragma unroll
                for (int x = 0; x < 4; ++x)
#pragma unroll
                    for (int y = 0; y < 4; ++y) {
                        B.A[4 * ti + x][4 * tj + y] = upd[x][y];
                        B.A[4 * tj + y][4 * ti + x] = upd[x][y];  // (diagonal tiles write their entries twice with the same value up to symmetry of U^T U)
                    }
            }
            __syncthreads();
            for (int idx = tid; idx < NB * NB; idx += T) Aw[idx] -= B.A[idx / NB][idx % NB];
            double* Rr = CW.Rr + (size_t)a * RR;
            for (int q = 0; q < 2; ++q) {
                const int idx = tid + q * T;
                if (idx < NB * NRHS) Rr[(idx / NRHS) * RLD + idx % NRHS] -= rupd[q];
            }
        }
        ct.mark(11);
        grid_sync(CW.bar, G);
        ct.mark(10);
    }
    // (2) root: frame 0
    if (cta == 0) {
        __syncthreads();
        blk_load<T>(B.A, CW.Aw);
        rhs_load<T>(B.R, CW.Rr);
        __syncthreads();
        const bool ok = blk_chol<T>(B.A, B.invd);
        if (!ok) { if (tid == 0) atomicExch(&CW.bar[2], 1u); }
        else {
            __syncthreads();
            blk_trsm<T>(B.A, B.invd, B.E1, B.E2, B.R, false);     // (E1 rides along: its content is not used)
            blk_trsm_t<T>(B.A, B.invd, B.R);
            rhs_store<T>(CW.Xs, B.R);
        }
    }
    grid_sync(CW.bar, G);
    ct.mark(12);
    // (3) back substitution, levels in reverse: X_f = L_f^-T (v_f - U_a X_a - U_c X_c)
    for (int lev = nlev - 1; lev >= 0; --lev) {
        const int s = 1 << lev;
        const int nodd = (F - s + 2 * s - 1) / (2 * s);
        for (int t = cta; t < nodd; t += G) {
            const int f = s * (2 * t + 1), a = f - s, c = f + s;
            const bool has_c = c < F;
            __syncthreads();
            blk_load<T>(B.A, CW.Aw + (size_t)f * BB);
            blk_load<T>(B.E1, CW.EL + (size_t)f * BB);
            if (has_c) blk_load<T>(B.E2, CW.ER + (size_t)f * BB);
            rhs_load<T>(B.R, CW.Rr + (size_t)f * RR);
            rhs_load<T>(B.X1, CW.Xs + (size_t)a * RR);
            if (has_c) rhs_load<T>(B.X2, CW.Xs + (size_t)c * RR);
            for (int i = tid; i < NB; i += T) B.invd[i] = W.invd[(size_t)f * NB + i];
            __syncthreads();
            double r[2] = {0.0, 0.0};
            for (int q = 0; q < 2; ++q) {
                const int idx = tid + q * T;
                if (idx < NB * NRHS) {
                    const int i = idx / NRHS, col = idx % NRHS;
                    double sacc = B.R[i][col];
                    for (int k = 0; k < NB; ++k) sacc -= B.E1[i][k] * B.X1[k][col];
                    if (has_c)
                        for (int k = 0; k < NB; ++k) sacc -= B.E2[i][k] * B.X2[k][col];
                    r[q] = sacc;
                }
            }
            __syncthreads();
            for (int q = 0; q < 2; ++q) {
                const int idx = tid + q * T;
                if (idx < NB * NRHS) B.R[idx / NRHS][idx % NRHS] = r[q];
            }
            __syncthreads();
            blk_trsm_t<T>(B.A, B.invd, B.R);
            rhs_store<T>(CW.Xs + (size_t)f * RR, B.R);
        }
        ct.mark(13);
        grid_sync(CW.bar, G);
        ct.mark(10);
    }
    // (4) the shape block: S_w = Hw + D_w^2 / radius - sum_f Ha_f X_f[:, 1:], z_w = gs_w - sum_f Ha_f X_f[:, 0]; partial sums per CTA
    {
        const int chunk = (F + G - 1) / G, fa = min(F, cta * chunk), fb = min(F, fa + chunk);
        double accw = 0.0;                       // thread (s, c): s < 10 rows of Ha, c < 11 columns of X
        const int sr = tid / NRHS, sc = tid % NRHS;
        if (tid < NS * NRHS)
            for (int f = fa; f < fb; ++f) {
                const double* Ha = W.Ha + (size_t)f * NS * NB + (size_t)sr * NB;
                const double* X = CW.Xs + (size_t)f * RR + sc;
                double sacc = 0.0;
                for (int k = 0; k < NB; ++k) sacc += Ha[k] * X[(size_t)k * RLD];
                accw += sacc;
            }
        if (tid < NS * NRHS) CW.part[(size_t)cta * CH_PART + 256 + tid] = accw;
    }
    grid_sync(CW.bar, G);
    const bool failed = *(volatile unsigned*)&CW.bar[2] != 0u;
    const size_t nw0 = (size_t)F * NB;
    if (tid < NS * NRHS) {
        double tot = 0.0;
        for (int g = 0; g < G; ++g) tot += CW.part[(size_t)g * CH_PART + 256 + tid];
        const int sr = tid / NRHS, sc = tid % NRHS;
        if (sc == 0) B.zw[sr] = W.gs[nw0 + sr] - tot;
        else {
            double v = W.Hw[sr * NS + (sc - 1)] - tot;
            if (sr == sc - 1) { const double l = sqrt(W.diag[nw0 + sr] / radius); v += l * l; }
            B.Sw[sr * NS + (sc - 1)] = v;
        }
    }
    __syncthreads();
    if (tid == 0) {     // 10 x 10 Cholesky by one thread, every CTA alike (win_arrow)
        int good = failed ? 0 : 1;
        for (int j = 0; j < NS && good; ++j) {
            double d = B.Sw[j * NS + j];
            for (int k = 0; k < j; ++k) d -= B.Sw[j * NS + k] * B.Sw[j * NS + k];
            if (!(d > 0.0)) { good = 0; break; }
            d = sqrt(d);
            B.Sw[j * NS + j] = d;
            for (int i = j + 1; i < NS; ++i) {
                double sacc = B.Sw[i * NS + j];
                for (int k = 0; k < j; ++k) sacc -= B.Sw[i * NS + k] * B.Sw[j * NS + k];
                B.Sw[i * NS + j] = sacc / d;
            }
        }
        if (good) {
            for (int i = 0; i < NS; ++i) { double sacc = B.zw[i]; for (int k = 0; k < i; ++k) sacc -= B.Sw[i * NS + k] * B.zw[k]; B.zw[i] = sacc / B.Sw[i * NS + i]; }
            for (int i = NS - 1; i >= 0; --i) { double sacc = B.zw[i]; for (int k = i + 1; k < NS; ++k) sacc -= B.Sw[k * NS + i] * B.yw[k]; B.yw[i] = sacc / B.Sw[i * NS + i]; }
        }
        B.flag = good;
    }
    __syncthreads();
    const bool ok = B.flag != 0;
    if (ok) {
        const int chunk = (F + G - 1) / G, fa = min(F, cta * chunk), fb = min(F, fa + chunk);
        for (int idx = fa * NB + tid; idx < fb * NB; idx += T) {
            const int f = idx / NB, i = idx % NB;
            const double* X = CW.Xs + (size_t)f * RR + (size_t)i * RLD;
            double v = X[0];
            for (int sidx = 0; sidx < NS; ++sidx) v -= X[1 + sidx] * B.yw[sidx];
            W.y[idx] = v;
        }
        if (cta == 0 && tid < NS) W.y[nw0 + tid] = B.yw[tid];
    }
    grid_sync(CW.bar, G);
    ct.mark(14);
    return ok;
}

// ---- the LM loop of one long chain on a cooperative grid of G CTAs ----
template <int T>
__global__ void __launch_bounds__(T, 1) fit_chain_kernel(const FitConst* __restrict__ C, WinProblem pr, int pb, const b3d_kp* __restrict__ kps_all,
                                                         const int32_t* __restrict__ nkp_all, const double* __restrict__ R0_all,
                                                         double* __restrict__ poses, double* __restrict__ Wshape, uint8_t* __restrict__ usable,
                                                         b3d_iter_record* __restrict__ trace, int trace_stride, int32_t* __restrict__ niters,
                                                         int32_t* __restrict__ term, double* __restrict__ workspace) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    ChainSmem& S = *reinterpret_cast<ChainSmem*>(smem_raw);
    const int tid = threadIdx.x, cta = blockIdx.x, G = gridDim.x;
    const int F = pr.count, n = F * NB + NS;
    const b3d_kp* kps = kps_all + (size_t)pr.first * MAXKP;
    const int32_t* nkp = nkp_all + pr.first;
    const double* R0 = R0_all + (size_t)pr.first * 9;
    double* xin = poses + (size_t)pr.first * NB;
    double* win = Wshape + (size_t)pr.first * NS;
    b3d_iter_record* tr = trace ? trace + (size_t)pb * trace_stride : nullptr;
    WinWs W;
    W.carve(workspace + pr.ws_offset, F);
    ChainWs CW;
    {
        double* p = W.xc + WIN_XC_DOUBLES;
        CW.Aw = W.Ld; CW.EL = W.Lo;
        CW.ER = p; p += (size_t)F * NB * NB;
        CW.Rr = p; p += (size_t)F * NB * RLD;
        CW.Xs = p; p += (size_t)F * NB * RLD;
        CW.part = p; p += (size_t)G * CH_PART;
        CW.red = p; p += (size_t)G * CH_RED;
        CW.bar = reinterpret_cast<unsigned*>(p);
    }
    WinWs Wl = W;                               // evaluation shares go to this CTA's private exchange area
    Wl.xc = CW.part + (size_t)cta * CH_PART;
    const int erank = cta == 0 ? 0 : 1;         // slot inside it; rank 0 carries the shape prior
    const double lam2 = C->lambda_temp * C->lambda_temp;
    const bool temporal = C->lambda_temp > 0.0 && F > 1;
    const int chunk = (F + G - 1) / G, fa = min(F, cta * chunk), fb = min(F, fa + chunk);   // this CTA's frames
    const size_t nw0 = (size_t)F * NB;
    if (tid < 16) win_phase_acc()[tid] = 0;
    __syncthreads();

    // sums / maxima over the grid: per-CTA values in CW.red (slot), summed in CTA order by every CTA
    auto grid_sum = [&](double v, int slot) -> double {
        const double b = block_sum<T>(v, S.u.b.red);
        if (tid == 0) CW.red[(size_t)cta * CH_RED + slot] = b;
        grid_sync(CW.bar, G);
        double tot = 0.0;
        for (int g = 0; g < G; ++g) tot += *(volatile double*)&CW.red[(size_t)g * CH_RED + slot];
        return tot;
    };
    auto grid_max = [&](double v, int slot) -> double {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_down_sync(0xffffffffu, v, o));
        __syncthreads();
        if ((tid & 31) == 0) S.u.b.red[tid >> 5] = v;
        __syncthreads();
        double m = 0.0;
        for (int i = 0; i < T / 32; ++i) m = fmax(m, S.u.b.red[i]);
        __syncthreads();
        if (tid == 0) CW.red[(size_t)cta * CH_RED + slot] = m;
        grid_sync(CW.bar, G);
        double tot = 0.0;
        for (int g = 0; g < G; ++g) tot = fmax(tot, *(volatile double*)&CW.red[(size_t)g * CH_RED + slot]);
        return tot;
    };
    // evaluation of the CTA's frames; the cost is summed over the grid.  With the Jacobian: CTA 0 then adds the shape-block shares
    auto evaluate_all = [&](const double* xvec, bool jac, bool scaled) -> double {
        const double part = jac ? win_evaluate<T, true>(C, F, xvec, kps, nkp, R0, Wl, S.u.w, fa, fb, erank, scaled)
                                : win_evaluate<T, false>(C, F, xvec, kps, nkp, R0, Wl, S.u.w, fa, fb, erank, false);
        __syncthreads();
        const double total = grid_sum(tid == 0 ? part : 0.0, jac ? 0 : 1);      // (the barrier inside also publishes the shares)
        if (jac && cta == 0) {
            const double bs2 = C->beta_shape * C->beta_shape;
            const bool shape_prior = C->beta_shape > 0.0;
            for (int idx = tid; idx < NS * NS; idx += T) {
                double a = 0.0;
                for (int g = 0; g < G; ++g) a += CW.part[(size_t)g * CH_PART + XC_HW + (g == 0 ? 0 : NS * NS) + idx];
                W.Hw[idx] = a + ((shape_prior && idx / NS == idx % NS) ? bs2 : 0.0);
            }
            if (tid < NS) {
                double a = 0.0;
                for (int g = 0; g < G; ++g) a += CW.part[(size_t)g * CH_PART + XC_GW + (g == 0 ? 0 : NS) + tid];
                W.g[nw0 + tid] = a + (shape_prior ? bs2 * xvec[nw0 + tid] : 0.0);
            }
        }
        if (jac) grid_sync(CW.bar, G);
        return total;
    };

    for (int k = fa * NB + tid; k < fb * NB; k += T) W.x[k] = xin[k];
    if (cta == 0 && tid < NS) W.x[nw0 + tid] = win[tid];
    grid_sync(CW.bar, G);

    double radius = wInitialRadius, decrease_factor = 2.0, x_cost = 0.0, cand_cost = 0.0, mcc = 0.0;
    bool reuse_diagonal = false, step_successful = true;
    int num_invalid = 0, iteration = 0, nrec = 0, termination = 0;
    b3d_iter_record it;
    it.cost = it.cost_change = it.gradient_max_norm = it.step_norm = it.relative_decrease = it.radius = 0.0;
    it.flags = 0; it.aux = -1;

    auto eval_grad_jac = [&](bool first) -> bool {
        x_cost = evaluate_all(W.x, true, !first);
        if (!isfinite(x_cost)) return false;
        it.cost = x_cost;
        if (first) {
            for (int k = fa * NB + tid; k < fb * NB; k += T) {
                const double d = W.Hd[(size_t)(k / NB) * NB * NB + (k % NB) * (NB + 1)];
                W.scale[k] = 1.0 / (1.0 + sqrt(d));
            }
            if (cta == 0 && tid < NS) W.scale[nw0 + tid] = 1.0 / (1.0 + sqrt(W.Hw[tid * (NS + 1)]));
            grid_sync(CW.bar, G);
            for (size_t idx = (size_t)fa * NB * NB + tid; idx < (size_t)fb * NB * NB; idx += T) {
                const int f = (int)(idx / (NB * NB)), r = (int)(idx % (NB * NB)) / NB, c = (int)(idx % NB);
                W.Hd[idx] *= W.scale[(size_t)f * NB + r] * W.scale[(size_t)f * NB + c];
            }
            for (size_t idx = (size_t)fa * NS * NB + tid; idx < (size_t)fb * NS * NB; idx += T) {
                const int f = (int)(idx / (NS * NB)), sidx = (int)(idx % (NS * NB)) / NB, c = (int)(idx % NB);
                W.Ha[idx] *= W.scale[nw0 + sidx] * W.scale[(size_t)f * NB + c];
            }
            for (int k = fa * NB + tid; k < fb * NB; k += T) W.gs[k] = W.g[k] * W.scale[k];
        }
        if (cta == 0) {
            if (tid < NS * NS) W.Hw[tid] *= W.scale[nw0 + tid / NS] * W.scale[nw0 + tid % NS];
            if (tid < NS) W.gs[nw0 + tid] = W.g[nw0 + tid] * W.scale[nw0 + tid];
        }
        double gm = 0.0;
        for (int k = fa * NB + tid; k < fb * NB; k += T) gm = fmax(gm, fabs(W.g[k]));
        if (cta == 0 && tid < NS) gm = fmax(gm, fabs(W.g[nw0 + tid]));
        it.gradient_max_norm = grid_max(gm, 2);     // (its barrier also publishes the scaled shape block)
        return true;
    };
    auto push = [&]() {
        it.radius = radius;
        it.aux = -1;
        if (tr && nrec < trace_stride && tid == 0 && cta == 0) tr[nrec] = it;
        ++nrec;
    };
    auto norm_of = [&](const double* v, int slot) -> double {
        double a = 0.0;
        for (int k = fa * NB + tid; k < fb * NB; k += T) a += v[k] * v[k];
        if (cta == 0 && tid < NS) a += v[nw0 + tid] * v[nw0 + tid];
        return sqrt(grid_sum(a, slot));
    };

    WinTimer kt;
    kt.start();
    double x_norm = norm_of(W.x, 3);
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
                for (int k = fa * NB + tid; k < fb * NB; k += T) {
                    const double d = W.Hd[(size_t)(k / NB) * NB * NB + (k % NB) * (NB + 1)];
                    W.diag[k] = fmin(fmax(d, wMinDiagonal), wMaxDiagonal);
                }
                if (cta == 0 && tid < NS) W.diag[nw0 + tid] = fmin(fmax(W.Hw[tid * (NS + 1)], wMinDiagonal), wMaxDiagonal);
            }
            grid_sync(CW.bar, G);
            reuse_diagonal = true;
            bool step_valid = false;
            const bool solved = chain_solve<T>(C, F, radius, W, CW, S.u.b, cta, G);
            if (solved) {
                // model_cost_change = y^T gs - y^T H_s y / 2 (step = -y), summed over the CTAs' frames
                double v = 0.0, fin = 0.0;
                for (size_t idx = (size_t)fa * NB * NB + tid; idx < (size_t)fb * NB * NB; idx += T) {
                    const int f = (int)(idx / (NB * NB)), r = (int)(idx % (NB * NB)) / NB, c = (int)(idx % NB);
                    v -= 0.5 * W.Hd[idx] * W.y[(size_t)f * NB + r] * W.y[(size_t)f * NB + c];
                }
                for (size_t idx = (size_t)fa * NS * NB + tid; idx < (size_t)fb * NS * NB; idx += T) {
                    const int f = (int)(idx / (NS * NB)), sidx = (int)(idx % (NS * NB)) / NB, c = (int)(idx % NB);
                    v -= W.Ha[idx] * W.y[nw0 + sidx] * W.y[(size_t)f * NB + c];
                }
                if (cta == 0 && tid < NS * NS) v -= 0.5 * W.Hw[tid] * W.y[nw0 + tid / NS] * W.y[nw0 + tid % NS];
                if (temporal) {      // the pair (f, f + 1) belongs to the owner of f
                    const int p_hi = min(fb, F - 1);
                    for (int idx = tid; idx < (p_hi - fa) * (NB - 1); idx += T) {
                        const int f = fa + idx / (NB - 1), k = 1 + idx % (NB - 1);
                        const size_t a = (size_t)f * NB + k, b = a + NB;
                        v += lam2 * W.scale[a] * W.scale[b] * W.y[a] * W.y[b];
                    }
                }
                for (int k = fa * NB + tid; k < fb * NB; k += T) { v += W.y[k] * W.gs[k]; fin += isfinite(W.y[k]) ? 0.0 : 1.0; }
                if (cta == 0 && tid < NS) { v += W.y[nw0 + tid] * W.gs[nw0 + tid]; fin += isfinite(W.y[nw0 + tid]) ? 0.0 : 1.0; }
                mcc = grid_sum(v, 4);
                const double bad = grid_sum(fin, 5);
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
            for (int k = fa * NB + tid; k < fb * NB + (cta == 0 ? 0 : 0); k += T) {
                const double d = -W.y[k] * W.scale[k];
                const double c = W.x[k] + d;
                W.cand[k] = c;
                const double e = W.x[k] - c;
                sn += e * e;
            }
            if (cta == 0 && tid < NS) {
                const size_t k = nw0 + tid;
                const double d = -W.y[k] * W.scale[k];
                const double c = W.x[k] + d;
                W.cand[k] = c;
                const double e = W.x[k] - c;
                sn += e * e;
            }
            it.flags = 1;
            it.step_norm = sqrt(grid_sum(sn, 6));          // (its barrier publishes the candidate)
            cand_cost = evaluate_all(W.cand, false, false);
            if (!isfinite(cand_cost)) cand_cost = DBL_MAX;
            if (it.step_norm <= wParameterTolerance * (x_norm + wParameterTolerance)) { termination = B3D_TERM_PARAM_TOL; break; }
            it.cost_change = x_cost - cand_cost;
            if (fabs(it.cost_change) <= wFunctionTolerance * x_cost) { termination = B3D_TERM_FUNC_TOL; break; }
            it.relative_decrease = it.cost_change / mcc;
            if (it.relative_decrease > wMinRelativeDecrease) {
                for (int k = fa * NB + tid; k < fb * NB; k += T) W.x[k] = W.cand[k];
                if (cta == 0 && tid < NS) W.x[nw0 + tid] = W.cand[nw0 + tid];
                x_norm = norm_of(W.x, 7);                   // (its barrier publishes x)
                if (!eval_grad_jac(false)) { termination = B3D_TERM_EVAL_FAILURE; break; }
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
    grid_sync(CW.bar, G);
    kt.mark(15);
    if (cta == 0 && tid < 16 && g_win_phase_on) g_win_phase[tid] += win_phase_acc()[tid];
    for (int k = fa * NB + tid; k < fb * NB; k += T) xin[k] = W.x[k];
    if (cta == 0) {
        if (tid < NS) win[tid] = W.x[nw0 + tid];
        if (tid == 0) {
            if (usable) usable[pb] = (termination >= B3D_TERM_PARAM_TOL && termination <= B3D_TERM_MIN_RADIUS) ? 1 : 0;
            if (niters) niters[pb] = nrec;
            if (term) term[pb] = termination;
        }
    }
}

}  // namespace b3d
