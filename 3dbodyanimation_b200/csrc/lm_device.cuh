// Device-side building blocks shared by the single-frame and window LM kernels.
//
// Residual model (reference: include/Sim3BA.h:34-88 ReprojCost, :126-227 ReprojCostShape; analytic Jacobians replace
// ceres::DynamicAutoDiffCostFunction, whose Jets give the exact derivative of the same expressions):
//   chain of keypoint joint jid:  a_0 = jid, a_i = parent(a_{i-1}) while that parent is neither -1 nor 0
//   q_0 = bone(a_0);  q_i = Rot(aa_{a_i}) q_{i-1} + bone(a_i);  pc = R0 q_d;  X' = Rot(rootAA) pc;  X = s X' + t
//   bone(j) = offset_j + (JSR_j - JSR_parent(j)) w          (shape term only when the functor was given jointShapeReg)
//   r = (fx X0/X2 + cx - u, fy X1/X2 + cy - v),  robustified by Huber(a) via the Corrector's sqrt(rho') scaling.
// Rot follows ceres::AngleAxisRotatePoint: Rodrigues for theta^2 > DBL_EPSILON, first-order p + aa x p otherwise.
#pragma once
#include <cfloat>

#include "internal.hpp"

namespace b3d {

constexpr int MAXCHAIN = 8;                    // a_0 .. a_7 (hands have 7 strict non-root ancestors)
constexpr int MAXNNZ = 7 + 3 * (MAXCHAIN - 1) + NS;   // 38 non-zero Jacobian columns per keypoint row pair
constexpr int NNZPAD = 40;

// Launch-constant data of a fit, built on the host, read-only on the device.
struct FitConst {
    int32_t parent[NJ];
    int32_t chain_len[NJ];                 // d + 1
    int32_t chain[NJ][MAXCHAIN];           // a_0 .. a_d
    double offset[NJ][3];
    double DB[NJ][3][NS];                  // d bone / d w; all zero when the shape term is off
    double cam[4];                         // fx fy cx cy
    double huber, beta_pose, beta_shape, lambda_temp, scale_lo, scale_hi;
    int32_t mode, max_iters, with_shape, has_w, use_gmm, n;   // n = number of free parameters
    int32_t free_idx[NFULL];               // free -> full (86-layout)
    int32_t col_of[NFULL];                 // full -> free, -1 = held constant
    int32_t gmm_ncomp, pad;
    int8_t chain_pos[NJ][NJ];              // chain_pos[jid][a] = i with chain[jid][i] == a, or -1 (dense Jacobian rows, eval_reproj_dense)
    const double* gmm_mean;                // ncomp x 69
    const double* gmm_L;                   // ncomp x 69 x 69, lower, row-major  (prec_cho)
    const double* gmm_L70;                 // the same with rows padded to 70 doubles
    const double* gmm_LT;                  // ncomp x 69 x 69, = L^T row-major
    const double* gmm_P;                   // ncomp x 69 x 69, = L L^T
    const double* gmm_consts;              // ncomp
    const double* gmm_sigma;               // ncomp: upper bound of the spectral norm of L_i
};

// host side (fit_single.cu): fills the launch-constant block
int build_fit_const(const b3d_model* m, const b3d_gmm* gmm, const b3d_fit_opts* o, bool multi, FitConst* fc);

// one multi-frame problem of a batched launch (fit_window.cu)
struct WinProblem {
    int32_t first, count;     // frames [first, first + count) of the frame arrays; the shared w block is row `first` of W
    int64_t ws_offset;        // doubles into the workspace
};
// exchange area between the two CTAs of a problem's cluster (fit_window.cu), at the end of the per-problem workspace
constexpr int WIN_XC_DOUBLES = 7424;
// per-problem global workspace, in doubles
// chains from CH_MIN_FRAMES frames on are solved by a cooperative grid (fit_chain.cuh: block cyclic reduction); their workspace also
// holds the reduction's working blocks and one exchange area per CTA (at most CH_MAX_GRID CTAs)
constexpr int CH_MIN_FRAMES = 48, CH_MAX_GRID = 160, CH_PART = 512, CH_RED = 16, CH_RLD = NS + 2;
__host__ __device__ inline size_t win_ws_base_doubles(int F) {
    const size_t n = (size_t)F * NPF + NS;
    return 10 * n + (size_t)F * (3 * NPF * NPF + 2 * NS * NPF) + 2 * NS * NS + (size_t)F * NPF + WIN_XC_DOUBLES;
}
__host__ __device__ inline size_t win_ws_doubles(int F) {
    size_t total = win_ws_base_doubles(F);
    if (F >= CH_MIN_FRAMES) total += (size_t)F * NPF * NPF + 2 * (size_t)F * NPF * CH_RLD + (size_t)CH_MAX_GRID * (CH_PART + CH_RED) + 64;
    return total;
}
// doubles from the start of a chain problem's workspace to its grid-barrier words, for a grid of G CTAs
__host__ __device__ inline size_t chain_bar_offset(int F, int G) {
    return win_ws_base_doubles(F) + (size_t)F * NPF * NPF + 2 * (size_t)F * NPF * CH_RLD + (size_t)G * (CH_PART + CH_RED);
}
// enqueue the LM solves of nprob problems (device-resident constants / problem table / workspace); no synchronisation
// hprobs (optional): the same problem table on the host; with it, launches made only of long chains take the grid kernel
int win_launch(const FitConst* dfc, const WinProblem* dprobs, int nprob, const b3d_kp* kps, const int32_t* nkp, const double* R0, double* poses,
               double* W, uint8_t* usable, b3d_iter_record* trace, int trace_stride, int32_t* niters, int32_t* term, double* ws, cudaStream_t st,
               const WinProblem* hprobs = nullptr);

// chain mode: cluster c solves the problems order[c * chain_len + 0 .. chain_len) (-1 = none) one after the other, each followed by the
// write-back of its frames into R0 / r / p (problems of one chain share a workspace slot; different chains must not)
int win_launch_chains(const FitConst* dfc, const WinProblem* dprobs, int nprob, const int32_t* order_d, int nchain, int chain_len, const b3d_kp* kps,
                      const int32_t* nkp, double* R0, double* poses, double* W, double* r, double* p, uint8_t* usable, int32_t* niters, int32_t* term,
                      double* ws, cudaStream_t st);

// ---------------------------------------------------------------------------------------------------------------
// rotations
// ---------------------------------------------------------------------------------------------------------------
// R (row-major 3x3) of ceres::AngleAxisRotatePoint(aa, .), and optionally dR[k] = dR/daa_k (3 row-major 3x3).
template <bool JAC>
__device__ __forceinline__ void aa_rotation(const double* aa, double* R, double* dR) {
    const double a0 = aa[0], a1 = aa[1], a2 = aa[2];
    const double th2 = a0 * a0 + a1 * a1 + a2 * a2;
    if (th2 > DBL_EPSILON) {
        const double th = sqrt(th2), ti = 1.0 / th;
        double s, c;
        sincos(th, &s, &c);
        const double w[3] = {a0 * ti, a1 * ti, a2 * ti};
        const double oc = 1.0 - c;
        R[0] = c + oc * w[0] * w[0];         R[1] = -s * w[2] + oc * w[0] * w[1];  R[2] = s * w[1] + oc * w[0] * w[2];
        R[3] = s * w[2] + oc * w[1] * w[0];  R[4] = c + oc * w[1] * w[1];          R[5] = -s * w[0] + oc * w[1] * w[2];
        R[6] = -s * w[1] + oc * w[2] * w[0]; R[7] = s * w[0] + oc * w[2] * w[1];   R[8] = c + oc * w[2] * w[2];
        if (JAC) {
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const double wk = w[k];
                // dw = (e_k - w w_k) / theta
                double dw[3];
#pragma unroll
                for (int i = 0; i < 3; ++i) dw[i] = ((i == k ? 1.0 : 0.0) - w[i] * wk) * ti;
                double* D = dR + 9 * k;
                const double cs = c * wk, sw = s * wk;
                // d/daa_k [ c I + s [w]x + (1-c) w w^T ]
#pragma unroll
                for (int i = 0; i < 3; ++i)
#pragma unroll
                    for (int j = 0; j < 3; ++j)
                        D[3 * i + j] = (i == j ? -sw : 0.0) + sw * w[i] * w[j] + oc * (dw[i] * w[j] + w[i] * dw[j]);
                // + c w_k [w]x + s [dw]x
                const double x0 = cs * w[0] + s * dw[0], x1 = cs * w[1] + s * dw[1], x2 = cs * w[2] + s * dw[2];
                D[1] -= x2; D[2] += x1; D[3] += x2; D[5] -= x0; D[6] -= x1; D[7] += x0;
            }
        }
    } else {
        R[0] = 1.0; R[1] = -a2; R[2] = a1;
        R[3] = a2;  R[4] = 1.0; R[5] = -a0;
        R[6] = -a1; R[7] = a0;  R[8] = 1.0;
        if (JAC) {
#pragma unroll
            for (int i = 0; i < 27; ++i) dR[i] = 0.0;
            // d(p + aa x p)/daa_k = e_k x p  =>  dR_k = [e_k]x
            dR[5] = -1.0; dR[7] = 1.0;            // k = 0
            dR[9 + 2] = 1.0; dR[9 + 6] = -1.0;    // k = 1
            dR[18 + 1] = -1.0; dR[18 + 3] = 1.0;  // k = 2
        }
    }
}

__device__ __forceinline__ void mat3_vec(const double* M, const double* v, double* o) {
    const double x = v[0], y = v[1], z = v[2];
    o[0] = M[0] * x + M[1] * y + M[2] * z;
    o[1] = M[3] * x + M[4] * y + M[5] * z;
    o[2] = M[6] * x + M[7] * y + M[8] * z;
}

// N (2x3 row-major) <- N * M (3x3 row-major)
__device__ __forceinline__ void n23_mul(const double* N, const double* M, double* o) {
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) o[3 * r + c] = N[3 * r] * M[c] + N[3 * r + 1] * M[3 + c] + N[3 * r + 2] * M[6 + c];
}

// ---------------------------------------------------------------------------------------------------------------
// block-wide sum in a fixed order (deterministic): every thread gets the result.  red: >= 33 doubles of shared memory.
// ---------------------------------------------------------------------------------------------------------------
template <int T>
__device__ __forceinline__ double block_sum(double v, double* red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();   // red may still be read from a previous call
    if (lane == 0) red[warp] = v;
    __syncthreads();
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < T / 32; ++i) s += red[i];
    return s;
}

// ---------------------------------------------------------------------------------------------------------------
// per-frame evaluation scratch (shared memory)
// ---------------------------------------------------------------------------------------------------------------
// temporaries of one evaluation: dead once eval_reproj returns, so callers may alias them with other storage
struct EvalTmp {
    double R[NJ][9];              // R[0] = Rot(rootAA), R[j] = Rot(jointAA_j)
    double dR[NJ][27];            // derivative matrices (JAC only)
    double bone[NJ][3];
    double q[MAXKP][MAXCHAIN][3]; // chain positions q_0 .. q_d
    double pc[MAXKP][3];          // R0 q_d
    double Xp[MAXKP][3];          // Rot(rootAA) pc
    double N[MAXKP][MAXCHAIN + 1][6];   // N[.][0] = d r / d X (2x3); N[.][1 + i] = d r / d q_i
};
// results of one evaluation
struct EvalOut {
    double res[MAXKP][2];         // robustified residuals
    double kcost[MAXKP];          // rho/2 per keypoint
    double J[MAXKP][2][NNZPAD];   // robustified Jacobian entries, local column slots (JAC only)
    int32_t jcol[MAXKP][NNZPAD];  // full-layout column of each slot
    int32_t jnnz[MAXKP];
};

// Evaluate all reprojection blocks of one frame.  x: 86-layout parameters in shared memory.  All T threads call this.
// After return (and a __syncthreads inside), O.res / O.kcost are valid; with JAC also O.J / O.jcol / O.jnnz.
template <int T, bool JAC>
__device__ void eval_reproj(const FitConst* __restrict__ C, const double* x, const b3d_kp* __restrict__ kps, int nkp,
                            const double* R0 /*col-major 3x3, shared or global*/, EvalTmp& S, EvalOut& O) {
    const int tid = threadIdx.x;
    // phase A: joint rotations (24 tasks) and bone vectors (72 tasks)
    for (int t = tid; t < NJ + 3 * NJ; t += T) {
        if (t < NJ) {
            const double* aa = t == 0 ? x + IDX_RAA : x + IDX_JAA + 3 * (t - 1);
            aa_rotation<JAC>(aa, S.R[t], S.dR[t]);
        } else {
            const int j = (t - NJ) / 3, c = (t - NJ) % 3;
            double b = C->offset[j][c];
            if (C->with_shape) {
#pragma unroll
                for (int s = 0; s < NS; ++s) b += C->DB[j][c][s] * x[IDX_W + s];
            }
            S.bone[j][c] = b;
        }
    }
    __syncthreads();
    // phase B1: one thread per keypoint — forward chain, residual, Huber, row-vector back-propagation
    if (tid < nkp) {
        const int k = tid;
        const int jid = clamp_jid(kps[k].jid);
        const int len = C->chain_len[jid];
        double q[3] = {S.bone[jid][0], S.bone[jid][1], S.bone[jid][2]};
        S.q[k][0][0] = q[0]; S.q[k][0][1] = q[1]; S.q[k][0][2] = q[2];
        for (int i = 1; i < len; ++i) {
            const int a = C->chain[jid][i];
            double o[3];
            mat3_vec(S.R[a], q, o);
            q[0] = o[0] + S.bone[a][0]; q[1] = o[1] + S.bone[a][1]; q[2] = o[2] + S.bone[a][2];
            S.q[k][i][0] = q[0]; S.q[k][i][1] = q[1]; S.q[k][i][2] = q[2];
        }
        double pc[3], Xp[3];
#pragma unroll
        for (int r = 0; r < 3; ++r) pc[r] = R0[r] * q[0] + R0[3 + r] * q[1] + R0[6 + r] * q[2];
        mat3_vec(S.R[0], pc, Xp);
        const double s = x[IDX_S];
        const double X0 = s * Xp[0] + x[IDX_T], X1 = s * Xp[1] + x[IDX_T + 1], X2 = s * Xp[2] + x[IDX_T + 2];
        const double iz = 1.0 / X2;
        double r0 = C->cam[0] * X0 * iz + C->cam[2] - kps[k].u;
        double r1 = C->cam[1] * X1 * iz + C->cam[3] - kps[k].v;
        // HuberLoss(a) + Corrector (rho'' <= 0 branch)
        const double sq = r0 * r0 + r1 * r1, b = C->huber * C->huber;
        double rho0, rho1;
        if (sq > b) { const double rr = sqrt(sq); rho0 = 2.0 * C->huber * rr - b; rho1 = fmax(DBL_MIN, C->huber / rr); }
        else { rho0 = sq; rho1 = 1.0; }
        const double sc = sqrt(rho1);
        O.res[k][0] = r0 * sc; O.res[k][1] = r1 * sc;
        O.kcost[k] = 0.5 * rho0;
        if (JAC) {
#pragma unroll
            for (int c = 0; c < 3; ++c) { S.pc[k][c] = pc[c]; S.Xp[k][c] = Xp[c]; }
            double N[6];
            N[0] = sc * C->cam[0] * iz; N[1] = 0.0;                  N[2] = -sc * C->cam[0] * X0 * iz * iz;
            N[3] = 0.0;                 N[4] = sc * C->cam[1] * iz;  N[5] = -sc * C->cam[1] * X1 * iz * iz;
#pragma unroll
            for (int e = 0; e < 6; ++e) S.N[k][0][e] = N[e];
            // d r / d q_d = s N Rot(rootAA) R0
            double A[6], Bm[6];
            n23_mul(N, S.R[0], A);
            const double R0rm[9] = {R0[0], R0[3], R0[6], R0[1], R0[4], R0[7], R0[2], R0[5], R0[8]};
            n23_mul(A, R0rm, Bm);
#pragma unroll
            for (int e = 0; e < 6; ++e) Bm[e] *= s;
            // S.N[k][1 + i] = d r / d q_i
            for (int i = len - 1; i >= 0; --i) {
#pragma unroll
                for (int e = 0; e < 6; ++e) S.N[k][1 + i][e] = Bm[e];
                if (i > 0) {
                    n23_mul(Bm, S.R[C->chain[jid][i]], A);
#pragma unroll
                    for (int e = 0; e < 6; ++e) Bm[e] = A[e];
                }
            }
            O.jnnz[k] = 7 + 3 * (len - 1) + (C->with_shape ? NS : 0);
        }
    }
    __syncthreads();
    if (!JAC) return;
    // phase B2: one task per (keypoint, non-zero column)
    for (int t = tid; t < nkp * NNZPAD; t += T) {
        const int k = t / NNZPAD, slot = t % NNZPAD;
        const int jid = clamp_jid(kps[k].jid);
        const int len = C->chain_len[jid];
        const int naa = 3 * (len - 1);
        if (slot >= 7 + naa + (C->with_shape ? NS : 0)) continue;
        const double* N0 = S.N[k][0];
        double j0, j1;
        int col;
        if (slot == 0) {                     // scale
            j0 = N0[0] * S.Xp[k][0] + N0[1] * S.Xp[k][1] + N0[2] * S.Xp[k][2];
            j1 = N0[3] * S.Xp[k][0] + N0[4] * S.Xp[k][1] + N0[5] * S.Xp[k][2];
            col = IDX_S;
        } else if (slot < 4) {               // root axis-angle
            double dv[3];
            mat3_vec(S.dR[0] + 9 * (slot - 1), S.pc[k], dv);
            const double s = x[IDX_S];
            j0 = s * (N0[0] * dv[0] + N0[1] * dv[1] + N0[2] * dv[2]);
            j1 = s * (N0[3] * dv[0] + N0[4] * dv[1] + N0[5] * dv[2]);
            col = IDX_RAA + slot - 1;
        } else if (slot < 7) {               // translation
            j0 = N0[slot - 4]; j1 = N0[3 + slot - 4];
            col = IDX_T + slot - 4;
        } else if (slot < 7 + naa) {         // axis-angle of chain ancestor a_i, i = 1..d
            const int i = (slot - 7) / 3 + 1, kk = (slot - 7) % 3;
            const int a = C->chain[jid][i];
            double dv[3];
            mat3_vec(S.dR[a] + 9 * kk, S.q[k][i - 1], dv);
            const double* Ni = S.N[k][1 + i];
            j0 = Ni[0] * dv[0] + Ni[1] * dv[1] + Ni[2] * dv[2];
            j1 = Ni[3] * dv[0] + Ni[4] * dv[1] + Ni[5] * dv[2];
            col = IDX_JAA + 3 * (a - 1) + kk;
        } else {                             // shape key
            const int s = slot - 7 - naa;
            j0 = 0.0; j1 = 0.0;
            for (int i = 0; i < len; ++i) {
                const int a = C->chain[jid][i];
                const double* Ni = S.N[k][1 + i];
                const double d0 = C->DB[a][0][s], d1 = C->DB[a][1][s], d2 = C->DB[a][2][s];
                j0 += Ni[0] * d0 + Ni[1] * d1 + Ni[2] * d2;
                j1 += Ni[3] * d0 + Ni[4] * d1 + Ni[5] * d2;
            }
            col = IDX_W + s;
        }
        O.J[k][0][slot] = j0;
        O.J[k][1][slot] = j1;
        O.jcol[k][slot] = col;
    }
    __syncthreads();
}

// packed lower-triangular index, i >= j
__device__ __forceinline__ int tri(int i, int j) { return i * (i + 1) / 2 + j; }

// ---------------------------------------------------------------------------------------------------------------
// Back substitution L^T y = z by ONE warp, four columns per step.  lane l holds z[l], z[l + 32], z[l + 64] in zreg.  Lget(r, i) reads
// L[r][i] (i < r < n), inv[j] = 1 / L_jj.  Per block of four columns (from the bottom): every lane gathers the block's four z by
// shuffles, solves the 4 x 4 triangle redundantly in registers, and subtracts the block's contribution from its own z entries.  The
// dependent chain per block is shuffle -> 4 x (multiply-add, multiply) -> 4 multiply-adds, about a third of one column at a time; the
// block's L entries are read before its chain starts.  n <= 96.
// ---------------------------------------------------------------------------------------------------------------
// Rows are addressed through a row pointer and a stride rule: row r starts at base + row_off(r).  Entries of a row at or right of
// its diagonal block are never meant to be used; they are read (valid memory, finite values) and only ever reach z entries that are
// dead by then, which keeps the loop free of predicates and index arithmetic — the chain is instruction-bound, not FMA-bound.
template <class RowPtr>
__device__ __forceinline__ void backsolve_warp_blocked(int n, RowPtr row_ptr, const double* inv, double (&zreg)[3], double* y) {
    const int lane = threadIdx.x & 31;
    int b = (n - 1) >> 2;
    // a partial top block (n not a multiple of 4): one column at a time
    for (int j = n - 1; j >= 4 * b && (n & 3); --j) {
        const int sl = j >> 5;
        const double v = sl == 0 ? zreg[0] : (sl == 1 ? zreg[1] : zreg[2]);
        const double yj = __shfl_sync(0xffffffffu, v, j & 31) * inv[j];
        if (lane == 0) y[j] = yj;
        const double* rj = row_ptr(j);
#pragma unroll
        for (int q = 0; q < 3; ++q) { const int i = lane + 32 * q; if (i < j) zreg[q] = fma(-rj[i], yj, zreg[q]); }
    }
    if (n & 3) --b;
    for (; b >= 0; --b) {
        const int j0 = 4 * b;
        const double *p0 = row_ptr(j0), *p1 = row_ptr(j0 + 1), *p2 = row_ptr(j0 + 2), *p3 = row_ptr(j0 + 3);
        const double l10 = p1[j0], l20 = p2[j0], l21 = p2[j0 + 1], l30 = p3[j0], l31 = p3[j0 + 1], l32 = p3[j0 + 2];
        const double i0 = inv[j0], i1 = inv[j0 + 1], i2 = inv[j0 + 2], i3 = inv[j0 + 3];
        double u0[3], u1[3], u2[3], u3[3];
#pragma unroll
        for (int q = 0; q < 3; ++q) { u0[q] = p0[lane + 32 * q]; u1[q] = p1[lane + 32 * q]; u2[q] = p2[lane + 32 * q]; u3[q] = p3[lane + 32 * q]; }
        // the block's z: its four entries sit in consecutive lanes of one register slot (j0 is a multiple of 4)
        const int sl = j0 >> 5, l0 = j0 & 31;
        const double v = sl == 0 ? zreg[0] : (sl == 1 ? zreg[1] : zreg[2]);
        const double z0 = __shfl_sync(0xffffffffu, v, l0), z1 = __shfl_sync(0xffffffffu, v, l0 + 1), z2 = __shfl_sync(0xffffffffu, v, l0 + 2),
                     z3 = __shfl_sync(0xffffffffu, v, l0 + 3);
        const double y3 = z3 * i3;
        const double y2 = (z2 - l32 * y3) * i2;
        const double y1 = (z1 - l21 * y2 - l31 * y3) * i1;
        const double y0 = (z0 - l10 * y1 - l20 * y2 - l30 * y3) * i0;
        if (lane == 0) { y[j0] = y0; y[j0 + 1] = y1; y[j0 + 2] = y2; y[j0 + 3] = y3; }
#pragma unroll
        for (int q = 0; q < 3; ++q) zreg[q] = zreg[q] - ((u0[q] * y0 + u1[q] * y1) + (u2[q] * y2 + u3[q] * y3));
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Line-search trial point: per keypoint the robustified cost rho / 2 at x and its derivative along the direction dx (both in the
// 86-layout), i.e. res^T (J dx) of the robustified block, by forward-mode propagation through the chain — no Jacobian is formed.
// What Ceres' Armijo search needs at a trial point is exactly the cost and this directional derivative.
// ---------------------------------------------------------------------------------------------------------------
template <int T>
__device__ void eval_reproj_dir(const FitConst* __restrict__ C, const double* x, const double* dx, const b3d_kp* __restrict__ kps, int nkp,
                                const double* R0, EvalTmp& S, double* kcost, double* kdphi) {
    const int tid = threadIdx.x;
    double (*Mj)[9] = reinterpret_cast<double (*)[9]>(&S.N[0][0][0]);     // d Rot(aa_j) along daa_j, one 3 x 3 per joint
    double (*dbone)[3] = reinterpret_cast<double (*)[3]>(&S.q[0][0][0]);   // d bone along dw
    static_assert(sizeof(S.N) >= sizeof(double) * NJ * 9 && sizeof(S.q) >= sizeof(double) * NJ * 3, "scratch aliases");
    for (int t = tid; t < NJ + 3 * NJ; t += T) {
        if (t < NJ) {
            const int o = t == 0 ? IDX_RAA : IDX_JAA + 3 * (t - 1);
            aa_rotation<true>(x + o, S.R[t], S.dR[t]);
            const double d0 = dx[o], d1 = dx[o + 1], d2 = dx[o + 2];
#pragma unroll
            for (int e = 0; e < 9; ++e) Mj[t][e] = d0 * S.dR[t][e] + d1 * S.dR[t][9 + e] + d2 * S.dR[t][18 + e];
        } else {
            const int j = (t - NJ) / 3, c = (t - NJ) % 3;
            double b = C->offset[j][c], db = 0.0;
            if (C->with_shape) {
#pragma unroll
                for (int s = 0; s < NS; ++s) { b += C->DB[j][c][s] * x[IDX_W + s]; db += C->DB[j][c][s] * dx[IDX_W + s]; }
            }
            S.bone[j][c] = b;
            dbone[j][c] = db;
        }
    }
    __syncthreads();
    if (tid < nkp) {
        const int k = tid;
        const int jid = clamp_jid(kps[k].jid);
        const int len = C->chain_len[jid];
        double q[3] = {S.bone[jid][0], S.bone[jid][1], S.bone[jid][2]};
        double dq[3] = {dbone[jid][0], dbone[jid][1], dbone[jid][2]};
        for (int i = 1; i < len; ++i) {
            const int a = C->chain[jid][i];
            double o[3], u[3], v[3];
            mat3_vec(S.R[a], q, o);
            mat3_vec(Mj[a], q, u);
            mat3_vec(S.R[a], dq, v);
#pragma unroll
            for (int c = 0; c < 3; ++c) { q[c] = o[c] + S.bone[a][c]; dq[c] = (u[c] + v[c]) + dbone[a][c]; }
        }
        double pc[3], dpc[3], Xp[3], dXp[3], tmp[3];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            pc[r] = R0[r] * q[0] + R0[3 + r] * q[1] + R0[6 + r] * q[2];
            dpc[r] = R0[r] * dq[0] + R0[3 + r] * dq[1] + R0[6 + r] * dq[2];
        }
        mat3_vec(S.R[0], pc, Xp);
        mat3_vec(Mj[0], pc, dXp);
        mat3_vec(S.R[0], dpc, tmp);
        const double s = x[IDX_S], ds = dx[IDX_S];
        const double X0 = s * Xp[0] + x[IDX_T], X1 = s * Xp[1] + x[IDX_T + 1], X2 = s * Xp[2] + x[IDX_T + 2];
        const double dX0 = ds * Xp[0] + s * (dXp[0] + tmp[0]) + dx[IDX_T], dX1 = ds * Xp[1] + s * (dXp[1] + tmp[1]) + dx[IDX_T + 1],
                     dX2 = ds * Xp[2] + s * (dXp[2] + tmp[2]) + dx[IDX_T + 2];
        const double iz = 1.0 / X2;
        const double r0 = C->cam[0] * X0 * iz + C->cam[2] - kps[k].u;
        const double r1 = C->cam[1] * X1 * iz + C->cam[3] - kps[k].v;
        const double dr0 = C->cam[0] * iz * (dX0 - X0 * iz * dX2), dr1 = C->cam[1] * iz * (dX1 - X1 * iz * dX2);
        const double sq = r0 * r0 + r1 * r1, b = C->huber * C->huber;
        double rho0, rho1;
        if (sq > b) { const double rr = sqrt(sq); rho0 = 2.0 * C->huber * rr - b; rho1 = fmax(DBL_MIN, C->huber / rr); }
        else { rho0 = sq; rho1 = 1.0; }
        kcost[k] = 0.5 * rho0;
        kdphi[k] = rho1 * (r0 * dr0 + r1 * dr1);
    }
    __syncthreads();
}

// ---------------------------------------------------------------------------------------------------------------
// Dense variant for the single-frame kernel: the robustified Jacobian of all reprojection blocks as one row-major
// [2 nkp][LDJ] matrix in the FREE column layout (held-constant parameters have no column), so that J^T J is a plain SYRK over
// 4 x 4 register tiles and J^T r a column sum.  Same arithmetic per entry as eval_reproj; structural zeros are stored.
// ---------------------------------------------------------------------------------------------------------------
constexpr int LDJ = 88;                       // 86 columns padded to whole 4-wide tiles (16-byte aligned rows)
// Position of column c inside a row: the first two columns of every 4-wide tile sit in the lower half of the row, the last two in
// the upper half.  A thread that owns tile t then reads two 16-byte pieces at [2 t] and [LDJ / 2 + 2 t]: consecutive tiles are 16
// bytes apart, so a warp's loads spread over all 32 banks (tiles as contiguous 32-byte groups put every load on half of them).
__host__ __device__ inline int jpos(int c) { return ((c >> 2) << 1) + (c & 1) + ((c & 2) ? LDJ / 2 : 0); }
struct EvalDense {
    double res[MAXKP][2];
    double kcost[MAXKP];
    alignas(16) double J[2 * MAXKP][LDJ];
};

template <int T, bool JAC>
__device__ void eval_reproj_dense(const FitConst* __restrict__ C, const double* x, const b3d_kp* __restrict__ kps, int nkp,
                                  const double* R0, EvalTmp& S, EvalDense& O) {
    const int tid = threadIdx.x;
    for (int t = tid; t < NJ + 3 * NJ; t += T) {
        if (t < NJ) {
            const double* aa = t == 0 ? x + IDX_RAA : x + IDX_JAA + 3 * (t - 1);
            aa_rotation<JAC>(aa, S.R[t], S.dR[t]);
        } else {
            const int j = (t - NJ) / 3, c = (t - NJ) % 3;
            double b = C->offset[j][c];
            if (C->with_shape) {
#pragma unroll
                for (int s = 0; s < NS; ++s) b += C->DB[j][c][s] * x[IDX_W + s];
            }
            S.bone[j][c] = b;
        }
    }
    __syncthreads();
    if (tid < nkp) {
        const int k = tid;
        const int jid = clamp_jid(kps[k].jid);
        const int len = C->chain_len[jid];
        double q[3] = {S.bone[jid][0], S.bone[jid][1], S.bone[jid][2]};
        S.q[k][0][0] = q[0]; S.q[k][0][1] = q[1]; S.q[k][0][2] = q[2];
        for (int i = 1; i < len; ++i) {
            const int a = C->chain[jid][i];
            double o[3];
            mat3_vec(S.R[a], q, o);
            q[0] = o[0] + S.bone[a][0]; q[1] = o[1] + S.bone[a][1]; q[2] = o[2] + S.bone[a][2];
            S.q[k][i][0] = q[0]; S.q[k][i][1] = q[1]; S.q[k][i][2] = q[2];
        }
        double pc[3], Xp[3];
#pragma unroll
        for (int r = 0; r < 3; ++r) pc[r] = R0[r] * q[0] + R0[3 + r] * q[1] + R0[6 + r] * q[2];
        mat3_vec(S.R[0], pc, Xp);
        const double s = x[IDX_S];
        const double X0 = s * Xp[0] + x[IDX_T], X1 = s * Xp[1] + x[IDX_T + 1], X2 = s * Xp[2] + x[IDX_T + 2];
        const double iz = 1.0 / X2;
        double r0 = C->cam[0] * X0 * iz + C->cam[2] - kps[k].u;
        double r1 = C->cam[1] * X1 * iz + C->cam[3] - kps[k].v;
        const double sq = r0 * r0 + r1 * r1, b = C->huber * C->huber;
        double rho0, rho1;
        if (sq > b) { const double rr = sqrt(sq); rho0 = 2.0 * C->huber * rr - b; rho1 = fmax(DBL_MIN, C->huber / rr); }
        else { rho0 = sq; rho1 = 1.0; }
        const double sc = sqrt(rho1);
        O.res[k][0] = r0 * sc; O.res[k][1] = r1 * sc;
        O.kcost[k] = 0.5 * rho0;
        if (JAC) {
#pragma unroll
            for (int c = 0; c < 3; ++c) { S.pc[k][c] = pc[c]; S.Xp[k][c] = Xp[c]; }
            double N[6];
            N[0] = sc * C->cam[0] * iz; N[1] = 0.0;                  N[2] = -sc * C->cam[0] * X0 * iz * iz;
            N[3] = 0.0;                 N[4] = sc * C->cam[1] * iz;  N[5] = -sc * C->cam[1] * X1 * iz * iz;
#pragma unroll
            for (int e = 0; e < 6; ++e) S.N[k][0][e] = N[e];
            double A[6], Bm[6];
            n23_mul(N, S.R[0], A);
            const double R0rm[9] = {R0[0], R0[3], R0[6], R0[1], R0[4], R0[7], R0[2], R0[5], R0[8]};
            n23_mul(A, R0rm, Bm);
#pragma unroll
            for (int e = 0; e < 6; ++e) Bm[e] *= s;
            for (int i = len - 1; i >= 0; --i) {
#pragma unroll
                for (int e = 0; e < 6; ++e) S.N[k][1 + i][e] = Bm[e];
                if (i > 0) {
                    n23_mul(Bm, S.R[C->chain[jid][i]], A);
#pragma unroll
                    for (int e = 0; e < 6; ++e) Bm[e] = A[e];
                }
            }
        }
    }
    __syncthreads();
    if (!JAC) return;
    // one task per (keypoint, padded free column): entries of columns the keypoint's chain does not touch are exact zeros
    const int n = C->n;
    for (int t = tid; t < nkp * LDJ; t += T) {
        const int k = t / LDJ, c = t - k * LDJ;
        double j0 = 0.0, j1 = 0.0;
        if (c < n) {
            const int fc = C->free_idx[c];
            const int jid = clamp_jid(kps[k].jid);
            const double* N0 = S.N[k][0];
            if (fc == IDX_S) {
                j0 = N0[0] * S.Xp[k][0] + N0[1] * S.Xp[k][1] + N0[2] * S.Xp[k][2];
                j1 = N0[3] * S.Xp[k][0] + N0[4] * S.Xp[k][1] + N0[5] * S.Xp[k][2];
            } else if (fc < IDX_T) {             // root axis-angle
                double dv[3];
                mat3_vec(S.dR[0] + 9 * (fc - IDX_RAA), S.pc[k], dv);
                const double s = x[IDX_S];
                j0 = s * (N0[0] * dv[0] + N0[1] * dv[1] + N0[2] * dv[2]);
                j1 = s * (N0[3] * dv[0] + N0[4] * dv[1] + N0[5] * dv[2]);
            } else if (fc < IDX_JAA) {           // translation
                j0 = N0[fc - IDX_T]; j1 = N0[3 + fc - IDX_T];
            } else if (fc < IDX_W) {             // axis-angle of joint a: non-zero iff a is a strict ancestor a_i, i >= 1, of the keypoint's joint
                const int a = (fc - IDX_JAA) / 3 + 1, kk = (fc - IDX_JAA) - 3 * (a - 1);
                const int i = C->chain_pos[jid][a];
                if (i >= 1) {
                    double dv[3];
                    mat3_vec(S.dR[a] + 9 * kk, S.q[k][i - 1], dv);
                    const double* Ni = S.N[k][1 + i];
                    j0 = Ni[0] * dv[0] + Ni[1] * dv[1] + Ni[2] * dv[2];
                    j1 = Ni[3] * dv[0] + Ni[4] * dv[1] + Ni[5] * dv[2];
                }
            } else if (C->with_shape) {          // shape key
                const int sidx = fc - IDX_W;
                const int len = C->chain_len[jid];
                for (int i = 0; i < len; ++i) {
                    const int a = C->chain[jid][i];
                    const double* Ni = S.N[k][1 + i];
                    const double d0 = C->DB[a][0][sidx], d1 = C->DB[a][1][sidx], d2 = C->DB[a][2][sidx];
                    j0 += Ni[0] * d0 + Ni[1] * d1 + Ni[2] * d2;
                    j1 += Ni[3] * d0 + Ni[4] * d1 + Ni[5] * d2;
                }
            }
        }
        O.J[2 * k][jpos(c)] = j0;
        O.J[2 * k + 1][jpos(c)] = j1;
    }
    __syncthreads();
}

}  // namespace b3d
