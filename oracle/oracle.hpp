// CPU ORACLE — TEST INFRASTRUCTURE ONLY (never shipped, never on the product path).
//
// Dependency-free C++17 / fp64 restatement of the reference's per-frame SMPL fitting hot path
// (jonH34400/3DBodyAnimation @ eb48afe).  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load this library.
//
// PARITY UNPINNED: the reference ships no tests or golden vectors for this path and its
// own build needs Eigen + Ceres 1.14, neither of which exists in this container, so the
// oracle cannot be diffed against the reference binary here.  It is pinned instead by self-checks
// (Jet autodiff vs analytic vs finite differences, QR vs Cholesky vs block-tridiagonal, numpy
// restatement of the forward pass) and by the two real fixtures (pose_prior.txt, 38 MediaPipe JSONs).
//
// Reference files followed (all relative to /root/reference):
//   external/avatar/AvatarModel.cpp:23-127   model derivation      -> smpl.cpp  Model::derive
//   external/avatar/Avatar.cpp:22-75         Avatar::update         -> smpl.cpp  forward
//   external/avatar/GaussianMixture.cpp:12-113 GMM load / residual  -> gmm.cpp
//   include/Sim3BA.h:22-346                  residual functors/priors -> resid.hpp
//   include/Sim3BA.h:348-683, include/MultiFrameBA.h:33-177 problem builders -> fit.cpp
//   src/main_multi_frame.cpp:109-201         anchor/window schedule -> fit.cpp  fit_sequence
//   Ceres Solver 1.14 (external, not in tree): TrustRegionMinimizer, LevenbergMarquardtStrategy,
//   Corrector, DenseQRSolver, ArmijoLineSearch, polynomial.cc  -> lm.cpp (restated from the
//   published algorithm)
#pragma once
#include <array>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <limits>
#include <string>
#include <vector>

namespace orc {

// ---------------------------------------------------------------------------------------------
// Forward-mode dual numbers that mirror ceres::Jet<double, N> (value + N partials).
// ---------------------------------------------------------------------------------------------
template <int N>
struct Jet {
    double a;
    double v[N];
    Jet() : a(0.0) { for (int i = 0; i < N; ++i) v[i] = 0.0; }
    Jet(double x) : a(x) { for (int i = 0; i < N; ++i) v[i] = 0.0; }  // NOLINT (implicit like ceres)
    Jet(double x, int k) : a(x) { for (int i = 0; i < N; ++i) v[i] = 0.0; v[k] = 1.0; }
};
template <int N> inline Jet<N> operator+(const Jet<N>& f, const Jet<N>& g) { Jet<N> h; h.a = f.a + g.a; for (int i = 0; i < N; ++i) h.v[i] = f.v[i] + g.v[i]; return h; }
template <int N> inline Jet<N> operator-(const Jet<N>& f, const Jet<N>& g) { Jet<N> h; h.a = f.a - g.a; for (int i = 0; i < N; ++i) h.v[i] = f.v[i] - g.v[i]; return h; }
template <int N> inline Jet<N> operator-(const Jet<N>& f) { Jet<N> h; h.a = -f.a; for (int i = 0; i < N; ++i) h.v[i] = -f.v[i]; return h; }
template <int N> inline Jet<N> operator*(const Jet<N>& f, const Jet<N>& g) { Jet<N> h; h.a = f.a * g.a; for (int i = 0; i < N; ++i) h.v[i] = f.a * g.v[i] + f.v[i] * g.a; return h; }
template <int N> inline Jet<N> operator/(const Jet<N>& f, const Jet<N>& g) {
    // ceres jet.h: h = f/g ; dh = (df - h*dg)/g
    Jet<N> h; const double gi = 1.0 / g.a; h.a = f.a * gi; for (int i = 0; i < N; ++i) h.v[i] = (f.v[i] - h.a * g.v[i]) * gi; return h;
}
template <int N> inline Jet<N>& operator+=(Jet<N>& f, const Jet<N>& g) { f = f + g; return f; }
template <int N> inline Jet<N> sqrt(const Jet<N>& f) { Jet<N> h; h.a = std::sqrt(f.a); const double t = 1.0 / (2.0 * h.a); for (int i = 0; i < N; ++i) h.v[i] = t * f.v[i]; return h; }
template <int N> inline Jet<N> sin(const Jet<N>& f) { Jet<N> h; h.a = std::sin(f.a); const double c = std::cos(f.a); for (int i = 0; i < N; ++i) h.v[i] = c * f.v[i]; return h; }
template <int N> inline Jet<N> cos(const Jet<N>& f) { Jet<N> h; h.a = std::cos(f.a); const double s = -std::sin(f.a); for (int i = 0; i < N; ++i) h.v[i] = s * f.v[i]; return h; }
template <int N> inline bool operator>(const Jet<N>& f, const Jet<N>& g) { return f.a > g.a; }
inline double sqrt(double x) { return std::sqrt(x); }
inline double sin(double x) { return std::sin(x); }
inline double cos(double x) { return std::cos(x); }
inline double value_of(double x) { return x; }
template <int N> inline double value_of(const Jet<N>& x) { return x.a; }

// ---------------------------------------------------------------------------------------------
// SMPL-shaped model (external/avatar/include/Avatar.h:92-147, AvatarModel.cpp:23-127)
// ---------------------------------------------------------------------------------------------
struct Model {
    int V = 0, J = 0, S = 0, P = 0;   // verts, joints, shape keys, pose-blend dims (0 = reference semantics, F5)
    std::vector<int> parent;          // J        parent[0] == -1
    std::vector<double> base;         // 3V       baseCloud, x1 y1 z1 x2 ...
    std::vector<double> key;          // 3V x S   keyClouds row-major: key[(3v+c)*S+s] = shapedirs[v][c][s]
    std::vector<int> wptr, wjoint;    // CSC of weights (J x V): column v = wptr[v]..wptr[v+1], joints ascending
    std::vector<double> wval;
    std::vector<double> J0;           // 3J       initialJointPos (3 x J col-major => J0[3j+c])
    std::vector<double> JSR;          // 3J x S   jointShapeReg row-major: JSR[(3j+c)*S+s]
    std::vector<double> posedirs;     // 3V x P   optional (standard SMPL; NOT in the reference)
    void derive(int V, int J, int S, const double* v_template, const double* shapedirs,
                const double* J_regressor, const double* weights, const uint32_t* kintree_row0,
                const double* posedirs_or_null, int P);
    // Avatar::update (external/avatar/Avatar.cpp:22-75). r: J col-major 3x3.  Outputs may be null.
    void forward(const double* w, const double* p, const double* r, double* cloud /*3V*/,
                 double* jointPos /*3J*/, double* jointTrans /*12J*/) const;
};

// ---------------------------------------------------------------------------------------------
// Gaussian mixture pose prior (external/avatar/GaussianMixture.cpp)
// ---------------------------------------------------------------------------------------------
struct GMM {
    int nComps = -1, nDims = 0;
    std::vector<double> weight, consts_log;
    std::vector<double> mean;                   // nComps x nDims row-major
    std::vector<std::vector<double>> cov;       // each nDims x nDims row-major
    std::vector<std::vector<double>> cov_cho;   // lower
    std::vector<std::vector<double>> prec_cho;  // lower: cov^-1 = L L^T
    // returns 0 on success, -1 missing file (nComps=-1, reference :15-19), -2 not PD (reference throws :55)
    int load_text(const char* path);
    int build(int nComps, int nDims, const double* weights, const double* means, const double* covs);
    // GaussianMixture::residual (:94-113): out has nDims+1 entries; returns component index
    int residual(const double* x, double* out) const;
};

// ---------------------------------------------------------------------------------------------
// Small dense linear algebra (row-major unless noted)
// ---------------------------------------------------------------------------------------------
bool cholesky_lower(double* A, int n);                              // in place, lower; false if not PD
void chol_solve(const double* L, int n, double* b);                 // solves L L^T x = b in place
// least squares min ||A x - b|| by Householder QR (A: m x n row-major, destroyed; b destroyed)
bool householder_qr_solve(double* A, int m, int n, double* b, double* x);
bool lu_fullpiv_solve(double* A, int n, double* b, double* x);      // small systems (line search)
void aa_to_matrix(const double aa[3], double R[9] /*col-major*/);   // write-back rule a15 (Sim3BA.h:481-507)
void matmul3_cm(const double A[9], const double B[9], double C[9]); // col-major 3x3

// ---------------------------------------------------------------------------------------------
// Sparse row Jacobian
// ---------------------------------------------------------------------------------------------
struct SparseRows {
    int n = 0;
    std::vector<int> rowptr{0};
    std::vector<int> col;
    std::vector<double> val;
    int rows() const { return (int)rowptr.size() - 1; }
    void clear(int n_) { n = n_; rowptr.assign(1, 0); col.clear(); val.clear(); }
    void push(int c, double v) { col.push_back(c); val.push_back(v); }
    void end_row() { rowptr.push_back((int)col.size()); }
};

// ---------------------------------------------------------------------------------------------
// LM (Ceres 1.14 TrustRegionMinimizer + LevenbergMarquardtStrategy)
// ---------------------------------------------------------------------------------------------
enum LinearSolverKind { LS_DENSE_QR = 0, LS_DENSE_CHOLESKY = 1, LS_BLOCK_TRIDIAG = 2 };

struct IterRecord {          // one pushed ceres::IterationSummary
    double cost, cost_change, gradient_max_norm, step_norm, relative_decrease, radius;
    int32_t flags;           // bit0 step_is_valid, bit1 step_is_successful, bit2 line-search contracted
    int32_t aux;             // GMM component active at x after this iteration (or -1)
};

enum Termination { T_CONVERGENCE_PARAM = 1, T_CONVERGENCE_FUNC = 2, T_CONVERGENCE_GRAD = 3, T_NO_CONVERGENCE_ITERS = 4,
                   T_NO_CONVERGENCE_RADIUS = 5, T_FAILURE_INVALID_STEPS = 6, T_FAILURE_EVAL = 7 };

struct BlockStructure {      // for LS_BLOCK_TRIDIAG: F diagonal blocks of size nb (tridiagonal coupling) + arrow of size na
    int F = 0, nb = 0, na = 0;
};

struct LMOptions {
    int max_iters = 100;
    int linear_solver = LS_DENSE_QR;
    BlockStructure blocks;
    double min_line_search_step_contraction = 0.6;   // Solver::Options default
    double max_line_search_step_contraction = 1e-3;
};

struct LMProblem {
    int n = 0;
    bool constrained = false;
    std::vector<double> lower, upper;    // size n when constrained (+-inf where unbounded)
    // evaluate at x: cost always; r/J/g when non-null (r and J are robustified, g = J^T r). false = failure.
    virtual bool evaluate(const double* x, double* cost, std::vector<double>* r, SparseRows* J, std::vector<double>* g) = 0;
    virtual int aux_state() const { return -1; }
    virtual ~LMProblem() {}
    void plus(const double* x, const double* d, double* out) const {
        for (int i = 0; i < n; ++i) {
            double v = x[i] + d[i];
            if (constrained) { if (v < lower[i]) v = lower[i]; if (v > upper[i]) v = upper[i]; }
            out[i] = v;
        }
    }
};

struct LMSummary {
    int termination = 0;
    bool usable = false;
    int num_iterations = 0;     // number of pushed iteration records (incl. iteration 0)
    double initial_cost = 0, final_cost = 0;
    int line_search_steps = 0;
};

LMSummary lm_minimize(LMProblem& prob, const LMOptions& opt, double* x, std::vector<IterRecord>* trace);

// ---------------------------------------------------------------------------------------------
// Fitting problems (Sim3BA.h / MultiFrameBA.h)
// ---------------------------------------------------------------------------------------------
struct KP { int32_t jid; int32_t pad; double u, v; };   // PixelKP (Sim3BA.h:9)

struct FitOptions {
    int mode = 0;                 // 0 = pose-only (OptimizePoseReprojection), 1 = pose+shape (OptimizePoseShapeReprojection)
    int max_iters = 100;
    double beta_pose = 0, beta_shape = 0, lambda_temp = 0;
    double fx = 0, fy = 0, cx = 0, cy = 0;
    double huber = 3.0;
    double scale_lo = 0.3, scale_hi = 3.0;
    int linear_solver = LS_DENSE_QR;
    int use_jets = 1;             // 1: Jet autodiff like ceres::DynamicAutoDiffCostFunction; 0: analytic Jacobians
};

constexpr int NJ = 24;            // SMPL joints
constexpr int NS = 10;            // shape keys
constexpr int NPF = 76;           // per-frame params: s, rootAA(3), rootT(3), jointAA[1..23](69)
constexpr int MAXKP = 17;

// single frame: sim3[7] in/out, w[S] in/out, jointAA[J*3] out (row j = joint j, row 0 unused/zero), R0[9] col-major in.
// returns usable flag; trace optional.
bool fit_single(const Model& M, const GMM* gmm, const FitOptions& o, const KP* kps, int nkp, const double* R0,
                double* sim3, double* w, double* jointAA, std::vector<IterRecord>* trace, LMSummary* summary);

// multi frame (OptimizeMultiFrame): poses F x 76 in/out, w[S] in/out (shared), R0 F x 9 (in; NOT rotated here),
// kps F x MAXKP, nkp F.
bool fit_window(const Model& M, const FitOptions& o, int F, const KP* kps, const int* nkp, const double* R0,
                double* poses, double* w, std::vector<IterRecord>* trace, LMSummary* summary);

// rest-pose offsets (Sim3BA.h:533-555): offset[3j+c] = J0_j - J0_parent (w=0); offset[0]=0
void rest_offsets(const Model& M, double* offset /*3J*/);

// ---------------------------------------------------------------------------------------------
// Vertex-level ICP residuals + analytic Jacobians of ark::AvatarOptimizer (icp.cpp; AvatarOptimizer.cpp:161-641)
// q: J quaternions, Eigen coefficient order (x, y, z, w).  corr e = (model point mpt[e], data point dpt[e]).
// resid 3 x ncorr, jac (3 ncorr) x (3 + 3J + S) row-major, grad 3 + 3J + S: any may be null.
// ---------------------------------------------------------------------------------------------
void icp_eval(const Model& M, const double* w, const double* p, const double* q, int ncorr, const int* mpt, const int* dpt,
              const double* data, int shape_enabled, double* resid, double* jac, double* cost, double* grad);
void quat_plus(const double* x, const double* delta, double* out);   // FakeQuaternionParameterization::Plus
// findNN with invert = true (AvatarOptimizer.cpp:841-907): out[i] = nearest visible model point of data point i's part, or -1
void icp_priors(const GMM* gmm, double beta_pose, double beta_shape, double total_residuals, int S, const double* w, const double* q,
                double* cost, double* grad);   // pose / shape prior blocks (:644-726, :1412-1428), added to cost / grad
void icp_visible(int V, int F, const uint32_t* faces, const double* cloud, uint8_t* visible);   // pointVisible (:1337-1352)
void icp_nearest(int V, const double* cloud, const int* model_part, const uint8_t* visible, int nd, const double* data,
                 const int* data_part, int* out);

// The minimiser of AvatarOptimizer::optimize (icp_solve.cpp): Ceres 1.14 LINE_SEARCH / BFGS / WOLFE / CUBIC on the ICP objective
// (data term + pose prior + shape prior).  w / p / q in and out.
enum { ICP_TERM_NO_CONVERGENCE = 0, ICP_TERM_GRADIENT_TOL = 1, ICP_TERM_PARAMETER_TOL = 2, ICP_TERM_FUNCTION_TOL = 3, ICP_TERM_FAILURE = 4 };
struct IcpSolveOptions {
    int max_iterations = 50;
    double function_tolerance = 1e-4;     // AvatarOptimizer.cpp:1334
    double gradient_tolerance = 1e-10;    // Ceres default
    double parameter_tolerance = 1e-8;    // Ceres default
};
struct IcpSolveRecord {
    double cost = 0, cost_change = 0, gradient_max_norm = 0, step_norm = 0, step_size = 0;
    int line_search_evals = 0, line_search_iters = 0;
};
struct IcpSolveSummary {
    double initial_cost = 0, final_cost = 0;
    int iterations = 0, evaluations = 0, restarts = 0, termination = 0;
};
int icp_minimize(const Model& M, const GMM* gmm, double beta_pose, double beta_shape, int ncorr, const int* mpt, const int* dpt,
                 const double* data, int shape_enabled, const IcpSolveOptions& o, double* w, double* p, double* q,
                 std::vector<IcpSolveRecord>* trace, IcpSolveSummary* summary);

}  // namespace orc
