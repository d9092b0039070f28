/* b3d.h — C-ABI of lib3dba_b200.so: the B200-native replacement for the per-frame SMPL fitting hot path of
 * jonH34400/3DBodyAnimation (reference @ eb48afe).  Plain pointers and sizes only; no C++/torch types.
 *
 * The reference has no FFI layer of its own: its boundary is three inline C++ functions plus the public fields of
 * ark::Avatar / ark::AvatarModel.  Every entry point below names the reference interface it
 * replaces (paths relative to the reference root).  include/b3d_shim.hpp offers the reference's own C++ signatures
 * on top of this header; INTEGRATION.md shows the binding a maintainer would add.
 *
 * Conventions
 *   - All fitting arithmetic is float64, as in the reference (Eigen double / ceres::Jet<double,4>).
 *   - 3x3 rotations are column-major (Eigen default): R[3*col + row].  jointTrans columns are col-major 3x4 [R|t].
 *   - Per-frame parameter vector ("pose", 76 doubles): [scale, rootAA(3), rootT(3), jointAA[1..23](69)]
 *     (FramePoseParams, include/MultiFrameBA.h:9-14; Sim3Params = first 7, include/Sim3BA.h:11-19).
 *   - Functions without a suffix take HOST pointers, copy in/out, and return when the result is in host memory.
 *     Functions ending in _dev take DEVICE pointers, enqueue on `stream` (a cudaStream_t passed as void*) and
 *     return without synchronising.
 *   - The library uses the calling thread's current CUDA device.  Handles belong to the device they were created on.
 *   - There is no CPU fallback: every entry point fails with B3D_ECUDA when no sm_100 device is usable.
 */
#ifndef B3D_H_
#define B3D_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B3D_NUM_JOINTS 24      /* SMPL joints (external/avatar/include/Avatar.h:27-59) */
#define B3D_NUM_SHAPE 10       /* shape keys */
#define B3D_POSE_DIM 76        /* per-frame parameters */
#define B3D_MAX_KP 17          /* USE_SMPL has 17 slots (include/Utils.h:22-23) */
#define B3D_POSEDIRS_DIM 207   /* optional pose blendshapes, 23*9 (NOT in the reference, AvatarModel.cpp:104-109) */

/* ---- status codes (reference: bool + std::exit / throw; SURVEY §8(b) "Errors") ---- */
enum {
    B3D_OK = 0,
    B3D_EINVAL = 1,    /* bad argument (null pointer, wrong size, kintree root parent != -1 — _ARK_ASSERT_EQ AvatarModel.cpp:41) */
    B3D_ECUDA = 2,     /* CUDA runtime error or no usable sm_100 device */
    B3D_ENOTPD = 3,    /* GMM covariance not positive definite (GaussianMixture.cpp:55 throws "Decomposition failed!") */
    B3D_ENOFILE = 4,   /* file missing / unreadable */
    B3D_ENOMEM = 5,
    B3D_EFORMAT = 6    /* malformed model / prior file */
};

/* Text of the last error raised on the calling thread ("" if none). */
const char* b3d_last_error(void);
/* Library version string and the SM architecture the kernels were compiled for ("sm_100a"). */
const char* b3d_version(void);
/* Number of kernel launches this library has issued in this process (all threads); for bench accounting. */
uint64_t b3d_launch_count(void);

/* ===================================================================================================== */
/* Model: replaces ark::AvatarModel (external/avatar/include/Avatar.h:64-147; AvatarModel.cpp:23-127)      */
/* ===================================================================================================== */
typedef struct b3d_model b3d_model;

enum { B3D_F32 = 0, B3D_F64 = 1 };

/* Raw arrays in the layout of SMPL's model.npz as the reference reads it (C order):
 *   v_template  V x 3           (AvatarModel.cpp:44-48)     shapedirs  V x 3 x S   (:93-102)
 *   J_regressor J x V  dense    (:60-64)                     weights    V x J dense (:66-71)
 *   kintree_row0 J  uint32, root = 0xFFFFFFFF -> parent -1  (:35-41)
 *   faces       F x 3  uint32 (optional; rendering only, :50-58)
 *   posedirs    V x 3 x 207 (optional; standard SMPL pose blendshapes — an extension, see b3d_smpl_forward flags) */
typedef struct b3d_model_desc {
    int32_t num_verts;       /* V (6890 for SMPL) */
    int32_t num_joints;      /* must be 24 */
    int32_t num_shape;       /* must be 10 */
    int32_t num_faces;       /* F or 0 */
    int32_t dtype;           /* B3D_F32 or B3D_F64: element type of the five float arrays below */
    int32_t reserved;
    const void* v_template;
    const void* shapedirs;
    const void* J_regressor;
    const void* weights;
    const uint32_t* kintree_row0;
    const uint32_t* faces;   /* may be NULL */
    const void* posedirs;    /* may be NULL */
} b3d_model_desc;

int b3d_model_create(const b3d_model_desc* desc, b3d_model** out);

/* ark::AvatarModel(model_dir)'s npz branch (external/avatar/AvatarModel.cpp:23-102 with cnpy.cpp:67-301 and
 * Util.cpp:249-309): reads v_template, kintree_table, f, J_regressor, weights, shapedirs from a NumPy .npz (stored or deflated
 * members, .npy format 1-3, little-endian float32/float64 and 32/64-bit integers, C or Fortran order) with the reference's
 * array views — v_template is read flat, shapedirs as a (3V x S) matrix, kintree_table row 0 = parents with the root's
 * 0xFFFFFFFF / -1 — and builds the model.  B3D_NPZ_WITH_POSEDIRS also loads posedirs (the reference has that commented out). */
enum { B3D_NPZ_WITH_POSEDIRS = 1 };
int b3d_model_load_npz(const char* path, uint32_t flags, b3d_model** out);
/* The loader's building blocks (host only, usable without a GPU): shape / element width / order of a member, and the
 * (rows x cols) views of Util.cpp's loadFloatMatrix / loadUintMatrix (row-major out). */
int b3d_npz_shape(const char* path, const char* name, int32_t* ndim, int64_t* shape4, int32_t* word_size, int32_t* fortran_order);
int b3d_npz_float_matrix(const char* path, const char* name, int64_t rows, int64_t cols, double* out);
int b3d_npz_uint_matrix(const char* path, const char* name, int64_t rows, int64_t cols, uint32_t* out);
void b3d_model_destroy(b3d_model* m);
int b3d_model_num_verts(const b3d_model* m);
/* parent[24], bit-exact indexing output (AvatarModel.cpp:35-41). */
int b3d_model_get_parents(const b3d_model* m, int32_t* parent);
/* Derived constants in float64 (any pointer may be NULL):
 *   J0 24x3 = initialJointPos (AvatarModel.cpp:113-117), JSR 24x3x10 = jointShapeReg (:121-127),
 *   offset 24x3 = rest bone vectors used by the fitting functors (include/Sim3BA.h:533-555). */
int b3d_model_get_constants(const b3d_model* m, double* J0, double* JSR, double* offset);

/* ===================================================================================================== */
/* GMM pose prior: replaces ark::GaussianMixture::load (external/avatar/GaussianMixture.cpp:12-75)         */
/* ===================================================================================================== */
typedef struct b3d_gmm b3d_gmm;

/* weights nComps; means nComps x nDims; covs nComps x nDims x nDims (row-major).  nDims must be 69. */
int b3d_gmm_create(int32_t nComps, int32_t nDims, const double* weights, const double* means, const double* covs,
                   b3d_gmm** out);
/* pose_prior.txt parser (format: scripts/convert_gmm_to_avatar.py:17-29).  A missing file returns B3D_ENOFILE and
 * *out = NULL; passing NULL as the gmm of a fit degrades to the L2 prior exactly like the reference's nComps = -1
 * (GaussianMixture.cpp:15-19, include/Sim3BA.h:266). */
int b3d_gmm_load(const char* path, b3d_gmm** out);
void b3d_gmm_destroy(b3d_gmm* g);
int b3d_gmm_dims(const b3d_gmm* g, int32_t* nComps, int32_t* nDims);
/* Derived arrays in float64 (any may be NULL): prec_cho nComps x nDims x nDims lower (chol of cov^-1), consts_log nComps. */
int b3d_gmm_get(const b3d_gmm* g, double* prec_cho, double* consts_log);
/* GaussianMixture::residual (GaussianMixture.cpp:94-113) for B pose vectors x (B x 69) evaluated on the GPU:
 * out B x 70, comp B (argmin component, strict '<', first wins). */
int b3d_gmm_residual(const b3d_gmm* g, int32_t B, const double* x, double* out, int32_t* comp);

/* ===================================================================================================== */
/* Entry (1) batched SMPL evaluator: replaces ark::Avatar::update (external/avatar/Avatar.cpp:22-75)       */
/* ===================================================================================================== */
enum {
    B3D_FWD_DEFAULT = 0,
    B3D_FWD_POSEDIRS = 1,  /* add posedirs . vec(R_j - I) (standard SMPL; the reference has this commented out) */
    /* float32 skinning kernel override (tests, profiling); without one of these the library picks by batch size:
     * tensor-core shape blend (tcgen05, lane = body) for large batches, tile-sorted lane = vertex kernel otherwise */
    B3D_FWD_LBS_TC = 0x10,
    B3D_FWD_LBS_SORTED = 0x20,
    B3D_FWD_LBS_BODYLANE = 0x40,
    B3D_FWD_LBS_GENERIC = 0x80,
    /* skinning only: jointTrans is an INPUT (the rest->posed transforms of an earlier call), the joint chain is not run and
     * p / r / jointPos are ignored.  Lets a caller re-skin with new shape-independent data, and lets bench.py time the skinning
     * kernel alone. */
    B3D_FWD_GIVEN_TRANSFORMS = 0x100
};

/* Inputs per body: w (10) = Avatar::w, p (3) = Avatar::p, r (24 x 9, col-major) = Avatar::r.
 * Outputs per body (any may be NULL): cloud V x 3 (= Avatar::cloud, 3 x V col-major), jointPos 24 x 3 (posed joints),
 * jointTrans 24 x 12 (col-major 3x4 rest->posed affine per joint).  dtype selects float or double for ALL six arrays. */
int b3d_smpl_forward(const b3d_model* m, int32_t B, int32_t dtype, const void* w, const void* p, const void* r,
                     void* cloud, void* jointPos, void* jointTrans, int32_t flags);
int b3d_smpl_forward_dev(const b3d_model* m, int32_t B, int32_t dtype, const void* w, const void* p, const void* r,
                         void* cloud, void* jointPos, void* jointTrans, int32_t flags, void* stream);

/* Allocation-free variant of b3d_smpl_forward_dev (cloud and jointPos only; no flags: shape blend + skinning, float32 through the
 * tile-sorted kernel, float64 through the generic one).  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  It calls no cudaMalloc / cudaMallocAsync / cudaFree and no allocating library routine: the B x 24 x 12 joint
 * transforms that b3d_smpl_forward_dev keeps in stream-ordered scratch live in a file-scope `static __device__` array sized for
 * B3D_NOALLOC_MAX_BODIES bodies of doubles (2.25 MiB, resident on every GPU that loads the library until exit).
 * B > B3D_NOALLOC_MAX_BODIES is rejected with B3D_EINVAL.  All pointers are DEVICE pointers; cloud and jointPos may be NULL.
 * Every call on a device shares the one array (it keeps the previous call's contents; the joint kernel overwrites every entry a
 * call reads, so nothing needs clearing): calls that overlap on the GPU — different streams, or two host threads — RACE and
 * must be serialised by the caller. */
#define B3D_NOALLOC_MAX_BODIES 1024
int b3d_smpl_forward_noalloc_dev(const b3d_model* m, int32_t B, int32_t dtype, const void* w, const void* p, const void* r,
                                 void* cloud, void* jointPos, void* stream);

/* ===================================================================================================== */
/* Entry (3) vertex-level ICP term of ark::AvatarOptimizer: cost and gradient of       */
/*   sum over correspondences 1/2 |cloud_i(w, p, q) - data_e|^2 for a batch of bodies.  Replaces            */
/*   AvatarEvaluationCommonData::PrepareForEvaluation (external/avatar/AvatarOptimizer.cpp:295-362),        */
/*   AvatarCostFunctorCache::updateData (:503-577), AvatarICPCostFunctor::Evaluate (:621-637) and the       */
/*   J^T r assembly of Ceres' line-search minimiser (:1318-1321).                                           */
/* ===================================================================================================== */
typedef struct b3d_icp b3d_icp;
enum {
    B3D_ICP_DEVICE_INPUT = 1,  /* b3d_icp_create: mpt / dpt / data are device pointers (cptr / dptr always host) */
    B3D_ICP_POSE_ONLY = 2      /* b3d_icp_eval: shapeEnabled = false (:176): the 10 shape entries of grad are 0 */
};
#define B3D_ICP_GRAD_DIM 85    /* root position (3) | rotation tangent of joint j at 3 + 3 j (FakeQuaternionParameterization,
                                  :118-153) | shape keys at 75 */

/* One set of correspondences per body (they stay fixed during a ceres::Solve, :1386-1412 / :1439): pairs e in
 * [cptr[b], cptr[b+1]) match model point mpt[e] with point dpt[e] of body b's data cloud, which starts at point dptr[b] of
 * `data` (3 values of `dtype` per point); dpt = NULL means pair e is data point e - cptr[b], and pairs with mpt[e] < 0 are
 * skipped (the output of b3d_icp_nearest).  The cloud is reduced to one (mean, count) record per model point; `dtype` is also the
 * arithmetic type of the evaluation kernel (B3D_F32: fast path; B3D_F64: accuracy path). */
int b3d_icp_create(const b3d_model* m, int32_t B, int32_t dtype, const int64_t* cptr, const int32_t* mpt, const int32_t* dpt,
                   const int64_t* dptr, const void* data, int32_t flags, b3d_icp** out);
void b3d_icp_destroy(b3d_icp* h);
/* Correspondence search: findNN with invert = true (external/avatar/AvatarOptimizer.cpp:841-907, called at :1372-1375).  For every
 * data point the nearest VISIBLE model point of the same part, -1 if the part has none.  cloud: B x V x 3 posed model clouds
 * (Avatar::cloud), model_part: V labels in [0, nparts) (host; modelPartLabels, :1296-1300), visible: B x V bytes or NULL
 * (pointVisible), data / data_part: the bodies' data points and their labels, body b at [dptr[b], dptr[b+1]) (dptr host).
 * nn_out has one entry per data point.  With B3D_ICP_DEVICE_INPUT cloud, visible, data, data_part and nn_out are device pointers.
 * The result is the pair list of b3d_icp_create: cptr = dptr, mpt = nn_out, dpt = NULL (pair e <-> data point e; mpt < 0 skipped). */
/* The prior blocks AvatarOptimizer::optimize adds to the ICP term (external/avatar/AvatarOptimizer.cpp:1412-1428):
 * AvatarPosePriorCostFunctor (:644-700, GMM pose prior on the 23 joint rotations) and AvatarShapePriorCostFunctor (:703-726), their
 * weights scaled by sqrt(number of ICP residual blocks of the body) / 15.  b3d_icp_add_priors_dev ADDS their cost and gradient to
 * cost / grad (device pointers, as left by b3d_icp_eval_dev on the same stream); b3d_icp_eval_full is the whole objective with host
 * pointers (data term + priors).  beta <= 0 switches a block off like the reference's `if (beta > 0.)`; gmm may be NULL only then. */
int b3d_icp_add_priors_dev(b3d_icp* h, const b3d_gmm* gmm, double beta_pose, double beta_shape, const double* w, const double* q,
                           double* cost, double* grad, void* stream);
int b3d_icp_eval_full(b3d_icp* h, const b3d_gmm* gmm, double beta_pose, double beta_shape, const double* w, const double* p,
                      const double* q, int32_t flags, double* cost, double* grad);
/* pointVisible of the ICP loop (external/avatar/AvatarOptimizer.cpp:1337-1352, enableOcclusion): visible[b][v] = 1 if vertex v of
 * body b belongs to a face with ((p2 - p1) x (p1 - p3)).z > 1e-4 in the posed cloud (B x V x 3), else 0.  The model must have
 * been created with its faces.  With B3D_ICP_DEVICE_INPUT cloud and visible are device pointers. */
int b3d_icp_visible(const b3d_model* m, int32_t B, int32_t dtype, const void* cloud, int32_t flags, uint8_t* visible);
int b3d_icp_nearest(const b3d_model* m, int32_t B, int32_t dtype, const void* cloud, int32_t nparts, const int32_t* model_part,
                    const uint8_t* visible, const int64_t* dptr, const void* data, const int32_t* data_part, int32_t flags,
                    int32_t* nn_out);
int b3d_icp_num_bodies(const b3d_icp* h);
/* One evaluation at w (B x 10), p (B x 3), q (B x 24 x 4 quaternions in Eigen's coefficient order x, y, z, w):
 * cost (B) and grad (B x 85).  Host pointers / device pointers + stream. */
int b3d_icp_eval(b3d_icp* h, const double* w, const double* p, const double* q, int32_t flags, double* cost, double* grad);
int b3d_icp_eval_dev(b3d_icp* h, const double* w, const double* p, const double* q, int32_t flags, double* cost, double* grad,
                     void* stream);

/* The solve of one ICP iteration for a batch of bodies: replaces the ceres::Solve call of AvatarOptimizer::optimize
 * (external/avatar/AvatarOptimizer.cpp:1486) with the options set at :1313-1339 — minimizer_type LINE_SEARCH, line_search_direction_type
 * BFGS, line_search_interpolation_type CUBIC, line_search_type WOLFE (Ceres default), function_tolerance 1e-4, max_num_iterations =
 * maxItersPerICP, every other option at its Ceres 1.14 default — on the objective b3d_icp_eval_dev + b3d_icp_add_priors_dev evaluate
 * (ICP blocks + pose prior + shape prior, :1439-1474).  w (B x 10), p (B x 3), q (B x 24 x 4) are the starting point on entry and
 * the minimiser's result on return (what :1491-1494 converts back to rotation matrices).  Optional outputs: trace (B x trace_cap x
 * 7 doubles per iteration record: cost, cost_change, gradient_max_norm, step_norm, step_size, line-search evaluations,
 * line-search iterations; record 0 is the starting point), nrec (B: records written or wanted), summary (B x 6 doubles: initial
 * cost, final cost, iterations, evaluations, direction restarts, termination B3D_ICP_TERM_*), evaluations_out (batched objective
 * evaluations launched = evaluations of the slowest body, rounded up to the host's polling interval). */
enum { B3D_ICP_TERM_NO_CONVERGENCE = 0, B3D_ICP_TERM_GRADIENT_TOL = 1, B3D_ICP_TERM_PARAMETER_TOL = 2, B3D_ICP_TERM_FUNCTION_TOL = 3,
       B3D_ICP_TERM_FAILURE = 4 };
typedef struct b3d_icp_solve_opts {
    int32_t max_iterations;        /* maxItersPerICP */
    int32_t flags;                 /* B3D_ICP_POSE_ONLY: shapeEnabled = false */
    double function_tolerance;     /* 1e-4 (:1334) */
    double gradient_tolerance;     /* 1e-10 (Ceres default) */
    double parameter_tolerance;    /* 1e-8 (Ceres default) */
    double beta_pose, beta_shape;  /* betaPose, betaShape (scaled by sqrt(#residual blocks) / 15 inside, :1457-1458) */
} b3d_icp_solve_opts;
int b3d_icp_solve(b3d_icp* h, const b3d_gmm* gmm, const b3d_icp_solve_opts* opts, double* w, double* p, double* q, double* trace,
                  int32_t trace_cap, int32_t* nrec, double* summary, int32_t* evaluations_out);
int b3d_icp_solve_dev(b3d_icp* h, const b3d_gmm* gmm, const b3d_icp_solve_opts* opts, double* w, double* p, double* q, double* trace,
                      int32_t trace_cap, int32_t* nrec, double* summary, int32_t* evaluations_out, void* stream);

/* ===================================================================================================== */
/* Entry (2) batched Levenberg-Marquardt driver: replaces the ceres::Problem + ceres::Solve inside          */
/*   OptimizePoseReprojection (include/Sim3BA.h:515-683), OptimizePoseShapeReprojection (:348-511)          */
/* ===================================================================================================== */
typedef struct b3d_kp {        /* PixelKP (include/Sim3BA.h:9) */
    int32_t jid;
    int32_t pad;
    double u, v;
} b3d_kp;

enum { B3D_FIT_POSE = 0, B3D_FIT_POSE_SHAPE = 1 };

typedef struct b3d_fit_opts {
    int32_t mode;              /* B3D_FIT_POSE (Sim3BA.h:515) or B3D_FIT_POSE_SHAPE (Sim3BA.h:348) */
    int32_t max_iters;         /* ceres max_num_iterations (default 100) */
    double beta_pose;          /* betaPose  */
    double beta_shape;         /* betaShape */
    double lambda_temp;        /* lambdaTemp (multi-frame only) */
    double fx, fy, cx, cy;     /* pinhole intrinsics */
    double huber;              /* ceres::HuberLoss(3.0) */
    double scale_lo, scale_hi; /* SetParameterLowerBound/UpperBound on scale: 0.3, 3.0 (Sim3BA.h:450-451) */
} b3d_fit_opts;

/* One ceres::IterationSummary as printed by minimizer_progress_to_stdout (Sim3BA.h:477). */
typedef struct b3d_iter_record {
    double cost, cost_change, gradient_max_norm, step_norm, relative_decrease, radius;
    int32_t flags;             /* bit0 step_is_valid, bit1 step_is_successful, bit2 line search contracted the step */
    int32_t aux;               /* GMM component active at x after this iteration, or -1 */
} b3d_iter_record;

/* ceres::TerminationType / message class of the solve */
enum {
    B3D_TERM_PARAM_TOL = 1, B3D_TERM_FUNC_TOL = 2, B3D_TERM_GRAD_TOL = 3, B3D_TERM_MAX_ITERS = 4,
    B3D_TERM_MIN_RADIUS = 5, B3D_TERM_INVALID_STEPS = 6, B3D_TERM_EVAL_FAILURE = 7
};

/* B independent single-frame fits (the frame loop of src/main_single_frame.cpp:192-249).
 *   kps   B x 17 (first nkp[b] valid)     nkp B            R0  B x 9 = avatar.r[0] on entry (Sim3BA.h:395/558)
 *   sim3  B x 7 in/out (initSim3)         w   B x 10 in/out (avatar.w)
 *   jointAA B x 24 x 3 out (row 0 zero)   usable B out = summary.IsSolutionUsable()
 *   trace  B x trace_stride records or NULL; niters / term B or NULL.
 * The caller applies the write-back (b3d_writeback) and the final update (b3d_smpl_forward), as the reference does. */
int b3d_fit_single(const b3d_model* m, const b3d_gmm* gmm, const b3d_fit_opts* opts, int32_t B, const b3d_kp* kps,
                   const int32_t* nkp, const double* R0, double* sim3, double* w, double* jointAA, uint8_t* usable,
                   b3d_iter_record* trace, int32_t trace_stride, int32_t* niters, int32_t* term);
int b3d_fit_single_dev(const b3d_model* m, const b3d_gmm* gmm, const b3d_fit_opts* opts, int32_t B, const b3d_kp* kps,
                       const int32_t* nkp, const double* R0, double* sim3, double* w, double* jointAA, uint8_t* usable,
                       b3d_iter_record* trace, int32_t trace_stride, int32_t* niters, int32_t* term, void* stream);

/* AA -> rotation write-back (include/Sim3BA.h:481-507, 649-679; include/MultiFrameBA.h:154-172) for F frames:
 *   R0 F x 9 in/out: R0 <- R(rootAA) * R0 (cumulative, in place);  r F x 24 x 9 out: r[0] = R0, r[j] = R(jointAA_j);
 *   p F x 3 out = rootT.  poses F x 76.  Scale is NOT applied (reference behaviour). */
int b3d_writeback(int32_t F, const double* poses, double* R0, double* r, double* p);
int b3d_writeback_dev(int32_t F, const double* poses, double* R0, double* r, double* p, void* stream);

/* mean_pixel_error (include/Utils.h:102-115): per frame mean over keypoints of hypot(fx*Jx/Jz+cx-u, fy*Jy/Jz+cy-v)
 * with jointPos (F x 24 x 3, float64) from the final update; frames with nkp = 0 give 0. */
int b3d_mean_pixel_error(int32_t F, const b3d_kp* kps, const int32_t* nkp, const double* jointPos, double fx, double fy,
                         double cx, double cy, double* err);

/* ===================================================================================================== */
/* Entry (2), multi-frame: replaces the ceres::Problem + ceres::Solve inside OptimizeMultiFrame            */
/*   (include/MultiFrameBA.h:33-177) and the anchor / window schedule of src/main_multi_frame.cpp:109-201   */
/* ===================================================================================================== */
/* One window or anchor chain of F frames (opts->mode is ignored; beta_pose is the L2 pose prior, beta_shape the shape
 * prior, lambda_temp the temporal weight; no bounds):
 *   kps F x 17, nkp F, R0 F x 9 = avatars[f]->r[0] (in; the caller applies b3d_writeback afterwards, as MultiFrameBA.h:154-172)
 *   poses F x 76 in/out (FramePoseParams), w 10 in/out (the shared block = avatars.front()->w, MultiFrameBA.h:67)
 *   usable 1, trace trace_cap records or NULL, niters / term 1 or NULL. */
int b3d_fit_window(const b3d_model* m, const b3d_fit_opts* opts, int32_t F, const b3d_kp* kps, const int32_t* nkp,
                   const double* R0, double* poses, double* w, uint8_t* usable, b3d_iter_record* trace, int32_t trace_cap,
                   int32_t* niters, int32_t* term);
/* nprob independent windows in one launch.  Problem p covers frames [first[p], first[p] + count[p]) of the DEVICE frame
 * arrays kps / nkp / R0 / poses / W (N x 10: row first[p] is problem p's shared shape block); first / count are HOST arrays.
 * usable / niters / term: nprob entries (device); trace: nprob x trace_stride (device) or NULL. */
int b3d_fit_windows_dev(const b3d_model* m, const b3d_fit_opts* opts, int32_t nprob, const int32_t* first, const int32_t* count,
                        const b3d_kp* kps, const int32_t* nkp, const double* R0, double* poses, double* W, uint8_t* usable,
                        b3d_iter_record* trace, int32_t trace_stride, int32_t* niters, int32_t* term, void* stream);
/* Chains of windows: chain c = problems [chain_ptr[c], chain_ptr[c + 1]) of the first / count tables, solved ONE AFTER THE OTHER on
 * the device by one thread-block cluster per chain — what `3dba_multi` does per sequence in its window loop
 * (src/main_multi_frame.cpp:165-201: OptimizeMultiFrame on the window, then the write-back of :154-172 before the next window, which
 * overlaps it) — with the write-back of each window's frames into R0 (rotated in place), r (F x 24 x 9) and p (F x 3) done between the
 * solves.  Different chains run side by side and must not share frames (the sharded schedule gives every chain private rows).  All
 * pointers but chain_ptr / first / count are device pointers; windows must be shorter than 48 frames. */
int b3d_fit_window_chains_dev(const b3d_model* m, const b3d_fit_opts* opts, int32_t nchain, const int32_t* chain_ptr, const int32_t* first,
                              const int32_t* count, const b3d_kp* kps, const int32_t* nkp, double* R0, double* poses, double* W, double* r,
                              double* p, uint8_t* usable, int32_t* niters, int32_t* term, void* stream);

/* The anchor / sliding-window schedule of 3dba_multi (src/main_multi_frame.cpp:109-201; argv defaults :29-36). */
typedef struct b3d_seq_opts {
    int32_t skip;            /* anchor_skip (10) */
    int32_t wsize;           /* WSIZE (20) */
    int32_t overlap;         /* OVERLAP (5) */
    int32_t it1;             /* max_iters of the anchor stage (1000) */
    int32_t it2;             /* max_iters of every window: the reference hard-codes 60 (:185) and ignores argv[6] */
    int32_t num_shards;      /* 1 = the reference's serial chain of windows; G > 1 = sharded schedule with halo hand-off */
    double beta_pose, beta_shape, lambda_temp;
    double beta_shape_lock;  /* beta_shape of stage 2: the reference hard-codes 1e5 (:183) */
} b3d_seq_opts;

/* One sequence of N frames, host pointers.  In/out per frame: poses N x 76 (FramePoseParams), R0 N x 9 (avatars[i]->r[0]),
 * W N x 10 (avatars[i]->w).  Out: r N x 24 x 9 and p N x 3 (avatar rotations / translation after the last write-back; may be
 * NULL); solve_iters / solve_term (may be NULL): one entry per LM solve in the order [anchor stage, windows in index order,
 * fix-up re-solves in shard order]; nfail = number of solves whose solution is not usable. */
int b3d_fit_sequence(const b3d_model* m, const b3d_fit_opts* base, const b3d_seq_opts* so, int32_t N, const b3d_kp* kps,
                     const int32_t* nkp, double* poses, double* R0, double* W, double* r, double* p, int32_t* solve_iters,
                     int32_t* solve_term, int32_t* nfail);
/* nseq independent sequences stored back to back in arrays of N frames: sequence q = frames [seq_first[q], + seq_len[q]).
 * Their chains are batched into the same kernel launches (one CTA per sequence and shard per wave). */
int b3d_fit_sequences(const b3d_model* m, const b3d_fit_opts* base, const b3d_seq_opts* so, int32_t nseq,
                      const int32_t* seq_first, const int32_t* seq_len, int32_t N, const b3d_kp* kps, const int32_t* nkp,
                      double* poses, double* R0, double* W, double* r, double* p, int32_t* solve_iters, int32_t* solve_term,
                      int32_t* nfail);

int b3d_mean_pixel_error_dev(int32_t F, const b3d_kp* kps, const int32_t* nkp, const double* jointPos, double fx, double fy,
                             double cx, double cy, double* err, void* stream);

/* ===================================================================================================== */
/* Overlay renderer: smpl::render::renderSMPLMesh (include/RenderSMPLMesh.h:16-110; call sites                */
/* src/main_single_frame.cpp:273-277, src/main_multi_frame.cpp:206-229)                                       */
/* ===================================================================================================== */
/* For each of B bodies: project the cloud (camera coordinates) with fx, fy, cx, cy, cull back faces, flat-shade each face
 * round(220 * clamp(n^ . v^)), and paint the faces far to near into that body's own H x W x 3 uint8 image (in/out: pixels no
 * face covers keep their value, so the caller's frame is the background).  fill = true, wireframe = false as at both call sites.
 *   cloud  B x V x 3 of dtype (what b3d_smpl_forward / b3d_fit_frames produce), images B x H x W x 3.
 * The model must have been created with faces.  Coverage per face is cv::fillConvexPoly's for LINE_8 (the reference passes
 * LINE_AA: its anti-aliased edge blend is not reproduced); equal-depth ties go to the higher face index; faces with a
 * projected pixel coordinate beyond +-2^20 are dropped (cv::Point would overflow / OpenCV would walk millions of rows). */
enum { B3D_RENDER_NO_CULL = 1 /* backface_cull = false */ };
int b3d_render_overlay(const b3d_model* m, int32_t B, int32_t dtype, const void* cloud, double fx, double fy, double cx,
                       double cy, int32_t W, int32_t H, uint32_t flags, uint8_t* images);
/* Same with device pointers, asynchronous on `stream` (scratch planes come from the stream-ordered allocator). */
int b3d_render_overlay_dev(const b3d_model* m, int32_t B, int32_t dtype, const void* cloud, double fx, double fy, double cx,
                           double cy, int32_t W, int32_t H, uint32_t flags, uint8_t* images, void* stream);

/* ===================================================================================================== */
/* The per-frame body of src/main_single_frame.cpp:205-255, batched: fit -> write-back -> update -> pixel error */
/* ===================================================================================================== */
/* Reusable device buffers + one CUDA stream, so that repeated calls do not allocate.  Not thread-safe per handle. */
typedef struct b3d_workspace b3d_workspace;
int b3d_workspace_create(b3d_workspace** out);
void b3d_workspace_destroy(b3d_workspace* ws);

/* Host pointers in, host pointers out (pinned host memory makes the copies asynchronous).  For each of B frames:
 *   OptimizePose[Shape]Reprojection (Sim3BA.h:348/515) from sim3 / w / R0, then the caller-side steps of the reference
 *   main: write-back of r, p (Sim3BA.h:481-507), Avatar::update (main_single_frame.cpp:254), mean_pixel_error (:255).
 * in : kps B x 17, nkp B, R0 B x 9 (avatar.r[0] before the solve), sim3 B x 7 and w B x 10 (both in/out)
 * out: jointAA B x 24 x 3; usable B; niters, term B (may be NULL); r B x 24 x 9, p B x 3, jointPos B x 24 x 3 (float64,
 *      may be NULL); px_err B (may be NULL); cloud B x V x 3 of cloud_dtype (may be NULL: the cloud is then computed and
 *      left in the workspace on the device — b3d_workspace_cloud returns its device address). */
int b3d_fit_frames(b3d_workspace* ws, const b3d_model* m, const b3d_gmm* gmm, const b3d_fit_opts* opts, int32_t B,
                   const b3d_kp* kps, const int32_t* nkp, const double* R0, double* sim3, double* w, double* jointAA,
                   uint8_t* usable, int32_t* niters, int32_t* term, double* r, double* p, double* jointPos, double* px_err,
                   void* cloud, int32_t cloud_dtype);
/* Device address and element type of the cloud left by the last b3d_fit_frames on this workspace (NULL before). */
const void* b3d_workspace_cloud(const b3d_workspace* ws, int32_t* dtype);

#ifdef __cplusplus
}
#endif
#endif /* B3D_H_ */
