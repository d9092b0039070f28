// Internal declarations of lib3dba_b200.so (not part of the C-ABI).
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/b3d.h"

namespace b3d {

constexpr int NJ = B3D_NUM_JOINTS;     // 24
constexpr int NS = B3D_NUM_SHAPE;      // 10
constexpr int NPF = B3D_POSE_DIM;      // 76
constexpr int MAXKP = B3D_MAX_KP;      // 17
constexpr int NFULL = NPF + NS;        // 86: [s, rootAA, rootT, jointAA[1..23], w]
constexpr int NPRI = 3 * (NJ - 1);     // 69 pose-prior dims
constexpr int IDX_S = 0, IDX_RAA = 1, IDX_T = 4, IDX_JAA = 7, IDX_W = NPF;

int set_error(int code, const char* fmt, ...);
void count_launch(int n = 1);

#define B3D_CUDA(expr)                                                                                   \
    do {                                                                                                 \
        cudaError_t e_ = (expr);                                                                         \
        if (e_ != cudaSuccess)                                                                           \
            return b3d::set_error(B3D_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

#define B3D_LAUNCH_CHECK()                                                                               \
    do {                                                                                                 \
        b3d::count_launch();                                                                             \
        cudaError_t e_ = cudaGetLastError();                                                             \
        if (e_ != cudaSuccess)                                                                           \
            return b3d::set_error(B3D_ECUDA, "kernel launch failed: %s (%s:%d)", cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

// element-wise double -> float conversion on the device (util.cu)
int b3d_convert_f64_f32_dev(const double* src, float* dst, size_t n, cudaStream_t st);

// Host-pointer entry points: every used keypoint must name a joint in [0, 24) (it indexes FitConst tables and jointPos on the
// device).  Returns B3D_EINVAL naming the first offender.  The _dev entry points cannot look: their kernels clamp jid instead.
int validate_keypoints(const b3d_kp* kps, const int32_t* nkp, size_t frames, const char* who);
__host__ __device__ inline int clamp_jid(int jid) { return jid < 0 ? 0 : (jid >= NJ ? NJ - 1 : jid); }

// Verifies that the current device can run sm_100a code.  No CPU fallback: callers propagate the error.
int require_device();

// Per-vertex model data in structure-of-arrays form so that thread = vertex loads are coalesced.
//   vert[k * Vpad + v], k in [0, 3)       : v_template component k        (baseCloud, AvatarModel.cpp:44-48)
//   vert[(3 + 10 c + s) * Vpad + v]        : shapedirs[v][c][s]            (keyClouds, AvatarModel.cpp:93-102)
//   wval[k * Vpad + v], wjoint[k*Vpad + v] : k-th non-zero skinning weight of vertex v and its joint, ascending
//                                            joint order, zero padded to `nw` slots (weights, AvatarModel.cpp:66-71)
template <typename T>
struct VertexTables {
    T* vert = nullptr;
    T* wval = nullptr;
    int32_t* wjoint = nullptr;
};

// Tile-sorted copy of the float tables for the float32 LBS kernel (forward.cu, smpl_lbs_sorted_kernel): inside every tile of
// LBS_TILE consecutive vertices the vertices are ordered by their joint tuple, so that the lanes of a warp mostly read the
// same joint transforms from shared memory (broadcast instead of a 32-address gather).  Slot t of a tile holds the vertex
// tile_base + loc[t]; vert / wval are stored in slot order; wjoint holds float offsets (12 * joint) into a body's transforms.
struct SortedTables {
    float* vert = nullptr;       // [(3 + 30)][Vpad]
    float* wval = nullptr;       // [4][Vpad]
    int32_t* wjoff = nullptr;    // [4][Vpad]
    int32_t* loc = nullptr;      // [Vpad]
    // body-lane kernel (smpl_lbs_bodylane_kernel): one 40-float record per slot, in slot order:
    //   [0..2] template xyz, [3] int 3 * loc, [4..13] [14..23] [24..33] shapedirs of x / y / z, [34..37] weights,
    //   [38] int 1 if the joint tuple differs from the previous slot's (or the slot opens a warp's 32-slot share), [39] 0
    float* rec = nullptr;        // [Vpad][40]
    int32_t* joff4 = nullptr;    // [Vpad][4]  float offsets 12 * joint of the four weight slots
};

// Tables of the tensor-core LBS kernel (lbs_tc.cu): the mesh is cut into sort tiles of TC_T = 64 consecutive vertices; a sort
// tile is two MMA tiles of 32 slots.  Per MMA tile one contiguous blob (one TMA bulk copy):
//   [0, 12288)      B operand of the shape blend, 96 columns (slot, coordinate) x K = 32, TF32 hi/lo split, in the canonical
//                   K-major / no-swizzle shared-memory layout tcgen05.mma reads
//   [12288, 13312)  32 slot records of 8 words: 4 skinning weights (aligned with the lane's 4 transform registers sets),
//                   3 * position in the sort tile, control word (reload flags + joint ids), 2 spare
struct TcTables {
    uint8_t* blob = nullptr;
    // pose blendshapes (extension) for the same kernel: per MMA tile 26 K = 8 steps of the [96 columns x 8] B operand, TF32,
    // canonical layout; columns in the tile's slot order (host copy of that order: slot_vertex)
    float* pose_b = nullptr;
    std::vector<int32_t> slot_vertex;   // [sort_tiles * TC_MPT][TC_SLOTS] vertex of every slot (-1: padding)
    int sort_tiles = 0;
    double mean_reloads = 0.0;   // joint transforms re-read from TMEM per slot (host statistic, reported by tools)
};
constexpr int TC_T = 64, TC_SLOTS = 32, TC_N = 96, TC_K = 32;
constexpr int TC_MPT = TC_T / TC_SLOTS;   // MMA tiles per sort tile
constexpr int TC_SEQS = 4;
constexpr int TC_POSE_KSTEPS = (B3D_POSEDIRS_DIM + 7) / 8;   // 26   // slot sequences per sort tile = epilogue warps per TMEM lane quadrant
constexpr int TC_OPER_BYTES = TC_N * TC_K * 4, TC_REC_WORDS = 8, TC_BLOB_BYTES = TC_OPER_BYTES + TC_SLOTS * TC_REC_WORDS * 4;

// small per-model constants passed to kernels by value
struct SkeletonConst {
    int32_t parent[NJ];
    int32_t depth[NJ];
    int32_t max_depth;
    int32_t order[NJ];          // joints in (depth, index) order: a level of the tree is a contiguous run
    int32_t lvstart[NJ + 2];    // lvstart[l] = number of joints of depth < l
    double J0[3 * NJ];          // initialJointPos, J0[3j + c]
    double JSR[3 * NJ * NS];    // jointShapeReg row-major (3j + c) x S
    double offset[3 * NJ];      // rest bone vectors (Sim3BA.h:533-555)
};

}  // namespace b3d

struct b3d_model;
namespace b3d {
// lbs_tc.cu
int build_tc_tables(b3d_model& m, const std::vector<double>& vert, const std::vector<double>& wval, const std::vector<int32_t>& wjoint);
int build_tc_pose_tables(b3d_model& m, const std::vector<float>& p32);   // p32[(k * 3 + c) * Vpad + v]
// posefeat: B x 207 row-major pose features (joints kernel) or null = reference semantics (no pose blendshapes)
int launch_lbs_tc(const b3d_model* m, int B, const float* wshape, const float* G, const float* posefeat, float* cloud, cudaStream_t st);
}  // namespace b3d

struct b3d_model {
    int device = -1;
    int V = 0, Vpad = 0, F = 0, nw = 0;        // nw = padded non-zeros per vertex (4, 8 or 24)
    bool has_posedirs = false;
    b3d::SkeletonConst sk;
    b3d::SkeletonConst* d_sk = nullptr;         // device copy
    b3d::VertexTables<float> t32;
    b3d::VertexTables<double> t64;
    b3d::SortedTables ts;                       // only when nw == 4
    b3d::TcTables tc;                           // only when nw == 4
    float* posedirs32 = nullptr;                // [207][3][Vpad] (k, c, v)
    double* posedirs64 = nullptr;
    std::vector<uint32_t> faces;                // host copy (rendering only)
    int32_t* d_faces = nullptr;                 // F x 3 on the device (render.cu)
};

struct b3d_gmm {
    int device = -1;
    int nComps = 0, nDims = 0;
    std::vector<double> prec_cho, consts_log, mean;   // host copies
    // device arrays
    double* d_mean = nullptr;        // nComps x 69
    double* d_prec_cho = nullptr;    // nComps x 69 x 69 lower, row-major (L[m][k])
    double* d_prec_cho70 = nullptr;  // nComps x 69 x 70: L with rows padded to 70 doubles (16-byte aligned column pairs)
    double* d_prec_choT = nullptr;   // nComps x 69 x 69, L^T row-major (coalesced column access)
    double* d_prec = nullptr;        // nComps x 69 x 69  P = L L^T (symmetric)
    double* d_consts = nullptr;      // nComps  consts_log
    double* d_sigma = nullptr;       // nComps  upper bound of ||L_i||_2
};
