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

// small per-model constants passed to kernels by value
struct SkeletonConst {
    int32_t parent[NJ];
    int32_t depth[NJ];
    int32_t max_depth;
    double J0[3 * NJ];          // initialJointPos, J0[3j + c]
    double JSR[3 * NJ * NS];    // jointShapeReg row-major (3j + c) x S
    double offset[3 * NJ];      // rest bone vectors (Sim3BA.h:533-555)
};

}  // namespace b3d

struct b3d_model {
    int device = -1;
    int V = 0, Vpad = 0, F = 0, nw = 0;        // nw = padded non-zeros per vertex (4, 8 or 24)
    bool has_posedirs = false;
    b3d::SkeletonConst sk;
    b3d::SkeletonConst* d_sk = nullptr;         // device copy
    b3d::VertexTables<float> t32;
    b3d::VertexTables<double> t64;
    float* posedirs32 = nullptr;                // [207][3][Vpad] (k, c, v)
    double* posedirs64 = nullptr;
    std::vector<uint32_t> faces;                // host copy (rendering only)
};

struct b3d_gmm {
    int device = -1;
    int nComps = 0, nDims = 0;
    std::vector<double> prec_cho, consts_log, mean;   // host copies
    // device arrays
    double* d_mean = nullptr;        // nComps x 69
    double* d_prec_cho = nullptr;    // nComps x 69 x 69 lower, row-major (L[m][k])
    double* d_prec_choT = nullptr;   // nComps x 69 x 69, L^T row-major (coalesced column access)
    double* d_prec = nullptr;        // nComps x 69 x 69  P = L L^T (symmetric)
    double* d_consts = nullptr;      // nComps  consts_log
};
