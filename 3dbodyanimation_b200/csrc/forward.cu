// Entry (1): batched SMPL evaluator — replaces ark::Avatar::update (external/avatar/Avatar.cpp:22-75).
//
//   smpl_joints_kernel : per body, jointPos = J0 + JSR w (Avatar.cpp:32-36), the 24-joint chain (:47-57), the
//                        rest->posed conversion (:59-64).  One warp per body, float64 arithmetic whatever the I/O type.
//   smpl_lbs_kernel    : per (vertex, body): shape blend (:26), linear-blend skinning (:69-73).  thread = vertex; the
//                        33 template/shapedir values and the vertex's weights stay in registers while the CTA streams
//                        bodies, whose 24 joint transforms are staged in shared memory.  HBM traffic is the cloud
//                        write (3 V sizeof(T) per body); everything else is L2 resident.
//   posedirs_kernel    : optional standard-SMPL pose blendshapes (an extension: the reference has them commented out,
//                        AvatarModel.cpp:104-109), CUDA-core path.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>

#include "internal.hpp"

namespace b3d {

constexpr int JW = 4;   // warps (= bodies) per CTA in the joints kernel
constexpr int TC_AUTO_MIN_B = 512;   // batches from which the tensor-core LBS kernel (lbs_tc.cu) is the default float32 path

template <typename T>
__global__ void __launch_bounds__(32 * JW) smpl_joints_kernel(const SkeletonConst* __restrict__ sk, int B,
                                                              const T* __restrict__ w, const T* __restrict__ p,
                                                              const T* __restrict__ r, T* __restrict__ jointPos,
                                                              T* __restrict__ jointTrans, T* __restrict__ posefeat, int bpw) {
    __shared__ double s_r[JW][NJ * 9];
    __shared__ double s_g[JW][NJ * 12];
    __shared__ double s_jp[JW][NJ * 3];
    __shared__ double s_w[JW][NS];
    // model constants of the CTA: jointShapeReg transposed ([s][3 j + c]: lanes = rows read consecutive words; straight from
    // global memory the row stride of 80 B made every one of these loads a 20-sector gather), initialJointPos, level order
    __shared__ double s_jsr[NS][NJ * 3], s_j0[NJ * 3];
    __shared__ int s_order[NJ], s_lvstart[NJ + 2], s_parent[NJ];
    for (int i = threadIdx.x; i < NJ * 3 * NS; i += 32 * JW) s_jsr[i % NS][i / NS] = sk->JSR[i];
    for (int i = threadIdx.x; i < NJ * 3; i += 32 * JW) s_j0[i] = sk->J0[i];
    if (threadIdx.x < NJ) { s_order[threadIdx.x] = sk->order[threadIdx.x]; s_parent[threadIdx.x] = sk->parent[threadIdx.x]; }
    if (threadIdx.x < NJ + 2) s_lvstart[threadIdx.x] = sk->lvstart[threadIdx.x];
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // a warp takes bpw consecutive bodies, so that the staging above is paid once per JW * bpw bodies
    // a body's 216 rotation entries travel through 7 registers per lane; the next body's are requested before this body's chain
    // starts, so their latency hides behind it
    constexpr int RV = (NJ * 9 + 31) / 32;
    T rv[RV];
    {
        const int bf = (blockIdx.x * JW + warp) * bpw;
#pragma unroll
        for (int u = 0; u < RV; ++u) rv[u] = (bf < B && lane + 32 * u < NJ * 9) ? r[(size_t)bf * NJ * 9 + lane + 32 * u] : (T)0;
    }
    for (int t = 0; t < bpw; ++t) {
    const int b = (blockIdx.x * JW + warp) * bpw + t;
    if (b >= B) break;
    double* R = s_r[warp];
    double* G = s_g[warp];
    double* jp = s_jp[warp];
#pragma unroll
    for (int u = 0; u < RV; ++u)
        if (lane + 32 * u < NJ * 9) R[lane + 32 * u] = (double)rv[u];
    if (t + 1 < bpw && b + 1 < B) {
#pragma unroll
        for (int u = 0; u < RV; ++u)
            if (lane + 32 * u < NJ * 9) rv[u] = r[(size_t)(b + 1) * NJ * 9 + lane + 32 * u];
    }
    if (lane < NS) s_w[warp][lane] = (double)w[(size_t)b * NS + lane];
    __syncwarp();
    // jointPos = initialJointPos + jointShapeReg * w
    for (int i = lane; i < NJ * 3; i += 32) {
        double acc = 0.0;
#pragma unroll
        for (int s = 0; s < NS; ++s) acc += s_jsr[s][i] * s_w[warp][s];
        jp[i] = s_j0[i] + acc;
    }
    __syncwarp();
    // chain, level by level: global = parent o [r_j | jp_j - jp_parent]   (mulAffine, Util.h:191-199)
    const int j = lane;
    const int depth = j < NJ ? sk->depth[j] : -1;
    const int pa = j < NJ ? sk->parent[j] : 0;
    const int max_depth = sk->max_depth;
    if (j == 0) {
#pragma unroll
        for (int k = 0; k < 9; ++k) G[k] = R[k];
#pragma unroll
        for (int c = 0; c < 3; ++c) G[9 + c] = (double)p[(size_t)b * 3 + c];
    }
    __syncwarp();
    const int* order = s_order;      // joints in depth order (model.cu): a level is the run order[lvstart[l] .. lvstart[l + 1])
    const int* lvstart = s_lvstart;
    // One lane per ELEMENT of a child's 3 x 4 matrix (not per joint): a level of n joints is n * 12 independent 3-term sums, so
    // the warp is busy whatever the width of the tree and a level costs 7 shared-memory instructions per pass instead of 39.
    for (int lvl = 1; lvl <= max_depth; ++lvl) {
        const int l0 = lvstart[lvl], work = (lvstart[lvl + 1] - l0) * 12;
        for (int base = 0; base < work; base += 32) {
            const int idx = base + lane;
            if (idx < work) {
                const int jj = order[l0 + idx / 12], e = idx % 12;
                const int pj = s_parent[jj];
                const double* A = G + 12 * pj;
                if (e < 9) {
                    const int col = e / 3, row = e - 3 * col;
                    const double* Rl = R + 9 * jj + 3 * col;
                    G[12 * jj + e] = A[row] * Rl[0] + A[3 + row] * Rl[1] + A[6 + row] * Rl[2];
                } else {
                    const int row = e - 9;
                    const double t0 = jp[3 * jj] - jp[3 * pj], t1 = jp[3 * jj + 1] - jp[3 * pj + 1], t2 = jp[3 * jj + 2] - jp[3 * pj + 2];
                    G[12 * jj + e] = A[9 + row] + (A[row] * t0 + A[3 + row] * t1 + A[6 + row] * t2);
                }
            }
        }
        __syncwarp();
    }
    // jointPos <- translation; translation <- t - R * Jinit
    if (j < NJ) {
        double* Bm = G + 12 * j;
        const double j0 = jp[3 * j], j1 = jp[3 * j + 1], j2 = jp[3 * j + 2];
        const double q0 = Bm[9], q1 = Bm[10], q2 = Bm[11];
#pragma unroll
        for (int row = 0; row < 3; ++row) Bm[9 + row] -= Bm[row] * j0 + Bm[3 + row] * j1 + Bm[6 + row] * j2;
        jp[3 * j] = q0; jp[3 * j + 1] = q1; jp[3 * j + 2] = q2;
    }
    __syncwarp();
    if (jointPos)
        for (int i = lane; i < NJ * 3; i += 32) jointPos[(size_t)b * NJ * 3 + i] = (T)jp[i];
    for (int i = lane; i < NJ * 12; i += 32) jointTrans[(size_t)b * NJ * 12 + i] = (T)G[i];
    if (posefeat)   // vec_rowmajor(R_j - I), j = 1..23
        for (int i = lane; i < B3D_POSEDIRS_DIM; i += 32) {
            const int jj = i / 9 + 1, a = (i % 9) / 3, c = i % 3;
            posefeat[(size_t)b * B3D_POSEDIRS_DIM + i] = (T)(R[9 * jj + 3 * c + a] - (a == c ? 1.0 : 0.0));
        }
    __syncwarp();   // the next body overwrites R / G / jp
    }
}

// ---------------------------------------------------------------------------------------------------------------
// posedirs (extension): off[b][c][v] = sum_k posedirs[k][c][v] * pf[b][k].  thread = vertex, PB bodies per pass.
// ---------------------------------------------------------------------------------------------------------------
constexpr int PB = 8;
template <typename T>
__global__ void __launch_bounds__(128) posedirs_kernel(const T* __restrict__ pd, int V, int Vpad, int B,
                                                       const T* __restrict__ posefeat, T* __restrict__ off) {
    __shared__ T s_pf[PB][B3D_POSEDIRS_DIM];
    const int v = blockIdx.x * 128 + threadIdx.x;
    const int b0 = blockIdx.y * PB;
    const int nb = min(PB, B - b0);
    for (int i = threadIdx.x; i < PB * B3D_POSEDIRS_DIM; i += 128) {
        const int bb = i / B3D_POSEDIRS_DIM;
        s_pf[bb][i % B3D_POSEDIRS_DIM] = bb < nb ? posefeat[(size_t)(b0 + bb) * B3D_POSEDIRS_DIM + i % B3D_POSEDIRS_DIM] : T(0);
    }
    __syncthreads();
    if (v >= V) return;
    T acc[PB][3];
#pragma unroll
    for (int bb = 0; bb < PB; ++bb) acc[bb][0] = acc[bb][1] = acc[bb][2] = T(0);
    for (int k = 0; k < B3D_POSEDIRS_DIM; ++k) {
        const T a0 = pd[((size_t)k * 3 + 0) * Vpad + v], a1 = pd[((size_t)k * 3 + 1) * Vpad + v], a2 = pd[((size_t)k * 3 + 2) * Vpad + v];
#pragma unroll
        for (int bb = 0; bb < PB; ++bb) {
            const T f = s_pf[bb][k];
            acc[bb][0] += a0 * f; acc[bb][1] += a1 * f; acc[bb][2] += a2 * f;
        }
    }
    for (int bb = 0; bb < nb; ++bb)
#pragma unroll
        for (int c = 0; c < 3; ++c) off[((size_t)(b0 + bb) * 3 + c) * Vpad + v] = acc[bb][c];
}

// ---------------------------------------------------------------------------------------------------------------
// LBS
// ---------------------------------------------------------------------------------------------------------------
constexpr int LBS_THREADS = 256;   // vertices per CTA tile
constexpr int LBS_NB = 8;          // bodies staged in shared memory per pass

template <typename T, int NW, bool HAS_OFF>
__global__ void __launch_bounds__(LBS_THREADS) smpl_lbs_kernel(const T* __restrict__ vert, const T* __restrict__ wval,
                                                               const int32_t* __restrict__ wjoint, int V, int Vpad, int B,
                                                               int bodies_per_cta, const T* __restrict__ wshape,
                                                               const T* __restrict__ G, const T* __restrict__ off,
                                                               T* __restrict__ cloud) {
    __shared__ __align__(16) T s_g[LBS_NB][NJ * 12];
    __shared__ T s_w[LBS_NB][NS];
    __shared__ T s_out[LBS_THREADS / 32][96];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int v = blockIdx.x * LBS_THREADS + tid;       // < Vpad by construction of the grid
    const int vw = blockIdx.x * LBS_THREADS + warp * 32;  // first vertex of this warp
    const int nout = 3 * max(0, min(32, V - vw));       // valid output scalars of this warp per body

    T base[3], sd[3][NS], ww[NW];
    int wj[NW];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        base[c] = vert[(size_t)c * Vpad + v];
#pragma unroll
        for (int s = 0; s < NS; ++s) sd[c][s] = vert[(size_t)(3 + NS * c + s) * Vpad + v];
    }
#pragma unroll
    for (int k = 0; k < NW; ++k) {
        ww[k] = wval[(size_t)k * Vpad + v];
        wj[k] = wjoint[(size_t)k * Vpad + v] * 12;
    }

    const int b_begin = blockIdx.y * bodies_per_cta;
    const int b_end = min(B, b_begin + bodies_per_cta);
    for (int b0 = b_begin; b0 < b_end; b0 += LBS_NB) {
        const int nb = min(LBS_NB, b_end - b0);
        __syncthreads();   // previous pass finished reading s_g / s_w
        for (int i = tid; i < nb * NJ * 12; i += LBS_THREADS) (&s_g[0][0])[i] = G[(size_t)b0 * NJ * 12 + i];
        for (int i = tid; i < nb * NS; i += LBS_THREADS) (&s_w[0][0])[i] = wshape[(size_t)b0 * NS + i];
        __syncthreads();
        for (int i = 0; i < nb; ++i) {
            const T* sw = s_w[i];
            T x = base[0], y = base[1], z = base[2];
#pragma unroll
            for (int s = 0; s < NS; ++s) {
                const T ws = sw[s];
                x += sd[0][s] * ws; y += sd[1][s] * ws; z += sd[2][s] * ws;
            }
            if (HAS_OFF) {
                const size_t o = (size_t)(b0 + i) * 3 * Vpad + v;
                x += off[o]; y += off[o + Vpad]; z += off[o + 2 * (size_t)Vpad];
            }
            T ax = T(0), ay = T(0), az = T(0);
#pragma unroll
            for (int k = 0; k < NW; ++k) {
                const T* g = &s_g[i][wj[k]];   // col-major 3x4: g[3 col + row]
                const T tx = g[0] * x + g[3] * y + g[6] * z + g[9];
                const T ty = g[1] * x + g[4] * y + g[7] * z + g[10];
                const T tz = g[2] * x + g[5] * y + g[8] * z + g[11];
                ax += ww[k] * tx; ay += ww[k] * ty; az += ww[k] * tz;
            }
            // transpose through shared memory so that the warp writes 3 x 32 consecutive scalars
            T* so = s_out[warp];
            so[3 * lane] = ax; so[3 * lane + 1] = ay; so[3 * lane + 2] = az;
            __syncwarp();
            T* dst = cloud + (size_t)(b0 + i) * 3 * V + (size_t)3 * vw;
#pragma unroll
            for (int q = 0; q < 3; ++q) {
                const int e = q * 32 + lane;
                if (e < nout) dst[e] = so[e];
            }
            __syncwarp();
        }
    }
}


// ---------------------------------------------------------------------------------------------------------------
// LBS, float32, <= 4 weights per vertex: the HBM-bound kernel of the path (82,680 B of cloud per body).
//
// The generic kernel above is bound by shared-memory bandwidth, not HBM: every lane gathers its own 4 x 12 joint-transform
// floats per body (6 KB per warp and body through a 128 B/clk crossbar).  Here a CTA still owns 256 consecutive vertices
// (so that its output is one contiguous 3 KB run per body), but inside the tile the vertices are processed in joint-tuple
// order (SortedTables): the lanes of a warp then share their joint transforms and each LDS.128 is a broadcast of a few
// addresses.  Results are scattered into a shared staging row at their original position and leave as coalesced 8-byte
// stores, two bodies per barrier, double buffered.
// ---------------------------------------------------------------------------------------------------------------
constexpr int LS_TILE = 256;                       // vertices per CTA = threads per CTA
constexpr int LS_NB = 8;                           // bodies staged per pass (even)
constexpr int LS_ROW = 3 * LS_TILE + 2 * (LS_TILE / 32);   // staging row: 2 pad floats after every 96 (bank spreading, keeps 8 B alignment)

__device__ __forceinline__ int ls_pos(int e) { return e + 2 * (e / 96); }   // element 3 * loc + c -> padded position

__global__ void __launch_bounds__(LS_TILE, 3) smpl_lbs_sorted_kernel(const float* __restrict__ vert, const float* __restrict__ wval,
                                                                      const int32_t* __restrict__ wjoff, const int32_t* __restrict__ locs,
                                                                      int V, int Vpad, int B, int bodies_per_cta,
                                                                      const float* __restrict__ wshape, const float* __restrict__ G,
                                                                      float* __restrict__ cloud) {
    __shared__ __align__(16) float s_g[LS_NB][NJ * 12];
    __shared__ __align__(16) float s_w[LS_NB][12];
    __shared__ __align__(16) float s_out[2][2][LS_ROW];
    const int tid = threadIdx.x;
    const int tile_base = blockIdx.x * LS_TILE;
    const int slot = tile_base + tid;                   // < Vpad by construction of the grid
    const int nvalid = min(LS_TILE, V - tile_base);     // vertices of this tile that exist
    const int nout2 = (3 * nvalid) / 2;                 // float2 per body (3 * nvalid is even: V and LS_TILE are even)

    float base[3], sd[3][NS], ww[4];
    int wj[4];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        base[c] = vert[(size_t)c * Vpad + slot];
#pragma unroll
        for (int s = 0; s < NS; ++s) sd[c][s] = vert[(size_t)(3 + NS * c + s) * Vpad + slot];
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        ww[k] = wval[(size_t)k * Vpad + slot];
        wj[k] = wjoff[(size_t)k * Vpad + slot];
    }
    const int opos = ls_pos(3 * locs[slot]);            // 3 * loc + c stays inside one 96-block for c = 0..2

    const int b_begin = blockIdx.y * bodies_per_cta;
    const int b_end = min(B, b_begin + bodies_per_cta);
    int buf = 0;
    for (int b0 = b_begin; b0 < b_end; b0 += LS_NB) {
        const int nb = min(LS_NB, b_end - b0);
        __syncthreads();   // previous pass finished reading s_g / s_w
        {
            const float4* src = reinterpret_cast<const float4*>(G + (size_t)b0 * NJ * 12);   // 1152 B per body: 16 B aligned
            float4* dst = reinterpret_cast<float4*>(&s_g[0][0]);
            for (int i = tid; i < nb * (NJ * 3); i += LS_TILE) dst[i] = src[i];
            for (int i = tid; i < nb * 12; i += LS_TILE) {
                const int bb = i / 12, s = i % 12;
                s_w[bb][s] = s < NS ? wshape[(size_t)(b0 + bb) * NS + s] : 0.f;
            }
        }
        __syncthreads();
        for (int i = 0; i < nb; i += 2) {
            const int npair = min(2, nb - i);
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if (h < npair) {
                    const float4* sw4 = reinterpret_cast<const float4*>(s_w[i + h]);
                    const float4 w0 = sw4[0], w1 = sw4[1], w2 = sw4[2];
                    const float wsv[NS] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w, w2.x, w2.y};
                    float x = base[0], y = base[1], z = base[2];
#pragma unroll
                    for (int s = 0; s < NS; ++s) {
                        x = fmaf(sd[0][s], wsv[s], x); y = fmaf(sd[1][s], wsv[s], y); z = fmaf(sd[2][s], wsv[s], z);
                    }
                    float ax = 0.f, ay = 0.f, az = 0.f;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const float4* g4 = reinterpret_cast<const float4*>(&s_g[i + h][wj[k]]);   // col-major 3x4: g[3 col + row]
                        const float4 ga = g4[0], gb = g4[1], gc = g4[2];
                        const float tx = fmaf(ga.x, x, fmaf(ga.w, y, fmaf(gb.z, z, gc.y)));
                        const float ty = fmaf(ga.y, x, fmaf(gb.x, y, fmaf(gb.w, z, gc.z)));
                        const float tz = fmaf(ga.z, x, fmaf(gb.y, y, fmaf(gc.x, z, gc.w)));
                        ax = fmaf(ww[k], tx, ax); ay = fmaf(ww[k], ty, ay); az = fmaf(ww[k], tz, az);
                    }
                    float* so = s_out[buf][h] + opos;
                    so[0] = ax; so[1] = ay; so[2] = az;
                }
            }
            __syncthreads();
            // copy-out: per body 3 * nvalid consecutive floats of the cloud, as float2 (body rows are 8 B aligned: 3 V is even)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if (h < npair) {
                    float2* dst = reinterpret_cast<float2*>(cloud + ((size_t)(b0 + i + h) * V + tile_base) * 3);
                    const float* row = s_out[buf][h];
                    for (int e2 = tid; e2 < nout2; e2 += LS_TILE)
                        dst[e2] = *reinterpret_cast<const float2*>(row + ls_pos(2 * e2));
                }
            }
            buf ^= 1;
        }
    }
}


// ---------------------------------------------------------------------------------------------------------------
// LBS, float32, large batches: lane = body.
//
// With lane = vertex every lane has to fetch its own 4 x 12 transform floats per body from shared memory: 48 register-fill
// wavefronts per warp and body, which is what bounds the kernels above (ncu: LSU wavefronts 72 % of peak, DRAM 15 %).
// Here the roles are swapped.  A CTA owns one 256-vertex range and streams batches of 32 bodies; a lane keeps ITS body's shape
// vector and the four transforms of the current joint tuple in registers, and the warp walks its 32 slots of the range in
// joint-tuple order: per slot the 38 per-vertex constants arrive as ten warp-uniform LDS.128 and the transforms are re-read only
// when the tuple changes.  Results go to a [body][3 x 256] staging tile (odd row stride: conflict free) and leave as full rows.
// The transforms of the next batch are fetched by TMA bulk copies (cp.async.bulk + mbarrier) while the current one computes.
// ---------------------------------------------------------------------------------------------------------------
constexpr int BL_R = 256, BL_THREADS = 256, BL_NB = 32, BL_REC = 40;
constexpr int BL_GROW = NJ * 12 + 4;     // 292 floats: rows stay 16 B aligned, lanes of a quarter warp hit distinct 16 B bank groups
constexpr int BL_OROW = 3 * BL_R + 1;    // 769

struct __align__(16) BlSmem {
    float rec[BL_R * BL_REC];
    int4 joff[BL_R];
    float g[2][BL_NB * BL_GROW];
    float beta[2][BL_NB * NS];
    float out[BL_NB * BL_OROW];
    unsigned long long mbar[2];
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// TMA bulk copy global -> shared, completion counted on an mbarrier (bytes % 16 == 0, both addresses 16 B aligned)
__device__ __forceinline__ void tma_bulk_g2s(void* dst, const void* src, uint32_t bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src),
                 "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

__global__ void __launch_bounds__(BL_THREADS, 1) smpl_lbs_bodylane_kernel(const float* __restrict__ rec, const int32_t* __restrict__ joff4,
                                                                           int V, int B, int bodies_per_cta, const float* __restrict__ wshape,
                                                                           const float* __restrict__ G, float* __restrict__ cloud) {
    extern __shared__ __align__(16) unsigned char bl_raw[];
    BlSmem& S = *reinterpret_cast<BlSmem*>(bl_raw);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int range_base = blockIdx.x * BL_R;
    const int nvalid = min(BL_R, V - range_base);
    const int b_begin = blockIdx.y * bodies_per_cta;
    const int b_end = min(B, b_begin + bodies_per_cta);
    if (b_begin >= b_end) return;

    // per-range constants -> shared memory (once per CTA)
    {
        const float4* src = reinterpret_cast<const float4*>(rec + (size_t)range_base * BL_REC);
        float4* dst = reinterpret_cast<float4*>(S.rec);
        for (int i = tid; i < BL_R * BL_REC / 4; i += BL_THREADS) dst[i] = src[i];
        S.joff[tid] = reinterpret_cast<const int4*>(joff4)[range_base + tid];
    }
    if (tid == 0) {
        mbar_init(&S.mbar[0], 1);
        mbar_init(&S.mbar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    auto load_beta = [&](int buf, int b0) {   // plain loads: 10 floats per body, read back by the owning lane
        const int nb = min(BL_NB, b_end - b0);
        for (int i = tid; i < nb * NS; i += BL_THREADS) S.beta[buf][i] = wshape[(size_t)b0 * NS + i];
    };
    auto issue_g = [&](int buf, int b0) {     // one thread: one bulk copy per body row (1152 B), all counted on mbar[buf]
        const int nb = min(BL_NB, b_end - b0);
        mbar_expect_tx(&S.mbar[buf], (uint32_t)nb * NJ * 12 * 4);
        for (int i = 0; i < nb; ++i) tma_bulk_g2s(&S.g[buf][i * BL_GROW], G + (size_t)(b0 + i) * NJ * 12, NJ * 12 * 4, &S.mbar[buf]);
    };
    load_beta(0, b_begin);
    __syncthreads();
    if (tid == 0) issue_g(0, b_begin);

    const int slot_begin = warp * 32, slot_end = min(slot_begin + 32, nvalid);
    int it = 0;
    for (int b0 = b_begin; b0 < b_end; b0 += BL_NB, ++it) {
        const int buf = it & 1;
        const int nb = min(BL_NB, b_end - b0);
        const bool has_next = b0 + BL_NB < b_end;
        if (has_next && tid == 0) issue_g(buf ^ 1, b0 + BL_NB);      // buffer buf^1 was last read before the barrier that closed batch it-1
        mbar_wait(&S.mbar[buf], (uint32_t)((it >> 1) & 1));

        float be[NS];
#pragma unroll
        for (int s = 0; s < NS; ++s) be[s] = S.beta[buf][lane * NS + s];
        const float* grow = &S.g[buf][lane * BL_GROW];
        float* orow = &S.out[lane * BL_OROW];
        float4 g[4][3];
#pragma unroll
        for (int k = 0; k < 4; ++k) g[k][0] = g[k][1] = g[k][2] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 2
        for (int slot = slot_begin; slot < slot_end; ++slot) {
            const float4* r4 = reinterpret_cast<const float4*>(&S.rec[slot * BL_REC]);
            const float4 q0 = r4[0], q1 = r4[1], q2 = r4[2], q3 = r4[3], q4 = r4[4], q5 = r4[5], q6 = r4[6], q7 = r4[7], q8 = r4[8], q9 = r4[9];
            if (__float_as_int(q9.z)) {   // warp-uniform: the joint tuple changed
                const int4 jo = S.joff[slot];
                const int o[4] = {jo.x, jo.y, jo.z, jo.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float4* g4 = reinterpret_cast<const float4*>(grow + o[k]);
                    g[k][0] = g4[0]; g[k][1] = g4[1]; g[k][2] = g4[2];
                }
            }
            // record: q0 = (bx, by, bz, 3 loc); sdx = q1 q2 q3.xy; sdy = q3.zw q4 q5; sdz = q6 q7 q8.xy; weights = q8.zw q9.xy
            float x = q0.x, y = q0.y, z = q0.z;
            x = fmaf(q1.x, be[0], x); x = fmaf(q1.y, be[1], x); x = fmaf(q1.z, be[2], x); x = fmaf(q1.w, be[3], x);
            x = fmaf(q2.x, be[4], x); x = fmaf(q2.y, be[5], x); x = fmaf(q2.z, be[6], x); x = fmaf(q2.w, be[7], x);
            x = fmaf(q3.x, be[8], x); x = fmaf(q3.y, be[9], x);
            y = fmaf(q3.z, be[0], y); y = fmaf(q3.w, be[1], y);
            y = fmaf(q4.x, be[2], y); y = fmaf(q4.y, be[3], y); y = fmaf(q4.z, be[4], y); y = fmaf(q4.w, be[5], y);
            y = fmaf(q5.x, be[6], y); y = fmaf(q5.y, be[7], y); y = fmaf(q5.z, be[8], y); y = fmaf(q5.w, be[9], y);
            z = fmaf(q6.x, be[0], z); z = fmaf(q6.y, be[1], z); z = fmaf(q6.z, be[2], z); z = fmaf(q6.w, be[3], z);
            z = fmaf(q7.x, be[4], z); z = fmaf(q7.y, be[5], z); z = fmaf(q7.z, be[6], z); z = fmaf(q7.w, be[7], z);
            z = fmaf(q8.x, be[8], z); z = fmaf(q8.y, be[9], z);
            const float wk[4] = {q8.z, q8.w, q9.x, q9.y};
            float ax = 0.f, ay = 0.f, az = 0.f;
#pragma unroll
            for (int k = 0; k < 4; ++k) {   // col-major 3x4 [R|t]: ga = (r00 r10 r20 r01), gb = (r11 r21 r02 r12), gc = (r22 t0 t1 t2)
                const float4 ga = g[k][0], gb = g[k][1], gc = g[k][2];
                const float tx = fmaf(ga.x, x, fmaf(ga.w, y, fmaf(gb.z, z, gc.y)));
                const float ty = fmaf(ga.y, x, fmaf(gb.x, y, fmaf(gb.w, z, gc.z)));
                const float tz = fmaf(ga.z, x, fmaf(gb.y, y, fmaf(gc.x, z, gc.w)));
                ax = fmaf(wk[k], tx, ax); ay = fmaf(wk[k], ty, ay); az = fmaf(wk[k], tz, az);
            }
            float* so = orow + __float_as_int(q0.w);
            so[0] = ax; so[1] = ay; so[2] = az;
        }
        __syncthreads();
        // copy-out: rows of 3 * nvalid consecutive floats; meanwhile stage the next batch's shape vectors
        if (has_next) load_beta(buf ^ 1, b0 + BL_NB);
        const int nfl = 3 * nvalid;
        for (int row = 0; row < nb; ++row) {
            float* dst = cloud + ((size_t)(b0 + row) * V + range_base) * 3;
            const float* src = &S.out[row * BL_OROW];
            for (int e = tid; e < nfl; e += BL_THREADS) dst[e] = src[e];
        }
        __syncthreads();
    }
}

template <typename T>
static int forward_impl(const b3d_model* m, int B, const T* w, const T* p, const T* r, T* cloud, T* jointPos, T* jointTrans,
                        int flags, cudaStream_t st) {
    const VertexTables<T>& t = sizeof(T) == 4 ? (const VertexTables<T>&)m->t32 : (const VertexTables<T>&)m->t64;
    const bool use_pd = (flags & B3D_FWD_POSEDIRS) != 0;
    if (use_pd && !m->has_posedirs) return set_error(B3D_EINVAL, "B3D_FWD_POSEDIRS requested but the model has no posedirs");
    if ((flags & B3D_FWD_GIVEN_TRANSFORMS) && (!jointTrans || !cloud || use_pd))   // validated before anything is allocated
        return set_error(B3D_EINVAL, "B3D_FWD_GIVEN_TRANSFORMS needs jointTrans and cloud, and excludes B3D_FWD_POSEDIRS");
    // stream-ordered scratch, released on every return path
    struct Scratch {
        cudaStream_t st;
        void* p[3] = {nullptr, nullptr, nullptr};
        ~Scratch() { for (void* q : p) if (q) cudaFreeAsync(q, st); }
    } scratch{st};
    T* G = jointTrans;
    T* pf = nullptr;
    T* off = nullptr;
    if (!G) { B3D_CUDA(cudaMallocAsync(&scratch.p[0], sizeof(T) * (size_t)B * NJ * 12, st)); G = (T*)scratch.p[0]; }
    // B3D_FWD_LBS_* flags (or env B3D_LBS = tc | sorted | bodylane | generic) force a float32 path (tests, profiling); default:
    // tensor-core kernel (lbs_tc.cu; with the pose blendshapes fused in when they are asked for) for batches that fill the
    // machine, tile-sorted lane = vertex kernel below that
    static const std::string want_env = [] { const char* e = getenv("B3D_LBS"); return std::string(e ? e : ""); }();
    const std::string want = (flags & B3D_FWD_LBS_TC) ? "tc" : (flags & B3D_FWD_LBS_SORTED) ? "sorted" : (flags & B3D_FWD_LBS_BODYLANE) ? "bodylane"
                             : (flags & B3D_FWD_LBS_GENERIC) ? "generic" : want_env;
    const bool g_al16 = (reinterpret_cast<uintptr_t>(G) & 15) == 0;
    const bool use_tc = cloud && sizeof(T) == 4 && m->nw == 4 && g_al16 && m->tc.blob && (!use_pd || m->tc.pose_b) &&
                        (want == "tc" || (want.empty() && B >= TC_AUTO_MIN_B));
    if (use_pd && cloud) {
        B3D_CUDA(cudaMallocAsync(&scratch.p[1], sizeof(T) * (size_t)B * B3D_POSEDIRS_DIM, st));
        pf = (T*)scratch.p[1];
        if (!use_tc) { B3D_CUDA(cudaMallocAsync(&scratch.p[2], sizeof(T) * (size_t)B * 3 * m->Vpad, st)); off = (T*)scratch.p[2]; }
    }
    if (!(flags & B3D_FWD_GIVEN_TRANSFORMS)) {
        const int bpw = B >= 4096 ? 4 : 1;   // bodies per warp: amortise the CTA's constant staging once the machine is full
        smpl_joints_kernel<T><<<(B + JW * bpw - 1) / (JW * bpw), 32 * JW, 0, st>>>(m->d_sk, B, w, p, r, jointPos, G, pf, bpw);
        B3D_LAUNCH_CHECK();
    }
    if (use_tc) {
        if (int rc = launch_lbs_tc(m, B, (const float*)w, (const float*)G, (const float*)pf, (float*)cloud, st)) return rc;
        B3D_LAUNCH_CHECK();
    } else if (cloud) {
        if (use_pd) {
            const T* pd = sizeof(T) == 4 ? (const T*)m->posedirs32 : (const T*)m->posedirs64;
            dim3 grid(m->Vpad / 128, (B + PB - 1) / PB);
            posedirs_kernel<T><<<grid, 128, 0, st>>>(pd, m->V, m->Vpad, B, pf, off);
            B3D_LAUNCH_CHECK();
        }
        const int tiles = (m->Vpad + LBS_THREADS - 1) / LBS_THREADS;
        // enough CTAs to fill 148 SMs several times over, but each CTA keeps its vertex registers for many bodies
        int groups = (148 * 8 + tiles - 1) / tiles;
        if (groups > B) groups = B;
        int per = (B + groups - 1) / groups;
        per = (per + LBS_NB - 1) / LBS_NB * LBS_NB;
        groups = (B + per - 1) / per;
        dim3 grid(tiles, groups);
#define B3D_LBS(NWV)                                                                                                     \
    do {                                                                                                                 \
        if (use_pd) smpl_lbs_kernel<T, NWV, true><<<grid, LBS_THREADS, 0, st>>>(t.vert, t.wval, t.wjoint, m->V, m->Vpad, B, per, w, G, off, cloud); \
        else smpl_lbs_kernel<T, NWV, false><<<grid, LBS_THREADS, 0, st>>>(t.vert, t.wval, t.wjoint, m->V, m->Vpad, B, per, w, G, nullptr, cloud);   \
    } while (0)
        const bool want_bodylane = want == "bodylane";
        if (false) {
        } else if (want != "generic" && want_bodylane && sizeof(T) == 4 && m->nw == 4 && !use_pd && m->ts.rec && g_al16 && B >= 256) {
            B3D_CUDA(cudaFuncSetAttribute(smpl_lbs_bodylane_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BlSmem)));   // per device: every call
            const int ranges = m->Vpad / BL_R;
            int bgroups = std::max(1, (148 * 4) / ranges);                      // ~4 CTAs per SM over the launch, 1 resident
            int bper = (B + bgroups - 1) / bgroups;
            bper = (bper + BL_NB - 1) / BL_NB * BL_NB;
            bgroups = (B + bper - 1) / bper;
            smpl_lbs_bodylane_kernel<<<dim3(ranges, bgroups), BL_THREADS, sizeof(BlSmem), st>>>(m->ts.rec, m->ts.joff4, m->V, B, bper, (const float*)w,
                                                                                                 (const float*)G, (float*)cloud);
        } else if (want != "generic" && sizeof(T) == 4 && m->nw == 4 && !use_pd && m->ts.vert && m->V % 2 == 0 && g_al16) {
            static_assert(LS_TILE == 256, "model.cu sorts the tables in 256-vertex tiles");
            smpl_lbs_sorted_kernel<<<grid, LS_TILE, 0, st>>>(m->ts.vert, m->ts.wval, m->ts.wjoff, m->ts.loc, m->V, m->Vpad, B, per,
                                                             (const float*)w, (const float*)G, (float*)cloud);
        } else if (m->nw == 4) B3D_LBS(4);
        else if (m->nw == 8) B3D_LBS(8);
        else B3D_LBS(NJ);
#undef B3D_LBS
        B3D_LAUNCH_CHECK();
    }
    return B3D_OK;
}

}  // namespace b3d

using namespace b3d;

extern "C" {

int b3d_smpl_forward_dev(const b3d_model* m, int32_t B, int32_t dtype, const void* w, const void* p, const void* r, void* cloud,
                         void* jointPos, void* jointTrans, int32_t flags, void* stream) {
    if (!m || !w || !p || !r) return set_error(B3D_EINVAL, "b3d_smpl_forward: null argument");
    if (B < 0) return set_error(B3D_EINVAL, "b3d_smpl_forward: B < 0");
    if (B == 0) return B3D_OK;
    if (dtype == B3D_F32)
        return forward_impl<float>(m, B, (const float*)w, (const float*)p, (const float*)r, (float*)cloud, (float*)jointPos,
                                   (float*)jointTrans, flags, (cudaStream_t)stream);
    if (dtype == B3D_F64)
        return forward_impl<double>(m, B, (const double*)w, (const double*)p, (const double*)r, (double*)cloud, (double*)jointPos,
                                    (double*)jointTrans, flags, (cudaStream_t)stream);
    return set_error(B3D_EINVAL, "b3d_smpl_forward: bad dtype %d", dtype);
}

int b3d_smpl_forward(const b3d_model* m, int32_t B, int32_t dtype, const void* w, const void* p, const void* r, void* cloud,
                     void* jointPos, void* jointTrans, int32_t flags) {
    if (!m || !w || !p || !r) return set_error(B3D_EINVAL, "b3d_smpl_forward: null argument");
    if (dtype != B3D_F32 && dtype != B3D_F64) return set_error(B3D_EINVAL, "b3d_smpl_forward: bad dtype %d", dtype);
    if (B <= 0) return B < 0 ? set_error(B3D_EINVAL, "b3d_smpl_forward: B < 0") : B3D_OK;
    const size_t es = dtype == B3D_F32 ? 4 : 8, nb = (size_t)B;
    auto al = [](size_t x) { return (x + 255) / 256 * 256; };   // sub-buffers stay 256 B aligned (the float kernels use 16 B loads)
    const size_t sz_w = al(es * nb * NS), sz_p = al(es * nb * 3), sz_r = al(es * nb * NJ * 9);
    const size_t sz_c = es * nb * 3 * m->V, sz_jp = al(es * nb * NJ * 3), sz_jt = al(es * nb * NJ * 12);
    char *dw = nullptr, *dc = nullptr;
    B3D_CUDA(cudaMalloc(&dw, sz_w + sz_p + sz_r + sz_jp + sz_jt));
    char *dp = dw + sz_w, *dr = dp + sz_p, *djp = dr + sz_r, *djt = djp + sz_jp;
    int rc = B3D_OK;
    cudaError_t e = cudaSuccess;
    if (cloud) e = cudaMalloc(&dc, sz_c);
    if (e == cudaSuccess) e = cudaMemcpy(dw, w, es * nb * NS, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(dp, p, es * nb * 3, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(dr, r, es * nb * NJ * 9, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && (flags & B3D_FWD_GIVEN_TRANSFORMS) && jointTrans) e = cudaMemcpy(djt, jointTrans, es * nb * NJ * 12, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        rc = b3d_smpl_forward_dev(m, B, dtype, dw, dp, dr, dc, djp, djt, flags, nullptr);
        if (!rc) {
            e = cudaStreamSynchronize(nullptr);
            if (e == cudaSuccess && cloud) e = cudaMemcpy(cloud, dc, sz_c, cudaMemcpyDeviceToHost);
            if (e == cudaSuccess && jointPos) e = cudaMemcpy(jointPos, djp, es * nb * NJ * 3, cudaMemcpyDeviceToHost);
            if (e == cudaSuccess && jointTrans) e = cudaMemcpy(jointTrans, djt, es * nb * NJ * 12, cudaMemcpyDeviceToHost);
        }
    }
    cudaFree(dw);
    cudaFree(dc);
    if (e != cudaSuccess) return set_error(B3D_ECUDA, "b3d_smpl_forward: %s", cudaGetErrorString(e));
    return rc;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------------
// b3d_smpl_forward_noalloc_dev: the evaluator for callers that may not allocate (CUDA graph capture, serving loops).  The
// rest->posed joint transforms between the joint-chain kernel and the skinning kernel — the one scratch buffer of
// b3d_smpl_forward_dev when the caller passes jointTrans = NULL — live in a file-scope device array sized for
// B3D_NOALLOC_MAX_BODIES bodies of doubles (1024 * 24 * 12 * 8 B = 2.25 MiB, held by every GPU that loads the library).
// The array's name is only a host-side handle: the device address is taken once per device with cudaGetSymbolAddress.
// ---------------------------------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_noalloc_trans[(size_t)B3D_NOALLOC_MAX_BODIES * NJ * 12 * sizeof(double)];

namespace b3d {

static int noalloc_scratch(void** out) {
    constexpr int MAXDEV = 64;
    static std::mutex mu;
    static void* addr[MAXDEV] = {};
    int dev = -1;
    B3D_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAXDEV) return set_error(B3D_EINVAL, "b3d_smpl_forward_noalloc_dev: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {   // once per device, with that device current
        void* q = nullptr;
        B3D_CUDA(cudaGetSymbolAddress(&q, g_noalloc_trans));
        if (!q) return set_error(B3D_ECUDA, "b3d_smpl_forward_noalloc_dev: cudaGetSymbolAddress returned a null address");
        addr[dev] = q;
    }
    *out = addr[dev];
    return B3D_OK;
}

template <typename T>
static int forward_noalloc_impl(const b3d_model* m, int B, const T* w, const T* p, const T* r, T* cloud, T* jointPos, cudaStream_t st) {
    void* q = nullptr;
    if (int rc = noalloc_scratch(&q)) return rc;
    T* G = static_cast<T*>(q);   // every entry of the B x 24 x 12 block is written by the joint kernel before the skinning kernel
                                 // reads it: nothing has to start at zero, and nothing of the previous call is read
    const VertexTables<T>& t = sizeof(T) == 4 ? (const VertexTables<T>&)m->t32 : (const VertexTables<T>&)m->t64;
    smpl_joints_kernel<T><<<(B + JW - 1) / JW, 32 * JW, 0, st>>>(m->d_sk, B, w, p, r, jointPos, G, nullptr, 1);
    B3D_LAUNCH_CHECK();
    if (!cloud) return B3D_OK;
    const int tiles = (m->Vpad + LBS_THREADS - 1) / LBS_THREADS;
    int groups = (148 * 8 + tiles - 1) / tiles;
    if (groups > B) groups = B;
    int per = (B + groups - 1) / groups;
    per = (per + LBS_NB - 1) / LBS_NB * LBS_NB;
    groups = (B + per - 1) / per;
    dim3 grid(tiles, groups);
    if (sizeof(T) == 4 && m->nw == 4 && m->ts.vert && m->V % 2 == 0) {
        smpl_lbs_sorted_kernel<<<grid, LS_TILE, 0, st>>>(m->ts.vert, m->ts.wval, m->ts.wjoff, m->ts.loc, m->V, m->Vpad, B, per,
                                                         (const float*)w, (const float*)G, (float*)cloud);
    } else if (m->nw == 4) {
        smpl_lbs_kernel<T, 4, false><<<grid, LBS_THREADS, 0, st>>>(t.vert, t.wval, t.wjoint, m->V, m->Vpad, B, per, w, G, nullptr, cloud);
    } else if (m->nw == 8) {
        smpl_lbs_kernel<T, 8, false><<<grid, LBS_THREADS, 0, st>>>(t.vert, t.wval, t.wjoint, m->V, m->Vpad, B, per, w, G, nullptr, cloud);
    } else {
        smpl_lbs_kernel<T, NJ, false><<<grid, LBS_THREADS, 0, st>>>(t.vert, t.wval, t.wjoint, m->V, m->Vpad, B, per, w, G, nullptr, cloud);
    }
    B3D_LAUNCH_CHECK();
    return B3D_OK;
}

}  // namespace b3d

extern "C" int b3d_smpl_forward_noalloc_dev(const b3d_model* m, int32_t B, int32_t dtype, const void* w, const void* p, const void* r,
                                            void* cloud, void* jointPos, void* stream) {
    if (!m || !w || !p || !r) return set_error(B3D_EINVAL, "b3d_smpl_forward_noalloc_dev: null argument");
    if (B < 0) return set_error(B3D_EINVAL, "b3d_smpl_forward_noalloc_dev: B < 0");
    if (B > B3D_NOALLOC_MAX_BODIES)
        return set_error(B3D_EINVAL, "b3d_smpl_forward_noalloc_dev: B = %d exceeds B3D_NOALLOC_MAX_BODIES = %d (the static scratch holds no more; "
                         "split the batch or call b3d_smpl_forward_dev)", B, B3D_NOALLOC_MAX_BODIES);
    if (B == 0) return B3D_OK;
    if (dtype == B3D_F32)
        return forward_noalloc_impl<float>(m, B, (const float*)w, (const float*)p, (const float*)r, (float*)cloud, (float*)jointPos, (cudaStream_t)stream);
    if (dtype == B3D_F64)
        return forward_noalloc_impl<double>(m, B, (const double*)w, (const double*)p, (const double*)r, (double*)cloud, (double*)jointPos, (cudaStream_t)stream);
    return set_error(B3D_EINVAL, "b3d_smpl_forward_noalloc_dev: bad dtype %d", dtype);
}
