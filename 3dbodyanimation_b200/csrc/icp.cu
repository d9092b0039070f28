// Entry (3): vertex-level ICP term of ark::AvatarOptimizer — cost and gradient of
//     sum over correspondences (i, e) of 1/2 | cloud_i(w, p, q) - data_e |^2
// in the tangent space of the reference's parameter blocks (root position | 24 quaternions under
// FakeQuaternionParameterization::Plus | 10 shape keys), for a batch of bodies.  Replaces, for this term,
//   AvatarEvaluationCommonData::PrepareForEvaluation   external/avatar/AvatarOptimizer.cpp:295-362
//   AvatarCostFunctorCache::updateData                 external/avatar/AvatarOptimizer.cpp:503-577
//   AvatarICPCostFunctor::Evaluate                     external/avatar/AvatarOptimizer.cpp:621-637
//   and the J^T r assembly Ceres does for its line-search (BFGS) minimiser (:1318-1321).
//
// Not a translation of the reference's tables.  With G_k / g_k the posed rotation / position of joint k and
// P_ik = G_k (x_i - J_k) + g_k the image of vertex i under joint k, the reference's per-ancestor Jacobian block
// R(-1, parent j) * dRot(v) * localJacobian_j equals -2 [d_ij]x G_parent(j), d_ij = sum_{k assigned to i below j} w_ik (P_ik - g_j)
// (the quaternion pseudo-Jacobian only ever meets tangent directions).  The gradient therefore needs no ancestor lists and no
// joint-pair tables: per assigned joint k two 3-vectors are accumulated over the vertices,
//     a_k = sum_i w_ik P_ik x rs_i,   b_k = sum_i w_ik rs_i          (rs_i = sum of vertex i's residuals)
// and a per-body epilogue sums them over subtrees: grad_rot_j = 2 G_parent(j)^T (A_j - g_j x B_j).  The shape gradient is
//     sum_i D_i^T (T_i^T rs_i) - sum_k S_k^T G_k^T b_k + sum_{j>=1} (S_j - S_parent(j))^T G_parent(j)^T B_j,
// T_i = sum_k w_ik G_k being the skinning matrix the forward pass forms anyway.
//
// Kernels:
//   icp_gather_kernel / icp_targets_kernel / icp_scatter_kernel : once per set of correspondences (they stay fixed over a whole
//       ceres::Solve, :1439).  A vertex with n matches d_e contributes n |P - mean|^2 + sum |d_e - mean|^2, so the data cloud
//       is reduced to one (mean, n) record per vertex and a per-body constant; the records are stored body-minor
//       [body group][vertex][32 bodies] so that a warp whose lanes are bodies reads them with one coalesced request.
//   icp_accumulate_kernel<T> : the fused shape blend + skinning + residual + gradient pass.  lane = body, the vertex is warp
//       uniform: model data are broadcast loads, a body's 24 transforms live in shared memory [entry][lane] (conflict free) and
//       are cached in registers while consecutive vertices keep the joint of a weight slot, all sums over vertices are per-lane
//       register / shared accumulators — no shuffles, no atomics inside the loop.  HBM traffic: 16 B (float) per vertex and body.
//   icp_finish_kernel : the per-body epilogue above, float64.
#include <algorithm>
#include <cmath>
#include <vector>

#include "internal.hpp"

struct b3d_icp {
    const b3d_model* m = nullptr;
    int B = 0, dtype = 0, groups = 0;
    void* tgt = nullptr;        // [groups][V][32] x (mean xyz, n) of T
    int32_t *vertex_of_pos = nullptr, *pos_of_vertex = nullptr;   // processing order: vertices sorted by their joint tuple
    void* rec = nullptr;        // [V + CH][IcpRec::WORDS] of T: per-vertex model records (icp_records_kernel)
    double* scatter = nullptr;  // [B] sum |d - mean|^2, then [B] number of residual blocks (matched pairs) per body
    double* acc = nullptr;      // [B][ICP_ACC]
    double* state = nullptr;    // device copies for the host-pointer call: w | p | q | cost | grad
    double *r = nullptr, *jp = nullptr, *jt = nullptr;
    int64_t npairs = 0;
};

namespace b3d {

constexpr int ICP_NC = 3 + 3 * NJ + NS;      // 85 tangent coordinates
constexpr int ICP_ACC = 16 + 6 * NJ;         // cost, grad_p(3), shape part (10), pad(2), bins [24][6]
constexpr int ICP_TPAD = 33;                 // row stride of the staged transforms (conflict-free staging and use)

// ---------------------------------------------------------------------------------------------------------------------
// correspondence reduction
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ int body_of_pair(const int64_t* __restrict__ cptr, int B, int64_t e) {
    int lo = 0, hi = B;   // cptr[lo] <= e < cptr[hi]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (cptr[mid] <= e) lo = mid; else hi = mid;
    }
    return lo;
}

template <typename T>
__global__ void __launch_bounds__(256) icp_gather_kernel(int B, int V, const int64_t* __restrict__ cptr, const int64_t* __restrict__ dptr,
                                                         const int32_t* __restrict__ mpt, const int32_t* __restrict__ dpt,
                                                         const T* __restrict__ data, double* __restrict__ tmp) {
    const int64_t n = cptr[B];
    for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < n; e += (int64_t)gridDim.x * blockDim.x) {
        if (mpt[e] < 0) continue;   // unmatched data point (b3d_icp_nearest)
        const int b = body_of_pair(cptr, B, e);
        const T* d = data + 3 * (dptr[b] + (dpt ? (int64_t)dpt[e] : e - cptr[b]));
        double* t = tmp + 4 * ((size_t)b * V + mpt[e]);
        atomicAdd(t + 0, (double)d[0]);
        atomicAdd(t + 1, (double)d[1]);
        atomicAdd(t + 2, (double)d[2]);
        atomicAdd(t + 3, 1.0);
    }
}

template <typename T>
__global__ void __launch_bounds__(256) icp_targets_kernel(int B, int V, const int32_t* __restrict__ pos_of_vertex, double* __restrict__ tmp,
                                                          T* __restrict__ tgt) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= (size_t)B * V) return;
    const int b = (int)(i / V), v = (int)(i % V);
    double* t = tmp + 4 * i;
    const double n = t[3], inv = n > 0.0 ? 1.0 / n : 0.0;
    t[0] *= inv; t[1] *= inv; t[2] *= inv;
    T* o = tgt + 4 * (((size_t)(b >> 5) * V + pos_of_vertex[v]) * 32 + (b & 31));
    o[0] = (T)t[0]; o[1] = (T)t[1]; o[2] = (T)t[2]; o[3] = (T)n;
}

template <typename T>
__global__ void __launch_bounds__(256) icp_scatter_kernel(int B, int V, const int64_t* __restrict__ cptr, const int64_t* __restrict__ dptr,
                                                          const int32_t* __restrict__ mpt, const int32_t* __restrict__ dpt,
                                                          const T* __restrict__ data, const double* __restrict__ tmp,
                                                          double* __restrict__ scatter) {
    const int64_t n = cptr[B];
    const int64_t span = (int64_t)gridDim.x * blockDim.x;
    for (int64_t e0 = blockIdx.x * (int64_t)blockDim.x; e0 < n; e0 += span) {
        const int64_t e = e0 + threadIdx.x;
        int b = -1;
        double s = 0.0, one = 0.0;
        if (e < n && mpt[e] >= 0) {
            one = 1.0;
            b = body_of_pair(cptr, B, e);
            const T* d = data + 3 * (dptr[b] + (dpt ? (int64_t)dpt[e] : e - cptr[b]));
            const double* mu = tmp + 4 * ((size_t)b * V + mpt[e]);   // float64 mean (icp_targets_kernel)
            const double ax = (double)d[0] - mu[0], ay = (double)d[1] - mu[1], az = (double)d[2] - mu[2];
            s = ax * ax + ay * ay + az * az;
        }
        // one atomic per warp when the warp's pairs belong to one body
        const int b0 = __shfl_sync(0xffffffffu, b, 0);
        if (__all_sync(0xffffffffu, b == b0)) {
#pragma unroll
            for (int o = 16; o; o >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, o); one += __shfl_xor_sync(0xffffffffu, one, o); }
            if ((threadIdx.x & 31) == 0 && b0 >= 0) { atomicAdd(scatter + b0, s); atomicAdd(scatter + B + b0, one); }
        } else if (b >= 0) {
            atomicAdd(scatter + b, s);
            atomicAdd(scatter + B + b, 1.0);
        }
    }
}

// quaternions (x, y, z, w) -> column-major rotation matrices (Eigen toRotationMatrix, as the optimiser applies them :326)
__global__ void __launch_bounds__(256) icp_quat_kernel(int n, const double* __restrict__ q, double* __restrict__ r) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double x = q[4 * i], y = q[4 * i + 1], z = q[4 * i + 2], w = q[4 * i + 3];
    const double tx = 2 * x, ty = 2 * y, tz = 2 * z;
    const double twx = tx * w, twy = ty * w, twz = tz * w, txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y, tyz = tz * y, tzz = tz * z;
    double* o = r + 9 * (size_t)i;
    o[0] = 1 - (tyy + tzz); o[3] = txy - twz;       o[6] = txz + twy;
    o[1] = txy + twz;       o[4] = 1 - (txx + tzz); o[7] = tyz - twx;
    o[2] = txz - twy;       o[5] = tyz + twx;       o[8] = 1 - (txx + tyy);
}

// ---------------------------------------------------------------------------------------------------------------------
// fused shape blend + skinning + residual + gradient accumulation.  grid = (body groups of 32, vertex splits)
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float fma_(float a, float b, float c) { return fmaf(a, b, c); }
__device__ __forceinline__ double fma_(double a, double b, double c) { return fma(a, b, c); }
// packed FP32 (FFMA2): two lanes per issue slot; a (v, v) pair becomes the instruction's broadcast operand (no copy)
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float lo, float hi) {
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpk2(f32x2 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ float hsum2(f32x2 v) { float lo, hi; unpk2(v, lo, hi); return lo + hi; }

template <typename T> struct Vec4T;
template <> struct Vec4T<float> { using type = float4; };
template <> struct Vec4T<double> { using type = double4; };

// One record per vertex, in the joint-tuple order of build_order() (built once per b3d_icp from the model's SoA tables):
//   [0..2] template xyz, [3] 0, [4 + 10 c + s] shapedirs, [34 + k] weight of slot k, [34 + NW + k] its joint (as a T value)
template <int NW> struct IcpRec { static constexpr int WORDS = (34 + 2 * NW + 3) / 4 * 4; };   // 44 (NW = 4), 52 (NW = 8)
template <typename T> struct IcpChunk { static constexpr int CH = sizeof(T) == 4 ? 8 : 4; };   // vertices per prefetched tile

template <typename T, int NW>
__global__ void __launch_bounds__(256) icp_records_kernel(const T* __restrict__ vert, const T* __restrict__ wval, const int32_t* __restrict__ wjoint,
                                                          int V, int Vpad, const int32_t* __restrict__ vertex_of_pos, T* __restrict__ rec) {
    const int pos = blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= V) return;
    const int v = vertex_of_pos[pos];
    T* o = rec + (size_t)pos * IcpRec<NW>::WORDS;
    for (int c = 0; c < 3; ++c) o[c] = vert[(size_t)c * Vpad + v];
    o[3] = 0;
    for (int i = 0; i < 3 * NS; ++i) o[4 + i] = vert[(size_t)(3 + i) * Vpad + v];
    for (int k = 0; k < NW; ++k) {
        const T wt = wval[(size_t)k * Vpad + v];
        o[34 + k] = (double)wt > 1e-12 ? wt : (T)0;   // assignedJoints keeps weights > 1e-12 only (AvatarModel.cpp:81)
        o[34 + NW + k] = (T)wjoint[(size_t)k * Vpad + v];
    }
    for (int i = 34 + 2 * NW; i < IcpRec<NW>::WORDS; ++i) o[i] = 0;
}

__device__ __forceinline__ uint32_t icp_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void icp_mbar_init(unsigned long long* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(icp_smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void icp_mbar_expect_tx(unsigned long long* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(icp_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void icp_mbar_wait(unsigned long long* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(icp_smem_u32(bar)), "r"(parity) : "memory");
}
// TMA bulk copy global -> shared, completion counted on an mbarrier (bytes % 16 == 0, both addresses 16 B aligned)
__device__ __forceinline__ void icp_bulk_g2s(void* dst, const void* src, uint32_t bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(icp_smem_u32(dst)), "l"(src),
                 "r"(bytes), "r"(icp_smem_u32(bar)) : "memory");
}

template <typename T, int NW, int NWARPS>
constexpr size_t icp_smem_bytes() {
    return sizeof(T) * ((size_t)NJ * 12 * ICP_TPAD + (size_t)NJ * 6 * 32 +
                        (size_t)NWARPS * 2 * IcpChunk<T>::CH * (IcpRec<NW>::WORDS + 128)) + sizeof(unsigned long long) * 2 * NWARPS;
}

template <typename T, int NW, int NWARPS>
__global__ void __launch_bounds__(32 * NWARPS, 1) icp_accumulate_kernel(const T* __restrict__ rec, int V, int B,
                                                                        const double* __restrict__ wshape, const double* __restrict__ jointTrans,
                                                                        const T* __restrict__ tgt, double* __restrict__ acc) {
    constexpr int CH = IcpChunk<T>::CH, RW = IcpRec<NW>::WORDS;
    constexpr uint32_t REC_BYTES = CH * RW * sizeof(T), TGT_BYTES = CH * 128 * sizeof(T);
    extern __shared__ __align__(128) unsigned char smem_raw[];
    T* sT = reinterpret_cast<T*>(smem_raw);                 // [NJ * 12][ICP_TPAD] transforms of the 32 bodies
    T* sBin = sT + NJ * 12 * ICP_TPAD;                      // [NJ * 6][32] per-joint sums of the CTA
    T* sRec = sBin + NJ * 6 * 32;                           // [NWARPS][2][CH][RW]
    T* sTgt = sRec + NWARPS * 2 * CH * RW;                  // [NWARPS][2][CH][32][4]
    unsigned long long* mbar = reinterpret_cast<unsigned long long*>(sTgt + NWARPS * 2 * CH * 128);   // [NWARPS][2]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int b0 = blockIdx.x * 32;
    const int b = min(b0 + lane, B - 1);

    // vertex range of this CTA, then of this warp
    const int per_cta = (V + gridDim.y - 1) / gridDim.y;
    const int c0 = blockIdx.y * per_cta, c1 = min(V, c0 + per_cta);
    const int per_warp = (max(c1 - c0, 0) + NWARPS - 1) / NWARPS;
    const int v0 = min(c1, c0 + warp * per_warp), v1 = min(c1, v0 + per_warp);
    const int nch = (v1 - v0 + CH - 1) / CH;
    T* myRec = sRec + (size_t)warp * 2 * CH * RW;
    T* myTgt = sTgt + (size_t)warp * 2 * CH * 128;
    unsigned long long* myBar = mbar + 2 * warp;
    const T* tgt_group = tgt + (size_t)blockIdx.x * V * 128;
    auto issue = [&](int c) {   // lane 0: model records and target records of chunk c into buffer c & 1
        const int vb = v0 + c * CH, s = c & 1;
        icp_mbar_expect_tx(myBar + s, REC_BYTES + TGT_BYTES);
        icp_bulk_g2s(myRec + s * CH * RW, rec + (size_t)vb * RW, REC_BYTES, myBar + s);
        icp_bulk_g2s(myTgt + s * CH * 128, tgt_group + (size_t)vb * 128, TGT_BYTES, myBar + s);
    };
    if (lane == 0) {
        icp_mbar_init(myBar, 1);
        icp_mbar_init(myBar + 1, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        if (nch > 0) issue(0);
    }
    // stage the group's transforms: sT[e][body]
    for (int i = threadIdx.x; i < 32 * NJ * 12; i += 32 * NWARPS) {
        const int bl = i / (NJ * 12), e = i - bl * (NJ * 12);
        sT[e * ICP_TPAD + bl] = (T)jointTrans[(size_t)min(b0 + bl, B - 1) * NJ * 12 + e];
    }
    for (int i = threadIdx.x; i < NJ * 6 * 32; i += 32 * NWARPS) sBin[i] = (T)0;
    constexpr bool PACKED = sizeof(T) == 4;   // float32: packed FFMA2 arithmetic (the float64 instance keeps the scalar chains)
    static_assert(NS % 2 == 0, "shape coefficients are processed in pairs");
    T ws[NS];
#pragma unroll
    for (int s = 0; s < NS; ++s) ws[s] = (T)wshape[(size_t)b * NS + s];
    f32x2 wsp[NS / 2], shp[NS / 2];
#pragma unroll
    for (int s = 0; s < NS / 2; ++s) { wsp[s] = pk2((float)ws[2 * s], (float)ws[2 * s + 1]); shp[s] = 0ull; }
    __syncthreads();

    T cost = 0, gp[3] = {0, 0, 0}, sh[NS];
#pragma unroll
    for (int s = 0; s < NS; ++s) sh[s] = 0;
    int curj[NW];
    T curjT[NW], G[NW][12], bin[NW][6];
#pragma unroll
    for (int k = 0; k < NW; ++k) {
        curj[k] = -1;
        curjT[k] = (T)-1;
#pragma unroll
        for (int c = 0; c < 6; ++c) bin[k][c] = 0;
#pragma unroll
        for (int c = 0; c < 12; ++c) G[k][c] = 0;
    }
    using V4 = typename Vec4T<T>::type;

    for (int c = 0; c < nch; ++c) {
        const int s = c & 1;
        if (lane == 0 && c + 1 < nch) issue(c + 1);   // its buffer was released by the __syncwarp that closed chunk c - 1
        icp_mbar_wait(myBar + s, (uint32_t)((c >> 1) & 1));
        const int nv = min(CH, v1 - (v0 + c * CH));
        for (int i = 0; i < nv; ++i) {
            // the vertex's record: warp-uniform 16-byte shared loads (the shapedirs are read again for the shape gradient
            // rather than kept in 30 registers)
            const V4* r4 = reinterpret_cast<const V4*>(myRec + (s * CH + i) * RW);
            const T* r1 = myRec + (s * CH + i) * RW;
            T wj[RW - 32];   // [2..] weights, then joints
#pragma unroll
            for (int u = 8; u < RW / 4; ++u) {
                const V4 t4 = r4[u];
                wj[4 * (u - 8)] = t4.x; wj[4 * (u - 8) + 1] = t4.y; wj[4 * (u - 8) + 2] = t4.z; wj[4 * (u - 8) + 3] = t4.w;
            }
            const V4 mu = reinterpret_cast<const V4*>(myTgt + (s * CH + i) * 128)[lane];
            // shape blend (AvatarOptimizer.cpp:263-264)
            T x[3];
            if constexpr (PACKED) {
                // words 4 .. 33 = shapedirs [coordinate][s]: ten consecutive words per coordinate, i.e. five aligned pairs;
                // products pair by pair against (w_s, w_s+1), the two halves added at the end
                const V4 t0 = r4[0];
                const float tpl[3] = {t0.x, t0.y, t0.z};
                f32x2 dp[15];
#pragma unroll
                for (int u = 1; u < 8; ++u) {
                    const V4 t4 = r4[u];
                    dp[2 * (u - 1)] = pk2(t4.x, t4.y);
                    dp[2 * (u - 1) + 1] = pk2(t4.z, t4.w);
                }
                dp[14] = pk2(wj[0], wj[1]);
#pragma unroll
                for (int cc = 0; cc < 3; ++cc) {
                    f32x2 a2 = mul2(dp[5 * cc], wsp[0]);
#pragma unroll
                    for (int q = 1; q < NS / 2; ++q) a2 = fma2(dp[5 * cc + q], wsp[q], a2);
                    x[cc] = tpl[cc] + hsum2(a2);
                }
            } else {
                T rv[32];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const V4 t4 = r4[u];
                    rv[4 * u] = t4.x; rv[4 * u + 1] = t4.y; rv[4 * u + 2] = t4.z; rv[4 * u + 3] = t4.w;
                }
                T hi0 = wj[0], hi1 = wj[1];   // words 32, 33 = shapedirs z, s = 8, 9
#pragma unroll
                for (int cc = 0; cc < 3; ++cc) {
                    T a = rv[cc];
#pragma unroll
                    for (int q = 0; q < NS; ++q) {
                        const int idx = 4 + NS * cc + q;
                        a += (idx < 32 ? rv[idx < 32 ? idx : 0] : (idx == 32 ? hi0 : hi1)) * ws[q];
                    }
                    x[cc] = a;
                }
            }
            // A weight slot keeps its joint over long runs of the sorted vertices: one (rarely taken) branch per vertex looks
            // after the changes; slots the reference does not assign carry weight 0 (icp_records_kernel), keep whatever
            // transform they hold and contribute exact zeros, so the arithmetic below is branch free.
            bool changed = false;
#pragma unroll
            for (int k = 0; k < NW; ++k) changed |= (wj[2 + k] > (T)0) & (wj[2 + NW + k] != curjT[k]);
            if (changed) {   // warp-uniform: the vertex is
#pragma unroll
                for (int k = 0; k < NW; ++k) {
                    if ((wj[2 + k] > (T)0) & (wj[2 + NW + k] != curjT[k])) {
                        if (curj[k] >= 0) {
#pragma unroll
                            for (int cc = 0; cc < 6; ++cc) { atomicAdd(&sBin[(curj[k] * 6 + cc) * 32 + lane], bin[k][cc]); bin[k][cc] = 0; }
                        }
                        curjT[k] = wj[2 + NW + k];
                        const int jk = (int)curjT[k];
                        curj[k] = jk;
#pragma unroll
                        for (int cc = 0; cc < 12; ++cc) G[k][cc] = sT[(jk * 12 + cc) * ICP_TPAD + lane];
                    }
                }
            }
            if constexpr (PACKED) {
                // rows x / y of every 3-vector as one packed value, row z scalar; dot products against (wr_x, wr_y) pair-wise
                // with the halves added once per vertex
                const f32x2 X0 = pk2(x[0], x[0]), X1 = pk2(x[1], x[1]), X2 = pk2(x[2], x[2]);
                f32x2 Pkxy[NW], Pxy = 0ull;
                float Pkz[NW], Pz = 0.f;
#pragma unroll
                for (int k = 0; k < NW; ++k) {
                    const float wt = wj[2 + k];
                    Pkxy[k] = fma2(pk2(G[k][0], G[k][1]), X0, fma2(pk2(G[k][3], G[k][4]), X1, fma2(pk2(G[k][6], G[k][7]), X2, pk2(G[k][9], G[k][10]))));
                    Pkz[k] = fmaf(G[k][2], x[0], fmaf(G[k][5], x[1], fmaf(G[k][8], x[2], G[k][11])));
                    Pxy = fma2(pk2(wt, wt), Pkxy[k], Pxy);
                    Pz = fmaf(wt, Pkz[k], Pz);
                }
                float Px, Py;
                unpk2(Pxy, Px, Py);
                const float n = mu.w;
                const float e0 = Px - mu.x, e1 = Py - mu.y, e2 = Pz - mu.z;
                const float rs0 = n * e0, rs1 = n * e1, rs2 = n * e2;
                cost = fmaf(n, fmaf(e0, e0, fmaf(e1, e1, e2 * e2)), cost);
                gp[0] += rs0; gp[1] += rs1; gp[2] += rs2;
                const f32x2 RS = pk2(rs0, rs1);
                f32x2 Y0 = 0ull, Y1 = 0ull, Y2 = 0ull;
                float y0z = 0.f, y1z = 0.f, y2z = 0.f;
#pragma unroll
                for (int k = 0; k < NW; ++k) {
                    const float wt = wj[2 + k];
                    const f32x2 WR = mul2(pk2(wt, wt), RS);
                    const float wr2 = wt * rs2;
                    float wr0, wr1, pk0, pk1;
                    unpk2(WR, wr0, wr1);
                    unpk2(Pkxy[k], pk0, pk1);
                    bin[k][0] = fmaf(pk1, wr2, fmaf(-Pkz[k], wr1, bin[k][0]));
                    bin[k][1] = fmaf(Pkz[k], wr0, fmaf(-pk0, wr2, bin[k][1]));
                    bin[k][2] = fmaf(pk0, wr1, fmaf(-pk1, wr0, bin[k][2]));
                    bin[k][3] += wr0; bin[k][4] += wr1; bin[k][5] += wr2;
                    Y0 = fma2(pk2(G[k][0], G[k][1]), WR, Y0); y0z = fmaf(G[k][2], wr2, y0z);
                    Y1 = fma2(pk2(G[k][3], G[k][4]), WR, Y1); y1z = fmaf(G[k][5], wr2, y1z);
                    Y2 = fma2(pk2(G[k][6], G[k][7]), WR, Y2); y2z = fmaf(G[k][8], wr2, y2z);
                }
                const float y0 = hsum2(Y0) + y0z, y1 = hsum2(Y1) + y1z, y2 = hsum2(Y2) + y2z;
                // shape: D^T y, shapedirs read again from the tile as pairs (s, s + 1)
                const f32x2 YB0 = pk2(y0, y0), YB1 = pk2(y1, y1), YB2 = pk2(y2, y2);
                f32x2 dq[15];
#pragma unroll
                for (int u = 1; u < 8; ++u) {
                    const V4 t4 = r4[u];
                    dq[2 * (u - 1)] = pk2(t4.x, t4.y);
                    dq[2 * (u - 1) + 1] = pk2(t4.z, t4.w);
                }
                dq[14] = pk2(wj[0], wj[1]);
#pragma unroll
                for (int q = 0; q < NS / 2; ++q) shp[q] = fma2(dq[q], YB0, fma2(dq[5 + q], YB1, fma2(dq[10 + q], YB2, shp[q])));
            } else {
                T Pk[NW][3], P[3] = {0, 0, 0};
    #pragma unroll
                for (int k = 0; k < NW; ++k) {
                    const T wt = wj[2 + k];
    #pragma unroll
                    for (int r = 0; r < 3; ++r) {
                        Pk[k][r] = fma_(G[k][r], x[0], fma_(G[k][3 + r], x[1], fma_(G[k][6 + r], x[2], G[k][9 + r])));
                        P[r] = fma_(wt, Pk[k][r], P[r]);
                    }
                }
                // sum of the vertex's residuals and its cost share
                const T n = mu.w;
                const T e0 = P[0] - mu.x, e1 = P[1] - mu.y, e2 = P[2] - mu.z;
                const T rs[3] = {n * e0, n * e1, n * e2};
                cost = fma_(n, fma_(e0, e0, fma_(e1, e1, e2 * e2)), cost);
    #pragma unroll
                for (int cc = 0; cc < 3; ++cc) gp[cc] += rs[cc];
                // per-joint sums, and y = T^T rs = sum_k G_k^T (w_k rs) for the shape gradient
                T y0 = 0, y1 = 0, y2 = 0;
    #pragma unroll
                for (int k = 0; k < NW; ++k) {
                    const T wt = wj[2 + k];
                    {
                        const T wr[3] = {wt * rs[0], wt * rs[1], wt * rs[2]};
                        // straight FMA chains into the accumulators (one instruction per product)
                        bin[k][0] = fma_(Pk[k][1], wr[2], fma_(-Pk[k][2], wr[1], bin[k][0]));
                        bin[k][1] = fma_(Pk[k][2], wr[0], fma_(-Pk[k][0], wr[2], bin[k][1]));
                        bin[k][2] = fma_(Pk[k][0], wr[1], fma_(-Pk[k][1], wr[0], bin[k][2]));
                        bin[k][3] += wr[0]; bin[k][4] += wr[1]; bin[k][5] += wr[2];
                        y0 = fma_(G[k][0], wr[0], fma_(G[k][1], wr[1], fma_(G[k][2], wr[2], y0)));
                        y1 = fma_(G[k][3], wr[0], fma_(G[k][4], wr[1], fma_(G[k][5], wr[2], y1)));
                        y2 = fma_(G[k][6], wr[0], fma_(G[k][7], wr[1], fma_(G[k][8], wr[2], y2)));
                    }
                }
                // shape: D^T y, shapedirs read again from the tile
    #pragma unroll
                for (int q = 0; q < NS; ++q) sh[q] = fma_(r1[4 + q], y0, fma_(r1[4 + NS + q], y1, fma_(r1[4 + 2 * NS + q], y2, sh[q])));
            }
            }
        __syncwarp();
    }
    if constexpr (PACKED) {
#pragma unroll
        for (int q = 0; q < NS / 2; ++q) { float lo, hi; unpk2(shp[q], lo, hi); sh[2 * q] = (T)lo; sh[2 * q + 1] = (T)hi; }
    }
#pragma unroll
    for (int k = 0; k < NW; ++k)
        if (curj[k] >= 0) {
#pragma unroll
            for (int cc = 0; cc < 6; ++cc) atomicAdd(&sBin[(curj[k] * 6 + cc) * 32 + lane], bin[k][cc]);
        }
    __syncthreads();
    // dense accumulators through the (now free) transform area: sD[warp][14][32]
    T* sD = sT;
    sD[(warp * 14 + 0) * 32 + lane] = cost;
#pragma unroll
    for (int cc = 0; cc < 3; ++cc) sD[(warp * 14 + 1 + cc) * 32 + lane] = gp[cc];
#pragma unroll
    for (int q = 0; q < NS; ++q) sD[(warp * 14 + 4 + q) * 32 + lane] = sh[q];
    __syncthreads();
    for (int i = threadIdx.x; i < (14 + NJ * 6) * 32; i += 32 * NWARPS) {
        const int e = i >> 5, l = i & 31;
        if (b0 + l >= B) continue;
        double sum = 0.0;
        if (e < 14) {
#pragma unroll
            for (int wq = 0; wq < NWARPS; ++wq) sum += (double)sD[(wq * 14 + e) * 32 + l];
        } else {
            sum = (double)sBin[(e - 14) * 32 + l];
        }
        if (sum != 0.0) atomicAdd(acc + (size_t)(b0 + l) * ICP_ACC + (e < 14 ? e : e + 2), sum);
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// per-body epilogue (float64): subtree sums, tangent-space gradient, cost
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) icp_finish_kernel(const SkeletonConst* __restrict__ sk, int B, int shape_enabled,
                                                         const double* __restrict__ acc, const double* __restrict__ scatter,
                                                         const double* __restrict__ jointPos, const double* __restrict__ jointTrans,
                                                         double* __restrict__ cost, double* __restrict__ grad) {
    // warp = body, lane = joint
    __shared__ double s_ab[4][NJ][6];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b = blockIdx.x * 4 + warp;
    if (b >= B) return;
    const double* a = acc + (size_t)b * ICP_ACC;
    const double* G = jointTrans + (size_t)b * NJ * 12;   // per joint col-major [G | t]
    const double* g = jointPos + (size_t)b * NJ * 3;
    double* out = grad + (size_t)b * ICP_NC;
    const int j = lane < NJ ? lane : 0;
    const bool act = lane < NJ;
    const int pa = sk->parent[j];
    // ancestors-or-self of the lane's joint as a bit mask (parents precede children: model.cu checks parent[j] < j)
    uint32_t anc = 0;
    for (int k = j; k >= 0; k = sk->parent[k]) anc |= 1u << k;
    double ab[6];
#pragma unroll
    for (int c = 0; c < 6; ++c) { ab[c] = act ? a[16 + 6 * j + c] : 0.0; if (act) s_ab[warp][j][c] = ab[c]; }
    __syncwarp();
    double sg[NS];
#pragma unroll
    for (int s = 0; s < NS; ++s) sg[s] = 0.0;
    double Gj[9];
#pragma unroll
    for (int c = 0; c < 9; ++c) Gj[c] = G[12 * j + c];
    if (shape_enabled && act) {   // - S_j^T (G_j^T b_j), b_j before the subtree sums
        double y[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) y[c] = Gj[3 * c] * ab[3] + Gj[3 * c + 1] * ab[4] + Gj[3 * c + 2] * ab[5];
#pragma unroll
        for (int s = 0; s < NS; ++s)
            sg[s] -= sk->JSR[(3 * j) * NS + s] * y[0] + sk->JSR[(3 * j + 1) * NS + s] * y[1] + sk->JSR[(3 * j + 2) * NS + s] * y[2];
    }
    // subtree sums A_j, B_j: every joint k that has j among its ancestors-or-self
    double AB[6] = {0, 0, 0, 0, 0, 0};
    for (int k = 0; k < NJ; ++k) {   // same trip count on every lane: the shuffle needs the whole warp
        const uint32_t ak = __shfl_sync(0xffffffffu, anc, k);
        if ((ak >> j) & 1u) {
#pragma unroll
            for (int c = 0; c < 6; ++c) AB[c] += s_ab[warp][k][c];
        }
    }
    const double* gj = g + 3 * j;
    const double tq[3] = {AB[0] - (gj[1] * AB[5] - gj[2] * AB[4]), AB[1] - (gj[2] * AB[3] - gj[0] * AB[5]),
                          AB[2] - (gj[0] * AB[4] - gj[1] * AB[3])};
    if (act) {
        if (j == 0) {
#pragma unroll
            for (int c = 0; c < 3; ++c) out[3 + c] = 2.0 * tq[c];
        } else {
            double Gp[9];
#pragma unroll
            for (int c = 0; c < 9; ++c) Gp[c] = G[12 * pa + c];
#pragma unroll
            for (int c = 0; c < 3; ++c) out[3 + 3 * j + c] = 2.0 * (Gp[3 * c] * tq[0] + Gp[3 * c + 1] * tq[1] + Gp[3 * c + 2] * tq[2]);
            if (shape_enabled) {   // + (S_j - S_parent)^T G_parent^T B_j
                double y[3];
#pragma unroll
                for (int c = 0; c < 3; ++c) y[c] = Gp[3 * c] * AB[3] + Gp[3 * c + 1] * AB[4] + Gp[3 * c + 2] * AB[5];
#pragma unroll
                for (int s = 0; s < NS; ++s)
#pragma unroll
                    for (int c = 0; c < 3; ++c) sg[s] += (sk->JSR[(3 * j + c) * NS + s] - sk->JSR[(3 * pa + c) * NS + s]) * y[c];
            }
        }
    }
#pragma unroll
    for (int s = 0; s < NS; ++s) {
#pragma unroll
        for (int o = 16; o; o >>= 1) sg[s] += __shfl_xor_sync(0xffffffffu, sg[s], o);
    }
    if (lane == 0) cost[b] = 0.5 * (a[0] + scatter[b]);
    if (lane < 3) out[lane] = a[1 + lane];
    if (lane < NS) {
        double v = 0.0;
#pragma unroll
        for (int s = 0; s < NS; ++s) if (s == lane) v = sg[s];
        out[3 + 3 * NJ + lane] = shape_enabled ? a[4 + lane] + v : 0.0;
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Prior blocks of AvatarOptimizer::optimize, added to cost / grad of the data term: AvatarPosePriorCostFunctor (:644-700; GMM
// residual of the 23 joint angle-axes times scaledBetaPose, Jacobian L.middleRows<3>(3 i)^T * 0.707106781186548 * scaledBetaPose)
// and AvatarShapePriorCostFunctor (:703-726; w * scaledBetaShape), with scaledBeta = beta * sqrt(#residual blocks) / 15
// (:1412-1414).  CTA = body.
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) icp_prior_kernel(int B, int nc, const double* __restrict__ mean, const double* __restrict__ L,
                                                        const double* __restrict__ consts, double beta_pose, double beta_shape,
                                                        const double* __restrict__ totals, const double* __restrict__ w,
                                                        const double* __restrict__ q, double* __restrict__ cost, double* __restrict__ grad) {
    __shared__ double s_x[NPRI], s_h[16][NPRI], s_p[16], s_cost[4];
    const int b = blockIdx.x, tid = threadIdx.x;
    const double root = sqrt(totals[b]) / 15.0, sbp = beta_pose * root, sbs = beta_shape * root;
    double add = 0.0;
    if (nc > 0 && beta_pose > 0.0) {
        if (tid < NJ - 1) {   // Eigen::AngleAxisd(q_j): axis * angle
            const double* qq = q + ((size_t)b * NJ + tid + 1) * 4;
            double n = sqrt(qq[0] * qq[0] + qq[1] * qq[1] + qq[2] * qq[2]);
            double ax = 0.0, ay = 0.0, az = 0.0;
            if (n != 0.0) {
                const double angle = 2.0 * atan2(n, fabs(qq[3]));
                if (qq[3] < 0.0) n = -n;
                ax = qq[0] / n * angle; ay = qq[1] / n * angle; az = qq[2] / n * angle;
            }
            s_x[3 * tid] = ax; s_x[3 * tid + 1] = ay; s_x[3 * tid + 2] = az;
        }
        __syncthreads();
        const double rh = sqrt(0.5);
        for (int t = tid; t < nc * NPRI; t += 128) {   // GaussianMixture::residual (GaussianMixture.cpp:94-113)
            const int i = t / NPRI, k = t % NPRI;
            const double* Li = L + (size_t)i * NPRI * NPRI;
            double acc = 0.0;
            for (int m = k; m < NPRI; ++m) acc += Li[m * NPRI + k] * (s_x[m] - mean[i * NPRI + m]);
            s_h[i][k] = acc * rh;
        }
        __syncthreads();
        if (tid < nc) {
            double sq = 0.0;
            for (int k = 0; k < NPRI; ++k) sq += s_h[tid][k] * s_h[tid][k];
            s_p[tid] = sq - consts[tid];
        }
        __syncthreads();
        int c = 0;
        double best = 1.7976931348623157e308;
        for (int i = 0; i < nc; ++i) if (s_p[i] < best) { best = s_p[i]; c = i; }
        // cost: 1/2 |sbp * [h; sqrt(-const)]|^2 ; gradient entry (joint i + 1, a) = 0.7071... sbp^2 * L[3 i + a][.] . h
        if (tid < NPRI) {
            const double r = s_h[c][tid] * sbp;
            add = 0.5 * r * r;
            const double* Lr = L + (size_t)c * NPRI * NPRI + (size_t)tid * NPRI;
            double acc = 0.0;
            for (int k = 0; k <= tid; ++k) acc += Lr[k] * 0.707106781186548 * sbp * (s_h[c][k] * sbp);
            grad[(size_t)b * ICP_NC + 6 + tid] += acc;   // 3 + 3 (i + 1) + a with tid = 3 i + a
        }
        if (tid == NPRI) { const double r = sqrt(-consts[c]) * sbp; add = 0.5 * r * r; }
    }
    if (beta_shape > 0.0 && tid >= 96 && tid < 96 + NS) {
        const double rs = w[(size_t)b * NS + tid - 96] * sbs;
        add += 0.5 * rs * rs;
        grad[(size_t)b * ICP_NC + 3 + 3 * NJ + tid - 96] += sbs * rs;
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) add += __shfl_xor_sync(0xffffffffu, add, o);
    if ((tid & 31) == 0) s_cost[tid >> 5] = add;
    __syncthreads();
    if (tid == 0) cost[b] += (s_cost[0] + s_cost[1]) + (s_cost[2] + s_cost[3]);
}

// The sums over vertices do not depend on the order, so the vertices are processed sorted by their tuple of assigned joints:
// a weight slot then keeps its joint over long runs and the accumulate kernel keeps that joint's transform and partial sums in
// registers (SMPL: a few hundred distinct tuples for 6890 vertices).
static int build_order(b3d_icp* h) {
    const b3d_model* m = h->m;
    const int V = m->V, nw = m->nw;
    std::vector<int32_t> wj((size_t)nw * m->Vpad);
    std::vector<double> wv((size_t)nw * m->Vpad);
    B3D_CUDA(cudaMemcpy(wj.data(), m->t64.wjoint, sizeof(int32_t) * wj.size(), cudaMemcpyDeviceToHost));
    B3D_CUDA(cudaMemcpy(wv.data(), m->t64.wval, sizeof(double) * wv.size(), cudaMemcpyDeviceToHost));
    auto key = [&](int v, int k) { return wv[(size_t)k * m->Vpad + v] > 1e-12 ? wj[(size_t)k * m->Vpad + v] : NJ; };
    std::vector<int32_t> order(V), inv(V);
    for (int v = 0; v < V; ++v) order[v] = v;
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) {
        for (int k = 0; k < nw; ++k) {
            const int ka = key(a, k), kb = key(b, k);
            if (ka != kb) return ka < kb;
        }
        return false;
    });
    for (int i = 0; i < V; ++i) inv[order[i]] = i;
    B3D_CUDA(cudaMalloc(&h->vertex_of_pos, sizeof(int32_t) * 2 * V));
    h->pos_of_vertex = h->vertex_of_pos + V;
    B3D_CUDA(cudaMemcpy(h->vertex_of_pos, order.data(), sizeof(int32_t) * V, cudaMemcpyHostToDevice));
    B3D_CUDA(cudaMemcpy(h->pos_of_vertex, inv.data(), sizeof(int32_t) * V, cudaMemcpyHostToDevice));
    return B3D_OK;
}

template <typename T>
static int icp_reduce(b3d_icp* h, const int64_t* d_cptr, const int64_t* d_dptr, const int32_t* d_mpt, const int32_t* d_dpt, const T* d_data) {
    const b3d_model* m = h->m;
    const int B = h->B, V = m->V;
    double* tmp = nullptr;
    const size_t tmp_bytes = sizeof(double) * 4 * (size_t)B * V, tgt_bytes = sizeof(T) * 128 * ((size_t)h->groups * V + 8);
    B3D_CUDA(cudaMalloc(&tmp, tmp_bytes));
    cudaError_t e = cudaMemsetAsync(tmp, 0, tmp_bytes, nullptr);
    if (e == cudaSuccess) e = cudaMemsetAsync(h->tgt, 0, tgt_bytes, nullptr);
    if (e == cudaSuccess) e = cudaMemsetAsync(h->scatter, 0, sizeof(double) * 2 * B, nullptr);
    if (e != cudaSuccess) { cudaFree(tmp); return set_error(B3D_ECUDA, "b3d_icp_create: %s", cudaGetErrorString(e)); }
    const int pair_blocks = (int)std::min<int64_t>(std::max<int64_t>((h->npairs + 255) / 256, 1), 148 * 16);
    icp_gather_kernel<T><<<pair_blocks, 256>>>(B, V, d_cptr, d_dptr, d_mpt, d_dpt, d_data, tmp);
    count_launch();
    icp_targets_kernel<T><<<(unsigned)(((size_t)B * V + 255) / 256), 256>>>(B, V, h->pos_of_vertex, tmp, (T*)h->tgt);
    count_launch();
    icp_scatter_kernel<T><<<pair_blocks, 256>>>(B, V, d_cptr, d_dptr, d_mpt, d_dpt, d_data, tmp, h->scatter);
    count_launch();
    e = cudaStreamSynchronize(nullptr);
    if (e == cudaSuccess) e = cudaGetLastError();
    cudaFree(tmp);
    if (e != cudaSuccess) return set_error(B3D_ECUDA, "b3d_icp_create: %s", cudaGetErrorString(e));
    return B3D_OK;
}

template <typename T, int NW>
static int build_records(b3d_icp* h, const VertexTables<T>& t) {
    const b3d_model* m = h->m;
    const size_t bytes = sizeof(T) * IcpRec<NW>::WORDS * ((size_t)m->V + IcpChunk<T>::CH);   // a partial last tile reads past V
    B3D_CUDA(cudaMalloc(&h->rec, bytes));
    B3D_CUDA(cudaMemsetAsync(h->rec, 0, bytes, nullptr));
    icp_records_kernel<T, NW><<<(m->V + 255) / 256, 256>>>(t.vert, t.wval, t.wjoint, m->V, m->Vpad, h->vertex_of_pos, (T*)h->rec);
    B3D_LAUNCH_CHECK();
    return B3D_OK;
}

template <typename T, int NW, int NWARPS>
static int launch_accumulate(const b3d_icp* h, const double* w, cudaStream_t st) {
    const b3d_model* m = h->m;
    constexpr size_t smem = icp_smem_bytes<T, NW, NWARPS>();
    static_assert(smem <= 232448, "icp_accumulate_kernel: shared memory over the sm_100 limit");
    B3D_CUDA(cudaFuncSetAttribute(icp_accumulate_kernel<T, NW, NWARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));   // per device: every call
    // enough CTAs for two waves of the 148 SMs when the batch is small; one CTA per body group when it is large
    int vsplit = 1;
    {   // the split whose grid fills whole waves of the 148 SMs best (ties: the smaller split)
        double best = 0.0;
        const int hi = h->groups >= 2 * 148 ? 4 : 32;
        for (int sp = 1; sp <= hi; ++sp) {
            // whole waves, and a CTA's fixed cost (transform staging, final reduction) taken as 60 vertices' worth of work
            const double waves = (double)h->groups * sp / 148.0, per = (double)m->V / sp;
            const double eff = waves / std::ceil(waves) * per / (per + 60.0);
            if (eff > best + 1e-9) { best = eff; vsplit = sp; }
        }
    }
    icp_accumulate_kernel<T, NW, NWARPS><<<dim3(h->groups, vsplit), 32 * NWARPS, smem, st>>>((const T*)h->rec, m->V, h->B, w, h->jt,
                                                                                             (const T*)h->tgt, h->acc);
    B3D_LAUNCH_CHECK();
    return B3D_OK;
}

}  // namespace b3d

using namespace b3d;

extern "C" {

int b3d_icp_create(const b3d_model* m, int32_t B, int32_t dtype, const int64_t* cptr, const int32_t* mpt, const int32_t* dpt,
                   const int64_t* dptr, const void* data, int32_t flags, b3d_icp** out) {
    if (!m || !cptr || !mpt || !dptr || !data || !out) return set_error(B3D_EINVAL, "b3d_icp_create: null argument");
    if (B <= 0) return set_error(B3D_EINVAL, "b3d_icp_create: B <= 0");
    if (dtype != B3D_F32 && dtype != B3D_F64) return set_error(B3D_EINVAL, "b3d_icp_create: bad dtype %d", dtype);
    if (m->nw != 4 && m->nw != 8) return set_error(B3D_EINVAL, "b3d_icp_create: models with more than 8 weights per vertex are not supported (nw = %d)", m->nw);
    if (int rc = require_device()) return rc;
    for (int b = 0; b < B; ++b)
        if (cptr[b + 1] < cptr[b] || dptr[b + 1] < dptr[b]) return set_error(B3D_EINVAL, "b3d_icp_create: offsets of body %d decrease", b);
    const int64_t np = cptr[B], nd = dptr[B];
    const size_t es = dtype == B3D_F32 ? 4 : 8;
    b3d_icp* h = new b3d_icp();
    h->m = m; h->B = B; h->dtype = dtype; h->groups = (B + 31) / 32; h->npairs = np;
    const bool dev_in = (flags & B3D_ICP_DEVICE_INPUT) != 0;
    int64_t *d_cptr = nullptr, *d_dptr = nullptr;
    int32_t *d_mpt = nullptr, *d_dpt = nullptr;
    void* d_data = nullptr;
    cudaError_t e = cudaMalloc(&h->tgt, es * 128 * ((size_t)h->groups * m->V + 8));   // + one tile: a partial last tile reads past V
    if (e == cudaSuccess) e = cudaMalloc(&h->scatter, sizeof(double) * 2 * B);
    if (e == cudaSuccess) e = cudaMalloc(&h->acc, sizeof(double) * (size_t)B * ICP_ACC);
    if (e == cudaSuccess) e = cudaMalloc(&h->state, sizeof(double) * (size_t)B * (NS + 3 + 4 * NJ + 1 + ICP_NC));
    if (e == cudaSuccess) e = cudaMalloc(&h->r, sizeof(double) * (size_t)B * NJ * 9);
    if (e == cudaSuccess) e = cudaMalloc(&h->jp, sizeof(double) * (size_t)B * NJ * 3);
    if (e == cudaSuccess) e = cudaMalloc(&h->jt, sizeof(double) * (size_t)B * NJ * 12);
    if (e == cudaSuccess) e = cudaMalloc(&d_cptr, sizeof(int64_t) * 2 * (B + 1));
    if (e == cudaSuccess) {
        d_dptr = d_cptr + B + 1;
        e = cudaMemcpy(d_cptr, cptr, sizeof(int64_t) * (B + 1), cudaMemcpyHostToDevice);
    }
    if (e == cudaSuccess) e = cudaMemcpy(d_dptr, dptr, sizeof(int64_t) * (B + 1), cudaMemcpyHostToDevice);
    if (e == cudaSuccess && !dev_in) {
        e = cudaMalloc(&d_mpt, sizeof(int32_t) * 2 * (size_t)std::max<int64_t>(np, 1));
        if (e == cudaSuccess) { d_dpt = d_mpt + std::max<int64_t>(np, 1); e = cudaMalloc(&d_data, es * 3 * (size_t)std::max<int64_t>(nd, 1)); }
        if (e == cudaSuccess) e = cudaMemcpy(d_mpt, mpt, sizeof(int32_t) * np, cudaMemcpyHostToDevice);
        if (e == cudaSuccess && dpt) e = cudaMemcpy(d_dpt, dpt, sizeof(int32_t) * np, cudaMemcpyHostToDevice);
        if (e == cudaSuccess) e = cudaMemcpy(d_data, data, es * 3 * nd, cudaMemcpyHostToDevice);
    }
    int rc = B3D_OK;
    if (e != cudaSuccess) rc = set_error(B3D_ECUDA, "b3d_icp_create: %s", cudaGetErrorString(e));
    if (!rc) rc = build_order(h);
    if (!rc) {
        const int32_t* pm = dev_in ? mpt : d_mpt;
        const int32_t* pd = !dpt ? nullptr : dev_in ? dpt : d_dpt;
        const void* pdat = dev_in ? data : d_data;
        rc = dtype == B3D_F32 ? icp_reduce<float>(h, d_cptr, d_dptr, pm, pd, (const float*)pdat)
                              : icp_reduce<double>(h, d_cptr, d_dptr, pm, pd, (const double*)pdat);
    }
    if (!rc) {
        if (dtype == B3D_F32) rc = m->nw == 4 ? build_records<float, 4>(h, m->t32) : build_records<float, 8>(h, m->t32);
        else rc = m->nw == 4 ? build_records<double, 4>(h, m->t64) : build_records<double, 8>(h, m->t64);
        if (!rc && cudaStreamSynchronize(nullptr) != cudaSuccess) rc = set_error(B3D_ECUDA, "b3d_icp_create: building the vertex records failed");
    }
    cudaFree(d_cptr); cudaFree(d_mpt); cudaFree(d_data);
    if (rc) { b3d_icp_destroy(h); return rc; }
    *out = h;
    return B3D_OK;
}

void b3d_icp_destroy(b3d_icp* h) {
    if (!h) return;
    cudaFree(h->tgt); cudaFree(h->rec); cudaFree(h->vertex_of_pos); cudaFree(h->scatter); cudaFree(h->acc); cudaFree(h->state); cudaFree(h->r); cudaFree(h->jp); cudaFree(h->jt);
    delete h;
}

int b3d_icp_eval_dev(b3d_icp* h, const double* w, const double* p, const double* q, int32_t flags, double* cost, double* grad,
                     void* stream) {
    if (!h || !w || !p || !q || !cost || !grad) return set_error(B3D_EINVAL, "b3d_icp_eval: null argument");
    cudaStream_t st = (cudaStream_t)stream;
    const b3d_model* m = h->m;
    const int B = h->B;
    icp_quat_kernel<<<(B * NJ + 255) / 256, 256, 0, st>>>(B * NJ, q, h->r);
    B3D_LAUNCH_CHECK();
    if (int rc = b3d_smpl_forward_dev(m, B, B3D_F64, w, p, h->r, nullptr, h->jp, h->jt, 0, st)) return rc;
    B3D_CUDA(cudaMemsetAsync(h->acc, 0, sizeof(double) * (size_t)B * ICP_ACC, st));
    int rc;
    if (h->dtype == B3D_F32)
        rc = m->nw == 4 ? launch_accumulate<float, 4, 12>(h, w, st) : launch_accumulate<float, 8, 8>(h, w, st);
    else
        rc = m->nw == 4 ? launch_accumulate<double, 4, 8>(h, w, st) : launch_accumulate<double, 8, 8>(h, w, st);
    if (rc) return rc;
    icp_finish_kernel<<<(B + 3) / 4, 128, 0, st>>>(m->d_sk, B, (flags & B3D_ICP_POSE_ONLY) ? 0 : 1, h->acc, h->scatter, h->jp, h->jt, cost, grad);
    B3D_LAUNCH_CHECK();
    return B3D_OK;
}

int b3d_icp_eval(b3d_icp* h, const double* w, const double* p, const double* q, int32_t flags, double* cost, double* grad) {
    if (!h || !w || !p || !q || !cost || !grad) return set_error(B3D_EINVAL, "b3d_icp_eval: null argument");
    const size_t B = (size_t)h->B;
    double *dw = h->state, *dp = dw + B * NS, *dq = dp + B * 3, *dc = dq + B * 4 * NJ, *dg = dc + B;
    B3D_CUDA(cudaMemcpy(dw, w, sizeof(double) * B * NS, cudaMemcpyHostToDevice));
    B3D_CUDA(cudaMemcpy(dp, p, sizeof(double) * B * 3, cudaMemcpyHostToDevice));
    B3D_CUDA(cudaMemcpy(dq, q, sizeof(double) * B * 4 * NJ, cudaMemcpyHostToDevice));
    if (int rc = b3d_icp_eval_dev(h, dw, dp, dq, flags, dc, dg, nullptr)) return rc;
    B3D_CUDA(cudaStreamSynchronize(nullptr));
    B3D_CUDA(cudaMemcpy(cost, dc, sizeof(double) * B, cudaMemcpyDeviceToHost));
    B3D_CUDA(cudaMemcpy(grad, dg, sizeof(double) * B * ICP_NC, cudaMemcpyDeviceToHost));
    return B3D_OK;
}

int b3d_icp_add_priors_dev(b3d_icp* h, const b3d_gmm* gmm, double beta_pose, double beta_shape, const double* w, const double* q,
                           double* cost, double* grad, void* stream) {
    if (!h || !w || !q || !cost || !grad) return set_error(B3D_EINVAL, "b3d_icp_add_priors: null argument");
    if (beta_pose > 0.0 && !gmm) return set_error(B3D_EINVAL, "b3d_icp_add_priors: beta_pose > 0 needs the pose prior");
    if (!(beta_pose > 0.0) && !(beta_shape > 0.0)) return B3D_OK;
    const bool pose = beta_pose > 0.0 && gmm->nComps > 0;   // a missing prior (nComps = -1) contributes nothing
    icp_prior_kernel<<<h->B, 128, 0, (cudaStream_t)stream>>>(h->B, pose ? gmm->nComps : 0, pose ? gmm->d_mean : nullptr, pose ? gmm->d_prec_cho : nullptr,
                                                             pose ? gmm->d_consts : nullptr, beta_pose, beta_shape, h->scatter + h->B, w, q, cost, grad);
    B3D_LAUNCH_CHECK();
    return B3D_OK;
}

int b3d_icp_eval_full(b3d_icp* h, const b3d_gmm* gmm, double beta_pose, double beta_shape, const double* w, const double* p, const double* q,
                      int32_t flags, double* cost, double* grad) {
    if (!h || !w || !p || !q || !cost || !grad) return set_error(B3D_EINVAL, "b3d_icp_eval_full: null argument");
    const size_t B = (size_t)h->B;
    double *dw = h->state, *dp = dw + B * NS, *dq = dp + B * 3, *dc = dq + B * 4 * NJ, *dg = dc + B;
    B3D_CUDA(cudaMemcpy(dw, w, sizeof(double) * B * NS, cudaMemcpyHostToDevice));
    B3D_CUDA(cudaMemcpy(dp, p, sizeof(double) * B * 3, cudaMemcpyHostToDevice));
    B3D_CUDA(cudaMemcpy(dq, q, sizeof(double) * B * 4 * NJ, cudaMemcpyHostToDevice));
    if (int rc = b3d_icp_eval_dev(h, dw, dp, dq, flags, dc, dg, nullptr)) return rc;
    if (int rc = b3d_icp_add_priors_dev(h, gmm, beta_pose, beta_shape, dw, dq, dc, dg, nullptr)) return rc;
    B3D_CUDA(cudaStreamSynchronize(nullptr));
    B3D_CUDA(cudaMemcpy(cost, dc, sizeof(double) * B, cudaMemcpyDeviceToHost));
    B3D_CUDA(cudaMemcpy(grad, dg, sizeof(double) * B * ICP_NC, cudaMemcpyDeviceToHost));
    return B3D_OK;
}

int b3d_icp_num_bodies(const b3d_icp* h) { return h ? h->B : 0; }

}  // extern "C"
