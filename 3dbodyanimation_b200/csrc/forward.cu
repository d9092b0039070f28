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
#include "internal.hpp"

namespace b3d {

constexpr int JW = 4;   // warps (= bodies) per CTA in the joints kernel

template <typename T>
__global__ void __launch_bounds__(32 * JW) smpl_joints_kernel(const SkeletonConst* __restrict__ sk, int B,
                                                              const T* __restrict__ w, const T* __restrict__ p,
                                                              const T* __restrict__ r, T* __restrict__ jointPos,
                                                              T* __restrict__ jointTrans, T* __restrict__ posefeat) {
    __shared__ double s_r[JW][NJ * 9];
    __shared__ double s_g[JW][NJ * 12];
    __shared__ double s_jp[JW][NJ * 3];
    __shared__ double s_w[JW][NS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b = blockIdx.x * JW + warp;
    if (b >= B) return;
    double* R = s_r[warp];
    double* G = s_g[warp];
    double* jp = s_jp[warp];
    for (int i = lane; i < NJ * 9; i += 32) R[i] = (double)r[(size_t)b * NJ * 9 + i];
    if (lane < NS) s_w[warp][lane] = (double)w[(size_t)b * NS + lane];
    __syncwarp();
    // jointPos = initialJointPos + jointShapeReg * w
    for (int i = lane; i < NJ * 3; i += 32) {
        double acc = 0.0;
#pragma unroll
        for (int s = 0; s < NS; ++s) acc += sk->JSR[i * NS + s] * s_w[warp][s];
        jp[i] = sk->J0[i] + acc;
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
    for (int lvl = 1; lvl <= max_depth; ++lvl) {
        if (depth == lvl) {
            const double* A = G + 12 * pa;
            const double* Rl = R + 9 * j;
            double* Bm = G + 12 * j;
            const double t0 = jp[3 * j] - jp[3 * pa], t1 = jp[3 * j + 1] - jp[3 * pa + 1], t2 = jp[3 * j + 2] - jp[3 * pa + 2];
#pragma unroll
            for (int col = 0; col < 3; ++col)
#pragma unroll
                for (int row = 0; row < 3; ++row)
                    Bm[3 * col + row] = A[row] * Rl[3 * col] + A[3 + row] * Rl[3 * col + 1] + A[6 + row] * Rl[3 * col + 2];
#pragma unroll
            for (int row = 0; row < 3; ++row) Bm[9 + row] = A[9 + row] + (A[row] * t0 + A[3 + row] * t1 + A[6 + row] * t2);
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

template <typename T>
static int forward_impl(const b3d_model* m, int B, const T* w, const T* p, const T* r, T* cloud, T* jointPos, T* jointTrans,
                        int flags, cudaStream_t st) {
    const VertexTables<T>& t = sizeof(T) == 4 ? (const VertexTables<T>&)m->t32 : (const VertexTables<T>&)m->t64;
    const bool use_pd = (flags & B3D_FWD_POSEDIRS) != 0;
    if (use_pd && !m->has_posedirs) return set_error(B3D_EINVAL, "B3D_FWD_POSEDIRS requested but the model has no posedirs");
    T* G = jointTrans;
    T* scratchG = nullptr;
    T* pf = nullptr;
    T* off = nullptr;
    if (!G) { B3D_CUDA(cudaMallocAsync(&scratchG, sizeof(T) * (size_t)B * NJ * 12, st)); G = scratchG; }
    if (use_pd && cloud) {
        B3D_CUDA(cudaMallocAsync(&pf, sizeof(T) * (size_t)B * B3D_POSEDIRS_DIM, st));
        B3D_CUDA(cudaMallocAsync(&off, sizeof(T) * (size_t)B * 3 * m->Vpad, st));
    }
    smpl_joints_kernel<T><<<(B + JW - 1) / JW, 32 * JW, 0, st>>>(m->d_sk, B, w, p, r, jointPos, G, pf);
    B3D_LAUNCH_CHECK();
    if (cloud) {
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
        if (m->nw == 4) B3D_LBS(4);
        else if (m->nw == 8) B3D_LBS(8);
        else B3D_LBS(NJ);
#undef B3D_LBS
        B3D_LAUNCH_CHECK();
    }
    if (scratchG) B3D_CUDA(cudaFreeAsync(scratchG, st));
    if (pf) B3D_CUDA(cudaFreeAsync(pf, st));
    if (off) B3D_CUDA(cudaFreeAsync(off, st));
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
    const size_t sz_w = es * nb * NS, sz_p = es * nb * 3, sz_r = es * nb * NJ * 9;
    const size_t sz_c = es * nb * 3 * m->V, sz_jp = es * nb * NJ * 3, sz_jt = es * nb * NJ * 12;
    char *dw = nullptr, *dc = nullptr;
    B3D_CUDA(cudaMalloc(&dw, sz_w + sz_p + sz_r + sz_jp + sz_jt));
    char *dp = dw + sz_w, *dr = dp + sz_p, *djp = dr + sz_r, *djt = djp + sz_jp;
    int rc = B3D_OK;
    cudaError_t e = cudaSuccess;
    if (cloud) e = cudaMalloc(&dc, sz_c);
    if (e == cudaSuccess) e = cudaMemcpy(dw, w, sz_w, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(dp, p, sz_p, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(dr, r, sz_r, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        rc = b3d_smpl_forward_dev(m, B, dtype, dw, dp, dr, dc, djp, djt, flags, nullptr);
        if (!rc) {
            e = cudaStreamSynchronize(nullptr);
            if (e == cudaSuccess && cloud) e = cudaMemcpy(cloud, dc, sz_c, cudaMemcpyDeviceToHost);
            if (e == cudaSuccess && jointPos) e = cudaMemcpy(jointPos, djp, sz_jp, cudaMemcpyDeviceToHost);
            if (e == cudaSuccess && jointTrans) e = cudaMemcpy(jointTrans, djt, sz_jt, cudaMemcpyDeviceToHost);
        }
    }
    cudaFree(dw);
    cudaFree(dc);
    if (e != cudaSuccess) return set_error(B3D_ECUDA, "b3d_smpl_forward: %s", cudaGetErrorString(e));
    return rc;
}

}  // extern "C"
