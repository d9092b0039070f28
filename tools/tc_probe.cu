// tc_probe: smallest possible check of the tcgen05 building blocks the LBS tensor-core kernel relies on, before they are used
// inside it: canonical K-major / no-swizzle shared-memory operand layout, UMMA shared-memory descriptors, the kind::tf32
// instruction descriptor (M = 128, N = 192), TMEM allocation, tcgen05.mma + commit to an mbarrier, tcgen05.ld 32x32b.
//   D[128 x 192] = A[128 x 16] * B[192 x 16]^T   (two K = 8 steps), inputs exactly representable in TF32 -> exact fp32 result.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/tc_probe tools/tc_probe.cu ; run on a B200.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>

constexpr int M = 128, N = 192, K = 16;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// canonical K-major, no swizzle: slab of 8 k-values; inside a slab [row group of 8][k chunk of 4][row in group][k in chunk]
__host__ __device__ inline int canon_off_floats(int rows, int r, int k) {
    const int slab = k >> 3, kc = (k >> 2) & 1;
    return slab * rows * 8 + (r >> 3) * 64 + kc * 32 + (r & 7) * 4 + (k & 3);
}

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);              // start address, bits [0,14)
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;    // leading-dimension byte offset, bits [16,30)
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;    // stride-dimension byte offset, bits [32,46)
    d |= (uint64_t)1 << 46;                              // descriptor version 1 (Blackwell)
    return d;                                            // base offset 0, layout type 0 = SWIZZLE_NONE
}

template <bool A_TMEM>
__global__ void __launch_bounds__(128, 1) probe(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ D, uint32_t lbo, uint32_t sbo) {
    __shared__ __align__(128) float sA[M * K];
    __shared__ __align__(128) float sB[N * K];
    __shared__ __align__(8) unsigned long long bar;
    __shared__ uint32_t tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < M * K; i += 128) sA[canon_off_floats(M, i / K, i % K)] = A[i];
    for (int i = tid; i < N * K; i += 128) sB[canon_off_floats(N, i / K, i % K)] = B[i];
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base)), "r"(256));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic-proxy smem writes -> visible to the tensor core
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = tmem_base;
    if (A_TMEM) {   // A[row = lane][k] -> TMEM columns [192, 192 + K) of the thread's lane
        const uint32_t ta = tm + ((uint32_t)(warp * 32) << 16) + 192u;
        for (int k = 0; k < K; k += 4)
            asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(ta + k), "f"(A[tid * K + k]), "f"(A[tid * K + k + 1]),
                         "f"(A[tid * K + k + 2]), "f"(A[tid * K + k + 3]) : "memory");
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    }
    if (tid == 0) {
        // instruction descriptor: D = F32 (bits 4-5 = 1), A = B = TF32 (bits 7-9, 10-12 = 2), K-major both, N >> 3 at 17, M >> 4 at 24
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
        for (int s = 0; s < K / 8; ++s) {
            const uint64_t da = make_desc(smem_u32(sA) + s * M * 32, lbo, sbo);
            const uint64_t db = make_desc(smem_u32(sB) + s * N * 32, lbo, sbo);
            const uint32_t acc = s > 0 ? 1u : 0u;
            if (A_TMEM)
                asm volatile(
                    "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                    "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tm), "r"(tm + 192u + (uint32_t)(8 * s)), "l"(db), "r"(idesc), "r"(acc)
                    : "memory");
            else
            asm volatile(
                "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tm), "l"(da), "l"(db), "r"(idesc), "r"(acc)
                : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    }
    // everybody waits for the MMAs
    asm volatile(
        "{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}" ::"r"(
            smem_u32(&bar))
        : "memory");
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    // warp w reads TMEM lanes 32w .. 32w+31 (rows), 8 columns at a time
    const int row = tid;
    for (int c = 0; c < N; c += 8) {
        uint32_t v[8];
        const uint32_t taddr = tm + ((uint32_t)(warp * 32) << 16) + (uint32_t)c;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                     : "r"(taddr));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int i = 0; i < 8; ++i) D[row * N + c + i] = __uint_as_float(v[i]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(256));
}

int main() {
    std::vector<float> A(M * K), B(N * K), D(M * N, -1.f), R(M * N);
    srand(7);
    auto tf = [] { return (float)((rand() % 2001) - 1000) / 256.0f; };   // multiples of 1/256, |x| < 4: exact in TF32
    for (auto& x : A) x = tf();
    for (auto& x : B) x = tf();
    for (int i = 0; i < M; ++i)
        for (int j = 0; j < N; ++j) {
            double s = 0;
            for (int k = 0; k < K; ++k) s += (double)A[i * K + k] * B[j * K + k];
            R[i * N + j] = (float)s;
        }
    float *dA, *dB, *dD;
    cudaMalloc(&dA, A.size() * 4); cudaMalloc(&dB, B.size() * 4); cudaMalloc(&dD, D.size() * 4);
    cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dD, D.data(), D.size() * 4, cudaMemcpyHostToDevice);
    int rc = 1;
    const uint32_t variants[2][2] = {{128, 256}, {256, 128}};      // {LBO, SBO}: the first is what the layout above means
    for (int v = 0; v < 2; ++v) {
        cudaMemset(dD, 0xff, D.size() * 4);
        if (v == 0) probe<false><<<1, 128>>>(dA, dB, dD, variants[v][0], variants[v][1]);
        else probe<true><<<1, 128>>>(dA, dB, dD, variants[0][0], variants[0][1]);   // variant 1: A operand read from TMEM
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("TC_PROBE CUDA ERROR: %s\n", cudaGetErrorString(e)); return 2; }
        cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
        double maxerr = 0; int bad = 0;
        for (int i = 0; i < M * N; ++i) { const double d = fabs((double)D[i] - R[i]); if (!(d <= 1e-4)) ++bad; if (d > maxerr) maxerr = d; }
        printf("TC_PROBE LBO=%u SBO=%u: max|D - ref| = %.3e, mismatches %d of %d; D[0][0..3] = %g %g %g %g ref %g %g %g %g; D[127][191] = %g ref %g\n",
               variants[v][0], variants[v][1], maxerr, bad, M * N, D[0], D[1], D[2], D[3], R[0], R[1], R[2], R[3], D[127 * N + 191], R[127 * N + 191]);
        if (bad == 0) { printf("TC_PROBE OK (variant %d: %s)\n", v, v ? "A from TMEM" : "A from shared memory"); if (v == 0) rc = 0; }
    }
    return rc;

}
