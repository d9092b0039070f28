// tmem_probe: tcgen05.ld / tcgen05.st throughput per SM on B200 (the tensor-core LBS kernel keeps the per-body joint transforms
// in TMEM, one body per lane, and re-reads 48 columns whenever the joint tuple changes).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/tmem_probe tools/tmem_probe.cu ; run on a B200.
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int NCOL>
__device__ __forceinline__ uint32_t ld_cols(uint32_t taddr);
template <>
__device__ __forceinline__ uint32_t ld_cols<4>(uint32_t taddr) {
    uint32_t v[4];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    return v[0] ^ v[1] ^ v[2] ^ v[3];
}
template <>
__device__ __forceinline__ uint32_t ld_cols<8>(uint32_t taddr) {
    uint32_t v[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    return v[0] ^ v[1] ^ v[2] ^ v[3] ^ v[4] ^ v[5] ^ v[6] ^ v[7];
}
template <>
__device__ __forceinline__ uint32_t ld_cols<16>(uint32_t taddr) {
    uint32_t v[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
                   "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]) : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    uint32_t x = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) x ^= v[i];
    return x;
}
// two x16 loads in flight before the wait
__device__ __forceinline__ uint32_t ld_2x16(uint32_t t0, uint32_t t1) {
    uint32_t v[32];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
                   "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]) : "r"(t0));
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]),
                   "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31]) : "r"(t1));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    uint32_t x = 0;
#pragma unroll
    for (int i = 0; i < 32; ++i) x ^= v[i];
    return x;
}

template <int MODE>
__global__ void probe(uint32_t* out, long long* cyc, int n) {
    __shared__ uint32_t tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
    // fill the lanes this warp owns (tcgen05.st), so the loads read defined data
    for (int c = 0; c < 512; c += 4) {
        const uint32_t a = tid * 7 + c, b = a + 1, d = a + 2, e = a + 3;
        asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(tm + c), "r"(a), "r"(b), "r"(d), "r"(e));
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    __syncthreads();
    uint32_t x = 0;
    const long long t0 = clock64();
    for (int i = 0; i < n; ++i) {
        const uint32_t col = (uint32_t)((i * 48) & 255);
        if (MODE == 4) x ^= ld_cols<4>(tm + col);
        if (MODE == 8) x ^= ld_cols<8>(tm + col);
        if (MODE == 16) x ^= ld_cols<16>(tm + col);
        if (MODE == 32) x ^= ld_2x16(tm + col, tm + col + 16);
    }
    __syncthreads();
    const long long t1 = clock64();
    out[tid] = x;
    if (tid == 0) cyc[0] = t1 - t0;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
}

int main() {
    uint32_t* o; long long* c; long long h;
    cudaMalloc(&o, 4 * 2048); cudaMalloc(&c, 8);
    const int n = 2048;
    for (int threads : {32, 128, 256, 512}) {
        probe<4><<<1, threads>>>(o, c, n); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
        printf("tcgen05.ld 32x32b.x4   %3d threads: %.1f cycles per load per warp, %.1f B/clk/SM\n", threads, (double)h / n, 4.0 * 128 * n * (threads / 32) / (double)h);
        probe<8><<<1, threads>>>(o, c, n); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
        printf("tcgen05.ld 32x32b.x8   %3d threads: %.1f cycles per load per warp, %.1f B/clk/SM\n", threads, (double)h / n, 8.0 * 128 * n * (threads / 32) / (double)h);
        probe<16><<<1, threads>>>(o, c, n); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
        printf("tcgen05.ld 32x32b.x16  %3d threads: %.1f cycles per load per warp, %.1f B/clk/SM\n", threads, (double)h / n, 16.0 * 128 * n * (threads / 32) / (double)h);
        probe<32><<<1, threads>>>(o, c, n); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
        printf("tcgen05.ld 2 x x16     %3d threads: %.1f cycles per pair per warp, %.1f B/clk/SM\n", threads, (double)h / n, 32.0 * 128 * n * (threads / 32) / (double)h);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
