// store_probe: how fast can one SM (and all 148 together) push rows of a staging tile from shared memory to global memory?
// The tensor-core LBS kernel leaves 128 rows of 1152 B per sort tile; variants: LDS.32 + STG.32 (what it did first), LDS.64/128 +
// STG.64/128, and TMA bulk stores (cp.async.bulk.global.shared::cta) of whole rows.  Rows go to dst + row * 82680 B like the cloud.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/store_probe tools/store_probe.cu ; run on a B200.
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#ifndef ROW_FLOATS
#define ROW_FLOATS 288
#endif
constexpr int ROWS = 128, PITCH = 292;   // 16 B aligned pitch for the vector / bulk variants
constexpr size_t BODY_STRIDE = 82688;                        // bytes, 16 B aligned stand-in for 6890 * 12 = 82680
constexpr int THREADS = 512;

__device__ __forceinline__ uint32_t s_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int MODE>
__global__ void __launch_bounds__(THREADS, 1) probe(unsigned char* dst, int tiles, long long* cyc) {
    extern __shared__ __align__(128) float sm[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < ROWS * PITCH; i += THREADS) sm[i] = (float)i;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    // CTA c owns bodies [128 c, 128 c + 128); tile t is the t-th 1152 B column block of their rows
    unsigned char* base = dst + (size_t)blockIdx.x * ROWS * BODY_STRIDE;
    const long long t0 = clock64();
    for (int t = 0; t < tiles; ++t) {
        unsigned char* tb = base + (size_t)t * ROW_FLOATS * 4;
        if (MODE == 0) {   // 16 warps x 8 rows, LDS.32 + STG.32
            for (int r = warp * 8; r < warp * 8 + 8; ++r) {
                float* d = reinterpret_cast<float*>(tb + (size_t)r * BODY_STRIDE) + lane;
                const float* s = sm + r * PITCH + lane;
                float v[ROW_FLOATS / 32];
#pragma unroll
                for (int i = 0; i < ROW_FLOATS / 32; ++i) v[i] = s[32 * i];
#pragma unroll
                for (int i = 0; i < ROW_FLOATS / 32; ++i) d[32 * i] = v[i];
            }
        } else if (MODE == 1) {   // LDS.128 + STG.128: 72 float4 per row
            for (int r = warp * 8; r < warp * 8 + 8; ++r) {
                float4* d = reinterpret_cast<float4*>(tb + (size_t)r * BODY_STRIDE);
                const float4* s = reinterpret_cast<const float4*>(sm + r * PITCH);
                float4 v[3];
#pragma unroll
                for (int i = 0; i < 3; ++i) if (lane + 32 * i < ROW_FLOATS / 4) v[i] = s[lane + 32 * i];
#pragma unroll
                for (int i = 0; i < 3; ++i) if (lane + 32 * i < ROW_FLOATS / 4) d[lane + 32 * i] = v[i];
            }
        } else {   // TMA bulk store: thread r < 128 stores row r (1152 B)
            if (tid < ROWS) {
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(tb + (size_t)tid * BODY_STRIDE),
                             "r"(s_u32(sm + tid * PITCH)), "r"(ROW_FLOATS * 4) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            }
        }
        __syncthreads();
    }
    if (MODE == 2 && tid < ROWS) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    __syncthreads();
    const long long t1 = clock64();
    if (tid == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
static void run(const char* name, unsigned char* d, long long* c, int ctas, int tiles) {
    cudaFuncSetAttribute(probe<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, ROWS * PITCH * 4);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    probe<MODE><<<ctas, THREADS, ROWS * PITCH * 4>>>(d, tiles, c);   // warm-up
    cudaEventRecord(e0);
    probe<MODE><<<ctas, THREADS, ROWS * PITCH * 4>>>(d, tiles, c);
    cudaEventRecord(e1);
    cudaError_t e = cudaDeviceSynchronize();
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    long long h[148];
    cudaMemcpy(h, c, sizeof(long long) * ctas, cudaMemcpyDeviceToHost);
    double mean = 0;
    for (int i = 0; i < ctas; ++i) mean += (double)h[i] / ctas;
    const double bytes = (double)ctas * tiles * ROWS * ROW_FLOATS * 4;
    printf("%-22s %3d CTAs x %2d tiles: %.1f B/clk/SM (in-kernel), %.0f GB/s aggregate (events)  [%s]\n", name, ctas, tiles,
           (double)tiles * ROWS * ROW_FLOATS * 4 / mean, bytes / ms / 1e6, cudaGetErrorString(e));
}

int main() {
    const int tiles = 82688 / (ROW_FLOATS * 4) - 1; printf("row bytes %d\n", ROW_FLOATS * 4);   // 71 * 1152 B <= 82688
    unsigned char* d; long long* c;
    cudaMalloc(&d, (size_t)148 * ROWS * BODY_STRIDE);
    cudaMalloc(&c, sizeof(long long) * 148);
    for (int ctas : {1, 148}) {
        run<0>("LDS.32 + STG.32", d, c, ctas, tiles);
        run<1>("LDS.128 + STG.128", d, c, ctas, tiles);
        run<2>("TMA bulk row stores", d, c, ctas, tiles);
    }
    return 0;
}
