// ffma_probe: FP32 issue rates that bound the CUDA-core half of the tensor-core LBS kernel on B200:
//   scalar FFMA (three register operands) vs packed FFMA2 (PTX fma.rn.f32x2), per SM, for 4 / 8 / 16 / 32 resident warps.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/ffma_probe tools/ffma_probe.cu ; run on a B200.
#include <cstdio>
#include <cuda_runtime.h>

__global__ void ffma_scalar(float* out, long long* cyc, int n) {
    float a[16];
    const float b = out[1], c = out[2];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = out[0] + i;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < n; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], b, c);
    }
    __syncthreads();
    const long long t1 = clock64();
    float s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += a[i];
    out[3 + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}

// distinct multiplier registers per accumulator (no operand reuse cache help)
__global__ void ffma_scalar_distinct(float* out, long long* cyc, int n) {
    float a[16], m[16];
    const float c = out[2];
#pragma unroll
    for (int i = 0; i < 16; ++i) { a[i] = out[0] + i; m[i] = out[1] + 1e-7f * i; }
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < n; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], m[i], a[(i + 5) & 15]);
    }
    __syncthreads();
    const long long t1 = clock64();
    float s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += a[i];
    out[3 + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}

__device__ __forceinline__ unsigned long long pack2(float lo, float hi) {
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ unsigned long long ffma2(unsigned long long a, unsigned long long b, unsigned long long c) {
    unsigned long long d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__global__ void ffma_packed(float* out, long long* cyc, int n) {
    unsigned long long a[8];
    const unsigned long long b = pack2(out[1], out[1]), c = pack2(out[2], out[2]);
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = pack2(out[0] + i, out[0] + i + 0.5f);
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < n; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = ffma2(a[i], b, c);
    }
    __syncthreads();
    const long long t1 = clock64();
    unsigned long long s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s ^= a[i];
    reinterpret_cast<unsigned long long*>(out + 4)[threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}

int main() {
    float* f; long long* c; long long h;
    cudaMalloc(&f, 4 * 8192); cudaMalloc(&c, 8);
    const float initf[3] = {1.f, 0.999999f, 1e-7f};
    cudaMemcpy(f, initf, 12, cudaMemcpyHostToDevice);
    const int n = 4096;
    for (int threads : {128, 256, 512, 1024}) {
        ffma_scalar<<<1, threads>>>(f, c, n); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
        printf("FFMA  (shared b, c)   %4d threads: %.1f FMA lanes/clk/SM (%.2f warp-instr/clk)\n", threads, 16.0 * n * threads / (double)h, 16.0 * n * threads / 32.0 / (double)h);
        ffma_scalar_distinct<<<1, threads>>>(f, c, n); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
        printf("FFMA  (3 distinct reg) %4d threads: %.1f FMA lanes/clk/SM (%.2f warp-instr/clk)\n", threads, 16.0 * n * threads / (double)h, 16.0 * n * threads / 32.0 / (double)h);
        ffma_packed<<<1, threads>>>(f, c, n); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
        printf("FFMA2 (f32x2)         %4d threads: %.1f FMA lanes/clk/SM (%.2f warp-instr/clk)\n", threads, 16.0 * n * threads / (double)h, 8.0 * n * threads / 32.0 / (double)h);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
