// Micro-benchmarks that ground the LM kernel design on B200: fp64 DFMA latency / throughput, LDS->DFMA chains, barrier cost.
#include <cstdio>
#include <cuda_runtime.h>

__global__ void dfma_latency(double* out, long long* cyc, int n) {
    double a = out[0], b = out[1], c = out[2];
    long long t0 = clock64();
    for (int i = 0; i < n; ++i) { a = fma(a, b, c); a = fma(a, b, c); a = fma(a, b, c); a = fma(a, b, c); }
    long long t1 = clock64();
    out[3] = a; if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void dfma_throughput(double* out, long long* cyc, int n) {
    double a0 = out[0], a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7, b = out[1], c = out[2];
    __syncthreads();
    long long t0 = clock64();
    for (int i = 0; i < n; ++i) {
        a0 = fma(a0, b, c); a1 = fma(a1, b, c); a2 = fma(a2, b, c); a3 = fma(a3, b, c);
        a4 = fma(a4, b, c); a5 = fma(a5, b, c); a6 = fma(a6, b, c); a7 = fma(a7, b, c);
    }
    __syncthreads();
    long long t1 = clock64();
    out[3 + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7; if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void ffma_throughput(float* out, long long* cyc, int n) {
    float a0 = out[0], a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7, b = out[1], c = out[2];
    __syncthreads();
    long long t0 = clock64();
    for (int i = 0; i < n; ++i) {
        a0 = fmaf(a0, b, c); a1 = fmaf(a1, b, c); a2 = fmaf(a2, b, c); a3 = fmaf(a3, b, c);
        a4 = fmaf(a4, b, c); a5 = fmaf(a5, b, c); a6 = fmaf(a6, b, c); a7 = fmaf(a7, b, c);
    }
    __syncthreads();
    long long t1 = clock64();
    out[3 + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7; if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void lds_dfma_chain(double* out, long long* cyc, int n) {
    __shared__ double s[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) s[i] = 1.0 + 1e-9 * i;
    __syncthreads();
    double a = out[0];
    long long t0 = clock64();
    for (int i = 0; i < n; ++i) { const int k = (i * 4 + threadIdx.x) & 1020; a = fma(s[k], s[k + 1], a); a = fma(s[k + 2], s[k + 3], a); }
    long long t1 = clock64();
    out[3 + threadIdx.x] = a; if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void barrier_cost(long long* cyc, int n) {
    __syncthreads();
    long long t0 = clock64();
    for (int i = 0; i < n; ++i) __syncthreads();
    long long t1 = clock64();
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void sqrt_div_latency(double* out, long long* cyc, int n) {
    double a = out[0] + 2.0;
    long long t0 = clock64();
    for (int i = 0; i < n; ++i) { a = 1.0 / sqrt(a) + 1.5; }
    long long t1 = clock64();
    out[3] = a; if (threadIdx.x == 0) cyc[0] = t1 - t0;
}

__global__ void rsqrt_latency(double* out, long long* cyc, int n) {
    double a = out[0] + 2.0;
    long long t0 = clock64();
    for (int i = 0; i < n; ++i) { a = rsqrt(a) + 1.5; }
    long long t1 = clock64();
    out[3] = a; if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void rcp_latency(double* out, long long* cyc, int n) {
    double a = out[0] + 2.0;
    long long t0 = clock64();
    for (int i = 0; i < n; ++i) { a = 1.0 / a + 1.5; }
    long long t1 = clock64();
    out[3] = a; if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void shfl_latency(double* out, long long* cyc, int n) {
    double a = out[0] + threadIdx.x;
    long long t0 = clock64();
    for (int i = 0; i < n; ++i) { a = __shfl_sync(0xffffffffu, a, (threadIdx.x + 1) & 31); }
    long long t1 = clock64();
    out[3 + threadIdx.x] = a; if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void sincos_latency(double* out, long long* cyc, int n) {
    double a = out[0] + 0.3;
    long long t0 = clock64();
    for (int i = 0; i < n; ++i) { double s, c; sincos(a, &s, &c); a = s + c; }
    long long t1 = clock64();
    out[3] = a; if (threadIdx.x == 0) cyc[0] = t1 - t0;
}

int main() {
    double* d; float* f; long long* c; long long h;
    cudaMalloc(&d, 8 * 4096); cudaMalloc(&f, 4 * 4096); cudaMalloc(&c, 8);
    double init[3] = {1.0, 0.999999, 1e-7}; float initf[3] = {1.f, 0.999999f, 1e-7f};
    cudaMemcpy(d, init, 24, cudaMemcpyHostToDevice); cudaMemcpy(f, initf, 12, cudaMemcpyHostToDevice);
    const int n = 4096;
    dfma_latency<<<1, 32>>>(d, c, n); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
    printf("DFMA dependent latency: %.2f cycles\n", (double)h / (4.0 * n));
    for (int threads : {32, 128, 256, 512, 1024}) {
        dfma_throughput<<<1, threads>>>(d, c, n); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
        printf("DFMA throughput, 1 CTA x %4d threads: %.2f DFMA/clk/SM\n", threads, 8.0 * n * threads / (double)h);
    }
    for (int threads : {128, 512, 1024}) {
        ffma_throughput<<<1, threads>>>(f, c, n); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
        printf("FFMA throughput, 1 CTA x %4d threads: %.2f FFMA/clk/SM\n", threads, 8.0 * n * threads / (double)h);
    }
    lds_dfma_chain<<<1, 32>>>(d, c, n); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
    printf("2x(2 LDS.64 + dependent DFMA): %.2f cycles per DFMA\n", (double)h / (2.0 * n));
    for (int threads : {128, 256, 512}) {
        barrier_cost<<<1, threads>>>(c, n); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
        printf("__syncthreads, %d threads: %.1f cycles\n", threads, (double)h / n);
    }
    sqrt_div_latency<<<1, 32>>>(d, c, n); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
    printf("1/sqrt(x) fp64 dependent: %.1f cycles\n", (double)h / n);
    rsqrt_latency<<<1, 32>>>(d, c, n); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
    printf("rsqrt(x) + add fp64 dependent: %.1f cycles\n", (double)h / n);
    rcp_latency<<<1, 32>>>(d, c, n); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
    printf("1/x + add fp64 dependent: %.1f cycles\n", (double)h / n);
    shfl_latency<<<1, 32>>>(d, c, n); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
    printf("shfl (64-bit) dependent: %.1f cycles\n", (double)h / n);
    sincos_latency<<<1, 32>>>(d, c, n); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
    printf("sincos fp64 + add dependent: %.1f cycles\n", (double)h / n);
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
