// chol_probe: what one phase of the register-tile Cholesky costs on B200, in isolation — 256 threads, per "panel": 16 LDS.128 of
// panel rows + 64 DFMA into a 4 x 4 register tile + __syncthreads; and a one-lane pivot chain (4 x rsqrt + updates) + __syncthreads.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/chol_probe tools/chol_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256, 2) update_phase(double* out, long long* cyc, int panels, int reps) {
    __shared__ __align__(16) double2 pan[4][2][22];
    const int tid = threadIdx.x;
    int ti = 0;
    while ((ti + 1) * (ti + 2) / 2 <= tid) ++ti;
    const int tj = tid - ti * (ti + 1) / 2;
    const bool ok = ti < 22;
    if (tid < 4 * 2 * 22) (&pan[0][0][0])[tid] = make_double2(1e-3 * tid, 2e-3 * tid);
    double a[4][4];
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) a[r][c] = out[(r * 4 + c) & 3];
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < reps; ++it)
        for (int k = 0; k < panels; ++k) {
            if (ok && tj > k) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const double2 r01 = pan[q][0][ti], r23 = pan[q][1][ti], c01 = pan[q][0][tj], c23 = pan[q][1][tj];
                    const double rv[4] = {r01.x, r01.y, r23.x, r23.y}, cv[4] = {c01.x, c01.y, c23.x, c23.y};
#pragma unroll
                    for (int r = 0; r < 4; ++r)
#pragma unroll
                        for (int c = 0; c < 4; ++c) a[r][c] = fma(-rv[r], cv[c], a[r][c]);
                }
            }
            __syncthreads();
        }
    const long long t1 = clock64();
    double s = 0;
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) s += a[r][c];
    out[8 + blockIdx.x * 256 + tid] = s;
    if (tid == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}

__global__ void __launch_bounds__(256, 2) factor_phase(double* out, long long* cyc, int panels, int reps) {
    __shared__ double dk[2][4][4];
    const int tid = threadIdx.x;
    double a[4][4];
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) a[r][c] = (r == c ? 4.0 : 0.1) + out[(r * 4 + c) & 3];
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < reps; ++it)
        for (int k = 0; k < panels; ++k) {
            if (tid == ((k * 37) & 255)) {
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const double inv = rsqrt(a[c][c]);
                    a[c][c] = inv;
#pragma unroll
                    for (int r = c + 1; r < 4; ++r) a[r][c] *= inv;
#pragma unroll
                    for (int c2 = c + 1; c2 < 4; ++c2)
#pragma unroll
                        for (int r = c2; r < 4; ++r) a[r][c2] -= a[r][c] * a[c2][c];
                }
#pragma unroll
                for (int r = 0; r < 4; ++r)
#pragma unroll
                    for (int c = 0; c <= r; ++c) { dk[k & 1][r][c] = a[r][c]; a[r][c] = (r == c ? 4.0 : 0.1) + 1e-9 * a[r][c]; }
            }
            __syncthreads();
        }
    const long long t1 = clock64();
    out[8 + blockIdx.x * 256 + tid] = a[3][3] + dk[0][1][1];
    if (tid == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}

int main() {
    double* d; long long* c; long long h;
    cudaMalloc(&d, 8 * (8 + 512 * 256)); cudaMalloc(&c, 8);
    cudaMemset(d, 0, 8 * (8 + 512 * 256));
    const int panels = 22, reps = 50;
    for (int blocks : {1, 296, 592}) {
        update_phase<<<blocks, 256>>>(d, c, panels, reps); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
        printf("update phase (16 LDS.128 + 64 DFMA + barrier), %3d CTAs: %.1f cycles per panel\n", blocks, (double)h / (panels * reps));
        factor_phase<<<blocks, 256>>>(d, c, panels, reps); cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
        printf("factor phase (one lane: 4 x rsqrt chain + publish + barrier), %3d CTAs: %.1f cycles per panel\n", blocks, (double)h / (panels * reps));
    }
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
