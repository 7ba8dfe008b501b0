// Dependent-chain latencies on one warp (sm_100a): what a sequential pivot chain is made of.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o lat lat.cu ; ./lat
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(double* out, long long* clk, double seed, int iters) {
    __shared__ double sm[64];
    const int lane = threadIdx.x & 31;
    sm[lane] = seed + lane; sm[lane + 32] = seed;
    __syncthreads();
    double x = seed + 1.0, y = 1.000001;
    long long t0, t1;
    // 0: DFMA chain
    t0 = clock64();
    for (int i = 0; i < iters; i++) x = fma(x, y, 1e-9);
    t1 = clock64(); if (lane == 0) clk[0] = t1 - t0;
    // 1: DMUL chain
    t0 = clock64();
    for (int i = 0; i < iters; i++) x = x * y;
    t1 = clock64(); if (lane == 0) clk[1] = t1 - t0;
    // 2: rsqrt(double) chain
    x = fabs(x) + 2.0;
    t0 = clock64();
    for (int i = 0; i < iters; i++) x = rsqrt(x) + 3.0;
    t1 = clock64(); if (lane == 0) clk[2] = t1 - t0;   // includes one DADD per step
    // 3: shfl of a double chain
    t0 = clock64();
    for (int i = 0; i < iters; i++) x = __shfl_sync(0xffffffffu, x, (lane + 1) & 31);
    t1 = clock64(); if (lane == 0) clk[3] = t1 - t0;
    // 4: shared-memory load chain (pointer chase through doubles)
    int idx = lane;
    t0 = clock64();
    for (int i = 0; i < iters; i++) { double v = sm[idx]; idx = ((int)v + i) & 31; }
    t1 = clock64(); if (lane == 0) clk[4] = t1 - t0;   // includes F2I + IADD + LOP per step
    // 5: float rsqrt seed + 2 Newton steps in double
    x = fabs(x) + 2.0 + idx;
    t0 = clock64();
    for (int i = 0; i < iters; i++) {
        double r = (double)rsqrtf((float)x);
        double e = fma(-x * r, r, 1.0); r = fma(0.5 * r, e, r);
        e = fma(-x * r, r, 1.0); r = fma(0.5 * r, e, r);
        x = r + 3.0;
    }
    t1 = clock64(); if (lane == 0) clk[5] = t1 - t0;
    // 6: 32 independent DFMAs per step (throughput of one warp)
    double a[32];
    for (int j = 0; j < 32; j++) a[j] = x + j;
    t0 = clock64();
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 32; j++) a[j] = fma(a[j], y, 1e-9);
    }
    t1 = clock64(); if (lane == 0) clk[6] = t1 - t0;
    for (int j = 0; j < 32; j++) x += a[j];
    // 7: 32 independent double shuffles per step
    t0 = clock64();
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 32; j++) a[j] = __shfl_sync(0xffffffffu, a[j], j);
    }
    t1 = clock64(); if (lane == 0) clk[7] = t1 - t0;
    for (int j = 0; j < 32; j++) x += a[j];
    out[threadIdx.x] = x + idx;
}
int main() {
    double* out; long long* clk; long long h[8];
    cudaMalloc(&out, 256 * 8); cudaMalloc(&clk, 64);
    const int iters = 2000;
    for (int rep = 0; rep < 2; rep++) k<<<1, 32>>>(out, clk, 1.5, iters);
    cudaDeviceSynchronize();
    cudaMemcpy(h, clk, 64, cudaMemcpyDeviceToHost);
    const char* names[] = {"DFMA dependent", "DMUL dependent", "rsqrt(double)+DADD dependent", "SHFL(double) dependent",
                           "LDS.64 + index arithmetic dependent", "rsqrtf seed + 2 Newton (double) + DADD", "32 independent DFMA (per DFMA)",
                           "32 independent SHFL(double) (per shuffle)"};
    for (int i = 0; i < 8; i++) printf("%-45s %8.1f cycles\n", names[i], (double)h[i] / iters / (i >= 6 ? 32 : 1));
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
