// FP64 tensor-pipe peak probe: a register-resident mma.sync.m8n8k4.f64 loop with no memory
// traffic.  Gives the issue-rate ceiling the DMMA GEMM engine is measured against (the
// driver-written MEASURED_PEAKS.json holds only bf16 and HBM figures).
#include <cstdio>
#include "internal.h"

namespace dpb {
__global__ void __launch_bounds__(256) dmma_peak_kernel(double* out, int iters) {
    double acc[32][2];
#pragma unroll
    for (int i = 0; i < 32; i++) acc[i][0] = acc[i][1] = 0.0;
    double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 32; i++) dmma884(acc[i][0], acc[i][1], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 32; i++) s += acc[i][0] + acc[i][1];
    if (s == 123.456) out[0] = s;  // keep the loop alive
}
}  // namespace dpb

extern "C" double dplasma_b200_dmma_peak_tflops(int ctas_per_sm, int iters, double* seconds_out) {
    using namespace dpb;
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    double* out = nullptr;
    if (cudaMalloc(&out, 8) != cudaSuccess) return -1.0;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int grid = sms * ctas_per_sm;
    dmma_peak_kernel<<<grid, 256>>>(out, 16);  // warm-up
    cudaEventRecord(e0);
    dmma_peak_kernel<<<grid, 256>>>(out, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    if (cudaGetLastError() != cudaSuccess) return -1.0;
    const double flops = 2.0 * 8 * 8 * 4 * 32.0 * iters * 8.0 /*warps*/ * grid;
    if (seconds_out) *seconds_out = ms * 1e-3;
    return flops / (ms * 1e-3) / 1e12;
}
