// dfma.cu — FP64 peak of this GPU: independent fma.rn.f64 chains on every SM (the denominator for the
// zone test's FP64-pipe figures; MEASURED_PEAKS.json has no FP64 entry).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/exp/dfma tools/exp/dfma.cu && tools/exp/dfma
// Prints one JSON line: {"dfma_tflops": ..., "dfma_per_clk_per_sm_at_max_clock": ..., ...}.
#include <cstdio>
#include <cuda_runtime.h>
#define CHAINS 8
#define ITERS 4096
__global__ void __launch_bounds__(256) dfma_kernel(double *out, double a, double b) {
    double x[CHAINS];
#pragma unroll
    for (int k = 0; k < CHAINS; k++) x[k] = threadIdx.x * 1e-3 + k;
    for (int i = 0; i < ITERS; i++)
#pragma unroll
        for (int k = 0; k < CHAINS; k++) x[k] = __fma_rn(x[k], a, b);  // CHAINS independent dependent-chains
    double s = 0;
#pragma unroll
    for (int k = 0; k < CHAINS; k++) s += x[k];
    if (s == 123.456) out[blockIdx.x * blockDim.x + threadIdx.x] = s;  // keeps the chains alive
}
int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int ctas = p.multiProcessorCount * 8, thr = 256;
    double *out;
    cudaMalloc(&out, sizeof(double) * ctas * thr);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 8; rep++) {
        cudaEventRecord(e0);
        dfma_kernel<<<ctas, thr>>>(out, 0.999999, 1e-7);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep >= 2 && ms < best) best = ms;
    }
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double fmas = (double)ctas * thr * CHAINS * ITERS;
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"dfma_per_s\": %.4e, \"dfma_tflops\": %.3f, \"ms\": %.4f, "
           "\"dfma_per_clk_per_sm_at_max_clock\": %.2f, \"max_clock_mhz\": %.0f, \"cuda_error\": \"%s\"}\n",
           p.name, p.multiProcessorCount, fmas / (best * 1e-3), 2.0 * fmas / (best * 1e-3) / 1e12, best,
           fmas / (best * 1e-3) / (khz * 1e3) / p.multiProcessorCount, khz / 1e3, cudaGetErrorString(cudaGetLastError()));
    return 0;
}
