// Experiment: what HBM rate does the decode kernel's ACCESS PATTERN reach with no parsing at all?
// Reads `payload` (63 B per record, 16-byte vectors, coalesced) and `rec_off` (8 B per record), writes
// nine 8-byte columns + a 1-byte and a 4-byte column per record, the way pb_decode_kernel<dense> does.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/exp/streams tools/exp/streams.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(256) pattern(const uint4 *payload, const uint64_t *off, uint64_t *cols, uint8_t *st,
                                               uint32_t *fr, uint32_t n, int ncol, int dep) {
    const uint32_t i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    uint64_t o = off[i];
    // the warp's 32 records: 2016 bytes = 126 vectors; each lane takes ~4
    const uint32_t w0 = (i & ~31u);
    uint64_t base = dep ? __shfl_sync(0xffffffffu, o, 0) / 16 : (uint64_t)w0 * 63 / 16;
    uint4 acc = make_uint4(0, 0, 0, 0);
    for (uint32_t k = (i & 31); k < 127; k += 32) {
        uint4 v = payload[base + k];
        acc.x ^= v.x; acc.y ^= v.y; acc.z ^= v.z; acc.w ^= v.w;
    }
    uint64_t x = o ^ acc.x ^ ((uint64_t)acc.y << 32) ^ acc.z ^ acc.w;
    for (int c = 0; c < ncol; c++) cols[(size_t)c * n + i] = x + c;
    st[i] = (uint8_t)x;
    fr[i] = (uint32_t)x;
}
int main() {
    const uint32_t n = 10000000;
    uint4 *p; uint64_t *off, *cols; uint8_t *st; uint32_t *fr;
    cudaMalloc(&p, (size_t)n * 63 + 4096); cudaMalloc(&off, (size_t)(n + 1) * 8); cudaMalloc(&cols, (size_t)n * 8 * 13);
    cudaMalloc(&st, n); cudaMalloc(&fr, (size_t)n * 4 + 4);
    cudaMemset(p, 1, (size_t)n * 63 + 4096);
    uint64_t *h = (uint64_t *)malloc((size_t)(n + 1) * 8);
    for (uint32_t i = 0; i <= n; i++) h[i] = (uint64_t)i * 63;
    cudaMemcpy(off, h, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int dep = 0; dep < 2; dep++)
        for (int ncol = 1; ncol <= 13; ncol += 4) {
            for (int it = 0; it < 3; it++) pattern<<<(n + 255) / 256, 256>>>(p, off, cols, st, fr, n, ncol, dep);
            cudaEventRecord(a);
            for (int it = 0; it < 10; it++) pattern<<<(n + 255) / 256, 256>>>(p, off, cols, st, fr, n, ncol, dep);
            cudaEventRecord(b); cudaEventSynchronize(b);
            float ms; cudaEventElapsedTime(&ms, a, b); ms /= 10;
            double bytes = (double)n * (63 + 8 + 8.0 * ncol + 1 + 4);
            printf("dependent_offsets=%d columns=%2d  %.3f ms  %.0f GB/s\n", dep, ncol, ms, bytes / ms / 1e6);
        }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
