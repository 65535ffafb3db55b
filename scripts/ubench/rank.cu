// Microbenchmark: cost of computing, for every lane, the mask of lanes holding the same
// 8-bit digit (the ranking step of a radix pass).  Cycles per warp-level item.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(512) k(const uint32_t *in, uint32_t *out, int iters)
{
    __shared__ uint32_t masks[16][256];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 16 * 256; i += blockDim.x) (&masks[0][0])[i] = 0;
    __syncthreads();
    uint32_t x = in[blockIdx.x * blockDim.x + threadIdx.x];
    uint32_t acc = 0;
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
        uint32_t d = (x >> 8) & 255u;
        uint32_t m;
        if (MODE == 0) {
            m = __match_any_sync(0xffffffffu, d);
        } else if (MODE == 1) {
            m = 0xffffffffu;
#pragma unroll
            for (int b = 0; b < 8; b++) {
                bool bit = (d >> b) & 1;
                uint32_t v = __ballot_sync(0xffffffffu, bit);
                m &= bit ? v : ~v;
            }
        } else if (MODE == 2) {
            atomicOr(&masks[warp][d], 1u << lane);
            __syncwarp();
            m = masks[warp][d];
            __syncwarp();
            masks[warp][d] = 0;
            __syncwarp();
        } else {  // 4-bit digit by ballots
            m = 0xffffffffu;
#pragma unroll
            for (int b = 0; b < 4; b++) {
                bool bit = (d >> b) & 1;
                uint32_t v = __ballot_sync(0xffffffffu, bit);
                m &= bit ? v : ~v;
            }
        }
        acc += __popc(m & ((1u << lane) - 1));
        x = x * 1664525u + 1013904223u + acc;
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = (uint32_t)((t1 - t0) / iters);
}

template <int MODE>
void run(const char *name, uint32_t *in, uint32_t *out)
{
    const int iters = 2000, grid = 148 * 4;
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    k<MODE><<<grid, 512>>>(in, out, 10);
    cudaEventRecord(a);
    k<MODE><<<grid, 512>>>(in, out, iters);
    cudaEventRecord(b);
    cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, a, b);
    uint32_t cyc; cudaMemcpy(&cyc, out, 4, cudaMemcpyDeviceToHost);
    double items = (double)grid * 512 * iters;
    printf("%-14s %8.3f ms  %7.2f Gitems/s  (%u cycles per iteration seen by one warp at 2x16 warps/SM)\n", name, ms,
           items / ms / 1e6, cyc);
}

int main()
{
    uint32_t *in, *out;
    const int n = 148 * 4 * 512;
    cudaMalloc(&in, n * 4); cudaMalloc(&out, n * 4);
    uint32_t *h = new uint32_t[n];
    for (int i = 0; i < n; i++) h[i] = (uint32_t)rand() * 2654435761u;
    cudaMemcpy(in, h, n * 4, cudaMemcpyHostToDevice);
    run<0>("match_any", in, out);
    run<1>("ballot8", in, out);
    run<2>("atomicOr", in, out);
    run<3>("ballot4", in, out);
    // skewed digits (all equal)
    cudaMemset(in, 0, n * 4);
    printf("all-equal digits:\n");
    run<0>("match_any", in, out);
    run<1>("ballot8", in, out);
    run<2>("atomicOr", in, out);
    cudaError_t e = cudaGetLastError();
    printf("status: %s\n", cudaGetErrorString(e));
    return 0;
}
