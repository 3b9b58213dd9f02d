// microbenchmark: DRAM bytes fetched per randomly gathered 64 B-aligned row, as a function of how many bytes of
// the row are read (16/32/48/64) and of cudaLimitMaxL2FetchGranularity.  Run under
//   ncu --metrics dram__bytes_read.sum,gpu__time_duration.sum
// and divide by 2^24 rows.  One launch per (limit, bytes) pair, in the order printed.
#include <cuda_runtime.h>
#include <cstdio>
#include <vector>
#include <numeric>
#include <algorithm>
#include <random>
template <int PARTS>
__global__ void k_gather(const float4* __restrict__ src, const uint32_t* __restrict__ idx, float4* __restrict__ dst, uint32_t n)
{
    const uint32_t total = n * PARTS;
    for (uint32_t j0 = blockIdx.x * blockDim.x * 4 + threadIdx.x; j0 < total; j0 += gridDim.x * blockDim.x * 4) {
        float4 r[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const uint32_t j = j0 + u * blockDim.x;
            if (j < total) { const uint32_t e = j / PARTS, part = j - e * PARTS; r[u] = __ldg(src + (size_t)__ldg(idx + e) * 4 + part); }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) { const uint32_t j = j0 + u * blockDim.x; if (j < total) dst[j] = r[u]; }
    }
}
template <int PARTS>
float run(const float4* src, const uint32_t* idx, float4* dst, uint32_t n, cudaEvent_t e0, cudaEvent_t e1)
{
    cudaEventRecord(e0);
    k_gather<PARTS><<<148 * 16, 256>>>(src, idx, dst, n);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}
int main()
{
    const uint32_t n = 1u << 24;
    float4 *src, *dst; uint32_t* idx;
    cudaMalloc(&src, (size_t)n * 64); cudaMalloc(&dst, (size_t)n * 64); cudaMalloc(&idx, (size_t)n * 4);
    cudaMemset(src, 1, (size_t)n * 64);
    std::vector<uint32_t> h(n); std::iota(h.begin(), h.end(), 0u);
    std::mt19937 g(1); std::shuffle(h.begin(), h.end(), g);
    cudaMemcpy(idx, h.data(), (size_t)n * 4, cudaMemcpyHostToDevice);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const size_t lims[4] = {0, 32, 64, 128};
    for (size_t lim : lims) {
        if (lim) { cudaError_t e = cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, lim); if (e) printf("set limit %zu: %s\n", lim, cudaGetErrorString(e)); }
        size_t got = 0; cudaDeviceGetLimit(&got, cudaLimitMaxL2FetchGranularity);
        float t1 = run<1>(src, idx, dst, n, e0, e1), t2 = run<2>(src, idx, dst, n, e0, e1), t3 = run<3>(src, idx, dst, n, e0, e1),
              t4 = run<4>(src, idx, dst, n, e0, e1);
        printf("limit asked %3zu got %3zu: 16 B/row %.3f ms, 32 B/row %.3f ms, 48 B/row %.3f ms, 64 B/row %.3f ms\n", lim, got, t1, t2, t3, t4);
    }
    return 0;
}
