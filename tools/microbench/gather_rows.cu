// microbenchmark: how fast can B200 gather 64 B-aligned rows by a random permutation and write them densely?
// (the access pattern of k_emit's view that sees almost everything)
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <numeric>
#include <algorithm>
#include <random>
template <int OUT4>  // float4 written per row (3 = 48 B rows out)
__global__ void k_gather(const float4* __restrict__ src, const uint32_t* __restrict__ idx, float4* __restrict__ dst, uint32_t n)
{
    const uint32_t total = n * OUT4;
    for (uint32_t j0 = blockIdx.x * blockDim.x * 4 + threadIdx.x; j0 < total; j0 += gridDim.x * blockDim.x * 4) {
        float4 r[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const uint32_t j = j0 + u * blockDim.x;
            if (j < total) { const uint32_t e = j / OUT4, part = j - e * OUT4; r[u] = __ldg(src + (size_t)__ldg(idx + e) * 4 + part); }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) { const uint32_t j = j0 + u * blockDim.x; if (j < total) dst[j] = r[u]; }
    }
}
__global__ void k_copy(const float4* __restrict__ src, float4* __restrict__ dst, size_t n4)
{
    for (size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x; j < n4; j += (size_t)gridDim.x * blockDim.x) dst[j] = __ldg(src + j);
}
int main()
{
    const uint32_t n = 1u << 24;
    float4 *src, *dst; uint32_t* idx;
    cudaMalloc(&src, (size_t)n * 64); cudaMalloc(&dst, (size_t)n * 64); cudaMalloc(&idx, (size_t)n * 4);
    cudaMemset(src, 1, (size_t)n * 64);
    std::vector<uint32_t> h(n); std::iota(h.begin(), h.end(), 0u);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int mode = 0; mode < 3; ++mode) {
        if (mode == 1) { std::mt19937 g(1); std::shuffle(h.begin(), h.end(), g); }
        if (mode == 2) { // zipf-like runs: sorted-by-key order of random keys over 4096 keys: idx increasing inside runs
            std::mt19937 g(2); std::vector<uint32_t> key(n); for (auto& k : key) { double u = (g() + 0.5) / 4294967296.0; k = (uint32_t)(4096 * u * u * u); }
            std::iota(h.begin(), h.end(), 0u); std::stable_sort(h.begin(), h.end(), [&](uint32_t a, uint32_t b) { return key[a] < key[b]; });
        }
        cudaMemcpy(idx, h.data(), (size_t)n * 4, cudaMemcpyHostToDevice);
        for (int blocks : {148 * 4, 148 * 8, 148 * 16}) {
            k_gather<3><<<blocks, 256>>>(src, idx, dst, n);
            cudaEventRecord(e0);
            for (int it = 0; it < 5; ++it) k_gather<3><<<blocks, 256>>>(src, idx, dst, n);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 5;
            printf("mode %d (%s) blocks %5d: %.3f ms  rows/s %.2f G  useful GB/s %.0f (48 in + 48 out + 4 idx)\n", mode,
                   mode == 0 ? "sequential" : mode == 1 ? "random perm" : "key-sorted runs", blocks, ms, n / ms / 1e6, n * 100.0 / ms / 1e6);
        }
    }
    cudaEventRecord(e0);
    for (int it = 0; it < 5; ++it) k_copy<<<148 * 16, 256>>>(src, dst, (size_t)n * 4);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 5;
    printf("plain copy of 1 GiB: %.3f ms, %.0f GB/s (read+write)\n", ms, 2.0 * n * 64 / ms / 1e6);
    return 0;
}
