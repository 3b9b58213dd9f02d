// microbenchmark: sequential read of 64 B rows, write each as one aligned 64 B row at a random position
// (the pattern an instance kernel would have if it wrote world rows straight into sorted order)
#include <cuda_runtime.h>
#include <cstdio>
#include <vector>
#include <numeric>
#include <algorithm>
#include <random>
__global__ void k_scatter(const float4* __restrict__ src, const uint32_t* __restrict__ pos, float4* __restrict__ dst, uint32_t n)
{
    // 4 lanes cooperate on one row: each writes one float4 of the 64 B row
    const uint32_t total = n * 4;
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < total; j += gridDim.x * blockDim.x) {
        const uint32_t e = j >> 2, part = j & 3;
        dst[(size_t)__ldg(pos + e) * 4 + part] = __ldg(src + j);
    }
}
__global__ void k_scatter_thread(const float4* __restrict__ src, const uint32_t* __restrict__ pos, float4* __restrict__ dst, uint32_t n)
{
    // one thread writes the whole 64 B row (what a one-instance-per-thread kernel does)
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x) {
        const float4 a = __ldg(src + (size_t)e * 4), b = __ldg(src + (size_t)e * 4 + 1), c = __ldg(src + (size_t)e * 4 + 2), d = __ldg(src + (size_t)e * 4 + 3);
        float4* q = dst + (size_t)__ldg(pos + e) * 4;
        q[0] = a; q[1] = b; q[2] = c; q[3] = d;
    }
}
int main()
{
    const uint32_t n = 1u << 24;
    float4 *src, *dst; uint32_t* pos;
    cudaMalloc(&src, (size_t)n * 64); cudaMalloc(&dst, (size_t)n * 64); cudaMalloc(&pos, (size_t)n * 4);
    cudaMemset(src, 1, (size_t)n * 64);
    std::vector<uint32_t> h(n); std::iota(h.begin(), h.end(), 0u);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int mode = 0; mode < 2; ++mode) {
        if (mode == 1) { std::mt19937 g(1); std::shuffle(h.begin(), h.end(), g); }
        cudaMemcpy(pos, h.data(), (size_t)n * 4, cudaMemcpyHostToDevice);
        for (int variant = 0; variant < 2; ++variant) {
            for (int blocks : {148 * 4, 148 * 16}) {
                auto run = [&]() { if (variant == 0) k_scatter<<<blocks, 256>>>(src, pos, dst, n); else k_scatter_thread<<<blocks, 256>>>(src, pos, dst, n); };
                run();
                cudaEventRecord(e0);
                for (int it = 0; it < 5; ++it) run();
                cudaEventRecord(e1); cudaEventSynchronize(e1);
                float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 5;
                printf("%s, %s, blocks %5d: %.3f ms  rows/s %.2f G  GB/s %.0f (64 in + 64 out + 4 idx)\n", mode == 0 ? "sequential" : "random perm",
                       variant == 0 ? "4 lanes/row" : "1 thread/row", blocks, ms, n / ms / 1e6, n * 132.0 / ms / 1e6);
            }
        }
    }
    return 0;
}
