// Draw-order-resident layout (SS_CONFIG_DRAW_ORDER_RESIDENT) for sm_100a: the kernels that build it.
//
// The sort keys of a scene are static between ss_set_* calls (in the reference the std::map nodes persist across
// frames as well, VC/Scene/SceneRenderer.cpp:1559-1591), so the library can sort ONCE when a static stream changes
// and keep its per-instance streams in (key, instance) order.  A frame is then two streaming passes and nothing else:
// k_instances_x2 over the permuted streams (it writes world rows and masks in draw order) and k_emit in its
// sequential mode (row j belongs to sorted word j: no sort, no random gather).
//   k_build_words    sort words [key | instance] from the static id streams (+ the sort's digit histograms)
//   k_permute        streams in draw order: TRS, box, slot, material word, parent of sorted position j
//   k_unpermute      world rows / masks of an instance range back in instance order (downloads, parity tests)
#include <algorithm>

#include "ss_kernels.h"

namespace ss {

namespace {

__global__ void __launch_bounds__(256) k_build_words(const __grid_constant__ WordsArgs a)
{
    __shared__ uint32_t h[3 * kMaxRadix];
    for (int i = threadIdx.x; i < a.hist_passes * kMaxRadix; i += blockDim.x) h[i] = 0;
    __syncthreads();
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += stride) {
        const uint32_t sl = __ldg(a.slot + i), mw = __ldg(a.matword + i);
        uint64_t key = 2ull << a.tess_shift;  // dropped: the bit above tess, behind every real key
        if (sl != SS_SLOT_NULL)
            key = __ldg(&a.slots[sl].key_hi) | (uint64_t)(mw & ~SS_MATERIAL_TESS_BIT) | ((uint64_t)(mw >> 31) << a.tess_shift);
        const uint64_t w = (key << a.idx_bits) | i;
        a.words[i] = w;
#pragma unroll
        for (int p = 0; p < 3; ++p)
            if (p < a.hist_passes) atomicAdd(&h[p * kMaxRadix + ((uint32_t)(w >> a.hist_shift[p]) & ((1u << a.hist_bits[p]) - 1u))], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < a.hist_passes * kMaxRadix; i += blockDim.x) {
        const uint32_t c = h[i];
        if (c) atomicAdd(a.hist + i, c);
    }
}

// One thread per (sorted position, float of the 36 B TRS record): the writes are coalesced, the reads gather whole
// records.  The first thread of a position also moves the static fields when asked to.
__global__ void __launch_bounds__(256) k_permute(const __grid_constant__ PermuteArgs a)
{
    const uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t j = (uint32_t)(g / 9u), f = (uint32_t)(g - (uint64_t)j * 9u);
    if (j >= a.n) return;
    const uint32_t i = (uint32_t)(__ldg(a.order + j) & a.idx_mask);
    a.trs_out[(size_t)j * 9 + f] = __ldg(a.trs + (size_t)i * 9 + f);
    if (a.trs_only) return;
    if (f < 6 && a.box) a.box_out[(size_t)j * 6 + f] = __ldg(a.box + (size_t)i * 6 + f);
    if (f == 6) a.slot_out[j] = __ldg(a.slot + i);
    if (f == 7) a.matword_out[j] = __ldg(a.matword + i);
    if (f == 8) {
        if (a.parent) a.parent_out[j] = __ldg(a.parent + i);
        a.pos_out[i] = j;
    }
}

__global__ void __launch_bounds__(256) k_unpermute(const uint64_t* __restrict__ words, uint64_t idx_mask, uint32_t n, uint32_t first,
                                                   uint32_t count, const float4* __restrict__ world_words, uint32_t stride4,
                                                   float4* __restrict__ rows_out, const uint8_t* __restrict__ mask_words,
                                                   uint8_t* __restrict__ mask_out)
{
    const uint32_t step = gridDim.x * blockDim.x;
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += step) {
        const uint32_t i = (uint32_t)(words[j] & idx_mask) - first;
        if (i < count) {
            if (rows_out) {
#pragma unroll
                for (int r = 0; r < 3; ++r) rows_out[(size_t)i * 3 + r] = world_words[(size_t)j * stride4 + r];
            }
            if (mask_out) mask_out[i] = mask_words[j];
        }
    }
}

}  // namespace

void launch_build_words(const WordsArgs& a, cudaStream_t st)
{
    if (!a.n) return;
    const uint32_t want = (a.n + 256 * 8 - 1) / (256 * 8);
    k_build_words<<<std::min<uint32_t>(want, 148 * 8), 256, 0, st>>>(a);
}

void launch_permute(const PermuteArgs& a, cudaStream_t st)
{
    if (!a.n) return;
    const uint64_t threads = (uint64_t)a.n * 9;
    k_permute<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(a);
}

void launch_unpermute(const uint64_t* words, uint64_t idx_mask, uint32_t n, uint32_t first, uint32_t count, const float4* world_words,
                      uint32_t stride4, float4* rows_out, const uint8_t* mask_words, uint8_t* mask_out, cudaStream_t st)
{
    if (!n || !count) return;
    const uint32_t want = (n + 255) / 256;
    k_unpermute<<<std::min<uint32_t>(want, 148 * 16), 256, 0, st>>>(words, idx_mask, n, first, count, world_words, stride4, rows_out,
                                                                     mask_words, mask_out);
}

}  // namespace ss
