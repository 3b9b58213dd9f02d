// Fused submission kernel for sm_100a: one pass that reads an instance's record,
// composes its world matrix, tests its box against every view, compacts the visible ones per view and writes the
// instance rows, keys and ids in draw order.  It replaces the three-kernel pipeline (k_instances_x2 -> sort ->
// k_emit) wherever that pipeline's fast path applied.
//
// Why one kernel.  The old pipeline wrote every world row to HBM in instance order (64 B), sorted, and then k_emit
// fetched the rows back in draw order: a random 64 B read costs a whole 128 B DRAM line, so the gather alone moved
// 2.1 GB per 2^24 instances, on top of the 1.1 GB the rows cost to write.  Here the sort runs FIRST, on words
// [key | instance] that do not depend on the frame's transforms, and the kernel walks the sorted words:
//   * the record of an instance ({TRS, box, slot}, 64 B, one aligned block) is fetched with one TMA bulk copy per
//     record into shared memory (cp.async.bulk + mbarrier, UBLKCP in SASS) while the previous tile computes; in
//     the draw-order-resident layout the records already lie in word order and a tile is ONE 16 KB bulk copy;
//   * each thread carries two instances on packed FP32 (FFMA2), same arithmetic as k_instances_x2 (ss_x2.cuh),
//     so rows and masks are the same bits;
//   * the rows never travel through HBM on their way to the lists: they are staged in shared memory, the
//     per-view ranks come from one block-wide scan of 16-bit packed counters plus a decoupled look-back over
//     tiles (a whole warp walks 32 predecessor tiles per round trip), and the visible rows leave as coalesced
//     128-bit stores in draw order; all rows (visible or not) also leave once, in word order, as one bulk store
//     shared -> global per tile (the "world rows" output);
//   * runs of equal keys become draws exactly as in k_emit (per (view, run) counts, k_draws finishes them).
#include <algorithm>

#include "ss_kernels.h"
#include "ss_x2.cuh"

namespace ss {

namespace {

constexpr int kSubWarps = kSubThreads / 32;

__device__ __forceinline__ void bulk_s2g(void* dst, const void* src_smem, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src_smem)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t* p)
{
    uint32_t v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u32(uint32_t* p, uint32_t v)
{
    asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__device__ __forceinline__ uint64_t shfl_up_u64(uint64_t v, int o)
{
    uint32_t lo = __shfl_up_sync(0xffffffffu, (uint32_t)v, o);
    uint32_t hi = __shfl_up_sync(0xffffffffu, (uint32_t)(v >> 32), o);
    return ((uint64_t)hi << 32) | lo;
}

// bit v of m -> 1 << 16v for v in 0..3
__device__ __forceinline__ uint64_t spread4(uint32_t m)
{
    return (uint64_t)(m & 1u) | ((uint64_t)(m & 2u) << 15) | ((uint64_t)(m & 4u) << 30) | ((uint64_t)(m & 8u) << 45);
}

// Decoupled look-back over a status array lb[c * stride + i]: value | flag (kLbAgg: element i's own count, kLbPrefix:
// inclusive prefix over elements <= i), for up to kLbPerWarp counters cid[] at once (cid < 0: unused).  Each round the
// 32 lanes read the 32 elements below `start` of every unresolved counter (one 128 B line per counter), take the longest
// run of published words that ends at the nearest prefix, and go on from there.  Elements below `floor_i` count as an
// empty prefix.  out[j] = the sum for counter cid[j] (the same in every lane).
constexpr int kLbPerWarp = 3;  // kSubWords counters over kSubWarps = 4 warps
__device__ __forceinline__ void warp_walk(const uint32_t* lb, uint32_t stride, int64_t first, int64_t floor_i, int lane,
                                          const int (&cid)[kLbPerWarp], uint32_t (&out)[kLbPerWarp])
{
    int64_t start[kLbPerWarp];
    uint32_t open = 0;
#pragma unroll
    for (int j = 0; j < kLbPerWarp; ++j) {
        out[j] = 0;
        start[j] = first;
        if (cid[j] >= 0) open |= 1u << j;
    }
    while (open) {
        uint32_t v[kLbPerWarp];
        bool progress = false;
#pragma unroll
        for (int j = 0; j < kLbPerWarp; ++j) {
            const int64_t t = start[j] - lane;
            v[j] = kLbPrefix;  // below the floor: an empty prefix
            if (((open >> j) & 1u) && t >= floor_i) v[j] = ld_volatile_u32(lb + (size_t)cid[j] * stride + (size_t)t);
        }
#pragma unroll
        for (int j = 0; j < kLbPerWarp; ++j) {
            if ((open >> j) & 1u) {
                const uint32_t has_p = __ballot_sync(0xffffffffu, (v[j] & kLbPrefix) != 0);
                const uint32_t has_any = __ballot_sync(0xffffffffu, (v[j] & (kLbPrefix | kLbAgg)) != 0);
                const int first_missing = (~has_any) ? __ffs(~has_any) - 1 : 32;
                const int first_prefix = has_p ? __ffs(has_p) - 1 : 32;
                const int take = min(first_missing, first_prefix + 1);  // lanes [0, take) are usable
                out[j] += __reduce_add_sync(0xffffffffu, lane < take ? (v[j] & kLbValue) : 0u);
                progress = progress || take > 0;
                if (first_prefix < first_missing)
                    open &= ~(1u << j);
                else
                    start[j] -= take;  // take may be 0: the nearest element has not published yet, poll again
            }
        }
        // nothing new was published: leave the issue slots to the warps that compute (a polling warp that spins at full
        // rate slows down the very CTAs it is waiting for)
        if (open && !progress) __nanosleep(256);
    }
}

// Exclusive prefixes over tiles come in two levels.  The persistent CTAs run in near lock-step, so at any moment
// several hundred consecutive tiles have published their counts but not yet their prefixes; walking them 32 at a time
// would cost dozens of L2 round trips per tile.  Instead tiles are grouped by kLbGroup = 32:
//   level 1  lb_tile[c][t]   = kLbAgg | count of tile t               (never upgraded)
//   level 2  lb_group[c][g]  = kLbAgg | count of group g, then kLbPrefix | inclusive prefix over groups <= g
// A tile sums the counts of the earlier tiles of its own group (one round), then walks the groups (32 groups = 1024
// tiles per round).  The last tile of a group publishes the group's count and, later, the group's prefix.  Every warp of
// the CTA handles its own share of the counters (see lb_publish / lb_resolve in submit_kernel.cuh).
constexpr int kLbGroup = 32;

#include "submit_kernel.cuh"

// ---- sort words from the static id streams -----------------------------------------------------------
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

// ---- instance records from the streams: four threads per record, one 16 B quarter each -----------------
__global__ void __launch_bounds__(256) k_pack_records(const __grid_constant__ PackArgs a)
{
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t r = a.lo + (g >> 2), q = g & 3u;
    if (r >= a.hi) return;
    uint32_t i = r;
    if (a.order) i = (uint32_t)(__ldg(a.order + r) & a.idx_mask);
    const float* t = a.trs + (size_t)i * 9;
    float4* dst = reinterpret_cast<float4*>(a.rec + r) + q;
    if (q == 0) {
        *dst = make_float4(t[0], t[1], t[2], t[3]);
        if (a.pos) a.pos[i] = r;
        if (a.parent_out) a.parent_out[r] = a.parent ? a.parent[i] : -1;
    } else if (q == 1) {
        *dst = make_float4(t[4], t[5], t[6], t[7]);
    } else if (q == 2) {
        if (a.trs_only) {
            reinterpret_cast<float*>(dst)[0] = t[8];
        } else {
            const float* b = a.inst_aabb ? a.inst_aabb + (size_t)i * 6 : nullptr;
            *dst = make_float4(t[8], b ? b[0] : 0.f, b ? b[1] : 0.f, b ? b[2] : 0.f);
        }
    } else if (!a.trs_only) {
        const float* b = a.inst_aabb ? a.inst_aabb + (size_t)i * 6 : nullptr;
        *dst = make_float4(b ? b[3] : 0.f, b ? b[4] : 0.f, b ? b[5] : 0.f, __uint_as_float(__ldg(a.slot + i)));
    }
}

__global__ void __launch_bounds__(256) k_unpermute(const uint64_t* __restrict__ words, uint64_t idx_mask, uint32_t n, uint32_t first,
                                                   uint32_t count, const float4* __restrict__ world_words, float4* __restrict__ rows_out,
                                                   const uint8_t* __restrict__ mask_words, uint8_t* __restrict__ mask_out)
{
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += stride) {
        const uint32_t i = (uint32_t)(words[j] & idx_mask) - first;
        if (i < count) {
            if (rows_out) {
#pragma unroll
                for (int r = 0; r < 3; ++r) rows_out[(size_t)i * 3 + r] = world_words[(size_t)j * 3 + r];
            }
            if (mask_out) mask_out[i] = mask_words[j];
        }
    }
}

}  // namespace

uint32_t submit_lookback_stride(uint64_t n)
{
    const uint64_t tiles = (n + kSubTile - 1) / kSubTile;
    return (uint32_t)(((tiles ? tiles : 1) + 31) & ~uint64_t(31));
}

// level 1: [kSubWords][stride] tile words, level 2: [kSubWords][stride / kLbGroup] group words
size_t submit_lookback_bytes(uint64_t n)
{
    const size_t stride = submit_lookback_stride(n);
    return (stride + stride / kLbGroup) * kSubWords * sizeof(uint32_t);
}

void launch_build_words(const WordsArgs& a, cudaStream_t st)
{
    if (!a.n) return;
    const uint32_t want = (a.n + 256 * 8 - 1) / (256 * 8);
    k_build_words<<<std::min<uint32_t>(want, 148 * 8), 256, 0, st>>>(a);
}

void launch_pack_records(const PackArgs& a, cudaStream_t st)
{
    if (a.hi <= a.lo) return;
    const uint64_t threads = (uint64_t)(a.hi - a.lo) * 4;
    k_pack_records<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(a);
}

void launch_unpermute(const uint64_t* words, uint64_t idx_mask, uint32_t n, uint32_t first, uint32_t count, const float4* world_words,
                      float4* rows_out, const uint8_t* mask_words, uint8_t* mask_out, cudaStream_t st)
{
    if (!n || !count) return;
    const uint32_t want = (n + 255) / 256;
    k_unpermute<<<std::min<uint32_t>(want, 148 * 16), 256, 0, st>>>(words, idx_mask, n, first, count, world_words, rows_out, mask_words,
                                                                     mask_out);
}

namespace {

struct SubmitLaunch {
    int sms = 0;
    int per_sm[4] = {0, 0, 0, 0};
};
SubmitLaunch g_submit_launch[64];  // per device ordinal: attribute set + occupancy looked up once

template <bool GATHER, bool HAS_PARENT>
cudaError_t launch_submit_t(const SubmitArgs& a, const CullParams& cp, const Neutral& nz, cudaStream_t st)
{
    using Smem = SubSmem<HAS_PARENT>;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    SubmitLaunch& L = g_submit_launch[dev];
    constexpr int which = (GATHER ? 1 : 0) + (HAS_PARENT ? 2 : 0);
    if (!L.per_sm[which]) {
        e = cudaDeviceGetAttribute(&L.sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(k_submit<GATHER, HAS_PARENT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem));
        if (e != cudaSuccess) return e;
        int per = 0;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, k_submit<GATHER, HAS_PARENT>, kSubThreads, sizeof(Smem));
        if (e != cudaSuccess) return e;
        L.per_sm[which] = std::max(per, 1);
    }
    const uint32_t n_tiles = (a.n + kSubTile - 1) / kSubTile;
    // persistent, and every CTA must be resident: a tile's look-back waits on tiles that other CTAs of the grid run
    const uint32_t grid = std::min<uint32_t>(n_tiles, (uint32_t)(L.sms * L.per_sm[which]));
    k_submit<GATHER, HAS_PARENT><<<grid, kSubThreads, sizeof(Smem), st>>>(a, cp, nz);
    return cudaGetLastError();
}

}  // namespace

cudaError_t launch_submit(const SubmitArgs& a, const CullParams& cp, cudaStream_t st)
{
    if (!a.n) return cudaSuccess;
    Neutral nz;
    const uint32_t one = 0x3f800000u, neg_one = 0xbf800000u, neg_zero = 0x80000000u;
    nz.one = ((uint64_t)one << 32) | one;
    nz.neg_one = ((uint64_t)neg_one << 32) | neg_one;
    nz.neg_zero = ((uint64_t)neg_zero << 32) | neg_zero;
    if (a.gather)
        return a.parent ? launch_submit_t<true, true>(a, cp, nz, st) : launch_submit_t<true, false>(a, cp, nz, st);
    return a.parent ? launch_submit_t<false, true>(a, cp, nz, st) : launch_submit_t<false, false>(a, cp, nz, st);
}

}  // namespace ss
