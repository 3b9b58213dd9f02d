// Hand-written onesweep LSD radix sort for sm_100a: 64-bit keys, with or without a 32-bit payload.
// No CUB / Thrust.  This is the device replacement of the grouping std::map<MeshKey, ...> performs
// on the CPU (VC/Scene/SceneRenderer.h:252-253, insert at SceneRenderer.cpp:1146-1148): a stable sort
// by key keeps entt iteration order inside each group (SURVEY.md §7 hard part 4).
//
// Structure (Adinets & Merrill, "Onesweep", 2022):
//   k_sort_hist  one read of the keys, the histogram of every digit position at once
//   k_sort_scan  exclusive bin offsets per digit; marks digits that are constant over all keys so
//                their pass is skipped (narrow key layouts leave most of the 64 bits constant)
//   k_onesweep   per digit: each CTA ranks a tile with warp match_any, publishes its per-digit counts,
//                resolves its global offsets by decoupled look-back over earlier tiles, and scatters
//                through shared memory so the global stores are coalesced per digit run.
// Digits are 9 bits wide when that brings a narrow key range down to <= 3 passes, else 8 bits.
// In the frame the key and the instance index travel packed in ONE 64-bit word (keys-only mode), so
// an executed pass reads and writes 8 B per element; the pairs mode moves 12 B.  HBM-bound.
#include <algorithm>
#include <mutex>
#include <string>

#include "ss_kernels.h"

namespace ss {

namespace {

constexpr uint32_t kFlagAgg = 1u << 30;     // tile's own count is published
constexpr uint32_t kFlagPrefix = 1u << 31;  // inclusive prefix over all earlier tiles is published
constexpr uint32_t kValueMask = kFlagAgg - 1;
constexpr int kSortWarps = kSortThreads / 32;
constexpr size_t kCtlBytes = (sizeof(SortControl) + 255) & ~size_t(255);

__host__ __device__ inline uint64_t num_tiles(uint64_t n) { return (n + kSortTile - 1) / kSortTile; }

__device__ __forceinline__ uint32_t digit_of(uint64_t k, int shift, uint32_t mask)
{
    return (uint32_t)(k >> shift) & mask;
}

struct HistArgs {
    int n_passes;
    int shift[kMaxPasses];
    int bits[kMaxPasses];
};

__global__ void __launch_bounds__(256) k_sort_hist(const uint64_t* __restrict__ keys, uint64_t n,
                                                   const __grid_constant__ HistArgs ha, SortControl* ctl)
{
    extern __shared__ uint32_t h[];  // [n_passes][kMaxRadix]
    for (int i = threadIdx.x; i < ha.n_passes * kMaxRadix; i += blockDim.x) h[i] = 0;
    __syncthreads();
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const uint64_t k = __ldg(keys + i);
        for (int p = 0; p < ha.n_passes; ++p)
            atomicAdd(&h[p * kMaxRadix + digit_of(k, ha.shift[p], (1u << ha.bits[p]) - 1u)], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < ha.n_passes * kMaxRadix; i += blockDim.x) {
        const uint32_t v = h[i];
        if (v) atomicAdd(&ctl->hist[0][0] + i, v);
    }
}

// one CTA of kMaxRadix threads
__global__ void __launch_bounds__(kMaxRadix) k_sort_scan(SortControl* ctl, uint64_t n, int n_passes)
{
    __shared__ uint32_t warp_sum[kMaxRadix / 32];
    __shared__ uint32_t s_skip[kMaxPasses];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    for (int p = 0; p < n_passes; ++p) {
        const uint32_t v = ctl->hist[p][tid];
        const int all_in_one = __syncthreads_or(n > 0 && (uint64_t)v == n);
        uint32_t x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_sum[wid] = x;
        __syncthreads();
        uint32_t base = 0;
        for (int w = 0; w < wid; ++w) base += warp_sum[w];
        ctl->hist[p][tid] = base + x - v;  // exclusive bin offset
        if (tid == 0) s_skip[p] = (all_in_one || n == 0) ? 1u : 0u;
        __syncthreads();
    }
    if (tid == 0) {
        uint32_t sel = 0, run = 0;
        for (int p = 0; p < n_passes; ++p) {
            ctl->skip[p] = s_skip[p];
            ctl->src_sel[p] = sel;
            if (!s_skip[p]) {
                sel ^= 1u;
                ++run;
            }
        }
        ctl->src_sel[n_passes] = sel;
        ctl->passes_run = run;
    }
}

// dynamic shared memory of k_onesweep:
//   uint64 keys[kSortTile] | uint32 vals[kSortTile] (pairs mode only) | uint32 warp_hist[kSortWarps][bins]
//   | uint32 off[bins] | uint32 dbase[bins] | uint32 warp_sum[kSortWarps] | uint32 tile
inline size_t onesweep_smem_bytes(int bins, bool has_vals)
{
    return (size_t)kSortTile * 8 + (has_vals ? (size_t)kSortTile * 4 : 0) + (size_t)(kSortWarps + 2) * bins * 4 +
           (kSortWarps + 4) * 4;
}

template <bool HAS_VALS>
__global__ void __launch_bounds__(kSortThreads, HAS_VALS ? 1 : 2)
k_onesweep(uint64_t* keys_a, uint32_t* vals_a, uint64_t* keys_b, uint32_t* vals_b, uint64_t n, int shift, int bits,
           int pass, SortControl* ctl, uint32_t* lookback)
{
    if (ctl->skip[pass]) return;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int bins = 1 << bits;
    uint64_t* s_keys = reinterpret_cast<uint64_t*>(smem_raw);
    uint32_t* s_vals = reinterpret_cast<uint32_t*>(smem_raw + (size_t)kSortTile * 8);
    uint32_t* s_warp_hist = s_vals + (HAS_VALS ? kSortTile : 0);  // [kSortWarps][bins]
    uint32_t* s_off = s_warp_hist + kSortWarps * bins;             // global destination of local position p is off[d] + p
    uint32_t* s_dbase = s_off + bins;                              // first local position of each digit
    uint32_t* s_warp_sum = s_dbase + bins;
    uint32_t* s_tile = s_warp_sum + kSortWarps;

    const bool a_to_b = ctl->src_sel[pass] == 0;
    const uint64_t* __restrict__ src_k = a_to_b ? keys_a : keys_b;
    const uint32_t* __restrict__ src_v = a_to_b ? vals_a : vals_b;
    uint64_t* __restrict__ dst_k = a_to_b ? keys_b : keys_a;
    uint32_t* __restrict__ dst_v = a_to_b ? vals_b : vals_a;

    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t dmask = (uint32_t)bins - 1u;

    if (tid == 0) *s_tile = atomicAdd(&ctl->ticket[pass], 1u);
    for (int i = tid; i < kSortWarps * bins; i += kSortThreads) s_warp_hist[i] = 0;
    __syncthreads();
    const uint32_t tile = *s_tile;
    const uint64_t tile_base = (uint64_t)tile * kSortTile;
    const uint32_t n_valid = (uint32_t)min((uint64_t)kSortTile, n - tile_base);

    // warp-striped load: element (wid, j, lane); padding keys are all ones and sort to the tile's end
    uint64_t key[kSortItems];
    const uint32_t warp_base = wid * (kSortItems * 32);
#pragma unroll
    for (int j = 0; j < kSortItems; ++j) {
        const uint32_t li = warp_base + j * 32 + lane;
        key[j] = li < n_valid ? __ldg(src_k + tile_base + li) : ~0ull;
    }

    // rank inside the warp, in element order (j major, lane minor) => stable
    uint16_t rank[kSortItems];
    const uint32_t lt_mask = (1u << lane) - 1u;
    uint32_t* my_hist = s_warp_hist + wid * bins;
#pragma unroll
    for (int j = 0; j < kSortItems; ++j) {
        const uint32_t d = digit_of(key[j], shift, dmask);
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const uint32_t pre = my_hist[d];
        __syncwarp();
        if ((peers & lt_mask) == 0) my_hist[d] = pre + __popc(peers);
        __syncwarp();
        rank[j] = (uint16_t)(pre + __popc(peers & lt_mask));
    }
    __syncthreads();

    // per digit: exclusive offsets of the warps, tile count, look-back.  bins is 2^k: with more bins than
    // threads every thread owns bins/threads consecutive digits.
    const int per_thread = bins > kSortThreads ? bins / kSortThreads : 1;
    const int d0 = tid * per_thread;
    uint32_t cnt[kMaxRadix / kSortThreads], excl[kMaxRadix / kSortThreads];
    uint32_t mine = 0;
#pragma unroll
    for (int k = 0; k < kMaxRadix / kSortThreads; ++k) {
        cnt[k] = 0;
        excl[k] = 0;
        const int d = d0 + k;
        if (k < per_thread && d < bins) {
            uint32_t c = 0;
#pragma unroll
            for (int w = 0; w < kSortWarps; ++w) {
                const uint32_t t = s_warp_hist[w * bins + d];
                s_warp_hist[w * bins + d] = c;
                c += t;
            }
            cnt[k] = c;
            mine += c;
            volatile uint32_t* lb = lookback + (size_t)tile * bins;
            if (tile == 0) {
                lb[d] = kFlagPrefix | c;
            } else {
                lb[d] = kFlagAgg | c;
            }
        }
    }
    if (tile != 0) {
#pragma unroll
        for (int k = 0; k < kMaxRadix / kSortThreads; ++k) {
            const int d = d0 + k;
            if (k < per_thread && d < bins) {
                const uint32_t e = lookback_exclusive(lookback + d, (size_t)bins, (int64_t)tile);
                excl[k] = e;
                *(volatile uint32_t*)(lookback + (size_t)tile * bins + d) = kFlagPrefix | (e + cnt[k]);
            }
        }
    }
    {
        // exclusive scan of the tile's digit counts (thread totals first, then inside the thread)
        uint32_t x = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) s_warp_sum[wid] = x;
        __syncthreads();
        uint32_t base = 0;
        for (int w = 0; w < wid; ++w) base += s_warp_sum[w];
        uint32_t dbase = base + x - mine;
#pragma unroll
        for (int k = 0; k < kMaxRadix / kSortThreads; ++k) {
            const int d = d0 + k;
            if (k < per_thread && d < bins) {
                s_dbase[d] = dbase;
                s_off[d] = ctl->hist[pass][d] + excl[k] - dbase;
                dbase += cnt[k];
            }
        }
    }
    __syncthreads();

    // scatter into shared memory in sorted-by-digit order
#pragma unroll
    for (int j = 0; j < kSortItems; ++j) {
        const uint32_t li = warp_base + j * 32 + lane;
        const uint32_t d = digit_of(key[j], shift, dmask);
        const uint32_t pos = s_dbase[d] + my_hist[d] + rank[j];
        s_keys[pos] = key[j];
        if (HAS_VALS) s_vals[pos] = li < n_valid ? __ldg(src_v + tile_base + li) : 0u;
    }
    __syncthreads();

    for (uint32_t p = tid; p < n_valid; p += kSortThreads) {
        const uint64_t k = s_keys[p];
        const uint32_t dst = s_off[digit_of(k, shift, dmask)] + p;
        dst_k[dst] = k;
        if (HAS_VALS) dst_v[dst] = s_vals[p];
    }
}

}  // namespace

SortPlan sort_plan(int begin_bit, int end_bit)
{
    SortPlan pl{};
    const int total = end_bit - begin_bit;
    if (total <= 0) return pl;
    if (total <= 3 * kMaxRadixBits) {  // narrow: 9-bit digits, spread evenly
        pl.n_passes = (total + kMaxRadixBits - 1) / kMaxRadixBits;
        const int base = total / pl.n_passes, rem = total % pl.n_passes;
        int shift = begin_bit;
        for (int p = 0; p < pl.n_passes; ++p) {
            pl.bits[p] = base + (p < rem ? 1 : 0);
            pl.shift[p] = shift;
            shift += pl.bits[p];
        }
    } else {  // wide: 8-bit digits
        pl.n_passes = (total + 7) / 8;
        for (int p = 0; p < pl.n_passes; ++p) {
            pl.shift[p] = begin_bit + 8 * p;
            pl.bits[p] = std::min(8, end_bit - pl.shift[p]);
        }
    }
    return pl;
}

// look-back words per tile over all passes never exceed 8 * 256 = 2048 (3 * 512 in the narrow plan)
constexpr size_t kLookbackWordsPerTile = 2048;

size_t sort_scratch_bytes(uint64_t n)
{
    return kCtlBytes + num_tiles(n ? n : 1) * kLookbackWordsPerTile * sizeof(uint32_t);
}

cudaError_t sort_prepare(void* scratch, uint64_t n, int begin_bit, int end_bit, cudaStream_t st)
{
    const SortPlan pl = sort_plan(begin_bit, end_bit);
    const uint64_t tiles = num_tiles(n ? n : 1);
    size_t words = 0;
    for (int p = 0; p < pl.n_passes; ++p) words += (size_t)1 << pl.bits[p];
    // the control block and the look-back words of the passes about to run must start at zero
    return cudaMemsetAsync(scratch, 0, kCtlBytes + tiles * words * sizeof(uint32_t), st);
}

cudaError_t sort_async(uint64_t* keys_a, uint32_t* vals_a, uint64_t* keys_b, uint32_t* vals_b, uint64_t n,
                       int begin_bit, int end_bit, void* scratch, cudaStream_t st, int* launches, bool prepared_with_hist)
{
    const bool has_vals = vals_a != nullptr;
    const SortPlan pl = sort_plan(begin_bit, end_bit);
    const uint64_t tiles = num_tiles(n ? n : 1);
    cudaError_t e = cudaSuccess;
    if (!prepared_with_hist) {
        e = sort_prepare(scratch, n, begin_bit, end_bit, st);
        if (e != cudaSuccess) return e;
    }
    SortControl* ctl = reinterpret_cast<SortControl*>(scratch);
    uint32_t* lookback = reinterpret_cast<uint32_t*>(static_cast<unsigned char*>(scratch) + kCtlBytes);
    int nl = 0;
    if (n > 0 && pl.n_passes > 0 && !prepared_with_hist) {
        HistArgs ha{};
        ha.n_passes = pl.n_passes;
        for (int p = 0; p < pl.n_passes; ++p) {
            ha.shift[p] = pl.shift[p];
            ha.bits[p] = pl.bits[p];
        }
        const uint64_t hist_want = (n + 256 * 8 - 1) / (256 * 8);
        const int hist_grid = (int)(hist_want < 148 * 8 ? hist_want : 148 * 8);
        k_sort_hist<<<hist_grid, 256, (size_t)pl.n_passes * kMaxRadix * 4, st>>>(keys_a, n, ha, ctl);
        ++nl;
    }
    k_sort_scan<<<1, kMaxRadix, 0, st>>>(ctl, n, pl.n_passes);
    ++nl;
    if (n > 0) {
        size_t lb_off = 0;
        for (int p = 0; p < pl.n_passes; ++p) {
            const int bins = 1 << pl.bits[p];
            const size_t smem = onesweep_smem_bytes(bins, has_vals);
            // per device and cheap: set on every call rather than cache a flag a second GPU would miss
            if (has_vals) {
                e = cudaFuncSetAttribute(k_onesweep<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
                if (e != cudaSuccess) return e;
                k_onesweep<true><<<(unsigned)tiles, kSortThreads, smem, st>>>(keys_a, vals_a, keys_b, vals_b, n,
                                                                              pl.shift[p], pl.bits[p], p, ctl,
                                                                              lookback + lb_off);
            } else {
                e = cudaFuncSetAttribute(k_onesweep<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
                if (e != cudaSuccess) return e;
                k_onesweep<false><<<(unsigned)tiles, kSortThreads, smem, st>>>(keys_a, nullptr, keys_b, nullptr, n,
                                                                               pl.shift[p], pl.bits[p], p, ctl,
                                                                               lookback + lb_off);
            }
            lb_off += tiles * bins;
            ++nl;
        }
    }
    if (launches) *launches += nl;
    return cudaGetLastError();
}

cudaError_t sort_pairs_async(uint64_t* keys_a, uint32_t* vals_a, uint64_t* keys_b, uint32_t* vals_b, uint64_t n,
                             int begin_bit, int end_bit, void* scratch, cudaStream_t st, int* launches)
{
    return sort_async(keys_a, vals_a, keys_b, vals_b, n, begin_bit, end_bit, scratch, st, launches, false);
}

}  // namespace ss

// ---- the sort without any allocation -------------------------------------------------------------
//
// Some hosts of this library are not allowed to allocate device memory when they need a sort: a
// render thread inside a frame whose memory budget was fixed at start-up, a plug-in running under a
// host application that owns the allocator, code that runs after the engine has handed all of HBM
// to its own pools.  For those callers the scratch that ss_sort_pairs takes from the caller lives
// here, in a fixed array inside the library image.  Every GPU that loads libscenesubmit.so carries
// this array (about 1 MiB) until the process exits, which is why its size is capped.
constexpr size_t kStaticSortScratchBytes =
    ss::kCtlBytes + (size_t)((SS_SORT_STATIC_MAX_N + ss::kSortTile - 1) / ss::kSortTile) * ss::kLookbackWordsPerTile *
                        sizeof(uint32_t);

static __device__ __align__(256) unsigned char g_static_sort_scratch[kStaticSortScratchBytes];

namespace {

constexpr int kStaticMaxDevices = 64;
void* g_static_scratch_ptr[kStaticMaxDevices];  // device address of the array, per device ordinal
bool g_static_scratch_ready[kStaticMaxDevices];
std::mutex g_static_scratch_mutex;

// In host code `g_static_sort_scratch` is only a handle for the symbol, not an address a kernel can
// use; the device address has to be asked for, once per device, with that device current.
cudaError_t static_scratch_for_current_device(void** out)
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kStaticMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_static_scratch_mutex);
    if (!g_static_scratch_ready[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_sort_scratch);
        if (e != cudaSuccess) return e;
        g_static_scratch_ptr[dev] = p;
        g_static_scratch_ready[dev] = true;
    }
    *out = g_static_scratch_ptr[dev];
    return cudaSuccess;
}

}  // namespace

extern "C" SS_API ss_status ss_sort_pairs_static(void* cuda_stream, uint64_t* keys_a, uint32_t* vals_a,
                                                 uint64_t* keys_b, uint32_t* vals_b, uint64_t n, int begin_bit,
                                                 int end_bit, int* sorted_in_b)
{
    if (!keys_a || !vals_a || !keys_b || !vals_b || begin_bit < 0 || end_bit > 64 || begin_bit > end_bit) {
        ss::set_global_error("ss_sort_pairs_static: invalid argument (null buffer or bad bit range)");
        return SS_ERR_INVALID;
    }
    if (n > SS_SORT_STATIC_MAX_N) {
        ss::set_global_error(
            "ss_sort_pairs_static: n exceeds SS_SORT_STATIC_MAX_N (the built-in scratch is fixed in size); "
            "use ss_sort_pairs with caller-provided scratch for larger sorts");
        return SS_ERR_CAPACITY;
    }
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    // the scratch address is looked up for the CURRENT device: a stream of another device would make the kernels
    // run where that address means nothing
    if (st != nullptr) {
        int cur = -1, owner = -1;
        if (cudaGetDevice(&cur) == cudaSuccess && cudaStreamGetDevice(st, &owner) == cudaSuccess && owner != cur) {
            ss::set_global_error("ss_sort_pairs_static: cuda_stream belongs to another device than the current one");
            return SS_ERR_INVALID;
        }
    }
    void* scratch = nullptr;
    cudaError_t e = static_scratch_for_current_device(&scratch);
    // sort_pairs_async first clears the control block and the look-back words of the passes it will run
    // (the array still holds whatever the previous call left there), then enqueues the kernels.
    if (e == cudaSuccess)
        e = ss::sort_pairs_async(keys_a, vals_a, keys_b, vals_b, n, begin_bit, end_bit, scratch, st, nullptr);
    uint32_t sel = 0;
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(&sel, &static_cast<ss::SortControl*>(scratch)->src_sel[ss::sort_num_passes(begin_bit, end_bit)],
                            sizeof sel, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) {
        ss::set_global_error((std::string("ss_sort_pairs_static: ") + cudaGetErrorString(e)).c_str());
        return SS_ERR_CUDA;
    }
    if (sorted_in_b) *sorted_in_b = (int)sel;
    return SS_OK;
}
