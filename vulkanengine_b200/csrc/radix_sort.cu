// Hand-written onesweep LSD radix sort of (uint64 key, uint32 value) pairs for sm_100a.
// No CUB / Thrust.  This is the device replacement of the grouping std::map<MeshKey, ...> performs
// on the CPU (VC/Scene/SceneRenderer.h:252-253, insert at SceneRenderer.cpp:1146-1148): a stable sort
// by key keeps entt iteration order inside each group (SURVEY.md §7 hard part 4).
//
// Structure (Adinets & Merrill, "Onesweep", 2022):
//   k_sort_hist  one read of the keys, a 256-bin histogram for every digit position at once
//   k_sort_scan  exclusive bin offsets per digit; marks digits that are constant over all keys so
//                their pass is skipped (narrow key layouts leave most of the 64 bits constant)
//   k_onesweep   per digit: each CTA ranks a tile with warp match/ballot, publishes its per-digit counts,
//                resolves its global offsets by decoupled look-back over earlier tiles, and scatters
//                through shared memory so the global stores are coalesced per digit run.
// HBM-bound: each executed pass reads and writes 12 B per element once.
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

__global__ void __launch_bounds__(256) k_sort_hist(const uint64_t* __restrict__ keys, uint64_t n, int begin_bit,
                                                   int end_bit, int n_passes, SortControl* ctl)
{
    __shared__ uint32_t h[kMaxPasses][kRadix];
    for (int i = threadIdx.x; i < n_passes * kRadix; i += blockDim.x) (&h[0][0])[i] = 0;
    __syncthreads();
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const uint64_t k = __ldg(keys + i);
        for (int p = 0; p < n_passes; ++p) {
            const int shift = begin_bit + p * kRadixBits;
            const int bits = min(kRadixBits, end_bit - shift);
            atomicAdd(&h[p][digit_of(k, shift, (1u << bits) - 1u)], 1u);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < n_passes * kRadix; i += blockDim.x) {
        const uint32_t v = (&h[0][0])[i];
        if (v) atomicAdd(&ctl->hist[0][0] + i, v);
    }
}

// one CTA of kRadix threads
__global__ void __launch_bounds__(kRadix) k_sort_scan(SortControl* ctl, uint64_t n, int n_passes)
{
    __shared__ uint32_t warp_sum[kRadix / 32];
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

struct OnesweepSmem {
    uint64_t keys[kSortTile];
    uint32_t vals[kSortTile];
    uint32_t warp_hist[kSortWarps][kRadix];
    uint32_t off[kRadix];    // global destination of local position p is off[digit] + p
    uint32_t dbase[kRadix];  // first local position of each digit
    uint32_t warp_sum[kSortWarps];
    uint32_t tile;
};

__global__ void __launch_bounds__(kSortThreads)
k_onesweep(const uint64_t* __restrict__ keys_a, const uint32_t* __restrict__ vals_a, uint64_t* __restrict__ keys_b,
           uint32_t* __restrict__ vals_b, uint64_t n, int shift, int bits, int pass, SortControl* ctl,
           uint32_t* lookback)
{
    if (ctl->skip[pass]) return;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    OnesweepSmem& s = *reinterpret_cast<OnesweepSmem*>(smem_raw);

    const bool a_to_b = ctl->src_sel[pass] == 0;
    const uint64_t* src_k = a_to_b ? keys_a : keys_b;
    const uint32_t* src_v = a_to_b ? vals_a : vals_b;
    uint64_t* dst_k = a_to_b ? keys_b : const_cast<uint64_t*>(keys_a);
    uint32_t* dst_v = a_to_b ? vals_b : const_cast<uint32_t*>(vals_a);

    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t dmask = (1u << bits) - 1u;

    if (tid == 0) s.tile = atomicAdd(&ctl->ticket[pass], 1u);
    for (int i = tid; i < kSortWarps * kRadix; i += kSortThreads) (&s.warp_hist[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = s.tile;
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
    uint32_t* my_hist = s.warp_hist[wid];
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

    // per digit: exclusive offsets of the warps, tile count, look-back
    uint32_t count = 0;
    {
        const int d = tid;  // kSortThreads == kRadix
#pragma unroll
        for (int w = 0; w < kSortWarps; ++w) {
            const uint32_t t = s.warp_hist[w][d];
            s.warp_hist[w][d] = count;
            count += t;
        }
        volatile uint32_t* lb = lookback + (size_t)tile * kRadix;
        uint32_t excl = 0;
        if (tile == 0) {
            lb[d] = kFlagPrefix | count;
        } else {
            lb[d] = kFlagAgg | count;
            int64_t t = (int64_t)tile - 1;
            while (true) {
                const uint32_t v = *(volatile uint32_t*)(lookback + (size_t)t * kRadix + d);
                if (v & kFlagPrefix) {
                    excl += v & kValueMask;
                    break;
                }
                if (v & kFlagAgg) {
                    excl += v & kValueMask;
                    --t;
                }
            }
            lb[d] = kFlagPrefix | (excl + count);
        }
        // exclusive scan of the tile's digit counts
        uint32_t x = count;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) s.warp_sum[wid] = x;
        __syncthreads();
        uint32_t base = 0;
        for (int w = 0; w < wid; ++w) base += s.warp_sum[w];
        const uint32_t dbase = base + x - count;
        s.dbase[d] = dbase;
        s.off[d] = ctl->hist[pass][d] + excl - dbase;
    }
    __syncthreads();

    // scatter into shared memory in sorted-by-digit order
#pragma unroll
    for (int j = 0; j < kSortItems; ++j) {
        const uint32_t li = warp_base + j * 32 + lane;
        const uint32_t d = digit_of(key[j], shift, dmask);
        const uint32_t pos = s.dbase[d] + my_hist[d] + rank[j];
        s.keys[pos] = key[j];
        s.vals[pos] = li < n_valid ? __ldg(src_v + tile_base + li) : 0u;
    }
    __syncthreads();

    for (uint32_t p = tid; p < n_valid; p += kSortThreads) {
        const uint64_t k = s.keys[p];
        const uint32_t dst = s.off[digit_of(k, shift, dmask)] + p;
        dst_k[dst] = k;
        dst_v[dst] = s.vals[p];
    }
}

}  // namespace

size_t sort_scratch_bytes(uint64_t n)
{
    return kCtlBytes + (size_t)kMaxPasses * num_tiles(n ? n : 1) * kRadix * sizeof(uint32_t);
}

static cudaError_t sort_reset_passes(void* scratch, uint64_t n, int n_passes, cudaStream_t st)
{
    const size_t bytes = kCtlBytes + (size_t)n_passes * num_tiles(n ? n : 1) * kRadix * sizeof(uint32_t);
    return cudaMemsetAsync(scratch, 0, bytes, st);
}

cudaError_t sort_reset(void* scratch, uint64_t n, cudaStream_t st) { return sort_reset_passes(scratch, n, kMaxPasses, st); }

cudaError_t sort_pairs_async(uint64_t* keys_a, uint32_t* vals_a, uint64_t* keys_b, uint32_t* vals_b, uint64_t n,
                             int begin_bit, int end_bit, void* scratch, cudaStream_t st, int* launches)
{
    // per device and cheap: set on every call rather than cache a flag that a second GPU would miss
    cudaError_t e = cudaFuncSetAttribute(k_onesweep, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)sizeof(OnesweepSmem));
    if (e != cudaSuccess) return e;
    const int n_passes = sort_num_passes(begin_bit, end_bit);
    e = sort_reset_passes(scratch, n, n_passes, st);
    if (e != cudaSuccess) return e;
    SortControl* ctl = reinterpret_cast<SortControl*>(scratch);
    uint32_t* lookback = reinterpret_cast<uint32_t*>(static_cast<unsigned char*>(scratch) + kCtlBytes);
    const uint64_t tiles = num_tiles(n);
    int nl = 0;
    if (n > 0 && n_passes > 0) {
        const uint64_t hist_want = (n + 256 * 8 - 1) / (256 * 8);
        const int hist_grid = (int)(hist_want < 148 * 8 ? hist_want : 148 * 8);
        k_sort_hist<<<hist_grid, 256, 0, st>>>(keys_a, n, begin_bit, end_bit, n_passes, ctl);
        ++nl;
    }
    k_sort_scan<<<1, kRadix, 0, st>>>(ctl, n, n_passes);
    ++nl;
    if (n > 0) {
        for (int p = 0; p < n_passes; ++p) {
            const int shift = begin_bit + p * kRadixBits;
            const int bits = min(kRadixBits, end_bit - shift);
            k_onesweep<<<(unsigned)tiles, kSortThreads, sizeof(OnesweepSmem), st>>>(
                keys_a, vals_a, keys_b, vals_b, n, shift, bits, p, ctl, lookback + (size_t)p * tiles * kRadix);
            ++nl;
        }
    }
    if (launches) *launches += nl;
    return cudaGetLastError();
}

}  // namespace ss

// ---- the sort without any allocation -------------------------------------------------------------
//
// Some hosts of this library are not allowed to allocate device memory when they need a sort: a
// render thread inside a frame whose memory budget was fixed at start-up, a plug-in running under a
// host application that owns the allocator, code that runs after the engine has handed all of HBM
// to its own pools.  For those callers the scratch that ss_sort_pairs takes from the caller lives
// here, in a fixed array inside the library image.  Every GPU that loads libscenesubmit.so carries
// this array (about 2 MiB) until the process exits, which is why its size is capped.
constexpr size_t kStaticSortScratchBytes =
    ss::kCtlBytes + (size_t)ss::kMaxPasses * ((SS_SORT_STATIC_MAX_N + ss::kSortTile - 1) / ss::kSortTile) * ss::kRadix *
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
    void* scratch = nullptr;
    cudaError_t e = static_scratch_for_current_device(&scratch);
    // sort_pairs_async first clears the control block and the look-back words of the passes it will run
    // (the array still holds whatever the previous call left there), then enqueues the kernels.
    if (e == cudaSuccess)
        e = ss::sort_pairs_async(keys_a, vals_a, keys_b, vals_b, n, begin_bit, end_bit, scratch, st, nullptr);
    uint32_t sel = 0;
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(&sel,
                            &static_cast<ss::SortControl*>(scratch)->src_sel[ss::sort_num_passes(begin_bit, end_bit)],
                            sizeof sel, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) {
        ss::set_global_error((std::string("ss_sort_pairs_static: ") + cudaGetErrorString(e)).c_str());
        return SS_ERR_CUDA;
    }
    if (sorted_in_b) *sorted_in_b = (int)sel;
    return SS_OK;
}
