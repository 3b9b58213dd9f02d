// Merged draw lists of several shards, behind the C ABI: ss_merge_create / ss_merge_views / ss_merge_map.
//
// Entity-wise shards are contiguous and in rank order, so the single-GPU stable order of a view is "run of key k: rank
// 0's elements, rank 1's, ...": merging is run placement, not a sort.  Everything runs on the device, in four launches
// for ALL views, with no host synchronisation and no library besides the CUDA runtime:
//   k_merge_fetch     every rank's draws, draw keys and counts of every view into local scratch (peer reads over
//                     NVLink when the pointers are peer mappings; plain coalesced loads, a few hundred KB)
//   k_merge_rank      each (rank, draw) finds its place in the union of the ranks' sorted draw-key lists by bisection
//                     (a K-way merge without a heap): segments in (key, rank) order
//   k_merge_segments  one CTA per view: exclusive prefix of the segment lengths, merged indirect draws per key
//   k_merge_place     persistent warps over the merged list: every output position finds its segment and reads its
//                     element from the owning rank's list WHERE IT LIES (peer memory), adding the rank's instance base
// The reference has one list on one CPU (VC/Scene/SceneRenderer.cpp:1336,1368,1381); this has no counterpart there.
#include <algorithm>
#include <new>
#include <string>
#include <vector>

#include "ss_kernels.h"

namespace ss {
namespace {

constexpr int kMergeMaxRanks = 16;

struct MergeTables {          // device scratch + outputs, [view] major
    uint32_t n_ranks, n_views, max_draws, cap_visible;
    // fetched per (view, rank): [v][r][max_draws]
    uint64_t* keys;
    ss_draw_cmd* draws;
    uint32_t* n_draws;        // [v][r]
    // segments per view: [v][n_ranks * max_draws], in (key, rank) order
    uint64_t* seg_key;
    uint32_t* seg_rank;
    uint32_t* seg_src;        // offset inside the rank's instance list
    uint32_t* seg_cnt;
    uint32_t* seg_out;        // exclusive prefix of seg_cnt, +1 entry
    ss_draw_cmd* seg_draw;    // the rank's draw record (index_count / first_index / vertex_offset)
    uint32_t* n_segs;         // [v]
    // outputs per view
    uint32_t* out_idx;        // [v][cap_visible]
    uint64_t* out_keys;       // [v][cap_visible]
    ss_draw_cmd* out_draws;   // [v][n_ranks * max_draws] (at most one per distinct key)
    uint64_t* out_draw_keys;
    uint32_t* out_counts;     // [v][2] {visible, draws}
    uint32_t* overflow;
};

struct PeerTable {
    ss_merge_peer p[SS_MAX_VIEWS][kMergeMaxRanks];
};

__global__ void __launch_bounds__(256) k_merge_fetch(const __grid_constant__ PeerTable pt, const __grid_constant__ MergeTables t)
{
    const uint32_t v = blockIdx.y, r = blockIdx.z;
    const ss_merge_peer& p = pt.p[v][r];
    const uint32_t nd = min(p.counts[1], t.max_draws);
    if (blockIdx.x == 0 && threadIdx.x == 0) t.n_draws[v * t.n_ranks + r] = nd;
    const size_t base = ((size_t)v * t.n_ranks + r) * t.max_draws;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nd; i += gridDim.x * blockDim.x) {
        t.keys[base + i] = p.draw_keys[i];
        t.draws[base + i] = p.draws[i];
    }
}

__device__ __forceinline__ uint32_t lower_bound_u64(const uint64_t* a, uint32_t n, uint64_t k)
{
    uint32_t lo = 0, hi = n;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (a[mid] < k) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// element (r, i) of view v -> its position among all (key, rank) pairs of the view
__global__ void __launch_bounds__(256) k_merge_rank(const __grid_constant__ MergeTables t)
{
    const uint32_t v = blockIdx.y;
    const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t r = e / t.max_draws, i = e - r * t.max_draws;
    if (r >= t.n_ranks) return;
    const uint32_t* nd = t.n_draws + v * t.n_ranks;
    if (i >= nd[r]) return;
    const size_t vb = (size_t)v * t.n_ranks * t.max_draws;
    const uint64_t key = t.keys[vb + (size_t)r * t.max_draws + i];
    uint32_t pos = i;  // own keys below this one (keys of one rank are distinct and ascending)
    for (uint32_t q = 0; q < t.n_ranks; ++q) {
        if (q == r) continue;
        const uint64_t* a = t.keys + vb + (size_t)q * t.max_draws;
        // ranks before r come first among equal keys: count their keys <= key, of later ranks the keys < key
        pos += q < r ? lower_bound_u64(a, nd[q], key + 1) : lower_bound_u64(a, nd[q], key);
    }
    const ss_draw_cmd d = t.draws[vb + (size_t)r * t.max_draws + i];
    const size_t o = vb + pos;
    t.seg_key[o] = key;
    t.seg_rank[o] = r;
    t.seg_src[o] = d.first_instance;
    t.seg_cnt[o] = d.instance_count;
    t.seg_draw[o] = d;
}

// one CTA per view: seg_out = exclusive prefix of seg_cnt; one merged draw per distinct key
constexpr int kSegThreads = 1024;
__global__ void __launch_bounds__(kSegThreads) k_merge_segments(const __grid_constant__ MergeTables t)
{
    __shared__ uint32_t warp_c[32], warp_h[32];
    __shared__ uint32_t carry_c, carry_h;
    const uint32_t v = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const size_t vb = (size_t)v * t.n_ranks * t.max_draws;
    uint32_t n_segs = 0;
    for (uint32_t q = 0; q < t.n_ranks; ++q) n_segs += t.n_draws[v * t.n_ranks + q];
    if (tid == 0) {
        carry_c = 0;
        carry_h = 0;
        t.n_segs[v] = n_segs;
    }
    __syncthreads();
    constexpr int kIt = 8;  // consecutive segments per thread and step
    for (uint32_t s0 = 0; s0 < n_segs; s0 += kSegThreads * kIt) {
        const uint32_t first = s0 + tid * kIt;
        uint32_t c[kIt];
        uint64_t key[kIt];
        bool head[kIt];
        uint64_t prev_key = (first > 0 && first <= n_segs) ? t.seg_key[vb + first - 1] : 0ull;
        uint32_t tc = 0, th = 0;
#pragma unroll
        for (int j = 0; j < kIt; ++j) {
            const uint32_t s = first + j;
            const bool ok = s < n_segs;
            c[j] = ok ? t.seg_cnt[vb + s] : 0u;
            key[j] = ok ? t.seg_key[vb + s] : 0ull;
            head[j] = ok && (s == 0 || key[j] != prev_key);
            prev_key = key[j];
            tc += c[j];
            th += head[j] ? 1u : 0u;
        }
        uint32_t xc = tc, xh = th;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t yc = __shfl_up_sync(0xffffffffu, xc, o), yh = __shfl_up_sync(0xffffffffu, xh, o);
            if (lane >= o) {
                xc += yc;
                xh += yh;
            }
        }
        if (lane == 31) {
            warp_c[wid] = xc;
            warp_h[wid] = xh;
        }
        __syncthreads();
        uint32_t bc = carry_c, bh = carry_h;
        for (int w = 0; w < wid; ++w) {
            bc += warp_c[w];
            bh += warp_h[w];
        }
        uint32_t excl_c = bc + xc - tc, draw_no = bh + xh - th;
#pragma unroll
        for (int j = 0; j < kIt; ++j) {
            const uint32_t s = first + j;
            if (s < n_segs) {
                t.seg_out[(size_t)v * (t.n_ranks * t.max_draws + 1) + s] = excl_c;
                if (head[j]) {  // the merged draw of this key: its instances start here; the count is filled below
                    ss_draw_cmd d = t.seg_draw[vb + s];
                    d.first_instance = excl_c;
                    d.instance_count = 0;
                    t.out_draws[vb + draw_no] = d;
                    t.out_draw_keys[vb + draw_no] = key[j];
                    ++draw_no;
                }
                excl_c += c[j];
            }
        }
        __syncthreads();
        if (tid == kSegThreads - 1) {
            carry_c = bc + xc;
            carry_h = bh + xh;
        }
        __syncthreads();
    }
    const uint32_t total = carry_c, n_draws = carry_h;
    if (tid == 0) {
        t.seg_out[(size_t)v * (t.n_ranks * t.max_draws + 1) + n_segs] = total;
        t.out_counts[v * 2 + 0] = min(total, t.cap_visible);
        t.out_counts[v * 2 + 1] = n_draws;
        if (total > t.cap_visible) *t.overflow = 1u;
    }
    // instance counts of the merged draws: next draw's first instance minus this one's
    __syncthreads();
    for (uint32_t d = tid; d < n_draws; d += kSegThreads) {
        const uint32_t next = d + 1 < n_draws ? t.out_draws[vb + d + 1].first_instance : total;
        t.out_draws[vb + d].instance_count = next - t.out_draws[vb + d].first_instance;
    }
}

// Persistent warps over the merged list of every view (grid.y = view), kChunk = 512 consecutive positions per warp and
// step.  A warp bisects the segment prefix once per chunk; every lane then keeps the segment it is in (its end and its
// source pointer) in registers and only re-reads the tables when it walks past the segment's end.  The peer reads of a
// whole chunk are issued at once as 4-byte cp.async copies into shared memory (16 per lane in flight, no registers
// held): over NVLink with every GPU reading every other one it is the bytes in flight that set the pace (four
// register loads per lane: 2.33 ms for 187 M elements on 8 GPUs).  Then the chunk is added up and stored, coalesced.
constexpr int kPlaceRounds = 16, kPlaceChunk = 32 * kPlaceRounds, kPlaceWarps = 8;

__global__ void __launch_bounds__(kPlaceWarps * 32) k_merge_place(const __grid_constant__ PeerTable pt, const __grid_constant__ MergeTables t)
{
    __shared__ uint32_t s_val[kPlaceWarps][kPlaceChunk];
    __shared__ uint32_t s_seg[kPlaceWarps][kPlaceChunk];
    const uint32_t v = blockIdx.y;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const size_t vb = (size_t)v * t.n_ranks * t.max_draws;
    const uint32_t n_segs = t.n_segs[v];
    const uint32_t* __restrict__ seg_out = t.seg_out + (size_t)v * (t.n_ranks * t.max_draws + 1);
    const uint32_t total = min(seg_out[n_segs], t.cap_visible);
    uint32_t* __restrict__ out_idx = t.out_idx + (size_t)v * t.cap_visible;
    uint64_t* __restrict__ out_keys = t.out_keys + (size_t)v * t.cap_visible;
    const uint32_t warps = (gridDim.x * blockDim.x) >> 5, warp_id = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    uint32_t* my_val = s_val[wid];
    uint32_t* my_seg = s_seg[wid];
    for (uint64_t g0 = (uint64_t)warp_id * kPlaceChunk; g0 < total; g0 += (uint64_t)warps * kPlaceChunk) {
        uint32_t lo = 0, hi = n_segs;
        while (hi - lo > 1) {
            const uint32_t mid = (lo + hi) >> 1;
            if (seg_out[mid] <= g0) lo = mid; else hi = mid;
        }
        // the lane's current segment, cached
        uint32_t cur = lo, seg_begin = 0, seg_end = 0;
        const uint32_t* src = nullptr;
        auto load_segment = [&](uint32_t sI) {
            seg_begin = seg_out[sI];
            seg_end = seg_out[sI + 1];
            src = pt.p[v][t.seg_rank[vb + sI]].idx + t.seg_src[vb + sI];
        };
        load_segment(cur);
#pragma unroll
        for (int r = 0; r < kPlaceRounds; ++r) {
            const uint64_t e = g0 + (uint64_t)r * 32 + lane;
            if (e < total) {
                while (e >= seg_end) load_segment(++cur);  // empty segments are stepped over
                const uint32_t dst = (uint32_t)__cvta_generic_to_shared(my_val + r * 32 + lane);
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src + ((uint32_t)e - seg_begin)) : "memory");
                my_seg[r * 32 + lane] = cur;
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncwarp();
#pragma unroll 4
        for (int r = 0; r < kPlaceRounds; ++r) {
            const uint64_t e = g0 + (uint64_t)r * 32 + lane;
            if (e < total) {
                const uint32_t sI = my_seg[r * 32 + lane];
                const uint32_t ibase = (uint32_t)pt.p[v][t.seg_rank[vb + sI]].instance_base;
                __stcs(out_idx + e, my_val[r * 32 + lane] + ibase);  // written once, read by the consumer: keep them out of L2's way
                __stcs(reinterpret_cast<unsigned long long*>(out_keys + e), (unsigned long long)t.seg_key[vb + sI]);
            }
        }
        __syncwarp();  // the chunk buffer is refilled on the next step
    }
}

}  // namespace
}  // namespace ss

using namespace ss;

struct ss_merge {
    int device = 0, sms = 148;
    MergeTables t{};
    std::vector<void*> allocs;
    std::string err;
    cudaEvent_t ev[5] = {};  // around the four kernels of the last ss_merge_views
    bool timed = false;
};

extern "C" {

SS_API ss_status ss_merge_create(int device, uint32_t n_ranks, uint32_t n_views, uint32_t max_draws_per_rank,
                                 uint64_t max_visible_per_view, ss_merge** out)
{
    if (!out) return SS_ERR_INVALID;
    *out = nullptr;
    if (n_ranks < 1 || n_ranks > (uint32_t)kMergeMaxRanks || n_views < 1 || n_views > SS_MAX_VIEWS || max_draws_per_rank == 0 ||
        max_visible_per_view == 0 || max_visible_per_view >= (1ull << 32)) {
        set_global_error("ss_merge_create: 1..16 ranks, 1..8 views, non-zero capacities below 2^32");
        return SS_ERR_INVALID;
    }
    ss_merge* m = new (std::nothrow) ss_merge();
    if (!m) return SS_ERR_NOMEM;
    m->device = device;
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&m->sms, cudaDevAttrMultiProcessorCount, device);
    MergeTables& t = m->t;
    t.n_ranks = n_ranks;
    t.n_views = n_views;
    t.max_draws = max_draws_per_rank;
    t.cap_visible = (uint32_t)max_visible_per_view;
    const size_t V = n_views, RD = (size_t)n_ranks * max_draws_per_rank, C = (size_t)t.cap_visible;
    auto alloc = [&](auto** p, size_t count) {
        if (e != cudaSuccess) return;
        void* q = nullptr;
        e = cudaMalloc(&q, std::max<size_t>(count * sizeof(**p), 256));
        if (e == cudaSuccess) {
            m->allocs.push_back(q);
            *p = static_cast<std::remove_reference_t<decltype(*p)>>(q);
        }
    };
    alloc(&t.keys, V * RD);
    alloc(&t.draws, V * RD);
    alloc(&t.n_draws, V * n_ranks);
    alloc(&t.seg_key, V * RD);
    alloc(&t.seg_rank, V * RD);
    alloc(&t.seg_src, V * RD);
    alloc(&t.seg_cnt, V * RD);
    alloc(&t.seg_out, V * (RD + 1));
    alloc(&t.seg_draw, V * RD);
    alloc(&t.n_segs, V);
    alloc(&t.out_idx, V * C);
    alloc(&t.out_keys, V * C);
    alloc(&t.out_draws, V * RD);
    alloc(&t.out_draw_keys, V * RD);
    alloc(&t.out_counts, V * 2);
    alloc(&t.overflow, 1);
    if (e == cudaSuccess) e = cudaMemset(t.overflow, 0, 4);
    if (e != cudaSuccess) {
        set_global_error((std::string("ss_merge_create: ") + cudaGetErrorString(e)).c_str());
        for (void* p : m->allocs) cudaFree(p);
        delete m;
        return e == cudaErrorMemoryAllocation ? SS_ERR_NOMEM : SS_ERR_CUDA;
    }
    *out = m;
    return SS_OK;
}

SS_API void ss_merge_destroy(ss_merge* m)
{
    if (!m) return;
    cudaSetDevice(m->device);
    for (void* p : m->allocs) cudaFree(p);
    for (cudaEvent_t e : m->ev)
        if (e) cudaEventDestroy(e);
    delete m;
}

SS_API ss_status ss_merge_views(ss_merge* m, void* cuda_stream, const ss_merge_peer* peers)
{
    if (!m || !peers) return SS_ERR_INVALID;
    const MergeTables& t = m->t;
    PeerTable pt;
    memset(&pt, 0, sizeof pt);
    for (uint32_t v = 0; v < t.n_views; ++v)
        for (uint32_t r = 0; r < t.n_ranks; ++r) {
            const ss_merge_peer& p = peers[v * t.n_ranks + r];
            if (!p.idx || !p.draws || !p.draw_keys || !p.counts) {
                m->err = "ss_merge_views: null pointer in the peer table";
                return SS_ERR_INVALID;
            }
            pt.p[v][r] = p;
        }
    if (cudaSetDevice(m->device) != cudaSuccess) return SS_ERR_CUDA;
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    const uint32_t chunks = (t.max_draws + 255) / 256;
    if (!m->ev[0])
        for (cudaEvent_t& e : m->ev)
            if (cudaEventCreate(&e) != cudaSuccess) return SS_ERR_CUDA;
    cudaEventRecord(m->ev[0], st);
    k_merge_fetch<<<dim3(std::min<uint32_t>(chunks, 8), t.n_views, t.n_ranks), 256, 0, st>>>(pt, t);
    cudaEventRecord(m->ev[1], st);
    k_merge_rank<<<dim3((t.n_ranks * t.max_draws + 255) / 256, t.n_views), 256, 0, st>>>(t);
    cudaEventRecord(m->ev[2], st);
    k_merge_segments<<<t.n_views, kSegThreads, 0, st>>>(t);
    cudaEventRecord(m->ev[3], st);
    k_merge_place<<<dim3((unsigned)m->sms * 4, t.n_views), kPlaceWarps * 32, 0, st>>>(pt, t);
    cudaEventRecord(m->ev[4], st);
    m->timed = true;
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        m->err = std::string("ss_merge_views: ") + cudaGetErrorString(e);
        return SS_ERR_CUDA;
    }
    return SS_OK;
}

SS_API ss_status ss_merge_stage_ms(ss_merge* m, float ms_out[4])
{
    if (!m || !ms_out) return SS_ERR_INVALID;
    if (!m->timed) {
        m->err = "ss_merge_stage_ms: no ss_merge_views yet";
        return SS_ERR_STATE;
    }
    if (cudaSetDevice(m->device) != cudaSuccess || cudaEventSynchronize(m->ev[4]) != cudaSuccess) return SS_ERR_CUDA;
    for (int i = 0; i < 4; ++i)
        if (cudaEventElapsedTime(&ms_out[i], m->ev[i], m->ev[i + 1]) != cudaSuccess) return SS_ERR_CUDA;
    return SS_OK;
}

SS_API ss_status ss_merge_map(ss_merge* m, void* cuda_stream, uint32_t view, ss_merged_view* out)
{
    if (!m || !out || view >= m->t.n_views) return SS_ERR_INVALID;
    const MergeTables& t = m->t;
    if (cudaSetDevice(m->device) != cudaSuccess) return SS_ERR_CUDA;
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    uint32_t counts[2] = {0, 0}, over = 0;
    cudaError_t e = cudaMemcpyAsync(counts, t.out_counts + view * 2, sizeof counts, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(&over, t.overflow, sizeof over, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) {
        m->err = std::string("ss_merge_map: ") + cudaGetErrorString(e);
        return SS_ERR_CUDA;
    }
    const size_t RD = (size_t)t.n_ranks * t.max_draws;
    out->instance_idx = t.out_idx + (size_t)view * t.cap_visible;
    out->keys = t.out_keys + (size_t)view * t.cap_visible;
    out->draws = t.out_draws + (size_t)view * RD;
    out->draw_keys = t.out_draw_keys + (size_t)view * RD;
    out->visible = counts[0];
    out->n_draws = counts[1];
    if (over) {
        m->err = "merged list exceeds max_visible_per_view (truncated)";
        return SS_ERR_CAPACITY;
    }
    return SS_OK;
}

SS_API ss_status ss_merge_download(ss_merge* m, void* cuda_stream, uint32_t view, uint32_t* instance_idx, uint64_t* keys,
                                   ss_draw_cmd* draws, uint64_t* draw_keys, uint32_t* visible_out, uint32_t* draws_out)
{
    ss_merged_view o;
    const ss_status st = ss_merge_map(m, cuda_stream, view, &o);
    if (st != SS_OK) return st;
    cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
    cudaError_t e = cudaSuccess;
    if (instance_idx && o.visible) e = cudaMemcpyAsync(instance_idx, o.instance_idx, (size_t)o.visible * 4, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess && keys && o.visible) e = cudaMemcpyAsync(keys, o.keys, (size_t)o.visible * 8, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess && draws && o.n_draws)
        e = cudaMemcpyAsync(draws, o.draws, (size_t)o.n_draws * sizeof(ss_draw_cmd), cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess && draw_keys && o.n_draws) e = cudaMemcpyAsync(draw_keys, o.draw_keys, (size_t)o.n_draws * 8, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) {
        m->err = std::string("ss_merge_download: ") + cudaGetErrorString(e);
        return SS_ERR_CUDA;
    }
    if (visible_out) *visible_out = o.visible;
    if (draws_out) *draws_out = o.n_draws;
    return SS_OK;
}

SS_API const char* ss_merge_last_error(const ss_merge* m) { return m ? m->err.c_str() : ""; }

}  // extern "C"
