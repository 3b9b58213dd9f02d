// Per-view stream compaction + instance-row gather + indirect-draw emission, sm_100a.
//
// Input: the (key, instance) list sorted by key (stable), the visibility byte of every instance and
// the world rows.  One pass over the sorted list serves every view at once:
//   - __ballot/__popc-free packed counting: the per-view visibility bits of an item are spread into
//     16-bit lanes of two 64-bit words, so ONE block-wide add-scan ranks the item in all views;
//   - decoupled look-back carries the running per-view counts (and the running number of key runs)
//     from tile to tile, so the output position of every visible item is known inside its tile;
//   - the 48 B TransformData rows (VC/Core/Components.h:95-98) are gathered through a shared-memory
//     index list so that the global stores are coalesced 128-bit writes in draw order;
//   - runs of equal keys become draws: per (view, run) visible counts are accumulated with one
//     atomic per run per tile, and k_draws turns them into VkDrawIndexedIndirectCommand records
//     (field meanings: Platform/Vulkan/VulkanRenderer.cpp:628) in MeshKey order.
#include <algorithm>

#include "ss_kernels.h"

namespace ss {

namespace {

constexpr uint32_t kFlagAgg = 1u << 30;
constexpr uint32_t kFlagPrefix = 1u << 31;
constexpr uint32_t kValueMask = kFlagAgg - 1;
constexpr int kEmitWarps = kEmitThreads / 32;
constexpr int kGatherBatch = 6;  // independent 16 B row gathers in flight per thread; with 64 registers four CTAs share an SM

struct EmitSmem {
    uint64_t key[kEmitTile];
    uint32_t idx[kEmitTile];
    union {  // a tile uses one of the two output orders
        uint16_t list[SS_MAX_VIEWS][kEmitTile];      // view-major: local item ids of the visible items, per view
        struct {
            uint64_t pos_a[kEmitTile], pos_b[kEmitTile];  // item-major: visible-before counts of views 0-3 / 4-7, 16 bits each
        };
    };
    uint8_t msk[kEmitTile];                          // per item: visibility byte
    uint16_t seg_start[kEmitTile + 2][SS_MAX_VIEWS]; // visible-before count at the start of each local run
    uint64_t last_key[kEmitThreads];
    uint64_t warp_a[kEmitWarps], warp_b[kEmitWarps];
    uint32_t warp_h[kEmitWarps];
    uint32_t base[kEmitWords];   // exclusive prefix over earlier tiles
    uint32_t total[kEmitWords];  // this tile
    uint32_t tile;
};

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

__global__ void __launch_bounds__(kEmitThreads, 4) k_emit(const __grid_constant__ EmitArgs a)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    EmitSmem& s = *reinterpret_cast<EmitSmem*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;

    if (tid == 0) s.tile = atomicAdd(a.ticket, 1u);
    __syncthreads();
    const uint32_t tile = s.tile;
    const uint32_t tile_base = tile * kEmitTile;
    const uint32_t n_valid = min((uint32_t)kEmitTile, a.n - tile_base);
    const uint32_t n_tiles = (a.n + kEmitTile - 1) / kEmitTile;

    const bool in_b = a.sort_ctl && a.sort_ctl->src_sel[a.sort_passes] != 0;
    const uint64_t* __restrict__ keys = in_b ? a.keys_b : a.keys_a;
    const uint32_t* __restrict__ vals = in_b ? a.vals_b : a.vals_a;

    // blocked arrangement: thread t owns local items 4t .. 4t+3
    uint64_t key[kEmitItems];
    uint32_t idx[kEmitItems], m[kEmitItems];
    const uint32_t l0 = tid * kEmitItems;
    const bool packed = a.idx_bits != 0;
    if (l0 + kEmitItems <= n_valid) {
        const ulonglong2* kp = reinterpret_cast<const ulonglong2*>(keys + tile_base + l0);
        ulonglong2 k01 = __ldg(kp), k23 = __ldg(kp + 1);
        key[0] = k01.x; key[1] = k01.y; key[2] = k23.x; key[3] = k23.y;
        if (!packed) {
            uint4 iv = __ldg(reinterpret_cast<const uint4*>(vals + tile_base + l0));
            idx[0] = iv.x; idx[1] = iv.y; idx[2] = iv.z; idx[3] = iv.w;
        }
    } else {
#pragma unroll
        for (int j = 0; j < kEmitItems; ++j) {
            const bool ok = l0 + j < n_valid;
            key[j] = ok ? __ldg(keys + tile_base + l0 + j) : ~0ull;
            idx[j] = (ok && !packed) ? __ldg(vals + tile_base + l0 + j) : 0u;
        }
    }
    const uint32_t key_shift = a.idx_bits + a.mask_bits;
    if (packed) {  // [key | mask | instance]
        const uint64_t imask = (1ull << a.idx_bits) - 1ull;
#pragma unroll
        for (int j = 0; j < kEmitItems; ++j) {
            const bool ok = l0 + j < n_valid;
            idx[j] = ok ? (uint32_t)(key[j] & imask) : 0u;
            m[j] = (ok && !a.gather_mask && !a.seq) ? (uint32_t)(key[j] >> a.idx_bits) & 0xffu : 0u;
            key[j] = ok ? key[j] >> key_shift : ~0ull;
        }
    }
    if (a.seq) {  // masks in word order: this thread's four bytes are one aligned word
        if (l0 + kEmitItems <= n_valid) {
            const uint32_t mw = __ldg(reinterpret_cast<const uint32_t*>(a.mask + tile_base + l0));
#pragma unroll
            for (int j = 0; j < kEmitItems; ++j) m[j] = (mw >> (8 * j)) & 0xffu;
        } else {
#pragma unroll
            for (int j = 0; j < kEmitItems; ++j) m[j] = (l0 + j < n_valid) ? (uint32_t)__ldg(a.mask + tile_base + l0 + j) : 0u;
        }
    }
#pragma unroll
    for (int j = 0; j < kEmitItems; ++j) {
        if (a.gather_mask && !a.seq) m[j] = (l0 + j < n_valid) ? (uint32_t)__ldg(a.mask + idx[j]) : 0u;
        s.key[l0 + j] = key[j];
        s.idx[l0 + j] = idx[j];
    }
    s.last_key[tid] = key[kEmitItems - 1];
    __syncthreads();

    uint64_t prev;
    bool first_is_head = false;
    if (tid > 0) {
        prev = s.last_key[tid - 1];
    } else if (tile > 0) {
        prev = __ldg(keys + tile_base - 1);
        if (packed) prev >>= key_shift;
    } else {
        prev = 0;
        first_is_head = true;
    }
    bool head[kEmitItems];
    uint64_t pa = 0, pb = 0;
    uint32_t ph = 0;
#pragma unroll
    for (int j = 0; j < kEmitItems; ++j) {
        const bool ok = l0 + j < n_valid;
        head[j] = ok && ((j == 0 && first_is_head) || key[j] != (j == 0 ? prev : key[j - 1]));
        pa += spread4(m[j] & 15u);
        pb += spread4(m[j] >> 4);
        ph += head[j] ? 1u : 0u;
    }

    // block-wide inclusive scans of (pa, pb, ph)
    uint64_t xa = pa, xb = pb;
    uint32_t xh = ph;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint64_t ya = shfl_up_u64(xa, o), yb = shfl_up_u64(xb, o);
        uint32_t yh = __shfl_up_sync(0xffffffffu, xh, o);
        if (lane >= o) {
            xa += ya;
            xb += yb;
            xh += yh;
        }
    }
    if (lane == 31) {
        s.warp_a[wid] = xa;
        s.warp_b[wid] = xb;
        s.warp_h[wid] = xh;
    }
    __syncthreads();
    uint64_t ba = 0, bb = 0;
    uint32_t bh = 0;
    for (int w = 0; w < wid; ++w) {
        ba += s.warp_a[w];
        bb += s.warp_b[w];
        bh += s.warp_h[w];
    }
    // exclusive prefixes of this thread's first item
    uint64_t ea = ba + xa - pa, eb = bb + xb - pb;
    uint32_t eh = bh + xh - ph;

    if (tid == kEmitThreads - 1) {
        const uint64_t ta = ba + xa, tb = bb + xb;
#pragma unroll
        for (int v = 0; v < 4; ++v) {
            s.total[v] = (uint32_t)(ta >> (16 * v)) & 0xffffu;
            s.total[4 + v] = (uint32_t)(tb >> (16 * v)) & 0xffffu;
        }
        s.total[SS_MAX_VIEWS] = bh + xh;
    }
    __syncthreads();

    // decoupled look-back, one thread per word
    if (tid < kEmitWords) {
        const uint32_t mine = s.total[tid];
        volatile uint32_t* lb = a.lookback + (size_t)tile * kEmitWords;
        uint32_t excl = 0;
        if (tile == 0) {
            lb[tid] = kFlagPrefix | mine;
        } else {
            lb[tid] = kFlagAgg | mine;
            excl = lookback_exclusive(a.lookback + tid, (size_t)kEmitWords, (int64_t)tile);
            lb[tid] = kFlagPrefix | (excl + mine);
        }
        s.base[tid] = excl;
        if (tile == n_tiles - 1) a.totals[tid] = excl + mine;
    }

    // local lists: where each visible item lands, and where each key run starts.
    // Two output orders, chosen per tile: view-major (a dense list per view; every store instruction is a full,
    // contiguous line, but a row seen by k views is fetched k times) when items are seen by few views, item-major
    // (each row fetched once and stored to every view that sees it) when they are seen by many.
    const uint32_t n_heads = s.total[SS_MAX_VIEWS];
    uint32_t rows_out = 0;
#pragma unroll
    for (int v = 0; v < (int)SS_MAX_VIEWS; ++v) rows_out += s.total[v];
    const bool item_major = rows_out > 3u * kEmitTile;
    if (tid < (int)SS_MAX_VIEWS) {
        s.seg_start[0][tid] = 0;
        s.seg_start[n_heads + 1][tid] = (uint16_t)s.total[tid];
    }
#pragma unroll
    for (int j = 0; j < kEmitItems; ++j) {
        if (head[j]) {
            const uint32_t lseg = eh + 1;  // 1-based among runs starting in this tile; 0 = run carried in
#pragma unroll
            for (int v = 0; v < 4; ++v) {
                s.seg_start[lseg][v] = (uint16_t)(ea >> (16 * v));
                s.seg_start[lseg][4 + v] = (uint16_t)(eb >> (16 * v));
            }
        }
        if (item_major) {
            s.pos_a[l0 + j] = ea;
            s.pos_b[l0 + j] = eb;
            s.msk[l0 + j] = (uint8_t)m[j];
        } else {
#pragma unroll
            for (int v = 0; v < (int)SS_MAX_VIEWS; ++v) {
                if ((m[j] >> v) & 1u) {
                    const uint32_t lpos = (uint32_t)((v < 4 ? ea >> (16 * v) : eb >> (16 * (v - 4))) & 0xffffu);
                    s.list[v][lpos] = (uint16_t)(l0 + j);
                }
            }
        }
        ea += spread4(m[j] & 15u);
        eb += spread4(m[j] >> 4);
        eh += head[j] ? 1u : 0u;
    }
    __syncthreads();

    // run bookkeeping: first key / first instance of each run, visible count per (view, run)
    const uint32_t seg_base = s.base[SS_MAX_VIEWS];
    {
        uint32_t ehh = bh + xh - ph;
#pragma unroll
        for (int j = 0; j < kEmitItems; ++j) {
            if (head[j]) {
                const uint32_t g = seg_base + ehh;
                if (g < a.max_draws) {
                    a.seg_key[g] = key[j];
                    a.seg_inst[g] = idx[j];
                } else {
                    *a.overflow = 1u;
                }
                ++ehh;
            }
        }
    }
    for (uint32_t k = tid; k <= n_heads; k += kEmitThreads) {
        const uint32_t g = seg_base + k - 1u;  // k == 0: run that started in an earlier tile
#pragma unroll
        for (int v = 0; v < (int)SS_MAX_VIEWS; ++v) {
            if (v < (int)a.n_views) {
                const uint32_t c = (uint32_t)s.seg_start[k + 1][v] - (uint32_t)s.seg_start[k][v];
                if (c) {
                    if (g < a.max_draws) atomicAdd(a.seg_count + (size_t)v * a.max_draws + g, c);
                }
            }
        }
    }

    if (!item_major) {
        // per view: keys + instance ids, then the gathered rows, all as coalesced stores
        for (uint32_t v = 0; v < a.n_views; ++v) {
            const uint32_t cnt = s.total[v];
            const uint32_t base = s.base[v];
            const ViewOut& o = a.out[v];
            if (base + cnt > a.cap_visible && tid == 0) *a.overflow = 1u;
            for (uint32_t e = tid; e < cnt; e += kEmitThreads) {
                const uint32_t li = s.list[v][e];
                const uint32_t g = base + e;
                if (g < a.cap_visible) {
                    __stcs(reinterpret_cast<unsigned long long*>(o.keys + g), (unsigned long long)s.key[li]);
                    __stcs(o.idx + g, s.idx[li]);
                }
            }
            const uint32_t total4 = cnt * 3u;
            for (uint32_t j0 = tid; j0 < total4; j0 += kEmitThreads * kGatherBatch) {
                float4 r[kGatherBatch];
#pragma unroll
                for (int u = 0; u < kGatherBatch; ++u) {
                    const uint32_t j = j0 + u * kEmitThreads;
                    if (j < total4) {
                        const uint32_t e = j / 3u, part = j - e * 3u;
                        const uint32_t li = s.list[v][e];
                        r[u] = __ldg(a.seq ? a.world + ((size_t)tile_base + li) * a.world_stride + part
                                           : a.world + (size_t)s.idx[li] * kWorldStride + part);
                    }
                }
#pragma unroll
                for (int u = 0; u < kGatherBatch; ++u) {
                    const uint32_t j = j0 + u * kEmitThreads;
                    if (j < total4) {
                        const uint32_t e = j / 3u;
                        if (base + e < a.cap_visible) __stcs(o.rows + (size_t)base * 3u + j, r[u]);  // written once, read by another kernel
                    }
                }
            }
        }
        return;
    }

    // Item-major output: every visible item's keys / id / row are read once and stored to each view that sees it,
    // at base[v] + (visible-before count of view v).  Consecutive threads hold consecutive items (and row
    // parts), so the stores of one view land next to each other.
    for (uint32_t v = tid; v < a.n_views; v += kEmitThreads)
        if (s.base[v] + s.total[v] > a.cap_visible) *a.overflow = 1u;
#pragma unroll
    for (int u = 0; u < kEmitItems; ++u) {
        const uint32_t e = tid + u * kEmitThreads;
        const uint32_t mm = s.msk[e];
        if (mm) {
            const uint64_t pa2 = s.pos_a[e], pb2 = s.pos_b[e], k = s.key[e];
            const uint32_t ix = s.idx[e];
#pragma unroll
            for (int v = 0; v < (int)SS_MAX_VIEWS; ++v) {
                if ((mm >> v) & 1u) {
                    const uint32_t g = s.base[v] + ((uint32_t)((v < 4 ? pa2 >> (16 * v) : pb2 >> (16 * (v - 4)))) & 0xffffu);
                    if (g < a.cap_visible) {
                        a.out[v].keys[g] = k;
                        a.out[v].idx[g] = ix;
                    }
                }
            }
        }
    }
    constexpr int kParts = kEmitTile * 3 / kEmitThreads;  // (item, row part) pairs per thread
    for (int u0 = 0; u0 < kParts; u0 += kGatherBatch) {
        float4 r[kGatherBatch];
#pragma unroll
        for (int u = 0; u < kGatherBatch; ++u) {
            const uint32_t j = tid + (u0 + u) * kEmitThreads;
            const uint32_t e = j / 3u, part = j - e * 3u;
            if (u0 + u < kParts && s.msk[e])
                r[u] = __ldg(a.seq ? a.world + ((size_t)tile_base + e) * a.world_stride + part : a.world + (size_t)s.idx[e] * kWorldStride + part);
        }
#pragma unroll
        for (int u = 0; u < kGatherBatch; ++u) {
            const uint32_t j = tid + (u0 + u) * kEmitThreads;
            const uint32_t e = j / 3u, part = j - e * 3u;
            const uint32_t mm = u0 + u < kParts ? s.msk[e] : 0u;
            if (mm) {
                const uint64_t pa2 = s.pos_a[e], pb2 = s.pos_b[e];
#pragma unroll
                for (int v = 0; v < (int)SS_MAX_VIEWS; ++v) {
                    if ((mm >> v) & 1u) {
                        const uint32_t g = s.base[v] + ((uint32_t)((v < 4 ? pa2 >> (16 * v) : pb2 >> (16 * (v - 4)))) & 0xffffu);
                        if (g < a.cap_visible) a.out[v].rows[(size_t)g * 3u + part] = r[u];
                    }
                }
            }
        }
    }
}

// one CTA per view: runs -> draws
constexpr int kDrawThreads = 256;
constexpr int kDrawItems = 4;  // consecutive runs per thread
constexpr int kDrawChunk = kDrawThreads * kDrawItems;
constexpr int kDrawWarps = kDrawThreads / 32;

// Grid (chunks of 1024 runs, views).  Run s of the sorted order becomes a draw of view v when its per-view count is
// non-zero; draws keep the run order (= MeshKey order) and firstInstance is the exclusive sum of the counts.  A CTA
// sums the counts of the chunks before its own itself (a few KB out of L2), so chunks need no hand-over; the
// records are staged in shared memory and leave as whole lines.  The per-view counters are double-buffered by
// frame: this frame's stay as they are, the other buffer is cleaned here for the next frame's k_emit.
__global__ void __launch_bounds__(kDrawThreads) k_draws(const __grid_constant__ DrawArgs a)
{
    __shared__ uint32_t sm_draw[kDrawChunk * 5];
    __shared__ uint64_t sm_key[kDrawChunk];
    __shared__ uint32_t warp_c[kDrawWarps], warp_f[kDrawWarps], pre_c[kDrawWarps], pre_f[kDrawWarps];
    const uint32_t v = blockIdx.y, c0 = blockIdx.x * kDrawChunk;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t n_raw = a.totals[SS_MAX_VIEWS];
    const uint32_t n_segs = min(n_raw, a.max_draws);
    const uint32_t n_dirty = min(*a.runs_next, a.max_draws);

    uint32_t* nxt = a.seg_count_next + (size_t)v * a.max_draws;
#pragma unroll
    for (int j = 0; j < kDrawItems; ++j) {
        const uint32_t i = c0 + j * kDrawThreads + tid;
        if (i < n_dirty) nxt[i] = 0u;
    }
    if (blockIdx.x == 0 && v == 0 && tid == 0) {
        const uint32_t over = (n_raw > a.max_draws) ? 1u : 0u;
        if (over) *a.overflow = 1u;
        *a.runs_cur = n_segs;
        a.host_stats[a.stats_views * 2] = *a.overflow | over;
        a.host_stats[a.stats_views * 2 + 1] = *a.passes_run;
    }
    // rows of views this frame does not use are cleaned all the same: the frame that filled the other buffer may
    // have run more views than this one (ss_frame takes any n_views up to max_views)
    if (v >= a.n_views) return;
    if (c0 >= n_segs && blockIdx.x != 0) return;

    // counts and draws of the chunks before this one
    const uint32_t* cnt = a.seg_count + (size_t)v * a.max_draws;
    uint32_t pc = 0, pf = 0;
#pragma unroll 4
    for (uint32_t i = tid; i < c0; i += kDrawThreads) {
        const uint32_t q = cnt[i];
        pc += q;
        pf += q ? 1u : 0u;
    }
    // this chunk: thread t owns runs c0 + 4t .. c0 + 4t + 3
    const uint32_t first = c0 + tid * kDrawItems;
    uint32_t c[kDrawItems], inst[kDrawItems];
#pragma unroll
    for (int j = 0; j < kDrawItems; ++j) {
        const bool ok = first + j < n_segs;
        c[j] = ok ? cnt[first + j] : 0u;
        inst[j] = ok ? a.seg_inst[first + j] : 0u;
    }
    uint32_t sl[kDrawItems];
#pragma unroll
    for (int j = 0; j < kDrawItems; ++j) sl[j] = c[j] ? a.slot[inst[j]] : 0u;  // c != 0: a live instance, slot valid
    uint32_t ic[kDrawItems], fi[kDrawItems];
    int32_t vo[kDrawItems];
    uint64_t sk[kDrawItems];
#pragma unroll
    for (int j = 0; j < kDrawItems; ++j) {
        ic[j] = fi[j] = 0u;
        vo[j] = 0;
        sk[j] = 0ull;
        if (c[j]) {
            const SlotDev* sd = a.slots + sl[j];
            ic[j] = sd->index_count;
            fi[j] = sd->first_index;
            vo[j] = sd->vertex_offset;
            sk[j] = a.seg_key[first + j];
        }
    }
#pragma unroll
    for (int ofs = 16; ofs > 0; ofs >>= 1) {
        pc += __shfl_xor_sync(0xffffffffu, pc, ofs);
        pf += __shfl_xor_sync(0xffffffffu, pf, ofs);
    }
    uint32_t tc = 0, tf = 0;
#pragma unroll
    for (int j = 0; j < kDrawItems; ++j) {
        tc += c[j];
        tf += c[j] ? 1u : 0u;
    }
    uint32_t xc = tc, xf = tf;
#pragma unroll
    for (int ofs = 1; ofs < 32; ofs <<= 1) {
        const uint32_t yc = __shfl_up_sync(0xffffffffu, xc, ofs), yf = __shfl_up_sync(0xffffffffu, xf, ofs);
        if (lane >= ofs) {
            xc += yc;
            xf += yf;
        }
    }
    if (lane == 31) {
        warp_c[wid] = xc;
        warp_f[wid] = xf;
    }
    if (lane == 0) {
        pre_c[wid] = pc;
        pre_f[wid] = pf;
    }
    __syncthreads();
    uint32_t base_c = 0, base_f = 0, loc_c = 0, loc_f = 0, all_f = 0;
#pragma unroll
    for (int w = 0; w < kDrawWarps; ++w) {
        base_c += pre_c[w];
        base_f += pre_f[w];
        if (w < wid) {
            loc_c += warp_c[w];
            loc_f += warp_f[w];
        }
        all_f += warp_f[w];
    }
    uint32_t run_c = base_c + loc_c + xc - tc;  // instances of this view before this thread's first run
    uint32_t run_f = loc_f + xf - tf;           // draws of this chunk before it
#pragma unroll
    for (int j = 0; j < kDrawItems; ++j) {
        if (c[j]) {
            uint32_t* d = sm_draw + run_f * 5;
            d[0] = ic[j];
            d[1] = c[j];
            d[2] = fi[j];
            d[3] = (uint32_t)vo[j];
            d[4] = run_c;
            sm_key[run_f] = sk[j];
            run_c += c[j];
            run_f += 1u;
        }
    }
    __syncthreads();
    const ViewOut& o = a.out[v];
    uint32_t* dst = reinterpret_cast<uint32_t*>(o.draws + base_f);
    for (uint32_t w = tid; w < all_f * 5; w += kDrawThreads) dst[w] = sm_draw[w];
    for (uint32_t k = tid; k < all_f; k += kDrawThreads) o.draw_keys[base_f + k] = sm_key[k];
    if (tid == 0 && c0 + kDrawChunk >= n_segs) {  // the chunk holding the last run (chunk 0 when there is none)
        const uint32_t vis = a.totals[v], nd = base_f + all_f;
        o.counts[0] = vis;
        o.counts[1] = nd;
        a.host_stats[v * 2 + 0] = vis;
        a.host_stats[v * 2 + 1] = nd;
    }
}

// Merged-list placement (multi-GPU): persistent warps over the output, 32 positions at a time; a warp bisects the
// segment prefix once for its first position and its lanes walk on from there (segments are mostly far longer than
// a warp, empty ones are stepped over).
__global__ void __launch_bounds__(256) k_merge_place(const int32_t* __restrict__ lists, const int64_t* __restrict__ seg_out,
                                                     const int64_t* __restrict__ seg_src, const int64_t* __restrict__ seg_add,
                                                     const int64_t* __restrict__ seg_key, uint32_t n_segs,
                                                     int32_t* __restrict__ out_idx, int64_t* __restrict__ out_keys)
{
    const int lane = threadIdx.x & 31;
    const int64_t total = seg_out[n_segs];
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5, warp_id = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    for (int64_t g0 = warp_id * 32; g0 < total; g0 += warps * 32) {
        uint32_t lo = 0, hi = n_segs;
        while (hi - lo > 1) {
            const uint32_t mid = (lo + hi) >> 1;
            if (seg_out[mid] <= g0) lo = mid; else hi = mid;
        }
        const int64_t e = g0 + lane;
        if (e < total) {
            uint32_t sI = lo;
            while (seg_out[sI + 1] <= e) ++sI;
            out_idx[e] = (int32_t)((int64_t)lists[seg_src[sI] + (e - seg_out[sI])] + seg_add[sI]);
            out_keys[e] = seg_key[sI];
        }
    }
}

// The same over peer memory: segment s lies in rank seg_rank[s]'s own output buffer, read across NVLink where it is.
// Four independent 128 B requests per warp are in flight per round (remote latency is several times the local one).
__global__ void __launch_bounds__(256) k_merge_place_peers(const int32_t* const* __restrict__ rank_lists,
                                                           const int32_t* __restrict__ seg_rank, const int64_t* __restrict__ seg_out,
                                                           const int64_t* __restrict__ seg_src, const int64_t* __restrict__ seg_add,
                                                           const int64_t* __restrict__ seg_key, uint32_t n_segs,
                                                           int32_t* __restrict__ out_idx, int64_t* __restrict__ out_keys)
{
    constexpr int kRounds = 4;
    const int lane = threadIdx.x & 31;
    const int64_t total = seg_out[n_segs];
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5, warp_id = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    for (int64_t g0 = warp_id * (32 * kRounds); g0 < total; g0 += warps * (32 * kRounds)) {
        uint32_t lo = 0, hi = n_segs;
        while (hi - lo > 1) {
            const uint32_t mid = (lo + hi) >> 1;
            if (seg_out[mid] <= g0) lo = mid; else hi = mid;
        }
        int32_t val[kRounds];
        int64_t add[kRounds], key[kRounds];
        uint32_t sI = lo;
#pragma unroll
        for (int r = 0; r < kRounds; ++r) {
            const int64_t e = g0 + r * 32 + lane;
            val[r] = 0;
            add[r] = key[r] = 0;
            if (e < total) {
                while (seg_out[sI + 1] <= e) ++sI;
                val[r] = rank_lists[seg_rank[sI]][seg_src[sI] + (e - seg_out[sI])];
                add[r] = seg_add[sI];
                key[r] = seg_key[sI];
            }
        }
#pragma unroll
        for (int r = 0; r < kRounds; ++r) {
            const int64_t e = g0 + r * 32 + lane;
            if (e < total) {
                out_idx[e] = (int32_t)((int64_t)val[r] + add[r]);
                out_keys[e] = key[r];
            }
        }
    }
}

}  // namespace

void launch_merge_place(const int32_t* lists, const int64_t* seg_out, const int64_t* seg_src, const int64_t* seg_add,
                        const int64_t* seg_key, uint32_t n_segs, int32_t* out_idx, int64_t* out_keys, cudaStream_t st);

size_t emit_lookback_bytes(uint64_t n)
{
    const uint64_t tiles = (n + kEmitTile - 1) / kEmitTile;
    return (size_t)(tiles ? tiles : 1) * kEmitWords * sizeof(uint32_t);
}

int device_sm_count()
{
    static int sms[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    int& n = sms[dev & 63];
    if (!n && cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) n = 148;
    return n;
}

void launch_emit(const EmitArgs& a, cudaStream_t st)
{
    static bool attr_set[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!attr_set[dev & 63]) {
        cudaFuncSetAttribute(k_emit, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(EmitSmem));
        attr_set[dev & 63] = true;
    }
    const uint32_t tiles = (a.n + kEmitTile - 1) / kEmitTile;
    if (tiles) k_emit<<<tiles, kEmitThreads, sizeof(EmitSmem), st>>>(a);
}

void launch_draws(const DrawArgs& a, cudaStream_t st)
{
    // grid.y covers every view row of the double-buffered run counters (stats_views = max_views), not only this frame's
    if (a.n_views)
        k_draws<<<dim3((a.max_draws + kDrawChunk - 1) / kDrawChunk, std::max(a.n_views, a.stats_views)), kDrawThreads, 0, st>>>(a);
}

void launch_merge_place(const int32_t* lists, const int64_t* seg_out, const int64_t* seg_src, const int64_t* seg_add,
                        const int64_t* seg_key, uint32_t n_segs, int32_t* out_idx, int64_t* out_keys, cudaStream_t st)
{
    if (n_segs) k_merge_place<<<device_sm_count() * 8, 256, 0, st>>>(lists, seg_out, seg_src, seg_add, seg_key, n_segs, out_idx, out_keys);
}

void launch_merge_place_peers(const int32_t* const* rank_lists, const int32_t* seg_rank, const int64_t* seg_out,
                              const int64_t* seg_src, const int64_t* seg_add, const int64_t* seg_key, uint32_t n_segs,
                              int32_t* out_idx, int64_t* out_keys, cudaStream_t st)
{
    if (n_segs)
        k_merge_place_peers<<<device_sm_count() * 8, 256, 0, st>>>(rank_lists, seg_rank, seg_out, seg_src, seg_add, seg_key, n_segs, out_idx, out_keys);
}

}  // namespace ss
