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

template <bool GATHER, bool HAS_PARENT>
struct SubSmem {
    static constexpr int kWordBufs = GATHER ? 3 : 2;
    InstRec rec[kSubTile];                             // stage: records of the tile about to be computed
    float4 rows[kSubTile * 3];                         // TransformData of the tile just computed, item-major
    uint64_t words[kWordBufs][kSubTile];               // sorted words: current tile, next (, the one after)
    int32_t parent[(HAS_PARENT && !GATHER) ? kSubTile : 4];
    f2 planes[SS_MAX_VIEWS * kViewWords];
    uint16_t seg_start[kSubTile + 2][SS_MAX_VIEWS];    // visible-before count at the start of each local run
    uint8_t list[SS_MAX_VIEWS][kSubTile];              // per view: local ids of its visible items, in order
    __align__(16) uint8_t msk[kSubTile];
    uint32_t flags[SS_MAX_VIEWS];
    uint64_t warp_a[kSubWarps], warp_b[kSubWarps];
    uint32_t warp_h[kSubWarps];
    uint32_t sink[kSubWarps];                          // see the unpack step
    uint32_t base[kSubWords];                          // exclusive prefix over earlier tiles
    uint32_t total[kSubWords];                         // this tile
    uint64_t rbar;
    uint64_t wbar[kWordBufs];
};

// Decoupled look-back by one warp, all kSubWords counters at once, over a status array lb[w * stride + i]: value | flag
// (kLbAgg: element i's own count, kLbPrefix: inclusive prefix over elements <= i).  Each round the 32 lanes read the
// 32 elements below `start` for every unresolved counter (one 128 B line per counter), take the longest run of
// published words that ends at the nearest prefix, and go on from there.  Elements below `floor_i` count as an empty
// prefix.  Returns the sum for counter `lane` (lanes >= kSubWords: 0).
__device__ __forceinline__ uint32_t warp_walk(const uint32_t* lb, uint32_t stride, int64_t first, int64_t floor_i, int lane)
{
    uint32_t sum[kSubWords];
    int64_t start[kSubWords];
    uint32_t open = (1u << kSubWords) - 1u;
#pragma unroll
    for (int w = 0; w < kSubWords; ++w) {
        sum[w] = 0;
        start[w] = first;
    }
    while (open) {
        uint32_t v[kSubWords];
#pragma unroll
        for (int w = 0; w < kSubWords; ++w) {
            const int64_t t = start[w] - lane;
            v[w] = kLbPrefix;  // below the floor: an empty prefix
            if (((open >> w) & 1u) && t >= floor_i) v[w] = ld_volatile_u32(lb + (size_t)w * stride + (size_t)t);
        }
#pragma unroll
        for (int w = 0; w < kSubWords; ++w) {
            if ((open >> w) & 1u) {
                const uint32_t has_p = __ballot_sync(0xffffffffu, (v[w] & kLbPrefix) != 0);
                const uint32_t has_any = __ballot_sync(0xffffffffu, (v[w] & (kLbPrefix | kLbAgg)) != 0);
                const int first_missing = (~has_any) ? __ffs(~has_any) - 1 : 32;
                const int first_prefix = has_p ? __ffs(has_p) - 1 : 32;
                const int take = min(first_missing, first_prefix + 1);  // lanes [0, take) are usable
                sum[w] += __reduce_add_sync(0xffffffffu, lane < take ? (v[w] & kLbValue) : 0u);
                if (first_prefix < first_missing)
                    open &= ~(1u << w);
                else
                    start[w] -= take;  // take may be 0: the nearest element has not published yet, poll again
            }
        }
    }
    uint32_t mine = 0;
#pragma unroll
    for (int w = 0; w < kSubWords; ++w)
        if (lane == w) mine = sum[w];
    return mine;
}

// Exclusive prefix of tile `tile` in two levels.  The persistent CTAs run in near lock-step, so at any moment several
// hundred consecutive tiles have published their counts but not yet their prefixes; walking them 32 at a time costs
// dozens of L2 round trips per tile.  Instead tiles are grouped by kLbGroup = 32:
//   level 1  lb_tile[w][t]   = kLbAgg | count of tile t               (never upgraded)
//   level 2  lb_group[w][g]  = kLbAgg | count of group g, then kLbPrefix | inclusive prefix over groups <= g
// A tile sums the counts of the earlier tiles of its own group (one round), then walks the groups (32 groups = 1024
// tiles per round).  The last tile of a group publishes the group's count as soon as its first walk is done and the
// group's prefix when its second is.
constexpr int kLbGroup = 32;
__device__ __forceinline__ uint32_t two_level_lookback(uint32_t* lb_tile, uint32_t* lb_group, uint32_t stride_t, uint32_t stride_g,
                                                       uint32_t tile, uint32_t mine, int lane, bool last_tile)
{
    const uint32_t g = tile / kLbGroup, q = tile % kLbGroup;
    const bool closes_group = q == kLbGroup - 1 || last_tile;
    if (lane < kSubWords) st_volatile_u32(lb_tile + (size_t)lane * stride_t + tile, kLbAgg | mine);
    const uint32_t within = warp_walk(lb_tile, stride_t, (int64_t)tile - 1, (int64_t)g * kLbGroup, lane);
    if (closes_group && lane < kSubWords) st_volatile_u32(lb_group + (size_t)lane * stride_g + g, kLbAgg | (within + mine));
    const uint32_t before = warp_walk(lb_group, stride_g, (int64_t)g - 1, 0, lane);
    if (closes_group && lane < kSubWords) st_volatile_u32(lb_group + (size_t)lane * stride_g + g, kLbPrefix | (before + within + mine));
    return before + within;
}

template <bool GATHER, bool HAS_PARENT>
__global__ void __launch_bounds__(kSubThreads, 5) k_submit(const __grid_constant__ SubmitArgs a, const __grid_constant__ CullParams cp,
                                                           const Neutral neutral)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    using Smem = SubSmem<GATHER, HAS_PARENT>;
    Smem& s = *reinterpret_cast<Smem*>(smem_raw);
    constexpr int kWB = Smem::kWordBufs;
    const P2 m{neutral};
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t n_tiles = (a.n + kSubTile - 1) / kSubTile;
    const uint64_t* __restrict__ words = (a.sort_ctl && a.sort_ctl->src_sel[a.sort_passes]) ? a.words_b : a.words_a;
    const uint64_t imask = (1ull << a.idx_bits) - 1ull;
    const bool has_box = a.has_box != 0;

    // Every array a full-tile bulk copy reads is padded by kSubPad elements, so the last, ragged tile is copied
    // like any other (what lies past n is never used as data or as an address).
    auto issue_words = [&](uint32_t tile, int b) {  // one thread
        mbar_expect_tx(&s.wbar[b], kSubTile * 8);
        bulk_g2s(s.words[b], words + (size_t)tile * kSubTile, kSubTile * 8, &s.wbar[b]);
    };
    auto issue_rec_seq = [&](uint32_t tile) {  // one thread; resident layout: the tile's records are contiguous
        mbar_expect_tx(&s.rbar, kSubTile * 64 + ((HAS_PARENT && !GATHER) ? kSubTile * 4 : 0));
        const uint32_t chunk = (kSubTile * 64) / a.seq_chunks;  // seq_chunks bulk copies per tile (see launch_submit)
        for (uint32_t c = 0; c < a.seq_chunks; ++c)
            bulk_g2s(reinterpret_cast<unsigned char*>(s.rec) + c * chunk,
                     reinterpret_cast<const unsigned char*>(a.rec + (size_t)tile * kSubTile) + c * chunk, chunk, &s.rbar);
        if (HAS_PARENT && !GATHER) bulk_g2s(s.parent, a.parent + (size_t)tile * kSubTile, kSubTile * 4, &s.rbar);
    };
    auto issue_rec_gather = [&](uint32_t tile, int b) {  // every thread, once the tile's words are in buffer b
        const uint32_t base = tile * kSubTile;
        const uint32_t nt = min((uint32_t)kSubTile, a.n - base);
        if (tid == 0) mbar_expect_tx(&s.rbar, nt * 64);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const uint32_t li = tid + h * kSubThreads;
            if (li < nt) bulk_g2s(&s.rec[li], a.rec + (size_t)(s.words[b][li] & imask), 64, &s.rbar);
        }
    };

    // ---- per-CTA set-up: barriers, duplicated plane constants
    if (tid == 0) {
        mbar_init(&s.rbar, 1);
#pragma unroll
        for (int b = 0; b < kWB; ++b) mbar_init(&s.wbar[b], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (uint32_t k = tid; k < cp.n_views * 3; k += kSubThreads) {
        const uint32_t v = k / 3, ax = k - v * 3;
        const float4 p0 = cp.v[v].pl[2 * ax], p1 = cp.v[v].pl[2 * ax + 1];
        f2* c = s.planes + v * kViewWords + ax * kAxisWords;
        c[0] = dup(p0.x); c[1] = dup(p0.y); c[2] = dup(p0.z);
        c[3] = dup(fabsf(p0.x)); c[4] = dup(fabsf(p0.y)); c[5] = dup(fabsf(p0.z));
        const bool skip_near = ax == 2 && (cp.v[v].flags & kViewSkipNear);
        c[6] = dup(skip_near ? __int_as_float(0x7f800000) : p0.w); c[7] = dup(p1.w);
        c[8] = dup(p1.x); c[9] = dup(p1.y); c[10] = dup(p1.z);
        c[11] = dup(fabsf(p1.x)); c[12] = dup(fabsf(p1.y)); c[13] = dup(fabsf(p1.z));
    }
    if (tid < (int)SS_MAX_VIEWS) s.flags[tid] = tid < (int)cp.n_views ? cp.v[tid].flags : 0u;
    __syncthreads();

    // ---- prologue
    {
        const uint32_t t0 = blockIdx.x, t1 = t0 + gridDim.x;
        if (tid == 0) {
            issue_words(t0, 0);
            if (GATHER && t1 < n_tiles) issue_words(t1, 1);
            if (!GATHER) issue_rec_seq(t0);
        }
        if (GATHER) {
            mbar_wait(&s.wbar[0], 0);
            issue_rec_gather(t0, 0);
        }
    }

    uint32_t k = 0;
    for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++k) {
        const int wb = k % kWB;
        const uint32_t base = tile * kSubTile;
        const int n_tile = min((uint32_t)kSubTile, a.n - base);
        const uint64_t* sw = s.words[wb];
        if (!GATHER) mbar_wait(&s.wbar[wb], (k / kWB) & 1u);  // gather: waited for when the gather was issued
        mbar_wait(&s.rbar, k & 1u);

        // ---- unpack the two instances of this thread (items tid and tid + kSubThreads of the tile)
        const int t0 = tid, t1 = tid + kSubThreads;
        const bool valid0 = t0 < n_tile, valid1 = t1 < n_tile;
        const float4* ra = reinterpret_cast<const float4*>(&s.rec[valid0 ? t0 : 0]);
        const float4* rb = reinterpret_cast<const float4*>(&s.rec[valid1 ? t1 : 0]);
        const float4 A0 = ra[0], A1 = ra[1], A2 = ra[2], A3 = ra[3];
        const float4 B0 = rb[0], B1 = rb[1], B2 = rb[2], B3 = rb[3];
        const uint32_t slot0 = __float_as_uint(A3.w), slot1 = __float_as_uint(B3.w);
        const bool live0 = valid0 && slot0 != SS_SLOT_NULL, live1 = valid1 && slot1 != SS_SLOT_NULL;
        const uint32_t sl0 = live0 ? slot0 : 0u, sl1 = live1 ? slot1 : 0u;
        int p0 = -1, p1 = -1;
        if (HAS_PARENT) {
            if (GATHER) {
                if (valid0) p0 = __ldg(a.parent + (size_t)(sw[t0] & imask));
                if (valid1) p1 = __ldg(a.parent + (size_t)(sw[t1] & imask));
            } else {
                if (valid0) p0 = s.parent[t0];
                if (valid1) p1 = s.parent[t1];
            }
        }
        const SlotDev* sd0 = a.slots + sl0;
        const SlotDev* sd1 = a.slots + sl1;
        const float4 x0 = ld_early_f4(&sd0->r0), x1 = ld_early_f4(&sd0->r1), x2 = ld_early_f4(&sd0->r2);
        const float4 y0 = ld_early_f4(&sd1->r0), y1 = ld_early_f4(&sd1->r1), y2 = ld_early_f4(&sd1->r2);
        float4 pr0[3], pr1[3];
        if (HAS_PARENT) {
            const float4* q0 = a.node_world + (size_t)max(p0, 0) * 3;
            const float4* q1 = a.node_world + (size_t)max(p1, 0) * 3;
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                pr0[r] = ld_early_f4(q0 + r);
                pr1[r] = ld_early_f4(q1 + r);
            }
        }
        f2 t9[9], mn[3], mx[3];
        t9[0] = pk(A0.x, B0.x); t9[1] = pk(A0.y, B0.y); t9[2] = pk(A0.z, B0.z); t9[3] = pk(A0.w, B0.w);
        t9[4] = pk(A1.x, B1.x); t9[5] = pk(A1.y, B1.y); t9[6] = pk(A1.z, B1.z); t9[7] = pk(A1.w, B1.w);
        t9[8] = pk(A2.x, B2.x);
        if (has_box) {
            mn[0] = pk(A2.y, B2.y); mn[1] = pk(A2.z, B2.z); mn[2] = pk(A2.w, B2.w);
            mx[0] = pk(A3.x, B3.x); mx[1] = pk(A3.y, B3.y); mx[2] = pk(A3.z, B3.z);
        }
        // the bulk stores of the previous tile (world rows, masks) must have read their shared-memory source
        // before this tile overwrites it
        // The stage is refilled by the async proxy (TMA) right after the barrier below, and a barrier only orders
        // loads that have been ISSUED: when every CTA of an SM unpacks at once (first tile of the launch) the
        // bank-conflicted 128-bit loads queue for longer than the next tile's copy takes to arrive, and a queued load
        // then reads the next tile's bytes.  A store that depends on every loaded word cannot issue before the
        // loads have returned, so the thread only reaches the barrier with its records really in registers.
        s.sink[wid] = __float_as_uint(A0.x) ^ __float_as_uint(A1.x) ^ __float_as_uint(A2.x) ^ slot0 ^ __float_as_uint(B0.x) ^
                      __float_as_uint(B1.x) ^ __float_as_uint(B2.x) ^ slot1 ^ (uint32_t)p0 ^ (uint32_t)p1;
        if (tid == 0) bulk_wait_read();
        __syncthreads();  // S1: records are in registers, the previous tile's output phase is over

        // ---- prefetch: records of this CTA's next tile, words of the one after
        {
            const uint32_t next = tile + gridDim.x, next2 = next + gridDim.x;
            if (GATHER) {
                if (next < n_tiles) {
                    mbar_wait(&s.wbar[(k + 1) % kWB], ((k + 1) / kWB) & 1u);
                    issue_rec_gather(next, (k + 1) % kWB);
                }
                if (tid == 0 && next2 < n_tiles) issue_words(next2, (k + 2) % kWB);
            } else if (tid == 0 && next < n_tiles) {
                issue_rec_seq(next);
                issue_words(next, (k + 1) % kWB);
            }
        }

        // ---- transform (ss_x2.cuh: same operations, same order as k_instances_x2)
        Aff2 node;
        trs_to_affine2(m, t9, node);
        if (HAS_PARENT) {  // N1: world = world[parent] * local
            Aff2 par, tmp;
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                const float4 x = pr0[r], y = pr1[r];
                par.r[r][0] = pk(x.x, y.x); par.r[r][1] = pk(x.y, y.y); par.r[r][2] = pk(x.z, y.z); par.r[r][3] = pk(x.w, y.w);
            }
            aff_mul2(m, par, node, tmp);
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    node.r[r][c] = pk(p0 >= 0 ? lo(tmp.r[r][c]) : lo(node.r[r][c]), p1 >= 0 ? hi(tmp.r[r][c]) : hi(node.r[r][c]));
        }
        Aff2 loc, w;
        loc.r[0][0] = pk(x0.x, y0.x); loc.r[0][1] = pk(x0.y, y0.y); loc.r[0][2] = pk(x0.z, y0.z); loc.r[0][3] = pk(x0.w, y0.w);
        loc.r[1][0] = pk(x1.x, y1.x); loc.r[1][1] = pk(x1.y, y1.y); loc.r[1][2] = pk(x1.z, y1.z); loc.r[1][3] = pk(x1.w, y1.w);
        loc.r[2][0] = pk(x2.x, y2.x); loc.r[2][1] = pk(x2.y, y2.y); loc.r[2][2] = pk(x2.z, y2.z); loc.r[2][3] = pk(x2.w, y2.w);
        aff_mul2(m, node, loc, w);  // VC/Scene/SceneRenderer.cpp:1150

        // ---- rows into the staging buffer (48 B per item: a lane's three float4 stores hit distinct banks)
        {
            const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                s.rows[t0 * 3 + r] = live0 ? make_float4(lo(w.r[r][0]), lo(w.r[r][1]), lo(w.r[r][2]), lo(w.r[r][3])) : zero;
                s.rows[t1 * 3 + r] = live1 ? make_float4(hi(w.r[r][0]), hi(w.r[r][1]), hi(w.r[r][2]), hi(w.r[r][3])) : zero;
            }
        }

        // ---- cull
        if (!has_box) {
#pragma unroll
            for (int q = 0; q < 3; ++q) {
                mn[q] = pk(__ldg(&sd0->mn[q]), __ldg(&sd1->mn[q]));
                mx[q] = pk(__ldg(&sd0->mx[q]), __ldg(&sd1->mx[q]));
            }
        }
        const uint2 cm = cull_box2(m, w, mn, mx, s.planes, s.flags, cp.n_views);
        s.msk[t0] = (uint8_t)(live0 ? cm.x : 0u);
        s.msk[t1] = (uint8_t)(live1 ? cm.y : 0u);
        fence_async_smem();  // rows / masks become visible to the bulk stores below
        __syncthreads();     // S2

        // ---- all rows of the tile, in word order: one bulk store shared -> global
        if (a.world_out) {
            if (tid == 0) {
                bulk_s2g(a.world_out + (size_t)base * 3, s.rows, (uint32_t)n_tile * 48u);
                if ((n_tile & 15) == 0) bulk_s2g(a.mask_out + base, s.msk, (uint32_t)n_tile);
                bulk_commit();
            }
            if (n_tile & 15)
                for (int i = tid; i < n_tile; i += kSubThreads) a.mask_out[base + i] = s.msk[i];
        }

        // ---- ranks: thread t now owns items 2t and 2t + 1 of the tile (word order)
        const uint32_t l0 = tid * 2;
        uint64_t key[2];
        uint32_t idx[2], mm[2];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const bool ok = l0 + j < (uint32_t)n_tile;
            const uint64_t wv = sw[l0 + j];
            key[j] = ok ? wv >> a.idx_bits : ~0ull;
            idx[j] = ok ? (uint32_t)(wv & imask) : 0u;
            mm[j] = ok ? (uint32_t)s.msk[l0 + j] : 0u;
        }
        uint64_t prev = 0;
        bool first_is_head = false;
        if (tid > 0) {
            prev = sw[l0 - 1] >> a.idx_bits;
        } else if (tile > 0) {
            prev = __ldg(words + base - 1) >> a.idx_bits;
        } else {
            first_is_head = true;
        }
        bool head[2];
        uint64_t pa = 0, pb = 0;
        uint32_t ph = 0;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const bool ok = l0 + j < (uint32_t)n_tile;
            head[j] = ok && ((j == 0 && first_is_head) || key[j] != (j == 0 ? prev : key[j - 1]));
            pa += spread4(mm[j] & 15u);
            pb += spread4(mm[j] >> 4);
            ph += head[j] ? 1u : 0u;
        }
        uint64_t xa = pa, xb = pb;
        uint32_t xh = ph;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint64_t ya = shfl_up_u64(xa, o), yb = shfl_up_u64(xb, o);
            const uint32_t yh = __shfl_up_sync(0xffffffffu, xh, o);
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
        __syncthreads();  // S3
        uint64_t ba = 0, bb = 0, ta = 0, tb = 0;
        uint32_t bh = 0, th = 0;
#pragma unroll
        for (int q = 0; q < kSubWarps; ++q) {
            if (q < wid) {
                ba += s.warp_a[q];
                bb += s.warp_b[q];
                bh += s.warp_h[q];
            }
            ta += s.warp_a[q];
            tb += s.warp_b[q];
            th += s.warp_h[q];
        }
        uint64_t ea = ba + xa - pa, eb = bb + xb - pb;  // exclusive prefixes of this thread's first item
        uint32_t eh = bh + xh - ph;
        // every thread knows the tile totals (no extra barrier): counter w of the tile
        auto tile_total = [&](int wq) -> uint32_t {
            return wq < 4 ? (uint32_t)(ta >> (16 * wq)) & 0xffffu : wq < 8 ? (uint32_t)(tb >> (16 * (wq - 4))) & 0xffffu : th;
        };

        // ---- decoupled look-back (warp 0) while the other warps build the local lists
        if (wid == 0) {
            uint32_t mine = 0;
#pragma unroll
            for (int wq = 0; wq < kSubWords; ++wq)
                if (lane == wq) mine = tile_total(wq);
            const uint32_t excl = two_level_lookback(a.lookback, a.lookback + (size_t)kSubWords * a.lb_stride, a.lb_stride,
                                                     a.lb_stride / kLbGroup, tile, mine, lane, tile == n_tiles - 1);
            if (lane < kSubWords) {
                s.base[lane] = excl;
                s.total[lane] = mine;
                if (tile == n_tiles - 1) a.totals[lane] = excl + mine;
            }
        }
        const uint32_t n_heads = th;
        if (tid >= kSubThreads - (int)SS_MAX_VIEWS) {  // last warp: the sentinels of the run table
            const int v = tid - (kSubThreads - (int)SS_MAX_VIEWS);
            s.seg_start[0][v] = 0;
            s.seg_start[n_heads + 1][v] = (uint16_t)tile_total(v);
        }
        {
            uint64_t fa = ea, fb = eb;
            uint32_t fh = eh;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                if (head[j]) {
                    const uint32_t lseg = fh + 1;  // 1-based among runs starting in this tile; 0 = run carried in
#pragma unroll
                    for (int v = 0; v < 4; ++v) {
                        s.seg_start[lseg][v] = (uint16_t)(fa >> (16 * v));
                        s.seg_start[lseg][4 + v] = (uint16_t)(fb >> (16 * v));
                    }
                }
#pragma unroll
                for (int v = 0; v < (int)SS_MAX_VIEWS; ++v) {
                    if ((mm[j] >> v) & 1u) {
                        const uint32_t lpos = (uint32_t)((v < 4 ? fa >> (16 * v) : fb >> (16 * (v - 4))) & 0xffffu);
                        s.list[v][lpos] = (uint8_t)(l0 + j);
                    }
                }
                fa += spread4(mm[j] & 15u);
                fb += spread4(mm[j] >> 4);
                fh += head[j] ? 1u : 0u;
            }
        }
        __syncthreads();  // S4: bases, lists and run table are complete

        // ---- run bookkeeping: first key / first instance of each run, visible count per (view, run)
        const uint32_t seg_base = s.base[SS_MAX_VIEWS];
        {
            uint32_t ehh = eh;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
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
        for (uint32_t q = tid; q <= n_heads; q += kSubThreads) {
            const uint32_t g = seg_base + q - 1u;  // q == 0: run that started in an earlier tile
#pragma unroll
            for (int v = 0; v < (int)SS_MAX_VIEWS; ++v) {
                if (v < (int)a.n_views) {
                    const uint32_t c = (uint32_t)s.seg_start[q + 1][v] - (uint32_t)s.seg_start[q][v];
                    if (c && g < a.max_draws) atomicAdd(a.seg_count + (size_t)v * a.max_draws + g, c);
                }
            }
        }

        // ---- per view: keys + instance ids, then the rows, all as coalesced stores in draw order
        for (uint32_t v = 0; v < a.n_views; ++v) {
            const uint32_t cnt = s.total[v];
            const uint32_t gbase = s.base[v];
            const ViewOut& o = a.out[v];
            if (gbase + cnt > a.cap_visible && tid == 0) *a.overflow = 1u;
            for (uint32_t e = tid; e < cnt; e += kSubThreads) {
                const uint64_t wv = sw[s.list[v][e]];
                const uint32_t g = gbase + e;
                if (g < a.cap_visible) {
                    o.keys[g] = wv >> a.idx_bits;
                    o.idx[g] = (uint32_t)(wv & imask);
                }
            }
            const uint32_t total4 = cnt * 3u;
            float4* dst = o.rows + (size_t)gbase * 3u;
            for (uint32_t j = tid; j < total4; j += kSubThreads) {
                const uint32_t e = j / 3u, part = j - e * 3u;
                const float4 r = s.rows[(uint32_t)s.list[v][e] * 3u + part];
                if (gbase + e < a.cap_visible) dst[j] = r;
            }
        }
    }
    if (tid == 0) bulk_wait_read();  // shared memory must outlive the last bulk store's reads
}

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
    using Smem = SubSmem<GATHER, HAS_PARENT>;
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
