// k_submit: the fused transform -> cull -> compact -> emit kernel.  Included by submit.cu only (inside namespace ss,
// anonymous namespace), after the look-back helpers.
//
// Structure.  Persistent CTAs of kSubThreads threads walk tiles of kSubTile = 2 * kSubThreads sorted words; thread t
// carries items t and t + kSubThreads of the tile on packed FP32.  Nothing in a tile's critical path waits for
// another CTA:
//   * a thread fetches ITS OWN two records of the next tile with asynchronous 16-byte copies (cp.async, LDGSTS)
//     into a thread-private, bank-conflict-free corner of the stage, and its words / parents of the next tiles into
//     registers, while it computes the current tile;
//   * the tile's per-view counts are published right after its scan, but its exclusive prefixes (two-level
//     look-back) are only resolved one tile LATER, after the next tile's arithmetic; by then every predecessor has
//     long published, so the walk never polls.  The tile's rows, lists and run table wait in shared memory
//     meanwhile, the next tile's matrices wait in registers, and the output of tile k-1 is written during tile k.
template <bool HAS_PARENT>
struct SubSmem {
    float4 rec[8][kSubThreads];                        // record quarters [half * 4 + q][thread]: thread-private
    float4 rows[kSubTile * 3];                         // TransformData of the staged tile, item-major
    uint64_t words[2][kSubTile];                       // sorted words of the staged tile and of the one before
    f2 planes[SS_MAX_VIEWS * kViewWords];
    uint16_t seg_start[kSubTile + 2][SS_MAX_VIEWS];    // visible-before count at the start of each local run
    uint8_t list[SS_MAX_VIEWS][kSubTile];              // per view: local ids of its visible items, in order
    __align__(16) uint8_t msk[kSubTile];
    uint8_t seg_item[kSubTile + 8];                    // local id of the first item of each run starting in the tile
    uint32_t flags[SS_MAX_VIEWS];
    uint64_t warp_a[kSubWarps], warp_b[kSubWarps];
    uint32_t warp_h[kSubWarps];
    uint32_t base[kSubWords];                          // exclusive prefix over earlier tiles (staged tile)
    uint32_t total[kSubWords];                         // counts of the staged tile
    uint32_t within[kSubWords];                        // counts of the earlier tiles of its look-back group
};

__device__ __forceinline__ void cp_async16(void* dst_smem, const void* src)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ uint64_t ldg_u64(const uint64_t* p)
{
    uint64_t r;
    asm volatile("ld.global.nc.u64 %0, [%1];" : "=l"(r) : "l"(p));
    return r;
}
__device__ __forceinline__ int ldg_s32(const int32_t* p)
{
    int r;
    asm volatile("ld.global.nc.s32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}

// look-back, first half (right after a tile's scan): publish the tile's counts; the tile that closes a group of
// kLbGroup also sums its group (its peers publish at about the same time) and publishes the group's count.
__device__ __forceinline__ void lb_publish(const SubmitArgs& a, uint32_t tile, bool last_tile, int lane, const int (&cid)[kLbPerWarp],
                                           const uint32_t (&mine)[kLbPerWarp], uint32_t (&within)[kLbPerWarp])
{
    uint32_t* lb_tile = a.lookback;
    uint32_t* lb_group = a.lookback + (size_t)kSubWords * a.lb_stride;
    const uint32_t stride_g = a.lb_stride / kLbGroup, g = tile / kLbGroup;
    const bool closes_group = tile % kLbGroup == kLbGroup - 1 || last_tile;
#pragma unroll
    for (int j = 0; j < kLbPerWarp; ++j) {
        within[j] = 0;
        if (lane == j && cid[j] >= 0) st_volatile_u32(lb_tile + (size_t)cid[j] * a.lb_stride + tile, kLbAgg | mine[j]);
    }
    if (closes_group) {
        warp_walk(lb_tile, a.lb_stride, (int64_t)tile - 1, (int64_t)g * kLbGroup, lane, cid, within);
#pragma unroll
        for (int j = 0; j < kLbPerWarp; ++j)
            if (lane == j && cid[j] >= 0) st_volatile_u32(lb_group + (size_t)cid[j] * stride_g + g, kLbAgg | (within[j] + mine[j]));
    }
}

// look-back, second half (one tile later): counts of the earlier tiles of the group, then the groups before it; the
// closing tile publishes the group's inclusive prefix.  excl[j] = exclusive prefix of counter cid[j].
__device__ __forceinline__ void lb_resolve(const SubmitArgs& a, uint32_t tile, bool last_tile, int lane, const int (&cid)[kLbPerWarp],
                                           const uint32_t (&mine)[kLbPerWarp], const uint32_t (&within_in)[kLbPerWarp],
                                           uint32_t (&excl)[kLbPerWarp])
{
    uint32_t* lb_tile = a.lookback;
    uint32_t* lb_group = a.lookback + (size_t)kSubWords * a.lb_stride;
    const uint32_t stride_g = a.lb_stride / kLbGroup, g = tile / kLbGroup;
    const bool closes_group = tile % kLbGroup == kLbGroup - 1 || last_tile;
    uint32_t within[kLbPerWarp], before[kLbPerWarp];
    if (closes_group) {
#pragma unroll
        for (int j = 0; j < kLbPerWarp; ++j) within[j] = within_in[j];
    } else {
        warp_walk(lb_tile, a.lb_stride, (int64_t)tile - 1, (int64_t)g * kLbGroup, lane, cid, within);
    }
    warp_walk(lb_group, stride_g, (int64_t)g - 1, 0, lane, cid, before);
    if (closes_group) {
#pragma unroll
        for (int j = 0; j < kLbPerWarp; ++j)
            if (lane == j && cid[j] >= 0)
                st_volatile_u32(lb_group + (size_t)cid[j] * stride_g + g, kLbPrefix | (before[j] + within[j] + mine[j]));
    }
#pragma unroll
    for (int j = 0; j < kLbPerWarp; ++j) excl[j] = before[j] + within[j];
}

// Output of the staged tile (tile index `tile`, words buffer sw): resolve its prefixes, then write run bookkeeping,
// keys, ids and rows of every view in draw order.  Called by the whole CTA; contains two barriers.
template <bool HAS_PARENT>
__device__ __forceinline__ void emit_staged_tile(const SubmitArgs& a, SubSmem<HAS_PARENT>& s, const uint64_t* sw, uint32_t tile,
                                                 uint32_t n_tiles, uint64_t imask, int tid, int lane, int wid)
{
    {
        int cid[kLbPerWarp];
        uint32_t mine[kLbPerWarp], within[kLbPerWarp], excl[kLbPerWarp];
#pragma unroll
        for (int j = 0; j < kLbPerWarp; ++j) {
            const uint32_t i = (uint32_t)wid + (uint32_t)kSubWarps * j;
            cid[j] = i < a.n_views ? (int)i : (i == a.n_views ? (int)SS_MAX_VIEWS : -1);
            mine[j] = cid[j] >= 0 ? s.total[cid[j]] : 0u;
            within[j] = cid[j] >= 0 ? s.within[cid[j]] : 0u;
        }
        lb_resolve(a, tile, tile == n_tiles - 1, lane, cid, mine, within, excl);
#pragma unroll
        for (int j = 0; j < kLbPerWarp; ++j) {
            if (lane == j && cid[j] >= 0) {
                s.base[cid[j]] = excl[j];
                if (tile == n_tiles - 1) a.totals[cid[j]] = excl[j] + mine[j];
            }
        }
    }
    __syncthreads();  // B1: bases visible

    // run bookkeeping: first key / first instance of each run, visible count per (view, run)
    const uint32_t n_heads = s.total[SS_MAX_VIEWS];
    const uint32_t seg_base = s.base[SS_MAX_VIEWS];
    for (uint32_t q = tid; q <= n_heads; q += kSubThreads) {
        const uint32_t g = seg_base + q - 1u;  // q == 0: run that started in an earlier tile
        if (q > 0) {
            if (g < a.max_draws) {
                const uint64_t wv = sw[s.seg_item[q]];
                a.seg_key[g] = wv >> a.idx_bits;
                a.seg_inst[g] = (uint32_t)(wv & imask);
            } else {
                *a.overflow = 1u;
            }
        }
#pragma unroll
        for (int v = 0; v < (int)SS_MAX_VIEWS; ++v) {
            if (v < (int)a.n_views) {
                const uint32_t c = (uint32_t)s.seg_start[q + 1][v] - (uint32_t)s.seg_start[q][v];
                if (c && g < a.max_draws) atomicAdd(a.seg_count + (size_t)v * a.max_draws + g, c);
            }
        }
    }
    // per view: keys + instance ids, then the rows, all as coalesced stores in draw order
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
    if (tid == 0) bulk_wait_read();  // the staged tile's bulk stores have read rows / masks
    __syncthreads();                 // B2: the staging buffers may be overwritten
}

template <bool GATHER, bool HAS_PARENT>
__global__ void __launch_bounds__(kSubThreads, 5) k_submit(const __grid_constant__ SubmitArgs a, const __grid_constant__ CullParams cp,
                                                           const Neutral neutral)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    using Smem = SubSmem<HAS_PARENT>;
    Smem& s = *reinterpret_cast<Smem*>(smem_raw);
    const P2 m{neutral};
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t n_tiles = (a.n + kSubTile - 1) / kSubTile, G = gridDim.x;
    const uint64_t* __restrict__ words = (a.sort_ctl && a.sort_ctl->src_sel[a.sort_passes]) ? a.words_b : a.words_a;
    const uint64_t imask = (1ull << a.idx_bits) - 1ull;
    const bool has_box = a.has_box != 0;

    // item h (0: tid, 1: tid + kSubThreads) of a tile: its position in word order, its word, its parent, its record
    auto item_pos = [&](uint32_t tile, int h) { return tile * (uint32_t)kSubTile + (uint32_t)tid + (uint32_t)h * kSubThreads; };
    auto load_word = [&](uint32_t tile, int h) -> uint64_t {
        const uint32_t j = item_pos(tile, h);
        return j < a.n ? ldg_u64(words + j) : 0ull;
    };
    auto load_parent = [&](uint32_t tile, int h, uint64_t w) -> int {
        const uint32_t j = item_pos(tile, h);
        if (j >= a.n) return -1;
        return ldg_s32(a.parent + (GATHER ? (size_t)(w & imask) : (size_t)j));
    };
    auto issue_records = [&](uint32_t tile, uint64_t w0, uint64_t w1) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const uint32_t j = item_pos(tile, h);
            if (j < a.n) {
                const float4* src = reinterpret_cast<const float4*>(a.rec + (GATHER ? (size_t)((h ? w1 : w0) & imask) : (size_t)j));
#pragma unroll
                for (int q = 0; q < 4; ++q) cp_async16(&s.rec[h * 4 + q][tid], src + q);
            }
        }
        cp_async_commit();
    };

    // ---- per-CTA set-up: duplicated plane constants
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

    // ---- prologue: words (and parents) of the first tile in registers, its records on their way
    uint64_t wc0 = load_word(blockIdx.x, 0), wc1 = load_word(blockIdx.x, 1);
    uint64_t wn0 = 0, wn1 = 0;
    int pc0 = -1, pc1 = -1, pn0 = -1, pn1 = -1;
    issue_records(blockIdx.x, wc0, wc1);
    if (HAS_PARENT) {
        pc0 = load_parent(blockIdx.x, 0, wc0);
        pc1 = load_parent(blockIdx.x, 1, wc1);
    }
    if (GATHER && blockIdx.x + G < n_tiles) {
        wn0 = load_word(blockIdx.x + G, 0);
        wn1 = load_word(blockIdx.x + G, 1);
    }

    uint32_t k = 0, staged_tile = 0;
    for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += G, ++k) {
        const uint32_t base = tile * kSubTile;
        const int n_tile = min((uint32_t)kSubTile, a.n - base);
        cp_async_wait_all();  // this thread's two records have landed in its own corner of the stage

        // ---- unpack
        const int t0 = tid, t1 = tid + kSubThreads;
        const bool valid0 = t0 < n_tile, valid1 = t1 < n_tile;
        const float4 A0 = s.rec[0][tid], A1 = s.rec[1][tid], A2 = s.rec[2][tid], A3 = s.rec[3][tid];
        const float4 B0 = s.rec[4][tid], B1 = s.rec[5][tid], B2 = s.rec[6][tid], B3 = s.rec[7][tid];
        const uint32_t slot0 = __float_as_uint(A3.w), slot1 = __float_as_uint(B3.w);
        const bool live0 = valid0 && slot0 != SS_SLOT_NULL, live1 = valid1 && slot1 != SS_SLOT_NULL;
        const uint32_t sl0 = live0 ? slot0 : 0u, sl1 = live1 ? slot1 : 0u;
        const int p0 = (HAS_PARENT && valid0) ? pc0 : -1, p1 = (HAS_PARENT && valid1) ? pc1 : -1;
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

        // ---- prefetch for this CTA's next tile(s).  The copies overwrite this thread's own corner of the stage: its
        // loads above were issued first and the load/store unit keeps a thread's shared-memory accesses in order.
        uint64_t wf0 = 0, wf1 = 0, prevw = 0;
        {
            const uint32_t next = tile + G;
            if (next < n_tiles) {
                if (!GATHER) {
                    wn0 = load_word(next, 0);
                    wn1 = load_word(next, 1);
                }
                issue_records(next, wn0, wn1);
                if (HAS_PARENT) {
                    pn0 = load_parent(next, 0, wn0);
                    pn1 = load_parent(next, 1, wn1);
                }
                if (GATHER && next + G < n_tiles) {
                    wf0 = load_word(next + G, 0);
                    wf1 = load_word(next + G, 1);
                }
            }
            if (tid == 0 && tile > 0) prevw = ldg_u64(words + base - 1);
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

        // ---- cull
        if (!has_box) {
#pragma unroll
            for (int q = 0; q < 3; ++q) {
                mn[q] = pk(__ldg(&sd0->mn[q]), __ldg(&sd1->mn[q]));
                mx[q] = pk(__ldg(&sd0->mx[q]), __ldg(&sd1->mx[q]));
            }
        }
        const uint2 cm = cull_box2(m, w, mn, mx, s.planes, s.flags, cp.n_views);
        const uint32_t mask0 = live0 ? cm.x : 0u, mask1 = live1 ? cm.y : 0u;

        // ---- the tile staged one iteration ago leaves now: its prefixes had this tile's arithmetic to become known
        if (a.rows_only) {
            if (tid == 0) bulk_wait_read();  // the previous tile's bulk stores have read the staging buffers
            __syncthreads();
        } else if (k > 0) {
            emit_staged_tile<HAS_PARENT>(a, s, s.words[(k - 1) & 1u], staged_tile, n_tiles, imask, tid, lane, wid);
        }

        // ---- stage this tile: rows (48 B per item: a lane's three float4 stores hit distinct banks), masks, words
        uint64_t* sw = s.words[k & 1u];
        {
            const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                s.rows[t0 * 3 + r] = live0 ? make_float4(lo(w.r[r][0]), lo(w.r[r][1]), lo(w.r[r][2]), lo(w.r[r][3])) : zero;
                s.rows[t1 * 3 + r] = live1 ? make_float4(hi(w.r[r][0]), hi(w.r[r][1]), hi(w.r[r][2]), hi(w.r[r][3])) : zero;
            }
            s.msk[t0] = (uint8_t)mask0;
            s.msk[t1] = (uint8_t)mask1;
            sw[t0] = wc0;
            sw[t1] = wc1;
        }
        fence_async_smem();  // rows / masks become visible to the bulk stores below
        __syncthreads();     // B3

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
        if (a.rows_only) {  // split pipeline: k_emit compacts the rows from here
            wc0 = wn0; wc1 = wn1;
            pc0 = pn0; pc1 = pn1;
            if (GATHER) {
                wn0 = wf0;
                wn1 = wf1;
            }
            continue;
        }

        // ---- ranks: thread t now owns items 2t and 2t + 1 of the tile (word order)
        const uint32_t l0 = tid * 2;
        uint64_t key[2];
        uint32_t mm[2];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const bool ok = l0 + j < (uint32_t)n_tile;
            key[j] = ok ? sw[l0 + j] >> a.idx_bits : ~0ull;
            mm[j] = ok ? (uint32_t)s.msk[l0 + j] : 0u;
        }
        uint64_t prev = 0;
        bool first_is_head = false;
        if (tid > 0) {
            prev = sw[l0 - 1] >> a.idx_bits;
        } else if (tile > 0) {
            prev = prevw >> a.idx_bits;
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
        __syncthreads();  // B4
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
        // every thread knows the tile totals: counter wq of the tile
        auto tile_total = [&](int wq) -> uint32_t {
            return wq < 4 ? (uint32_t)(ta >> (16 * wq)) & 0xffffu : wq < 8 ? (uint32_t)(tb >> (16 * (wq - 4))) & 0xffffu : th;
        };

        // ---- publish the tile's counts (look-back, first half).  The active counters (n_views visible counts + the
        // run count) are dealt out over the warps: slot i = wid + kSubWarps * j -> counter i (i < n_views), the run
        // counter (i == n_views), none.
        {
            int cid[kLbPerWarp];
            uint32_t mine[kLbPerWarp], within[kLbPerWarp];
#pragma unroll
            for (int j = 0; j < kLbPerWarp; ++j) {
                const uint32_t i = (uint32_t)wid + (uint32_t)kSubWarps * j;
                cid[j] = i < a.n_views ? (int)i : (i == a.n_views ? (int)SS_MAX_VIEWS : -1);
                mine[j] = cid[j] >= 0 ? tile_total(cid[j]) : 0u;
            }
            lb_publish(a, tile, tile == n_tiles - 1, lane, cid, mine, within);
#pragma unroll
            for (int j = 0; j < kLbPerWarp; ++j) {
                if (lane == j && cid[j] >= 0) {
                    s.total[cid[j]] = mine[j];
                    s.within[cid[j]] = within[j];
                }
            }
            __syncwarp();  // read back by every lane of this warp when the tile is emitted
        }

        // ---- local lists: where each visible item lands inside the tile, and where each key run starts
        const uint32_t n_heads = th;
        if (tid >= kSubThreads - (int)SS_MAX_VIEWS) {  // the sentinels of the run table
            const int v = tid - (kSubThreads - (int)SS_MAX_VIEWS);
            s.seg_start[0][v] = 0;
            s.seg_start[n_heads + 1][v] = (uint16_t)tile_total(v);
        }
        {
            uint64_t fa = ba + xa - pa, fb = bb + xb - pb;  // exclusive prefixes of this thread's first item
            uint32_t fh = bh + xh - ph;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                if (head[j]) {
                    const uint32_t lseg = fh + 1;  // 1-based among runs starting in this tile; 0 = run carried in
                    s.seg_item[lseg] = (uint8_t)(l0 + j);
#pragma unroll
                    for (int v = 0; v < 4; ++v) {
                        s.seg_start[lseg][v] = (uint16_t)(fa >> (16 * v));
                        s.seg_start[lseg][4 + v] = (uint16_t)(fb >> (16 * v));
                    }
                }
#pragma unroll
                for (int v = 0; v < (int)SS_MAX_VIEWS; ++v) {
                    if (v < (int)a.n_views && ((mm[j] >> v) & 1u)) {
                        const uint32_t lpos = (uint32_t)((v < 4 ? fa >> (16 * v) : fb >> (16 * (v - 4))) & 0xffffu);
                        s.list[v][lpos] = (uint8_t)(l0 + j);
                    }
                }
                fa += spread4(mm[j] & 15u);
                fb += spread4(mm[j] >> 4);
                fh += head[j] ? 1u : 0u;
            }
        }

        // ---- rotate the registers that run ahead of the tile
        staged_tile = tile;
        wc0 = wn0; wc1 = wn1;
        pc0 = pn0; pc1 = pn1;
        if (GATHER) {
            wn0 = wf0;
            wn1 = wf1;
        }
    }
    // the last staged tile
    if (a.rows_only) {
        if (tid == 0) bulk_wait_read();
        return;
    }
    if (k > 0) emit_staged_tile<HAS_PARENT>(a, s, s.words[(k - 1) & 1u], staged_tile, n_tiles, imask, tid, lane, wid);
}
