// Fast path of the transform -> cull -> key kernel for sm_100a:
// persistent CTAs, TMA bulk copies of the next tile into shared memory, packed FP32 arithmetic.
//
// Why.  The one-instance-per-thread kernel (transform_cull.cu) is issue-bound and latency-bound, not
// HBM-bound: ~450 of its ~1000 warp instructions per 32 instances are FMUL/FADD that -fmad=false forbids
// to fuse, and every CTA first waits a full DRAM round trip for its tile.  Here
//   * each CTA lives for the whole launch and walks tiles of 128 instances; one thread issues
//     cp.async.bulk (TMA, UBLKCP in SASS) copies of the NEXT tile's records (TRS 36 B, AABB 24 B, ids)
//     into the stage buffer as soon as the current tile has been unpacked into registers; the arrival
//     is tracked by an mbarrier, so HBM latency is hidden behind the arithmetic of the current tile;
//   * each thread carries TWO instances in the halves of 64-bit registers and computes with FFMA2
//     (PTX fma.rn.f32x2): one issue slot does the arithmetic of both, and the uniform work (plane
//     constants, addressing, control) is paid once per pair;
//   * the duplicated frustum-plane constants sit in shared memory (written once per CTA);
//   * the digit histograms the radix sort needs are accumulated in shared memory on the way and flushed
//     once per CTA, which removes the sort's own pass over the keys.
//
// Exactness.  The reference arithmetic is unfused multiplies and adds.  ptxas contracts mul.rn.f32x2 +
// add.rn.f32x2 into FFMA2 even under --fmad=false, so those two instructions cannot be used.  Instead every
// operation is ONE fma with a neutral third operand, which no later pass can fuse further:
//     a*b  = fma(a, b, -0.0)      (x + -0.0 == x for every x, including both zeros)
//     a+b  = fma(a, 1.0, b)       (a*1.0 is exact)
//     a-b  = fma(b, -1.0, a)
// The neutral constants come from kernel parameters (values unknown to ptxas), so they cannot be folded
// back into mul/add either.  Each result is the correctly rounded product/sum/difference: bit-identical to
// the scalar kernel and to the CPU oracle; tests/test_gpu_parity.py runs both kernels against the oracle.
#include <algorithm>

#include "ss_kernels.h"
#include "ss_x2.cuh"

namespace ss {

namespace {

constexpr int kThreads2 = 64;             // threads per CTA: two warps, ten CTAs per SM at 96 registers (128 threads x 5 and
                                          // 128 x 6 at 80 registers were measured slower, 32 x 20 too)
constexpr int kTile2 = 2 * kThreads2;     // instances per tile: thread t owns instances t and t + kThreads2
// One stage is enough: a tile's records are unpacked into registers first thing, the stage is refilled with
// the CTA's next tile right after, and that copy has the whole arithmetic of the current tile to land.
constexpr int kStages = 1;

struct Stage {
    float trs[kTile2 * 9];
    float box[kTile2 * 6];
    uint32_t slot[kTile2];
    uint32_t matword[kTile2];
    int32_t parent[kTile2];
};

// swizzled slot of float4 number k (= instance*4 + row) in a warp's transpose buffer: conflict-free for both
// the per-lane row writes (stride 4 float4) and the linear read-out
__device__ __forceinline__ int row_slot2(int k) { return (k & ~3) | ((k ^ (k >> 3)) & 3); }

struct Smem2 {
    Stage st[kStages];
    float4 tr[kThreads2 / 32][32 * kWorldStride];  // per-warp transpose buffers for the row stores
    f2 planes[SS_MAX_VIEWS * kViewWords];
    uint32_t flags[SS_MAX_VIEWS];
    uint32_t hist[3][kMaxRadix];
    uint64_t bar[kStages];
};

template <bool HAS_PARENT>
__global__ void __launch_bounds__(kThreads2, HAS_PARENT ? 8 : 10) k_instances_x2(const __grid_constant__ InstanceArgs a,
                                                            const __grid_constant__ CullParams cp, const Neutral neutral)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem2& s = *reinterpret_cast<Smem2*>(smem_raw);
    const P2 m{neutral};
    const int tid = threadIdx.x;
    const uint32_t n_tiles = (a.n + kTile2 - 1) / kTile2;
    const bool has_box = a.inst_aabb != nullptr;

    // one thread: bulk copies of tile `tile` into stage `sg`; the stage's mbarrier counts the bytes
    auto issue = [&](uint32_t tile, int sg) {
        const uint32_t base = tile * kTile2;
        uint32_t bytes = kTile2 * 36 + kTile2 * 8 + (has_box ? kTile2 * 24 : 0) + (HAS_PARENT ? kTile2 * 4 : 0);
        mbar_expect_tx(&s.bar[sg], bytes);
        bulk_g2s(s.st[sg].trs, a.trs + ((size_t)a.node_base + base) * 9, kTile2 * 36, &s.bar[sg]);
        if (has_box) bulk_g2s(s.st[sg].box, a.inst_aabb + (size_t)base * 6, kTile2 * 24, &s.bar[sg]);
        bulk_g2s(s.st[sg].slot, a.slot + base, kTile2 * 4, &s.bar[sg]);
        bulk_g2s(s.st[sg].matword, a.matword + base, kTile2 * 4, &s.bar[sg]);
        if (HAS_PARENT) bulk_g2s(s.st[sg].parent, a.parent + a.node_base + base, kTile2 * 4, &s.bar[sg]);
    };
    auto is_full = [&](uint32_t tile) { return (tile + 1) * (uint32_t)kTile2 <= a.n; };

    // ---- per-CTA set-up: barriers, duplicated plane constants, histograms
    if (tid == 0) {
#pragma unroll
        for (int g = 0; g < kStages; ++g) mbar_init(&s.bar[g], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (uint32_t k = tid; k < cp.n_views * 3; k += kThreads2) {
        const uint32_t v = k / 3, ax = k - v * 3;
        const float4 p0 = cp.v[v].pl[2 * ax], p1 = cp.v[v].pl[2 * ax + 1];
        f2* c = s.planes + v * kViewWords + ax * kAxisWords;
        c[0] = dup(p0.x); c[1] = dup(p0.y); c[2] = dup(p0.z);
        c[3] = dup(fabsf(p0.x)); c[4] = dup(fabsf(p0.y)); c[5] = dup(fabsf(p0.z));
        // near plane of a view flagged SS_VIEW_SKIP_NEAR: d = +inf makes the test pass for every finite box
        const bool skip_near = ax == 2 && (cp.v[v].flags & kViewSkipNear);
        c[6] = dup(skip_near ? __int_as_float(0x7f800000) : p0.w); c[7] = dup(p1.w);
        c[8] = dup(p1.x); c[9] = dup(p1.y); c[10] = dup(p1.z);
        c[11] = dup(fabsf(p1.x)); c[12] = dup(fabsf(p1.y)); c[13] = dup(fabsf(p1.z));
    }
    if (tid < (int)SS_MAX_VIEWS) s.flags[tid] = tid < (int)cp.n_views ? cp.v[tid].flags : 0u;
    for (int k = tid; k < 3 * kMaxRadix; k += kThreads2) (&s.hist[0][0])[k] = 0;
    __syncthreads();

    // ---- prologue: fill both stages
    if (tid == 0) {
#pragma unroll
        for (int g = 0; g < kStages; ++g) {
            const uint32_t tile = blockIdx.x + g * gridDim.x;
            if (tile < n_tiles && is_full(tile)) issue(tile, g);
        }
    }

    uint32_t it = 0;
    for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
        const int sg = it % kStages;
        const uint32_t base = tile * kTile2;
        const int n_tile = min((uint32_t)kTile2, a.n - base);
        Stage& st = s.st[sg];
        if (n_tile == kTile2) {
            mbar_wait(&s.bar[sg], (it / kStages) & 1u);
        } else {  // the ragged last tile: plain guarded loads (its sizes are not multiples of 16 B)
            for (int i = tid; i < n_tile * 9; i += kThreads2) st.trs[i] = __ldg(a.trs + ((size_t)a.node_base + base) * 9 + i);
            if (has_box)
                for (int i = tid; i < n_tile * 6; i += kThreads2) st.box[i] = __ldg(a.inst_aabb + (size_t)base * 6 + i);
            for (int i = tid; i < n_tile; i += kThreads2) {
                st.slot[i] = __ldg(a.slot + base + i);
                st.matword[i] = __ldg(a.matword + base + i);
                if (HAS_PARENT) st.parent[i] = __ldg(a.parent + a.node_base + base + i);
            }
            __syncthreads();
        }

        // ---- unpack the two instances of this thread from the stage
        const int t0 = tid, t1 = tid + kThreads2;
        const bool valid0 = t0 < n_tile, valid1 = t1 < n_tile;
        const uint32_t i0 = base + t0, i1 = base + t1;
        const uint32_t slot0 = valid0 ? st.slot[t0] : SS_SLOT_NULL, slot1 = valid1 ? st.slot[t1] : SS_SLOT_NULL;
        const uint32_t mw0 = valid0 ? st.matword[t0] : 0u, mw1 = valid1 ? st.matword[t1] : 0u;
        const bool live0 = valid0 && slot0 != SS_SLOT_NULL, live1 = valid1 && slot1 != SS_SLOT_NULL;
        // dead halves run on the record of instance 0 of the tile (always present): finite numbers, result ignored
        const int u0 = live0 ? t0 : 0, u1 = live1 ? t1 : 0;
        const uint32_t sl0 = live0 ? slot0 : 0u, sl1 = live1 ? slot1 : 0u;
        int p0 = -1, p1 = -1;
        if (HAS_PARENT) {
            p0 = valid0 ? st.parent[t0] : -1;
            p1 = valid1 ? st.parent[t1] : -1;
        }
        // The submesh-table rows (and parent rows) come from L2: issue those loads NOW, in program order, so
        // their latency runs under the ~300 cycles of quaternion/matrix arithmetic that does not need them.
        const SlotDev* sd0 = a.slots + sl0;
        const SlotDev* sd1 = a.slots + sl1;
        const float4 x0 = ld_early_f4(&sd0->r0), x1 = ld_early_f4(&sd0->r1), x2 = ld_early_f4(&sd0->r2);
        const float4 y0 = ld_early_f4(&sd1->r0), y1 = ld_early_f4(&sd1->r1), y2 = ld_early_f4(&sd1->r2);
        const uint64_t kh0 = ld_early_u64(&sd0->key_hi), kh1 = ld_early_u64(&sd1->key_hi);
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
#pragma unroll
        for (int k = 0; k < 9; ++k) t9[k] = pk(st.trs[u0 * 9 + k], st.trs[u1 * 9 + k]);
        if (has_box) {
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                mn[k] = pk(st.box[u0 * 6 + k], st.box[u1 * 6 + k]);
                mx[k] = pk(st.box[u0 * 6 + 3 + k], st.box[u1 * 6 + 3 + k]);
            }
        }
        __syncthreads();  // every thread holds its records in registers: the stage can be refilled
        if (tid == 0) {
            const uint32_t next = tile + kStages * gridDim.x;
            if (next < n_tiles && is_full(next)) issue(next, sg);
        }

        // ---- transform
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
                for (int c = 0; c < 4; ++c)  // instances without a parent keep their own matrix
                    node.r[r][c] = pk(p0 >= 0 ? lo(tmp.r[r][c]) : lo(node.r[r][c]), p1 >= 0 ? hi(tmp.r[r][c]) : hi(node.r[r][c]));
        }
        Aff2 loc, w;
        {
            loc.r[0][0] = pk(x0.x, y0.x); loc.r[0][1] = pk(x0.y, y0.y); loc.r[0][2] = pk(x0.z, y0.z); loc.r[0][3] = pk(x0.w, y0.w);
            loc.r[1][0] = pk(x1.x, y1.x); loc.r[1][1] = pk(x1.y, y1.y); loc.r[1][2] = pk(x1.z, y1.z); loc.r[1][3] = pk(x1.w, y1.w);
            loc.r[2][0] = pk(x2.x, y2.x); loc.r[2][1] = pk(x2.y, y2.y); loc.r[2][2] = pk(x2.z, y2.z); loc.r[2][3] = pk(x2.w, y2.w);
        }
        aff_mul2(m, node, loc, w);  // VC/Scene/SceneRenderer.cpp:1150

        // ---- world rows to HBM, 64 B (or, world_stride == 3, 48 B) per instance.  A lane's rows are contiguous but
        // the lanes of a warp sit a whole instance apart, so they are transposed through a warp-private shared buffer
        // (no block barrier) and leave as fully coalesced 512 B stores per group of 32 instances.
        {
            float4* tr = s.tr[tid >> 5];
            const int lane = tid & 31;
            const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f), last = make_float4(0.f, 0.f, 0.f, 1.f);
            const bool packed_rows = a.world_stride == 3;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const bool live = h ? live1 : live0;
                const int first = (h ? t1 : t0) - lane;  // first instance of this warp's group inside the tile
#pragma unroll
                for (int r = 0; r < 3; ++r) {
                    const float4 row = !live ? zero
                                             : (h ? make_float4(hi(w.r[r][0]), hi(w.r[r][1]), hi(w.r[r][2]), hi(w.r[r][3]))
                                                  : make_float4(lo(w.r[r][0]), lo(w.r[r][1]), lo(w.r[r][2]), lo(w.r[r][3])));
                    // 48 B stride: 12 words per lane, eight consecutive lanes cover all 32 banks as they are
                    tr[packed_rows ? lane * 3 + r : row_slot2(lane * 4 + r)] = row;
                }
                if (!packed_rows) tr[row_slot2(lane * 4 + 3)] = last;
                __syncwarp();
                const int per = packed_rows ? 3 : kWorldStride;
                float4* dst = a.world + ((size_t)base + first) * per;
                const int n_rows = min(max(n_tile - first, 0), 32) * per;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int k = j * 32 + lane;
                    if (k < n_rows) __stcs(dst + k, tr[packed_rows ? k : row_slot2(k)]);  // read once, by a later kernel
                }
                __syncwarp();
            }
        }

        // ---- cull + key
        if (!has_box) {
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                mn[k] = pk(__ldg(&sd0->mn[k]), __ldg(&sd1->mn[k]));
                mx[k] = pk(__ldg(&sd0->mx[k]), __ldg(&sd1->mx[k]));
            }
        }
        const uint2 cm = cull_box2(m, w, mn, mx, s.planes, s.flags, cp.n_views);
        const uint32_t mask0 = live0 ? cm.x : 0u, mask1 = live1 ? cm.y : 0u;
        uint64_t key0 = 2ull << a.tess_shift, key1 = key0;  // dropped: the bit above tess
        if (live0) key0 = kh0 | (uint64_t)(mw0 & ~SS_MATERIAL_TESS_BIT) | ((uint64_t)(mw0 >> 31) << a.tess_shift);
        if (live1) key1 = kh1 | (uint64_t)(mw1 & ~SS_MATERIAL_TESS_BIT) | ((uint64_t)(mw1 >> 31) << a.tess_shift);

        if (a.keys == nullptr) {  // SS_FRAME_REUSE_SORT: masks only
            if (valid0) a.mask[i0] = (uint8_t)mask0;
            if (valid1) a.mask[i1] = (uint8_t)mask1;
            continue;
        }
        const uint32_t sh = a.idx_bits + a.mask_bits;
        if (valid0) {
            a.mask[i0] = (uint8_t)mask0;
            if (a.idx_bits) {
                a.keys[i0] = (key0 << sh) | ((uint64_t)(a.mask_bits ? mask0 : 0u) << a.idx_bits) | i0;
            } else {
                a.keys[i0] = key0;
                a.vals[i0] = i0;
            }
        }
        if (valid1) {
            a.mask[i1] = (uint8_t)mask1;
            if (a.idx_bits) {
                a.keys[i1] = (key1 << sh) | ((uint64_t)(a.mask_bits ? mask1 : 0u) << a.idx_bits) | i1;
            } else {
                a.keys[i1] = key1;
                a.vals[i1] = i1;
            }
        }
        // ---- digit histograms for the sort (digits of the key part only)
#pragma unroll
        for (int p = 0; p < 3; ++p) {
            if (p < a.hist_passes) {
                const uint32_t dm = (1u << a.hist_bits[p]) - 1u;
                if (valid0) atomicAdd(&s.hist[p][(uint32_t)(key0 >> a.hist_shift[p]) & dm], 1u);
                if (valid1) atomicAdd(&s.hist[p][(uint32_t)(key1 >> a.hist_shift[p]) & dm], 1u);
            }
        }
    }

    if (a.hist_passes) {
        __syncthreads();
        for (int k = tid; k < a.hist_passes * kMaxRadix; k += kThreads2) {
            const uint32_t c = (&s.hist[0][0])[k];
            if (c) atomicAdd(a.hist + k, c);
        }
    }
}

}  // namespace

bool instances_x2_eligible(const InstanceArgs& a)
{
    // TRS-driven (no explicit node map, no given world rows), deterministic trig; every bulk copy source must
    // be 16 B aligned: the 36 B / 4 B node streams need node_base % 4 == 0 (tiles start at multiples of kTile2 = 128)
    return !a.world_given && a.node == nullptr && a.trig_mode == SS_TRIG_DETERMINISTIC && a.n > 0 &&
           (a.node_base & 3u) == 0 && (reinterpret_cast<uintptr_t>(a.trs) & 15u) == 0 &&
           (reinterpret_cast<uintptr_t>(a.slot) & 15u) == 0 && (reinterpret_cast<uintptr_t>(a.matword) & 15u) == 0 &&
           (a.parent == nullptr || (reinterpret_cast<uintptr_t>(a.parent) & 15u) == 0) &&
           (a.inst_aabb == nullptr || (reinterpret_cast<uintptr_t>(a.inst_aabb) & 15u) == 0);
}

int instances_x2_hist_capacity() { return 3; }

void launch_instances_x2(const InstanceArgs& a, const CullParams& cp, cudaStream_t st)
{
    Neutral nz;
    const uint32_t one = 0x3f800000u, neg_one = 0xbf800000u, neg_zero = 0x80000000u;
    nz.one = ((uint64_t)one << 32) | one;
    nz.neg_one = ((uint64_t)neg_one << 32) | neg_one;
    nz.neg_zero = ((uint64_t)neg_zero << 32) | neg_zero;
    const uint32_t n_tiles = (a.n + kTile2 - 1) / kTile2;
    // attribute + occupancy are looked up once per device and kernel variant, not per frame
    struct PerDevice {
        int sms = 0, per_sm[2] = {0, 0};
    };
    static PerDevice cache[64];
    int dev = 0;
    cudaGetDevice(&dev);
    PerDevice& pd = cache[dev & 63];
    const size_t smem = sizeof(Smem2);
    const int which = a.parent ? 1 : 0;
    if (!pd.per_sm[which]) {
        int sms_q = 148, per = 1;
        cudaDeviceGetAttribute(&sms_q, cudaDevAttrMultiProcessorCount, dev);
        if (a.parent) {
            cudaFuncSetAttribute(k_instances_x2<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, k_instances_x2<true>, kThreads2, smem);
        } else {
            cudaFuncSetAttribute(k_instances_x2<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, k_instances_x2<false>, kThreads2, smem);
        }
        pd.sms = sms_q;
        pd.per_sm[which] = std::max(per, 1);
    }
    const int sms = pd.sms, per_sm = pd.per_sm[which];
    // persistent: one resident wave; a multiple of the SM count
    const uint32_t grid = std::min<uint32_t>(n_tiles, (uint32_t)(sms * std::max(per_sm, 1)));
    if (a.parent)
        k_instances_x2<true><<<grid, kThreads2, smem, st>>>(a, cp, nz);
    else
        k_instances_x2<false><<<grid, kThreads2, smem, st>>>(a, cp, nz);
}

}  // namespace ss
