// Consumer of the compacted lists: depth + id rasterisation of one view's draws into a device
// buffer, as CUDA.  It stands where the reference binds the instance buffer and issues its draws:
//   * ShadowPass  (VulkanCore/assets/shaders/ShadowDepth.vert + .geom: world = transform * position,
//     clip = LightSpaceMatrices[cascade] * world; depth only, DepthClamp on — SceneRenderer.cpp:188)
//   * SelectionPass (CoreEditor.vert / .frag: clip = Projection * View * world, the fragment stores a_EntityID;
//     the editor reads ONE pixel back, SceneRenderer.cpp:1449-1456)
// Both passes are the same work per (instance row, triangle); the target word is [depth bits : 32 | id : 32] and the
// depth test is one 64-bit atomicMin, so the result does not depend on the order triangles arrive in.
//
// Per triangle, the fixed-function steps of that pipeline:
//   1. clip = view_proj * (transform * position)
//   2. clip against the near plane z >= 0 (with DepthClamp: against w >= 1e-5, depth then clamped per sample); new
//      vertices are interpolated from the inside vertex towards the outside one, so neighbours agree on them
//   3. perspective divide, viewport (y down), snap to 1/256 pixel; coverage by exact integer edge functions at pixel
//      centres, a shared edge belongs to exactly one of its triangles; depth = screen-linear z/w
//   3b. a triangle that cannot be snapped (w <= 0 left after clipping, or a vertex farther than 2^20 pixels out) is
//      rasterised in homogeneous float coordinates (edge functions = adjugate rows of the (X, Y, W) matrix) over the
//      whole target
// Every float operation is written out in a fixed order (-fmad=false), the order the tests' CPU restatement uses:
// the two agree bit for bit.  The reference's own rasteriser is the Vulkan device: unavailable on this box.
#include <algorithm>
#include <cmath>

#include "ss_kernels.h"

namespace ss {

namespace {

constexpr int kRasterThreads = 128;
constexpr int kCoopSamples = 192;             // bounds holding more samples than this get the whole warp to themselves
constexpr long long kHugeSamples = 16384;      // beyond this a triangle is queued and spread over the grid, tile by tile
constexpr int kWorkWords = 20;                // queue record: Work::w[13], x0, x1, y0, y1, mode, instance, first tile entry
constexpr float kSubpixel = 256.0f;
constexpr float kSnapLimit = 268435456.0f;    // 2^28 sub-pixel units = 2^20 pixels

// A triangle ready to be sampled.  mode 0 (snapped): w = x[3], y[3] as int32 in 1/256 pixel, zn[3] = z/w per vertex.
// mode 1 (homogeneous): w = a[3], b[3], c[3] edge functions, z[3], det.
struct Work {
    uint32_t w[13];
    int x0, x1, y0, y1;  // inclusive sample bounds inside the target
    int mode;            // -1: nothing to draw
};

__device__ __forceinline__ void clip_of(const float* vp, float4 r0, float4 r1, float4 r2, const float* p, float out[4])
{
    // world = transform * (p, 1); clip = view_proj * (world, 1): column-major m * v as (c0 x + c1 y) + (c2 z + c3)
    const float wx = (r0.x * p[0] + r0.y * p[1]) + (r0.z * p[2] + r0.w);
    const float wy = (r1.x * p[0] + r1.y * p[1]) + (r1.z * p[2] + r1.w);
    const float wz = (r2.x * p[0] + r2.y * p[1]) + (r2.z * p[2] + r2.w);
#pragma unroll
    for (int j = 0; j < 4; ++j) out[j] = (vp[j] * wx + vp[4 + j] * wy) + (vp[8 + j] * wz + vp[12 + j]);
}

__device__ __forceinline__ Work setup(const float c[3][4], int width, int height)
{
    Work t;
    t.mode = -1;
    t.x0 = t.y0 = 0;
    t.x1 = t.y1 = -1;
    const float half_w = 0.5f * (float)width, half_h = 0.5f * (float)height;
    int x[3], y[3];
    float zn[3];
    bool snap = true;
#pragma unroll
    for (int v = 0; v < 3; ++v) {
        if (!(c[v][3] > 0.f)) { snap = false; break; }
        const float iw = 1.0f / c[v][3];
        const float sx = ((c[v][0] * iw) * half_w + half_w) * kSubpixel;
        const float sy = ((c[v][1] * iw) * half_h + half_h) * kSubpixel;
        if (!(sx > -kSnapLimit && sx < kSnapLimit && sy > -kSnapLimit && sy < kSnapLimit)) { snap = false; break; }
        x[v] = (int)__float2ll_rn(sx);
        y[v] = (int)__float2ll_rn(sy);
        zn[v] = c[v][2] * iw;
    }
    if (snap) {
        const long long area = (long long)(x[1] - x[0]) * (y[2] - y[0]) - (long long)(y[1] - y[0]) * (x[2] - x[0]);
        if (area == 0) return t;
        const int minx = min(min(x[0], x[1]), x[2]), maxx = max(max(x[0], x[1]), x[2]);
        const int miny = min(min(y[0], y[1]), y[2]), maxy = max(max(y[0], y[1]), y[2]);
        // pixel i is sampled at 256 i + 128; >> 8 is the floor division
        t.x0 = max((minx - 128 + 255) >> 8, 0);
        t.x1 = min((maxx - 128) >> 8, width - 1);
        t.y0 = max((miny - 128 + 255) >> 8, 0);
        t.y1 = min((maxy - 128) >> 8, height - 1);
        if (t.x0 > t.x1 || t.y0 > t.y1) return t;  // covers no sample
#pragma unroll
        for (int v = 0; v < 3; ++v) {
            t.w[v] = (uint32_t)x[v];
            t.w[3 + v] = (uint32_t)y[v];
            t.w[6 + v] = __float_as_uint(zn[v]);
        }
        t.mode = 0;
        return t;
    }
    // homogeneous pixel-space coordinates: X / W is the pixel abscissa, Y / W the ordinate
    const float X0 = (c[0][0] + c[0][3]) * half_w, X1 = (c[1][0] + c[1][3]) * half_w, X2 = (c[2][0] + c[2][3]) * half_w;
    const float Y0 = (c[0][1] + c[0][3]) * half_h, Y1 = (c[1][1] + c[1][3]) * half_h, Y2 = (c[2][1] + c[2][3]) * half_h;
    const float W0 = c[0][3], W1 = c[1][3], W2 = c[2][3];
    float a[3], b[3], cc[3];
    a[0] = Y1 * W2 - Y2 * W1; b[0] = X2 * W1 - X1 * W2; cc[0] = X1 * Y2 - X2 * Y1;
    a[1] = Y2 * W0 - Y0 * W2; b[1] = X0 * W2 - X2 * W0; cc[1] = X2 * Y0 - X0 * Y2;
    a[2] = Y0 * W1 - Y1 * W0; b[2] = X1 * W0 - X0 * W1; cc[2] = X0 * Y1 - X1 * Y0;
    const float det = (X0 * a[0] + X1 * a[1]) + X2 * a[2];
    if (!(det > 0.f || det < 0.f)) return t;                  // degenerate, or NaN
    if (!(W0 > 0.f || W1 > 0.f || W2 > 0.f)) return t;        // wholly behind
#pragma unroll
    for (int q = 0; q < 3; ++q) {
        t.w[q] = __float_as_uint(a[q]);
        t.w[3 + q] = __float_as_uint(b[q]);
        t.w[6 + q] = __float_as_uint(cc[q]);
        t.w[9 + q] = __float_as_uint(c[q][2]);
    }
    t.w[12] = __float_as_uint(det);
    t.x0 = 0; t.x1 = width - 1; t.y0 = 0; t.y1 = height - 1;
    t.mode = 1;
    return t;
}

__device__ __forceinline__ void put(unsigned long long* target, uint32_t width, int i, int j, float depth, bool clamp_depth,
                                    uint32_t inst)
{
    if (clamp_depth) depth = depth < 0.f ? 0.f : (depth > 1.f ? 1.f : depth);  // a NaN stays one and is dropped
    if (!(depth >= 0.f && depth <= 1.f)) return;
    atomicMin(target + (size_t)j * width + i, ((unsigned long long)__float_as_uint(depth) << 32) | inst);
}

// Conservative lower bound of every depth a snapped triangle can store, as float bits (0: no bound).  A sample's depth
// is fl(fl(fl(z0 l0) + fl(z1 l1)) + fl(z2 l2)) with l_q = fl(fl(e_q) / fl(area)), e_q >= 0 and sum e_q = area: for
// z_q >= 0 that is at least min z * (1 - 9 * 2^-24).  With DepthClamp the bound stops at 1, the largest depth stored.
__device__ __forceinline__ uint32_t depth_floor_bits(float z0, float z1, float z2, bool clamp_depth)
{
    const float lo = fminf(fminf(z0, z1), z2);
    if (!(lo > 0.f)) return 0u;
    float zc = lo * 0.99999905f;  // 1 - 2^-20
    if (clamp_depth) zc = fminf(zc, 1.0f);
    return __float_as_uint(zc);
}

constexpr int kCellShift = 3;  // coarse depth cells of 8 x 8 samples

// what the sample loop needs, derived once per triangle
struct Prep {
    int x[3], y[3], dx[3], dy[3];  // mode 0; edge q runs from vertex q+1 to vertex q+2 (opposite vertex q); dx, dy carry the
                                   // triangle's orientation sign, so that the inside is the positive side
    float fa, zn[3];
    float a[3], b[3], c[3], z[3], det, sf;  // mode 1
};

__device__ __forceinline__ uint32_t work_depth_floor(const Work& t, bool clamp_depth)
{
    return t.mode == 0 ? depth_floor_bits(__uint_as_float(t.w[6]), __uint_as_float(t.w[7]), __uint_as_float(t.w[8]), clamp_depth) : 0u;
}

__device__ __forceinline__ Prep prepare(const Work& t)
{
    Prep p;
    if (t.mode == 0) {
#pragma unroll
        for (int v = 0; v < 3; ++v) {
            p.x[v] = (int)t.w[v];
            p.y[v] = (int)t.w[3 + v];
            p.zn[v] = __uint_as_float(t.w[6 + v]);
        }
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            const int p0 = (q + 1) % 3, p1 = (q + 2) % 3;
            p.dx[q] = p.x[p1] - p.x[p0];
            p.dy[q] = p.y[p1] - p.y[p0];
        }
        const long long area = (long long)p.dx[2] * (p.y[2] - p.y[0]) - (long long)p.dy[2] * (p.x[2] - p.x[0]);
        if (area < 0) {
#pragma unroll
            for (int q = 0; q < 3; ++q) {
                p.dx[q] = -p.dx[q];
                p.dy[q] = -p.dy[q];
            }
        }
        p.fa = (float)(area < 0 ? -area : area);
    } else {
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            p.a[q] = __uint_as_float(t.w[q]);
            p.b[q] = __uint_as_float(t.w[3 + q]);
            p.c[q] = __uint_as_float(t.w[6 + q]);
            p.z[q] = __uint_as_float(t.w[9 + q]);
        }
        p.det = __uint_as_float(t.w[12]);
        p.sf = p.det > 0.f ? 1.f : -1.f;
    }
    return p;
}

__device__ __forceinline__ void shade(const Prep& p, int mode, int i, int j, bool clamp_depth, uint32_t inst,
                                      unsigned long long* target, uint32_t width)
{
    if (mode == 0) {
        const int px = 256 * i + 128, py = 256 * j + 128;
        long long e[3];
        bool inside = true;
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            const int p0 = (q + 1) % 3;
            e[q] = (long long)p.dx[q] * (py - p.y[p0]) - (long long)p.dy[q] * (px - p.x[p0]);
            // a sample exactly on an edge belongs to the triangle for which that edge runs down, or level to the right
            inside = inside && (e[q] > 0 || (e[q] == 0 && (p.dy[q] > 0 || (p.dy[q] == 0 && p.dx[q] > 0))));
        }
        if (!inside) return;
        const float l0 = (float)e[0] / p.fa, l1 = (float)e[1] / p.fa, l2 = (float)e[2] / p.fa;
        put(target, width, i, j, (p.zn[0] * l0 + p.zn[1] * l1) + p.zn[2] * l2, clamp_depth, inst);
    } else {
        const float px = (float)i + 0.5f, py = (float)j + 0.5f;
        float e[3];
        bool inside = true;
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            e[q] = (p.a[q] * px + p.b[q] * py) + p.c[q];
            const float es = p.sf * e[q], as = p.sf * p.a[q], bs = p.sf * p.b[q];
            inside = inside && (es > 0.f || (es == 0.f && (as > 0.f || (as == 0.f && bs > 0.f))));
        }
        if (!inside) return;
        if (!(p.sf * ((e[0] + e[1]) + e[2]) > 0.f)) return;  // 1 / w at the sample must be positive
        put(target, width, i, j, ((p.z[0] * e[0] + p.z[1] * e[1]) + p.z[2] * e[2]) / p.det, clamp_depth, inst);
    }
}

// True when the samples x0..x1 x y0..y1 all lie outside one edge of a snapped triangle.  Exact: an edge function is
// linear, so it is negative over the rectangle iff it is negative at the corner it favours (a sample ON an edge may
// belong to the triangle, hence < and not <=).
__device__ __forceinline__ bool rect_outside(const Prep& p, int x0, int x1, int y0, int y1)
{
    bool out = false;
#pragma unroll
    for (int q = 0; q < 3; ++q) {
        const int p0 = (q + 1) % 3;
        // E = dx (py - y0) - dy (px - x0) is largest where dx favours py and -dy favours px
        const int py = 256 * (p.dx[q] >= 0 ? y1 : y0) + 128, px = 256 * (p.dy[q] <= 0 ? x1 : x0) + 128;
        const long long best = (long long)p.dx[q] * (py - p.y[p0]) - (long long)p.dy[q] * (px - p.x[p0]);
        out = out || best < 0;
    }
    return out;
}

// One warp samples the rectangle x0..x1 x y0..y1 of one triangle, coarse cell by coarse cell: the lanes first sort out
// which cells can matter (a cell whose coarse depth lies in front of everything the triangle can store, or that lies
// outside an edge, is skipped whole), then take the samples of each remaining cell two apiece.  `zc` = 0 or no coarse
// buffer: every cell is sampled.  Skipping never changes the target: a skipped sample could not have won its cell.
__device__ __forceinline__ void warp_sample_rect(const Prep& p, int mode, int x0, int x1, int y0, int y1, uint32_t zc,
                                                 bool clamp_depth, uint32_t inst, const RasterArgs& a, int lane)
{
    const int cx0 = x0 >> kCellShift, cy0 = y0 >> kCellShift;
    const int ncx = (x1 >> kCellShift) - cx0 + 1, ncy = (y1 >> kCellShift) - cy0 + 1;
    const int n_cells = ncx * ncy;
    for (int base = 0; base < n_cells; base += 32) {
        const int c = base + lane;
        bool alive = c < n_cells;
        int cx = 0, cy = 0;
        if (alive) {
            cy = c / ncx;
            cx = cx0 + (c - cy * ncx);
            cy += cy0;
            if (a.hiz && zc > __ldg(a.hiz + (size_t)cy * a.hiz_w + cx)) alive = false;
            if (alive && mode == 0)
                alive = !rect_outside(p, max(x0, cx << kCellShift), min(x1, (cx << kCellShift) + 7), max(y0, cy << kCellShift),
                                      min(y1, (cy << kCellShift) + 7));
        }
        uint32_t todo = __ballot_sync(0xffffffffu, alive);
        while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            const int ccx = __shfl_sync(0xffffffffu, cx, src), ccy = __shfl_sync(0xffffffffu, cy, src);
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int sidx = lane + 32 * h;
                const int i = (ccx << kCellShift) + (sidx & 7), j = (ccy << kCellShift) + (sidx >> 3);
                if (i >= x0 && i <= x1 && j >= y0 && j <= y1) shade(p, mode, i, j, clamp_depth, inst, a.target, a.width);
            }
        }
    }
}

// hiz[cell] = the largest depth bits among the cell's samples (all ones while any of them is empty).  The target's
// words only ever fall, so the value stays an upper bound for as long as it is used.
__global__ void __launch_bounds__(256) k_raster_hiz(const unsigned long long* __restrict__ target, uint32_t width, uint32_t height,
                                                    uint32_t* __restrict__ hiz, uint32_t hiz_w, uint32_t hiz_h)
{
    const uint32_t n = hiz_w * hiz_h;
    for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < n; c += gridDim.x * blockDim.x) {
        const uint32_t cy = c / hiz_w, cx = c - cy * hiz_w;
        const uint32_t x0 = cx << kCellShift, y0 = cy << kCellShift;
        uint32_t m = 0;
        const uint32_t x1 = min(x0 + 8u, width), y1 = min(y0 + 8u, height);
        for (uint32_t j = y0; j < y1; ++j)
            for (uint32_t i = x0; i < x1; ++i) m = max(m, (uint32_t)(__ldcg(target + (size_t)j * width + i) >> 32));
        hiz[c] = m;
    }
}

// pairs[d] = number of (instance, triangle) pairs of the draws before d; pairs[n_draws] = all of them
__global__ void __launch_bounds__(1024) k_raster_pairs(const ss_draw_cmd* draws, const uint32_t* counts, uint32_t max_draws,
                                                       unsigned long long* pairs)
{
    __shared__ unsigned long long warp_sum[32];
    __shared__ unsigned long long carry;
    const uint32_t n_draws = min(counts[1], max_draws);
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (uint32_t d0 = 0; d0 < n_draws; d0 += 1024) {
        const uint32_t d = d0 + tid;
        const unsigned long long mine = d < n_draws ? (unsigned long long)draws[d].instance_count * (draws[d].index_count / 3u) : 0ull;
        unsigned long long x = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_sum[wid] = x;
        __syncthreads();
        unsigned long long base = carry;
        for (int w = 0; w < wid; ++w) base += warp_sum[w];
        if (d < n_draws) pairs[d] = base + x - mine;
        __syncthreads();
        if (tid == 1023) carry = base + x;
        __syncthreads();
    }
    if (tid == 0) pairs[n_draws] = carry;
}

// A triangle its thread does not sample itself: queued for k_raster_tiles (one entry per 64 x 64-sample tile of its
// bounds, spread over the whole grid), or, when the queue is full, walked here sample by sample.  Kept out of line:
// it is the rare path and must not cost the common one its registers.
__device__ __noinline__ void queue_or_walk(const Work& t, uint32_t inst, const RasterArgs& a)
{
    const bool clamp_depth = (a.flags & SS_RASTER_DEPTH_CLAMP) != 0;
    if (a.queue_works) {
        const int tx0 = t.x0 >> 6, ty0 = t.y0 >> 6;
        const uint32_t n_tiles = (uint32_t)(((t.x1 >> 6) - tx0 + 1) * ((t.y1 >> 6) - ty0 + 1));
        const uint32_t slot_w = atomicAdd(a.queue_counts + 0, 1u);
        if (slot_w < a.queue_work_cap) {
            const uint32_t base_t = atomicAdd(a.queue_counts + 1, n_tiles);
            if ((unsigned long long)base_t + n_tiles <= a.queue_tile_cap) {
                uint32_t* rec = a.queue_works + (size_t)slot_w * kWorkWords;
#pragma unroll
                for (int q = 0; q < 13; ++q) rec[q] = t.w[q];
                rec[13] = (uint32_t)t.x0; rec[14] = (uint32_t)t.x1; rec[15] = (uint32_t)t.y0; rec[16] = (uint32_t)t.y1;
                rec[17] = (uint32_t)t.mode; rec[18] = inst; rec[19] = base_t;
                for (uint32_t k = 0; k < n_tiles; ++k) a.queue_tiles[base_t + k] = slot_w;
                return;
            }
            // the tile list is full: the part of the range that exists is marked as nobody's
            for (unsigned long long k = base_t; k < a.queue_tile_cap; ++k) a.queue_tiles[k] = 0xFFFFFFFFu;
        }
    }
    const Prep p = prepare(t);
    for (int j = t.y0; j <= t.y1; ++j)
        for (int i = t.x0; i <= t.x1; ++i) shade(p, t.mode, i, j, clamp_depth, inst, a.target, a.width);
}

// Step 2 for a triangle the near plane cuts: 0, 1 or 2 triangles, each set up and handed on.  Out of line for the
// same reason (the polygon lives in local memory).
__device__ __noinline__ void cut_and_queue(const float c[3][4], const float dd[3], uint32_t inst, const RasterArgs& a)
{
    float poly[4][4];
    int n_out = 0;
#pragma unroll
    for (int v = 0; v < 3; ++v) {
        const int nx = (v + 1) % 3;
        const bool in_v = dd[v] >= 0.f, in_n = dd[nx] >= 0.f;
        if (in_v) {
#pragma unroll
            for (int q = 0; q < 4; ++q) poly[n_out][q] = c[v][q];
            ++n_out;
        }
        if (in_v != in_n) {
            // from the inside vertex towards the outside one
            const float da = in_v ? dd[v] : dd[nx], db = in_v ? dd[nx] : dd[v];
            const float tt = da / (da - db);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float ca = in_v ? c[v][q] : c[nx][q], cb = in_v ? c[nx][q] : c[v][q];
                poly[n_out][q] = ca + tt * (cb - ca);
            }
            ++n_out;
        }
    }
    for (int u = 0; u + 2 < n_out; ++u) {
        float c3[3][4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            c3[0][q] = poly[0][q];
            c3[1][q] = poly[u + 1][q];
            c3[2][q] = poly[u + 2][q];
        }
        const Work t = setup(c3, (int)a.width, (int)a.height);
        if (t.mode >= 0) queue_or_walk(t, inst, a);
    }
}

// One thread per (instance, triangle) pair of the view, pairs numbered draw by draw, instance by instance: a warp
// holds 32 consecutive triangles (of one instance of a large mesh, or of a few instances of a small one), whatever the
// mix of meshes.  A warp finds the draw of its first pair by bisection over the draws' prefix, its lanes walk on from
// there.  What a triangle costs decides who samples it: a few samples (the common case at scene scale) — its thread;
// up to 16 K samples — its warp, lanes striding over the bounds; more, or anything cut by the near plane or not
// snappable — the queue, so that a view whose instances cover thousands of pixels each (the near shadow cascades)
// is spread over the whole grid.
__global__ void __launch_bounds__(kRasterThreads, 4) k_raster(const __grid_constant__ RasterArgs a)
{
    __shared__ __align__(16) uint32_t s_rec[kRasterThreads / 32][32][20];  // per lane: x, y, dx, dy, zn, fa, x0, y0, width, instance
    __shared__ uint32_t s_pre[kRasterThreads / 32][32];                     // samples before each lane's
    const int lane = threadIdx.x & 31;
    const uint32_t n_draws = min(a.counts[1], a.max_draws);
    const unsigned long long all_pairs = a.pairs[n_draws];
    const bool clamp_depth = (a.flags & SS_RASTER_DEPTH_CLAMP) != 0;
    const unsigned long long warps = ((unsigned long long)gridDim.x * kRasterThreads) >> 5;
    const unsigned long long warp_id = ((unsigned long long)blockIdx.x * kRasterThreads + threadIdx.x) >> 5;
    // this launch's part of the pairs (the pass is drawn in phases, the coarse depth refreshed in between)
    const unsigned long long first = (unsigned long long)(((unsigned __int128)all_pairs * a.phase_lo) >> 16);
    const unsigned long long total = (unsigned long long)(((unsigned __int128)all_pairs * a.phase_hi) >> 16);
    for (unsigned long long g0 = first + warp_id * 32ull; g0 < total; g0 += warps * 32ull) {
        // the last draw whose first pair is <= g0
        uint32_t lo = 0, hi = n_draws;
        while (hi - lo > 1) {
            const uint32_t mid = (lo + hi) >> 1;
            if (a.pairs[mid] <= g0) lo = mid; else hi = mid;
        }
        const unsigned long long g = g0 + (unsigned long long)lane;
        Work t;
        t.mode = -1;
        t.x0 = t.y0 = 0;
        t.x1 = t.y1 = -1;
        uint32_t inst = 0;
        if (g < total) {
            uint32_t d = lo;
            while (a.pairs[d + 1] <= g) ++d;  // draws without pairs are stepped over
            const ss_draw_cmd dc = a.draws[d];
            const uint32_t tris = dc.index_count / 3u;
            const unsigned long long local = g - a.pairs[d];
            const uint32_t ki = (uint32_t)(local / tris), tI = (uint32_t)(local - (unsigned long long)ki * tris);
            const uint32_t k = dc.first_instance + ki;
            const float4 r0 = a.rows[(size_t)k * 3 + 0], r1 = a.rows[(size_t)k * 3 + 1], r2 = a.rows[(size_t)k * 3 + 2];
            inst = a.idx[k];
            // 1: the triangle in clip space
            float c[3][4], dd[3];
            bool in_range = true;
#pragma unroll
            for (int v = 0; v < 3; ++v) {
                const uint64_t ii = (uint64_t)dc.first_index + (uint64_t)tI * 3u + (uint64_t)v;
                if (ii >= a.n_indices) { in_range = false; break; }
                const int64_t vi = (int64_t)dc.vertex_offset + (int64_t)a.indices[ii];
                if (vi < 0 || vi >= (int64_t)a.n_vertices) { in_range = false; break; }
                clip_of(a.vp, r0, r1, r2, a.positions + (size_t)vi * 3, c[v]);
            }
            if (in_range) {  // a table that points outside the geometry draws nothing
#pragma unroll
                for (int v = 0; v < 3; ++v) dd[v] = clamp_depth ? c[v][3] - 1e-5f : c[v][2];
                if (dd[0] >= 0.f && dd[1] >= 0.f && dd[2] >= 0.f) {
                    t = setup(c, (int)a.width, (int)a.height);
                    if (t.mode == 1) {  // not snappable: never sampled here
                        queue_or_walk(t, inst, a);
                        t.mode = -1;
                    }
                } else {
                    cut_and_queue(c, dd, inst, a);
                }
            }
        }
        // from here on: uncut, snapped triangles only
        long long samples = t.mode == 0 ? (long long)(t.x1 - t.x0 + 1) * (t.y1 - t.y0 + 1) : 0;
        if (samples > kHugeSamples) {
            queue_or_walk(t, inst, a);
            t.mode = -1;
            samples = 0;
        }
        // a small triangle that lies behind the coarse depth of every cell it touches is dropped here (larger ones are
        // sorted out cell by cell where they are sampled)
        const uint32_t zc = work_depth_floor(t, clamp_depth);
        if (a.hiz && zc && samples > 0 && samples <= kCoopSamples) {
            const int cx0 = t.x0 >> kCellShift, cx1 = t.x1 >> kCellShift, cy0 = t.y0 >> kCellShift, cy1 = t.y1 >> kCellShift;
            if ((cx1 - cx0 + 1) * (cy1 - cy0 + 1) <= 16) {
                uint32_t far = 0;
                for (int cy = cy0; cy <= cy1; ++cy)
                    for (int cx = cx0; cx <= cx1; ++cx) far = max(far, __ldg(a.hiz + (size_t)cy * a.hiz_w + cx));
                if (zc > far) {
                    t.mode = -1;
                    samples = 0;
                }
            }
        }
        // The samples of the warp's 32 triangles, dealt out evenly: each lane leaves its triangle's record in shared
        // memory and the lanes then take the samples of all bounds 32 at a time, whoever's they are.  (Sampling one's
        // own triangle would cost the warp as many rounds as its largest triangle has samples, with most lanes idle:
        // measured 14 of 32 lanes busy and 10 K instructions per 32 triangles on the camera view of cfg 3.)
        // A triangle of more than kCoopSamples keeps the warp to itself instead (its record is broadcast once and the
        // lanes stride over its bounds: no owner search and no division per sample).
        const bool coop = t.mode == 0 && samples > kCoopSamples;
        const uint32_t count = (t.mode == 0 && !coop) ? (uint32_t)samples : 0u;
        uint32_t todo = __ballot_sync(0xffffffffu, coop);
        while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            Work w;
#pragma unroll
            for (int q = 0; q < 9; ++q) w.w[q] = __shfl_sync(0xffffffffu, t.w[q], src);
            w.x0 = __shfl_sync(0xffffffffu, t.x0, src);
            w.x1 = __shfl_sync(0xffffffffu, t.x1, src);
            w.y0 = __shfl_sync(0xffffffffu, t.y0, src);
            w.y1 = __shfl_sync(0xffffffffu, t.y1, src);
            w.mode = 0;
            const uint32_t w_inst = __shfl_sync(0xffffffffu, inst, src);
            const Prep p = prepare(w);
            warp_sample_rect(p, 0, w.x0, w.x1, w.y0, w.y1, __shfl_sync(0xffffffffu, zc, src), clamp_depth, w_inst, a, lane);
        }
        const uint32_t most = __reduce_max_sync(0xffffffffu, count);
        if (most <= 2u) {  // nothing to balance
            if (count) {
                const Prep p = prepare(t);
                for (int j = t.y0; j <= t.y1; ++j)
                    for (int i = t.x0; i <= t.x1; ++i) shade(p, 0, i, j, clamp_depth, inst, a.target, a.width);
            }
            continue;
        }
        uint32_t incl = count;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += y;
        }
        const uint32_t all = __shfl_sync(0xffffffffu, incl, 31);
        uint32_t* rec = s_rec[threadIdx.x >> 5][lane];
        uint32_t* pre = s_pre[threadIdx.x >> 5];
        pre[lane] = incl - count;
        if (count) {
            const Prep p = prepare(t);
#pragma unroll
            for (int q = 0; q < 3; ++q) {
                rec[q] = (uint32_t)p.x[q];
                rec[3 + q] = (uint32_t)p.y[q];
                rec[6 + q] = (uint32_t)p.dx[q];
                rec[9 + q] = (uint32_t)p.dy[q];
                rec[12 + q] = __float_as_uint(p.zn[q]);
            }
            rec[15] = __float_as_uint(p.fa);
            rec[16] = (uint32_t)t.x0;
            rec[17] = (uint32_t)t.y0;
            rec[18] = (uint32_t)(t.x1 - t.x0 + 1);
            rec[19] = inst;
        }
        __syncwarp();
        for (uint32_t sIdx = lane; sIdx < all; sIdx += 32) {
            // the last lane whose first sample is <= sIdx (lanes without samples share their successor's start)
            int owner = 0;
#pragma unroll
            for (int step = 16; step > 0; step >>= 1)
                if (pre[owner + step] <= sIdx) owner += step;
            const uint4* r4 = reinterpret_cast<const uint4*>(s_rec[threadIdx.x >> 5][owner]);
            const uint4 q0 = r4[0], q1 = r4[1], q2 = r4[2], q3 = r4[3], q4 = r4[4];
            Prep p;
            p.x[0] = (int)q0.x; p.x[1] = (int)q0.y; p.x[2] = (int)q0.z;
            p.y[0] = (int)q0.w; p.y[1] = (int)q1.x; p.y[2] = (int)q1.y;
            p.dx[0] = (int)q1.z; p.dx[1] = (int)q1.w; p.dx[2] = (int)q2.x;
            p.dy[0] = (int)q2.y; p.dy[1] = (int)q2.z; p.dy[2] = (int)q2.w;
            p.zn[0] = __uint_as_float(q3.x); p.zn[1] = __uint_as_float(q3.y); p.zn[2] = __uint_as_float(q3.z);
            p.fa = __uint_as_float(q3.w);
            const uint32_t local = sIdx - pre[owner], bw = q4.z;
            const uint32_t row = local / bw;
            shade(p, 0, (int)(q4.x + (local - row * bw)), (int)(q4.y + row), clamp_depth, q4.w, a.target, a.width);
        }
        __syncwarp();  // the records are rewritten on the next round
    }
}

// The queued huge triangles: one warp per (triangle, 64 x 64-sample tile) entry.  A tile that lies wholly outside one
// edge is dropped after three corner evaluations (exact for the snapped mode: an edge function is linear, so it is
// negative over the tile iff it is negative at the tile corner it favours).
__global__ void __launch_bounds__(kRasterThreads) k_raster_tiles(const __grid_constant__ RasterArgs a)
{
    const int lane = threadIdx.x & 31;
    const bool clamp_depth = (a.flags & SS_RASTER_DEPTH_CLAMP) != 0;
    const uint32_t n_entries = min(a.queue_counts[1], a.queue_tile_cap);
    const uint32_t warps = (gridDim.x * kRasterThreads) >> 5;
    const uint32_t warp_id = (blockIdx.x * kRasterThreads + threadIdx.x) >> 5;
    for (uint32_t e = warp_id; e < n_entries; e += warps) {
        const uint32_t owner = a.queue_tiles[e];
        if (owner == 0xFFFFFFFFu) continue;
        const uint32_t* rec = a.queue_works + (size_t)owner * kWorkWords;
        Work w;
#pragma unroll
        for (int q = 0; q < 13; ++q) w.w[q] = rec[q];
        w.x0 = (int)rec[13]; w.x1 = (int)rec[14]; w.y0 = (int)rec[15]; w.y1 = (int)rec[16];
        w.mode = (int)rec[17];
        const uint32_t inst = rec[18];
        const int tx0 = w.x0 >> 6, ty0 = w.y0 >> 6, tnx = (w.x1 >> 6) - tx0 + 1;
        const int local = (int)(e - rec[19]);
        const int tx = tx0 + local % tnx, ty = ty0 + local / tnx;
        // the tile's samples that lie inside the triangle's bounds
        const int x0 = max(w.x0, tx << 6), x1 = min(w.x1, (tx << 6) + 63), y0 = max(w.y0, ty << 6), y1 = min(w.y1, (ty << 6) + 63);
        const Prep p = prepare(w);
        warp_sample_rect(p, w.mode, x0, x1, y0, y1, work_depth_floor(w, clamp_depth), clamp_depth, inst, a, lane);
    }
}

__global__ void k_raster_clear(unsigned long long* p, size_t n)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = ~0ull;
}

}  // namespace

void launch_raster_clear(unsigned long long* target, size_t n, cudaStream_t st)
{
    if (n) k_raster_clear<<<148 * 4, 256, 0, st>>>(target, n);
}

size_t raster_hiz_words(uint32_t width, uint32_t height)
{
    return (size_t)((width + 7) >> kCellShift) * ((height + 7) >> kCellShift);
}

// The pairs are drawn in phases whose cumulative share grows by sqrt(2): after each, the coarse depth is rebuilt from
// the target, so that a later phase skips what the earlier ones already hide (draw order is by key, i.e. any prefix of
// the pairs is a spatially uniform sample of the view).  A view of a few instances is drawn in one phase.
void launch_raster(const RasterArgs& a_in, cudaStream_t st)
{
    if (!a_in.visible) return;
    RasterArgs a = a_in;
    k_raster_pairs<<<1, 1024, 0, st>>>(a.draws, a.counts, a.max_draws, a.pairs);
    constexpr int kPhases = 12;
    const int phases = (a.hiz && a.visible >= 256) ? kPhases : 1;
    if (phases == 1) a.hiz = nullptr;
    a.hiz_w = (a.width + 7) >> kCellShift;
    a.hiz_h = (a.height + 7) >> kCellShift;
    const size_t cells = (size_t)a.hiz_w * a.hiz_h;
    if (a.hiz) cudaMemsetAsync(a.hiz, 0xFF, cells * sizeof(uint32_t), st);
    const int sms = device_sm_count();
    uint32_t lo = 0;
    for (int ph = 0; ph < phases; ++ph) {
        const uint32_t hi = ph + 1 == phases ? 65536u : (uint32_t)(65536.0 * exp2(-0.5 * (phases - 1 - ph)));
        a.phase_lo = lo;
        a.phase_hi = hi;
        lo = hi;
        if (a.queue_works) cudaMemsetAsync(a.queue_counts, 0, 2 * sizeof(uint32_t), st);
        k_raster<<<sms * 12, kRasterThreads, 0, st>>>(a);  // persistent warps striding over the pairs
        if (a.queue_works) k_raster_tiles<<<sms * 12, kRasterThreads, 0, st>>>(a);  // ... and over the queued tiles
        if (a.hiz && ph + 1 < phases) {
            const size_t want = (cells + 255) / 256;
            k_raster_hiz<<<(unsigned)std::min<size_t>(want, (size_t)sms * 8), 256, 0, st>>>(a.target, a.width, a.height, a.hiz, a.hiz_w, a.hiz_h);
        }
    }
}

}  // namespace ss
