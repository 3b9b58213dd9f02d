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
#include <cstdio>
#include <vector>

#include "ss_kernels.h"

namespace ss {

namespace {

constexpr int kRasterThreads = 128;
constexpr int kCoopSamples = 192;             // bounds holding more samples than this get the whole warp to themselves
constexpr long long kHugeSamples = 16384;      // beyond this a triangle is queued and spread over the grid, tile by tile
constexpr int kChunkPairs = 1024;             // pairs a warp takes at a time, at most (a short phase is cut finer)
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

// -DSS_RASTER_STATS: counters of what the pass did, printed by launch_raster (a debugging build, never the shipped one)
#ifdef SS_RASTER_STATS
__device__ unsigned long long g_rstats[12];
#define RSTAT(i, v) atomicAdd(&g_rstats[i], (unsigned long long)(v))
#else
#define RSTAT(i, v)
#endif

__device__ __forceinline__ void put(unsigned long long* target, uint32_t width, int i, int j, float depth, bool clamp_depth,
                                    uint32_t inst)
{
    if (clamp_depth) depth = depth < 0.f ? 0.f : (depth > 1.f ? 1.f : depth);  // a NaN stays one and is dropped
    if (!(depth >= 0.f && depth <= 1.f)) return;
    RSTAT(7, 1);
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
// outside an edge, is skipped whole), then take the samples of each remaining cell two apiece.  `floor_word` bounds
// every word the triangle can store from below ([depth floor | instance]); without a coarse buffer every cell is
// sampled.  Skipping never changes the target: a skipped sample could not have won its place.
__device__ __forceinline__ void warp_sample_rect(const Prep& p, int mode, int x0, int x1, int y0, int y1, unsigned long long floor_word,
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
            if (a.hiz && floor_word > __ldg(a.hiz + (size_t)cy * a.hiz_w + cx)) alive = false;
            if (alive && mode == 0)
                alive = !rect_outside(p, max(x0, cx << kCellShift), min(x1, (cx << kCellShift) + 7), max(y0, cy << kCellShift),
                                      min(y1, (cy << kCellShift) + 7));
        }
        uint32_t todo = __ballot_sync(0xffffffffu, alive);
        RSTAT(5, c < n_cells);
        RSTAT(6, alive);
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

// hiz[cell] = the largest word among the cell's samples (all ones while any of them is empty).  The target's words
// only ever fall, so the value stays an upper bound for as long as it is used.
// a warp takes four cells next to each other: lane = 4 row + cell, so the lanes of a row read 256 contiguous bytes
__global__ void __launch_bounds__(256) k_raster_hiz(const unsigned long long* __restrict__ target, uint32_t width, uint32_t height,
                                                    unsigned long long* __restrict__ hiz, uint32_t hiz_w, uint32_t hiz_h)
{
    const uint32_t groups_x = (hiz_w + 3) / 4, n_groups = groups_x * hiz_h;
    const uint32_t lane = threadIdx.x & 31u, warps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t row = lane >> 2, cl = lane & 3u;
    for (uint32_t g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; g < n_groups; g += warps) {
        const uint32_t cy = g / groups_x, cx = (g - cy * groups_x) * 4 + cl;
        const uint32_t x0 = cx << kCellShift, y = (cy << kCellShift) + row;
        unsigned long long m = 0;
        if (cx < hiz_w && y < height) {
            const uint32_t x1 = min(x0 + 8u, width);
            for (uint32_t i = x0; i < x1; ++i) m = max(m, __ldcg(target + (size_t)y * width + i));
        }
        m = max(m, __shfl_xor_sync(0xffffffffu, m, 4));
        m = max(m, __shfl_xor_sync(0xffffffffu, m, 8));
        m = max(m, __shfl_xor_sync(0xffffffffu, m, 16));
        if (row == 0 && cx < hiz_w) hiz[(size_t)cy * hiz_w + cx] = m;
    }
}

// one warp per tile: its 8 x 8 cells two per lane
__global__ void __launch_bounds__(256) k_raster_hiz2(const unsigned long long* __restrict__ hiz, uint32_t hiz_w, uint32_t hiz_h,
                                                     unsigned long long* __restrict__ hiz2, uint32_t hiz2_w, uint32_t hiz2_h)
{
    const uint32_t n = hiz2_w * hiz2_h;
    const uint32_t lane = threadIdx.x & 31u, warps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; c < n; c += warps) {
        const uint32_t ty = c / hiz2_w, tx = c - ty * hiz2_w;
        unsigned long long m = 0;
#pragma unroll
        for (uint32_t h = 0; h < 2; ++h) {
            const uint32_t q = lane + 32u * h, cy = ty * 8 + (q >> 3), cx = tx * 8 + (q & 7u);
            if (cy < hiz_h && cx < hiz_w) m = max(m, hiz[(size_t)cy * hiz_w + cx]);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (lane == 0) hiz2[c] = m;
    }
}

// ---- the order the pass is drawn in: front to back ---------------------------------------------------------------
// Every row of the view gets the depth of its transform's origin as a sort key; the rows are sorted by it (the library's
// own radix sort) and the triangles before every rank are counted, so that a (rank, triangle) pair has a number and any
// range of those numbers can be found again by bisection.  The order only decides how much the coarse depth can skip.
constexpr uint32_t kNoDraw = 0xFFFFFFFFu;
constexpr int kScanThreads = 1024, kScanItems = 4, kScanTile = kScanThreads * kScanItems;

__device__ __forceinline__ const unsigned long long* sorted_order(const RasterArgs& a)
{
    const SortControl* ctl = static_cast<const SortControl*>(a.sort_scratch);
    return ctl->src_sel[a.sort_passes] ? a.order_b : a.order_a;
}

// Bounds of the vertices every draw's index range names (one CTA per draw at a time).  A vertex or an index outside the
// geometry is left out: the triangle it belongs to is never drawn.
__global__ void __launch_bounds__(128) k_raster_draw_bounds(const __grid_constant__ RasterArgs a)
{
    __shared__ float s_red[4][6];
    const uint32_t n_draws = min(a.counts[1], a.max_draws);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (uint32_t d = blockIdx.x; d < n_draws; d += gridDim.x) {
        const ss_draw_cmd dc = a.draws[d];
        float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (uint32_t i = threadIdx.x; i < dc.index_count; i += blockDim.x) {
            const uint64_t ii = (uint64_t)dc.first_index + i;
            if (ii >= a.n_indices) continue;
            const int64_t vi = (int64_t)dc.vertex_offset + (int64_t)a.indices[ii];
            if (vi < 0 || vi >= (int64_t)a.n_vertices) continue;
#pragma unroll
            for (int q = 0; q < 3; ++q) {
                const float p = a.positions[(size_t)vi * 3 + q];
                lo[q] = fminf(lo[q], p);
                hi[q] = fmaxf(hi[q], p);
            }
        }
#pragma unroll
        for (int q = 0; q < 3; ++q)
            for (int o = 16; o > 0; o >>= 1) {
                lo[q] = fminf(lo[q], __shfl_xor_sync(0xffffffffu, lo[q], o));
                hi[q] = fmaxf(hi[q], __shfl_xor_sync(0xffffffffu, hi[q], o));
            }
        __syncthreads();  // s_red of the previous draw has been read
        if (lane == 0)
            for (int q = 0; q < 3; ++q) {
                s_red[wid][q] = lo[q];
                s_red[wid][3 + q] = hi[q];
            }
        __syncthreads();
        if (threadIdx.x < 6) {
            float v = s_red[0][threadIdx.x];
            for (int w = 1; w < 4; ++w) v = threadIdx.x < 3 ? fminf(v, s_red[w][threadIdx.x]) : fmaxf(v, s_red[w][threadIdx.x]);
            a.draw_aabb[(size_t)d * 6 + threadIdx.x] = v;
        }
    }
}

// Where on the target, and how near at the nearest, anything of one instance can land: the eight corners of its draw's
// bounds through the same two products the triangles go through, with the rounding of every step accounted for.
// A vertex is a convex combination of the corners; its clip coordinates, as k_raster computes them, lie within `err`
// of the exact ones (each is a sum of products whose magnitudes are bounded by the corners': 2^-19 of that sum covers
// the dozen roundings on the way twice over), so they lie in the hull of the corners' computed coordinates widened by
// 2 err, and with w > 0 throughout a ratio x / w is extreme at an extreme point of that set.  The snap to 1/256 pixel
// and the float evaluation of the viewport transform add at most 1/64 pixel + 2^-20 |x|.  No bound (whole target,
// depth 0) when a corner may reach w <= 0 or lies beyond 2^19 pixels (triangles that cannot be snapped are sampled
// with float edge functions, whose coverage is not exact).
__device__ __forceinline__ uint4 instance_bound(const RasterArgs& a, float4 r0, float4 r1, float4 r2, const float* bb, uint32_t inst)
{
    const bool clamp_depth = (a.flags & SS_RASTER_DEPTH_CLAMP) != 0;
    const int cxmax = (int)((a.width - 1) >> kCellShift), cymax = (int)((a.height - 1) >> kCellShift);
    const uint4 whole = make_uint4((uint32_t)cxmax << 16, (uint32_t)cymax << 16, 0u, inst);
    if (!(bb[0] <= bb[3])) return make_uint4(1u, 0u, 0u, inst);  // no valid vertex: nothing of it is ever drawn
    const float half_w = 0.5f * (float)a.width, half_h = 0.5f * (float)a.height;
    float sx_lo = INFINITY, sx_hi = -INFINITY, sy_lo = INFINITY, sy_hi = -INFINITY, zn_lo = INFINITY;
    const float k = 1.9073486e-6f;  // 2^-19
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        const float px = bb[(c & 1) ? 3 : 0], py = bb[(c & 2) ? 4 : 1], pz = bb[(c & 4) ? 5 : 2];
        const float wx = (r0.x * px + r0.y * py) + (r0.z * pz + r0.w);
        const float wy = (r1.x * px + r1.y * py) + (r1.z * pz + r1.w);
        const float wz = (r2.x * px + r2.y * py) + (r2.z * pz + r2.w);
        const float mx = (fabsf(r0.x * px) + fabsf(r0.y * py)) + (fabsf(r0.z * pz) + fabsf(r0.w));
        const float my = (fabsf(r1.x * px) + fabsf(r1.y * py)) + (fabsf(r1.z * pz) + fabsf(r1.w));
        const float mz = (fabsf(r2.x * px) + fabsf(r2.y * py)) + (fabsf(r2.z * pz) + fabsf(r2.w));
        float cl[4], er[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            cl[j] = (a.vp[j] * wx + a.vp[4 + j] * wy) + (a.vp[8 + j] * wz + a.vp[12 + j]);
            er[j] = 2.f * k * ((fabsf(a.vp[j]) * mx + fabsf(a.vp[4 + j]) * my) + (fabsf(a.vp[8 + j]) * mz + fabsf(a.vp[12 + j])));
        }
        const float w_lo = cl[3] - er[3], w_hi = cl[3] + er[3];
        if (!(w_lo > 0.f)) return whole;
        const float xl = cl[0] - er[0], xh = cl[0] + er[0], yl = cl[1] - er[1], yh = cl[1] + er[1];
        const float rxl = xl >= 0.f ? xl / w_hi : xl / w_lo, rxh = xh >= 0.f ? xh / w_lo : xh / w_hi;
        const float ryl = yl >= 0.f ? yl / w_hi : yl / w_lo, ryh = yh >= 0.f ? yh / w_lo : yh / w_hi;
        sx_lo = fminf(sx_lo, rxl * half_w + half_w);
        sx_hi = fmaxf(sx_hi, rxh * half_w + half_w);
        sy_lo = fminf(sy_lo, ryl * half_h + half_h);
        sy_hi = fmaxf(sy_hi, ryh * half_h + half_h);
        const float zl = cl[2] - er[2];
        zn_lo = fminf(zn_lo, zl > 0.f ? zl / w_hi : 0.f);
    }
    const float far_out = 524288.f;  // 2^19 pixels
    if (!(sx_lo > -far_out && sx_hi < far_out && sy_lo > -far_out && sy_hi < far_out)) return whole;
    const float mgx = 0.015625f + fmaxf(fabsf(sx_lo), fabsf(sx_hi)) * 9.5367432e-7f;
    const float mgy = 0.015625f + fmaxf(fabsf(sy_lo), fabsf(sy_hi)) * 9.5367432e-7f;
    // sample i sits at i + 0.5
    const int i0 = max((int)ceilf(sx_lo - mgx - 0.5f), 0), i1 = min((int)floorf(sx_hi + mgx - 0.5f), (int)a.width - 1);
    const int j0 = max((int)ceilf(sy_lo - mgy - 0.5f), 0), j1 = min((int)floorf(sy_hi + mgy - 0.5f), (int)a.height - 1);
    if (i0 > i1 || j0 > j1) return make_uint4(1u, 0u, 0u, inst);  // covers no sample
    float zf = zn_lo > 0.f ? zn_lo * (1.f - k) : 0.f;
    if (clamp_depth) zf = fminf(zf, 1.0f);
    return make_uint4((uint32_t)(i0 >> kCellShift) | ((uint32_t)(i1 >> kCellShift) << 16),
                      (uint32_t)(j0 >> kCellShift) | ((uint32_t)(j1 >> kCellShift) << 16), __float_as_uint(zf), inst);
}

__global__ void __launch_bounds__(256) k_raster_keys(const __grid_constant__ RasterArgs a)
{
    const uint32_t n_draws = min(a.counts[1], a.max_draws);
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < a.visible; k += gridDim.x * blockDim.x) {
        // the last draw that starts at or before row k (the draws of a view lie in row order)
        uint32_t d = kNoDraw;
        if (n_draws) {
            uint32_t lo = 0, hi = n_draws;
            while (hi - lo > 1) {
                const uint32_t mid = (lo + hi) >> 1;
                if (a.draws[mid].first_instance <= k) lo = mid; else hi = mid;
            }
            const uint32_t fi = a.draws[lo].first_instance;
            if (k >= fi && k - fi < a.draws[lo].instance_count) d = lo;
        }
        const float tx = a.rows[(size_t)k * 3 + 0].w, ty = a.rows[(size_t)k * 3 + 1].w, tz = a.rows[(size_t)k * 3 + 2].w;
        const float z = (a.vp[2] * tx + a.vp[6] * ty) + (a.vp[10] * tz + a.vp[14]);
        const float w = (a.vp[3] * tx + a.vp[7] * ty) + (a.vp[11] * tz + a.vp[15]);
        float key = w > 0.f ? z / w : -1.f;  // at or behind the eye: first
        uint32_t low = k;
        if (a.flags & SS_RASTER_DEPTH_CLAMP) {
            // everything in front of the near plane is stored at depth 0 and the lowest instance wins there: those rows
            // are drawn in instance order (the sort then runs over the instance bits too)
            key = key > 0.f ? fminf(key, 1.0f) : 0.f;
            low = a.idx[k];
            a.row_of[low] = k;
        }
        uint32_t sortable;
        if (a.flags & SS_RASTER_DEPTH_CLAMP) {
            sortable = __float_as_uint(key);  // in [0, 1]: the bits order like the values
        } else {
            // 16 bits of key are enough for an order that only decides how much can be skipped: the upper half of the
            // float bits of 1 - depth, reversed — for a perspective view that is 1 / distance in steps of 0.8 %.  Two
            // sort passes instead of four.  (An orthographic DepthClamp pass keeps all 32 bits: with 16, steps of 0.4 %
            // of the depth range, cascade 2 of cfg 3 went from 3.6 to 4.6 ms.)
            const float gap = fmaxf(1.0f - key, 1e-30f);  // > 1 in front of the near plane, tiny at and beyond the far one
            sortable = 0xFFFFu - (__float_as_uint(gap) >> 16);
        }
        a.order_a[k] = ((unsigned long long)sortable << 32) | low;
        a.draw_of[k] = d;
        const uint32_t inst = a.idx[k];
        if (d == kNoDraw) {
            a.bound[k] = make_uint4(1u, 0u, 0u, inst);
        } else {
            float bb[6];
#pragma unroll
            for (int q = 0; q < 6; ++q) bb[q] = a.draw_aabb[(size_t)d * 6 + q];
            a.bound[k] = instance_bound(a, a.rows[(size_t)k * 3 + 0], a.rows[(size_t)k * 3 + 1], a.rows[(size_t)k * 3 + 2], bb, inst);
        }
    }
}

__device__ __forceinline__ uint32_t row_at_rank(const RasterArgs& a, const unsigned long long* order, uint32_t r)
{
    const uint32_t low = (uint32_t)order[r];
    return (a.flags & SS_RASTER_DEPTH_CLAMP) ? a.row_of[low] : low;
}

__device__ __forceinline__ uint32_t triangles_at_rank(const RasterArgs& a, const unsigned long long* order, uint32_t r)
{
    const uint32_t d = a.draw_of[row_at_rank(a, order, r)];
    return d == kNoDraw ? 0u : a.draws[d].index_count / 3u;
}

__device__ __forceinline__ unsigned long long block_exclusive_scan(unsigned long long mine, unsigned long long* total)
{
    __shared__ unsigned long long warp_sum[kScanThreads / 32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned long long x = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
    }
    __syncthreads();  // warp_sum may still be read by the previous call
    if (lane == 31) warp_sum[wid] = x;
    __syncthreads();
    unsigned long long base = 0, all = 0;
    for (int w = 0; w < kScanThreads / 32; ++w) {
        const unsigned long long v = warp_sum[w];
        if (w < wid) base += v;
        all += v;
    }
    *total = all;
    return base + x - mine;
}

__global__ void __launch_bounds__(kScanThreads) k_raster_scan_sums(const __grid_constant__ RasterArgs a)
{
    const unsigned long long* order = sorted_order(a);
    const uint32_t r0 = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
    unsigned long long mine = 0;
#pragma unroll
    for (int q = 0; q < kScanItems; ++q)
        if (r0 + q < a.visible) mine += triangles_at_rank(a, order, r0 + q);
    unsigned long long all;
    block_exclusive_scan(mine, &all);
    if (threadIdx.x == 0) a.block_sums[blockIdx.x] = all;
}

__global__ void __launch_bounds__(kScanThreads) k_raster_scan_top(unsigned long long* block_sums, uint32_t n_blocks)
{
    unsigned long long carry = 0;
    for (uint32_t b0 = 0; b0 < n_blocks; b0 += kScanThreads) {
        const uint32_t b = b0 + threadIdx.x;
        const unsigned long long mine = b < n_blocks ? block_sums[b] : 0ull;
        unsigned long long all;
        const unsigned long long ex = block_exclusive_scan(mine, &all);
        if (b < n_blocks) block_sums[b] = carry + ex;
        carry += all;
    }
}

__global__ void __launch_bounds__(kScanThreads) k_raster_scan_apply(const __grid_constant__ RasterArgs a)
{
    const unsigned long long* order = sorted_order(a);
    const uint32_t r0 = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
    uint32_t cnt[kScanItems];
    unsigned long long mine = 0;
#pragma unroll
    for (int q = 0; q < kScanItems; ++q) {
        cnt[q] = r0 + q < a.visible ? triangles_at_rank(a, order, r0 + q) : 0u;
        mine += cnt[q];
    }
    unsigned long long all;
    unsigned long long at = a.block_sums[blockIdx.x] + block_exclusive_scan(mine, &all);
#pragma unroll
    for (int q = 0; q < kScanItems; ++q) {
        if (r0 + q < a.visible) a.tri_prefix[r0 + q] = at;
        at += cnt[q];
        if (r0 + q + 1 == a.visible) a.tri_prefix[a.visible] = at;
    }
}

// A triangle its thread does not sample itself: queued for k_raster_tiles (one entry per 64 x 64-sample tile of its
// bounds, spread over the whole grid), or, when the queue is full, walked here sample by sample.  Kept out of line:
// it is the rare path and must not cost the common one its registers.
__device__ __noinline__ void queue_or_walk(const Work& t, uint32_t inst, const RasterArgs& a)
{
    const bool clamp_depth = (a.flags & SS_RASTER_DEPTH_CLAMP) != 0;
    const Prep p = prepare(t);
    if (a.queue_works) {
        // Only the tiles that can matter get an entry: not those that already lie in front of everything the triangle
        // can store (64 x 64 level of the coarse depth), nor those wholly outside one of its edges.
        const int tx0 = t.x0 >> 6, ty0 = t.y0 >> 6;
        const uint32_t tnx = (uint32_t)((t.x1 >> 6) - tx0 + 1);
        const uint32_t n_tiles = tnx * (uint32_t)((t.y1 >> 6) - ty0 + 1);
        const unsigned long long floor_word = ((unsigned long long)work_depth_floor(t, clamp_depth) << 32) | inst;
        auto matters = [&](uint32_t k) {
            const int ty = ty0 + (int)(k / tnx), tx = tx0 + (int)(k % tnx);
            if (a.hiz2 && floor_word > __ldg(a.hiz2 + (size_t)ty * a.hiz2_w + tx)) return false;
            if (t.mode == 0 && rect_outside(p, max(t.x0, tx << 6), min(t.x1, (tx << 6) + 63), max(t.y0, ty << 6), min(t.y1, (ty << 6) + 63)))
                return false;
            return true;
        };
        uint32_t n_alive = 0;
        for (uint32_t k = 0; k < n_tiles; ++k) n_alive += matters(k) ? 1u : 0u;
        if (n_alive == 0) return;
        const uint32_t slot_w = atomicAdd(a.queue_counts + 0, 1u);
        if (slot_w < a.queue_work_cap) {
            const uint32_t base_t = atomicAdd(a.queue_counts + 1, n_alive);
            if ((unsigned long long)base_t + n_alive <= a.queue_tile_cap) {
                uint32_t* rec = a.queue_works + (size_t)slot_w * kWorkWords;
#pragma unroll
                for (int q = 0; q < 13; ++q) rec[q] = t.w[q];
                rec[13] = (uint32_t)t.x0; rec[14] = (uint32_t)t.x1; rec[15] = (uint32_t)t.y0; rec[16] = (uint32_t)t.y1;
                rec[17] = (uint32_t)t.mode; rec[18] = inst; rec[19] = 0u;
                uint32_t at = base_t;
                for (uint32_t k = 0; k < n_tiles; ++k)
                    if (matters(k)) {
                        a.queue_tiles[2 * (size_t)at] = slot_w;
                        a.queue_tiles[2 * (size_t)at + 1] = k;
                        ++at;
                    }
                return;
            }
            // the tile list is full: the part of the range that exists is marked as nobody's
            for (unsigned long long k = base_t; k < a.queue_tile_cap; ++k) a.queue_tiles[2 * k] = 0xFFFFFFFFu;
        }
    }
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

// One thread per (instance, triangle) pair of the view, pairs numbered rank by rank of the front-to-back order: a warp
// takes a chunk of the phase's pairs, finds its first rank by bisection, and goes through the chunk's ranks 32 at a time
// — one instance-level test per lane — before the triangles of the ranks that survive are set up, one per thread, 32
// at a time whoever's they are (so a 12-triangle box scene and a 9 K-triangle vase load the lanes alike, and a phase
// in which nearly everything is hidden costs a few coalesced loads per instance: id pass of cfg 3 3.3 -> 2.4 ms,
// far cascade 21.3 -> 16.3 ms against testing pair by pair).  What a triangle costs decides who samples it: a few
// samples (the common case at scene scale) — its thread, balanced over the warp; up to 16 K samples — its warp, cell by
// cell; more, or anything cut by the near plane or not snappable — the queue, so that a view whose instances cover
// thousands of pixels each (the near shadow cascades) is spread over the whole grid.
// 8 CTAs of 128 threads per SM (64 registers): the per-triangle set-up is a chain of dependent loads, and with most
// triangles dropped by the coarse depth it is what the pass spends its time on; measured on cfg 3's far cascade
// (199 M triangles) 31.8 ms at 4 CTAs, 26.4 at 6, 23.4 at 8.
constexpr int kRasterCtasPerSm = 8;
__global__ void __launch_bounds__(kRasterThreads, kRasterCtasPerSm) k_raster(const __grid_constant__ RasterArgs a)
{
    __shared__ __align__(16) uint32_t s_rec[kRasterThreads / 32][32][20];  // per lane: x, y, dx, dy, zn, fa, x0, y0, width, instance
    __shared__ uint32_t s_pre[kRasterThreads / 32][32];                     // samples before each lane's
    __shared__ uint4 s_ent[kRasterThreads / 32][32];                         // the round's ranks: row, instance, first triangle, draw
    __shared__ uint32_t s_epre[kRasterThreads / 32][32];                    // live pairs before each of them
    const int lane = threadIdx.x & 31;
    const unsigned long long* order = sorted_order(a);
    const unsigned long long all_pairs = a.tri_prefix[a.visible];
    const bool clamp_depth = (a.flags & SS_RASTER_DEPTH_CLAMP) != 0;
    // this launch's part of the pairs (the pass is drawn in phases, the coarse depth rebuilt in between), handed out to
    // the warps in chunks: one bisection per chunk, the lanes walk on from there
    const unsigned long long first = (unsigned long long)(((unsigned __int128)all_pairs * a.phase_lo) >> 16);
    const unsigned long long last = (unsigned long long)(((unsigned __int128)all_pairs * a.phase_hi) >> 16);
    // at least four chunks per warp of the grid, so that a short phase (or a view of few, large triangles) still
    // spreads over all of it
    const unsigned long long grid_warps = ((unsigned long long)gridDim.x * kRasterThreads) >> 5;
    const unsigned long long chunk_pairs =
        max(32ull, min((unsigned long long)kChunkPairs, ((last - first) / (grid_warps * 4ull)) & ~31ull));
    const unsigned long long n_chunks = (last - first + chunk_pairs - 1) / chunk_pairs;
    // a phase too short to occupy the grid (the nearest, largest triangles come first) hands everything a warp would
    // have to itself to the queue instead, which spreads it tile by tile
    const long long huge_samples = (last - first) < grid_warps * 64ull ? (long long)kCoopSamples : kHugeSamples;
    for (;;) {
        unsigned long long chunk = 0;
        if (lane == 0) chunk = atomicAdd(a.queue_counts + 2, 1u);
        chunk = __shfl_sync(0xffffffffu, chunk, 0);
        if (chunk >= n_chunks) break;
        const unsigned long long c0 = first + chunk * chunk_pairs;
        const unsigned long long total = min(c0 + chunk_pairs, last);
        // the last rank whose first pair is <= c0
        uint32_t r_warp = 0;
        {
            uint32_t lo = 0, hi = a.visible;
            while (hi - lo > 1) {
                const uint32_t mid = (lo + hi) >> 1;
                if (a.tri_prefix[mid] <= c0) lo = mid; else hi = mid;
            }
            r_warp = lo;
        }
        // The ranks of the chunk 32 at a time, one per lane: the instance first — where any of its triangles can land and how
        // near (k_raster_keys); one that covers no sample, or that lies behind the coarse depth of everything it can
        // touch, is done with here, before a single index or vertex of it is read.  The triangles of the ranks that are
        // left are numbered through (exclusive prefix in shared memory) and taken 32 at a time, whoever's they are.
        for (uint32_t rb = r_warp;; rb += 32u) {
            const uint32_t r = rb + (uint32_t)lane;
            const unsigned long long p0 = r < a.visible ? a.tri_prefix[r] : all_pairs;
            const unsigned long long p1 = r < a.visible ? a.tri_prefix[r + 1] : all_pairs;
            uint32_t cnt = 0;
            uint4 ent = make_uint4(0u, 0u, 0u, 0u);  // row, instance, first triangle of the rank in this chunk, draw
            if (r < a.visible && p0 < total && p1 > c0) {
                const unsigned long long lo = max(p0, c0), hi = min(p1, total);
                const uint32_t k = row_at_rank(a, order, r);
                const uint4 bd = a.bound[k];
                const int bx0 = (int)(bd.x & 0xFFFFu), bx1 = (int)(bd.x >> 16), by0 = (int)(bd.y & 0xFFFFu), by1 = (int)(bd.y >> 16);
                bool hidden = bx0 > bx1;
                if (!hidden && a.hiz) {
                    const unsigned long long inst_floor = ((unsigned long long)bd.z << 32) | bd.w;
                    if ((bx1 - bx0 + 1) * (by1 - by0 + 1) <= 16) {
                        unsigned long long far = 0;
                        for (int cy = by0; cy <= by1; ++cy)
                            for (int cx = bx0; cx <= bx1; ++cx) far = max(far, __ldg(a.hiz + (size_t)cy * a.hiz_w + cx));
                        hidden = inst_floor > far;
                    } else if (((bx1 >> 3) - (bx0 >> 3) + 1) * ((by1 >> 3) - (by0 >> 3) + 1) <= 16) {
                        unsigned long long far = 0;
                        for (int ty = by0 >> 3; ty <= (by1 >> 3); ++ty)
                            for (int tx = bx0 >> 3; tx <= (bx1 >> 3); ++tx) far = max(far, __ldg(a.hiz2 + (size_t)ty * a.hiz2_w + tx));
                        hidden = inst_floor > far;
                    }
                }
                RSTAT(9, hidden ? hi - lo : 0);
                if (!hidden) {
                    cnt = (uint32_t)(hi - lo);
                    ent = make_uint4(k, bd.w, (uint32_t)(lo - p0), a.draw_of[k]);
                }
            }
            uint32_t incl_p = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xffffffffu, incl_p, o);
                if (lane >= o) incl_p += y;
            }
            const uint32_t live = __shfl_sync(0xffffffffu, incl_p, 31);
            const bool chunk_done = __any_sync(0xffffffffu, r >= a.visible || p1 >= total);
            __syncwarp();  // the entries of the round before have been read
            s_ent[threadIdx.x >> 5][lane] = ent;
            s_epre[threadIdx.x >> 5][lane] = incl_p - cnt;
            __syncwarp();
            for (uint32_t j0 = 0; j0 < live; j0 += 32u) {
                const uint32_t jp = j0 + (uint32_t)lane;
                const bool g_ok = jp < live;
                Work t;
                t.mode = -1;
                t.x0 = t.y0 = 0;
                t.x1 = t.y1 = -1;
                uint32_t inst = 0;
                if (g_ok) {
                    // the last entry whose first pair is <= jp (entries without pairs share their successor's start)
                    int owner = 0;
#pragma unroll
                    for (int step = 16; step > 0; step >>= 1)
                        if (s_epre[threadIdx.x >> 5][owner + step] <= jp) owner += step;
                    const uint4 en = s_ent[threadIdx.x >> 5][owner];
                    const uint32_t k = en.x;
                    inst = en.y;
                    {
                    const ss_draw_cmd dc = a.draws[en.w];
                    const uint32_t tI = en.z + (jp - s_epre[threadIdx.x >> 5][owner]);
                    const float4 r0 = a.rows[(size_t)k * 3 + 0], r1 = a.rows[(size_t)k * 3 + 1], r2 = a.rows[(size_t)k * 3 + 2];
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
                }
                // from here on: uncut, snapped triangles only
                long long samples = t.mode == 0 ? (long long)(t.x1 - t.x0 + 1) * (t.y1 - t.y0 + 1) : 0;
                const unsigned long long floor_word = ((unsigned long long)work_depth_floor(t, clamp_depth) << 32) | inst;
                // a large triangle that lies behind the coarse depth of every 64 x 64 tile it touches is dropped here
                if (a.hiz2 && samples > kCoopSamples) {
                    const int tx0 = t.x0 >> 6, tx1 = t.x1 >> 6, ty0 = t.y0 >> 6, ty1 = t.y1 >> 6;
                    if ((tx1 - tx0 + 1) * (ty1 - ty0 + 1) <= 16) {
                        unsigned long long far = 0;
                        for (int ty = ty0; ty <= ty1; ++ty)
                            for (int tx = tx0; tx <= tx1; ++tx) far = max(far, __ldg(a.hiz2 + (size_t)ty * a.hiz2_w + tx));
                        if (floor_word > far) {
                            t.mode = -1;
                            samples = 0;
                        }
                    }
                }
                if (samples > huge_samples) {
                    queue_or_walk(t, inst, a);
                    t.mode = -1;
                    samples = 0;
                }
                // a small triangle that lies behind the coarse depth of every cell it touches is dropped here (the others are
                // sorted out cell by cell where they are sampled)
                if (a.hiz && samples > 0 && samples <= kCoopSamples) {
                    const int cx0 = t.x0 >> kCellShift, cx1 = t.x1 >> kCellShift, cy0 = t.y0 >> kCellShift, cy1 = t.y1 >> kCellShift;
                    if ((cx1 - cx0 + 1) * (cy1 - cy0 + 1) <= 16) {
                        unsigned long long far = 0;
                        for (int cy = cy0; cy <= cy1; ++cy)
                            for (int cx = cx0; cx <= cx1; ++cx) far = max(far, __ldg(a.hiz + (size_t)cy * a.hiz_w + cx));
                        RSTAT(2, floor_word > far);
                        if (floor_word > far) {
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
                RSTAT(0, g_ok);
                RSTAT(1, t.mode == 0 && samples > 0);
                if (t.mode == 0 && !coop) RSTAT(3, samples);
                RSTAT(4, coop);
                RSTAT(8, t.mode == 0 && samples > 0 && (floor_word >> 32) == 0u);
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
                    warp_sample_rect(p, 0, w.x0, w.x1, w.y0, w.y1, __shfl_sync(0xffffffffu, floor_word, src), clamp_depth, w_inst, a, lane);
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
            if (chunk_done) break;
        }
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
        const uint32_t owner = a.queue_tiles[2 * (size_t)e];
        if (owner == 0xFFFFFFFFu) continue;
        const uint32_t* rec = a.queue_works + (size_t)owner * kWorkWords;
        Work w;
#pragma unroll
        for (int q = 0; q < 13; ++q) w.w[q] = rec[q];
        w.x0 = (int)rec[13]; w.x1 = (int)rec[14]; w.y0 = (int)rec[15]; w.y1 = (int)rec[16];
        w.mode = (int)rec[17];
        const uint32_t inst = rec[18];
        const int tx0 = w.x0 >> 6, ty0 = w.y0 >> 6, tnx = (w.x1 >> 6) - tx0 + 1;
        const int local = (int)a.queue_tiles[2 * (size_t)e + 1];
        const int tx = tx0 + local % tnx, ty = ty0 + local / tnx;
        // the tile's samples that lie inside the triangle's bounds
        const int x0 = max(w.x0, tx << 6), x1 = min(w.x1, (tx << 6) + 63), y0 = max(w.y0, ty << 6), y1 = min(w.y1, (ty << 6) + 63);
        const Prep p = prepare(w);
        warp_sample_rect(p, w.mode, x0, x1, y0, y1, ((unsigned long long)work_depth_floor(w, clamp_depth) << 32) | inst, clamp_depth, inst, a, lane);
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
    return (size_t)((width + 7) >> kCellShift) * ((height + 7) >> kCellShift) + (size_t)((width + 63) >> 6) * ((height + 63) >> 6);
}

size_t raster_order_bytes(uint32_t n, uint32_t max_draws)
{
    auto up = [](size_t x) { return (x + 255) & ~size_t(255); };
    const size_t n1 = (size_t)(n ? n : 1);
    return 2 * up(n1 * 8) + up((n1 + 1) * 8) + 2 * up(n1 * 4) + up(n1 * 16) + up(((size_t)max_draws + 1) * 24) + up(((n1 + kScanTile - 1) / kScanTile) * 8) + up(sort_scratch_bytes(n1));
}

void raster_order_carve(void* blob, uint32_t n, uint32_t max_draws, RasterArgs* a)
{
    auto up = [](size_t x) { return (x + 255) & ~size_t(255); };
    const size_t n1 = (size_t)(n ? n : 1);
    unsigned char* p = static_cast<unsigned char*>(blob);
    a->order_a = reinterpret_cast<unsigned long long*>(p); p += up(n1 * 8);
    a->order_b = reinterpret_cast<unsigned long long*>(p); p += up(n1 * 8);
    a->tri_prefix = reinterpret_cast<unsigned long long*>(p); p += up((n1 + 1) * 8);
    a->draw_of = reinterpret_cast<uint32_t*>(p); p += up(n1 * 4);
    a->row_of = reinterpret_cast<uint32_t*>(p); p += up(n1 * 4);
    a->bound = reinterpret_cast<uint4*>(p); p += up(n1 * 16);
    a->draw_aabb = reinterpret_cast<float*>(p); p += up(((size_t)max_draws + 1) * 24);
    a->block_sums = reinterpret_cast<unsigned long long*>(p); p += up(((n1 + kScanTile - 1) / kScanTile) * 8);
    a->sort_scratch = p;
}

// The pairs are drawn front to back in phases whose cumulative share doubles; after each, the coarse depth is
// rebuilt from the target, so that a later phase skips what the earlier, nearer ones already hide.  A view of a few
// instances is drawn in one phase.
void launch_raster(const RasterArgs& a_in, cudaStream_t st)
{
    if (!a_in.visible) return;
    RasterArgs a = a_in;
    const int sms = device_sm_count();
    const unsigned row_grid = (unsigned)std::min<size_t>(((size_t)a.visible + 255) / 256, (size_t)sms * 8);
    const unsigned scan_grid = (a.visible + kScanTile - 1) / kScanTile;
    const int sort_begin = (a.flags & SS_RASTER_DEPTH_CLAMP) ? 0 : 32;
    const int sort_end = (a.flags & SS_RASTER_DEPTH_CLAMP) ? 64 : 48;
    a.sort_passes = sort_num_passes(sort_begin, sort_end);
    k_raster_draw_bounds<<<(unsigned)std::min<uint32_t>(std::max<uint32_t>(a.max_draws, 1u), (uint32_t)sms * 16u), 128, 0, st>>>(a);
    k_raster_keys<<<row_grid, 256, 0, st>>>(a);
    int sort_launches = 0;
    sort_async(reinterpret_cast<uint64_t*>(a.order_a), nullptr, reinterpret_cast<uint64_t*>(a.order_b), nullptr, a.visible, sort_begin, sort_end,
               a.sort_scratch, st, &sort_launches, false);
    k_raster_scan_sums<<<scan_grid, kScanThreads, 0, st>>>(a);
    k_raster_scan_top<<<1, kScanThreads, 0, st>>>(a.block_sums, scan_grid);
    k_raster_scan_apply<<<scan_grid, kScanThreads, 0, st>>>(a);
    constexpr int kPhases = 12;
    const int phases = (a.hiz && a.visible >= 256) ? kPhases : 1;
    if (phases == 1) a.hiz = nullptr;
    a.hiz_w = (a.width + 7) >> kCellShift;
    a.hiz_h = (a.height + 7) >> kCellShift;
    const size_t cells = (size_t)a.hiz_w * a.hiz_h;
    a.hiz2_w = (a.width + 63) >> 6;
    a.hiz2_h = (a.height + 63) >> 6;
    const size_t tiles = (size_t)a.hiz2_w * a.hiz2_h;
    a.hiz2 = a.hiz ? a.hiz + cells : nullptr;
    if (a.hiz) cudaMemsetAsync(a.hiz, 0xFF, (cells + tiles) * sizeof(unsigned long long), st);
    uint32_t lo = 0;
    for (int ph = 0; ph < phases; ++ph) {
        const uint32_t hi = ph + 1 == phases ? 65536u : (uint32_t)(65536u >> (phases - 1 - ph));
        a.phase_lo = lo;
        a.phase_hi = hi;
        lo = hi;
        cudaMemsetAsync(a.queue_counts, 0, 4 * sizeof(uint32_t), st);  // queue fill and the next chunk of pairs
        k_raster<<<sms * kRasterCtasPerSm, kRasterThreads, 0, st>>>(a);  // persistent warps taking chunks of the phase's pairs
        if (a.queue_works) k_raster_tiles<<<sms * 8, kRasterThreads, 0, st>>>(a);  // ... and striding over the queued tiles
        if (a.hiz && ph + 1 < phases) {
            const size_t want = ((size_t)((a.hiz_w + 3) / 4) * a.hiz_h + 7) / 8;  // a warp per four cells
            k_raster_hiz<<<(unsigned)std::min<size_t>(want, (size_t)sms * 8), 256, 0, st>>>(a.target, a.width, a.height, a.hiz, a.hiz_w, a.hiz_h);
            k_raster_hiz2<<<(unsigned)std::min<size_t>((tiles + 7) / 8, (size_t)sms * 8), 256, 0, st>>>(a.hiz, a.hiz_w, a.hiz_h, a.hiz2, a.hiz2_w, a.hiz2_h);
        }
#ifdef SS_RASTER_STATS
        {
            unsigned long long h[12], z[12] = {};
            uint32_t qc[4] = {};
            cudaStreamSynchronize(st);
            cudaMemcpyFromSymbol(h, g_rstats, sizeof h);
            cudaMemcpyToSymbol(g_rstats, z, sizeof z);
            cudaMemcpy(qc, a.queue_counts, 16, cudaMemcpyDeviceToHost);
            if (a.hiz) {
                std::vector<unsigned long long> hz(cells);
                cudaMemcpy(hz.data(), a.hiz, cells * 8, cudaMemcpyDeviceToHost);
                size_t ones = 0, zeros = 0, one_f = 0;
                for (unsigned long long v : hz) { ones += v == ~0ull; zeros += (v >> 32) == 0u; one_f += (v >> 32) == 0x3F800000u; }
                fprintf(stderr, "  hiz before the next phase: %zu cells, %zu open, %zu at depth 0, %zu at depth 1; no floor %llu\n", cells, ones, zeros, one_f, h[8]);
            }
            fprintf(stderr, "raster phase %2d [%5u, %5u): pairs %llu with_samples %llu small_rejected %llu balanced_samples %llu coop_tris %llu "
                            "cells %llu alive %llu puts %llu hidden_at_instance_level %llu queued %u tiles %u\n", ph, a.phase_lo, a.phase_hi, h[0], h[1],
                    h[2], h[3], h[4], h[5], h[6], h[7], h[9], qc[0], qc[1]);
        }
#endif
    }
}

}  // namespace ss
