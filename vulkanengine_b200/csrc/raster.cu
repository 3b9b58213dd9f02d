// Consumer of the compacted lists: depth + id rasterisation of one view's draws into a device
// buffer, as CUDA.  It stands where the reference binds the instance buffer and issues its draws:
//   * ShadowPass  (VulkanCore/assets/shaders/ShadowDepth.vert + .geom: world = transform * position,
//     clip = LightSpaceMatrices[cascade] * world; depth only, DepthClamp on — SceneRenderer.cpp:188)
//   * SelectionPass (CoreEditor.vert / .frag: clip = Projection * View * world, the fragment stores a_EntityID;
//     the editor reads ONE pixel back, SceneRenderer.cpp:1449-1456)
// Both passes are the same arithmetic over (instance row, triangle); the target word is [depth bits : 32 | id : 32]
// and the depth test is one 64-bit atomicMin, so the result does not depend on the order triangles arrive in.
//
// Rasterisation is in homogeneous coordinates (edge functions from the adjugate of the clip-space (x, y, w)
// matrix), so triangles that cross w = 0 need no clipper.  Every float operation is written out in a fixed order
// (-fmad=false), the same order the tests' CPU restatement uses: the two must agree bit for bit.
// The reference's own rasteriser is the Vulkan device: unavailable on this box.
#include "ss_kernels.h"

namespace ss {

namespace {

constexpr int kRasterThreads = 128;

struct Tri {
    float a[3], b[3], c[3];  // edge functions E_i(px, py) = (a_i * px + b_i * py) + c_i
    float z[3];
    float det;
    int x0, x1, y0, y1;      // inclusive pixel bounds, already clamped to the target
    bool ok;
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

__device__ __forceinline__ Tri setup(const float c0[4], const float c1[4], const float c2[4], float half_w, float half_h,
                                     int width, int height)
{
    Tri t;
    // pixel-space homogeneous coordinates: X / W is the pixel abscissa, Y / W the ordinate (y down, as Vulkan)
    const float X0 = (c0[0] + c0[3]) * half_w, X1 = (c1[0] + c1[3]) * half_w, X2 = (c2[0] + c2[3]) * half_w;
    const float Y0 = (c0[1] + c0[3]) * half_h, Y1 = (c1[1] + c1[3]) * half_h, Y2 = (c2[1] + c2[3]) * half_h;
    const float W0 = c0[3], W1 = c1[3], W2 = c2[3];
    t.a[0] = Y1 * W2 - Y2 * W1; t.b[0] = X2 * W1 - X1 * W2; t.c[0] = X1 * Y2 - X2 * Y1;
    t.a[1] = Y2 * W0 - Y0 * W2; t.b[1] = X0 * W2 - X2 * W0; t.c[1] = X2 * Y0 - X0 * Y2;
    t.a[2] = Y0 * W1 - Y1 * W0; t.b[2] = X1 * W0 - X0 * W1; t.c[2] = X0 * Y1 - X1 * Y0;
    t.det = (X0 * t.a[0] + X1 * t.a[1]) + X2 * t.a[2];
    t.z[0] = c0[2]; t.z[1] = c1[2]; t.z[2] = c2[2];
    t.ok = (t.det > 0.f || t.det < 0.f) && (W0 > 0.f || W1 > 0.f || W2 > 0.f);  // not degenerate (nor NaN), not wholly behind
    float lx = 0.f, hx = (float)width, ly = 0.f, hy = (float)height;
    if (W0 > 0.f && W1 > 0.f && W2 > 0.f) {  // all in front: the projected bounding box; otherwise the whole target
        const float sx0 = X0 / W0, sx1 = X1 / W1, sx2 = X2 / W2, sy0 = Y0 / W0, sy1 = Y1 / W1, sy2 = Y2 / W2;
        lx = fmaxf(fminf(fminf(sx0, sx1), sx2), 0.f);
        hx = fminf(fmaxf(fmaxf(sx0, sx1), sx2), (float)width);
        ly = fmaxf(fminf(fminf(sy0, sy1), sy2), 0.f);
        hy = fminf(fmaxf(fmaxf(sy0, sy1), sy2), (float)height);
        if (!(lx <= hx && ly <= hy)) t.ok = false;  // off the target (or NaN)
    }
    // a pixel (i, j) is sampled at (i + 0.5, j + 0.5); one pixel of slack each side, the edge tests decide
    t.x0 = max((int)floorf(lx) - 1, 0);
    t.x1 = min((int)ceilf(hx), width - 1);
    t.y0 = max((int)floorf(ly) - 1, 0);
    t.y1 = min((int)ceilf(hy), height - 1);
    return t;
}

__global__ void __launch_bounds__(kRasterThreads) k_raster(const __grid_constant__ RasterArgs a)
{
    const uint32_t k = blockIdx.x * kRasterThreads + threadIdx.x;
    if (k >= a.visible) return;
    const float4 r0 = a.rows[(size_t)k * 3 + 0], r1 = a.rows[(size_t)k * 3 + 1], r2 = a.rows[(size_t)k * 3 + 2];
    const uint32_t inst = a.idx[k];
    const SlotDev& sd = a.slots[a.slot[inst]];
    const uint32_t n_tri = sd.index_count / 3u;
    const float half_w = 0.5f * (float)a.width, half_h = 0.5f * (float)a.height;
    const bool clamp_depth = (a.flags & SS_RASTER_DEPTH_CLAMP) != 0;
    for (uint32_t tI = 0; tI < n_tri; ++tI) {
        float c[3][4];
        bool in_range = true;
#pragma unroll
        for (int v = 0; v < 3; ++v) {
            const uint64_t ii = (uint64_t)sd.first_index + (uint64_t)tI * 3u + (uint64_t)v;
            if (ii >= a.n_indices) { in_range = false; break; }
            const int64_t vi = (int64_t)sd.vertex_offset + (int64_t)a.indices[ii];
            if (vi < 0 || vi >= (int64_t)a.n_vertices) { in_range = false; break; }
            clip_of(a.vp, r0, r1, r2, a.positions + (size_t)vi * 3, c[v]);
        }
        if (!in_range) continue;  // a table that points outside the geometry draws nothing
        const Tri t = setup(c[0], c[1], c[2], half_w, half_h, (int)a.width, (int)a.height);
        if (!t.ok) continue;
        const float s = t.det > 0.f ? 1.f : -1.f;
        for (int j = t.y0; j <= t.y1; ++j) {
            const float py = (float)j + 0.5f;
            for (int i = t.x0; i <= t.x1; ++i) {
                const float px = (float)i + 0.5f;
                float e[3];
                bool inside = true;
#pragma unroll
                for (int q = 0; q < 3; ++q) {
                    e[q] = (t.a[q] * px + t.b[q] * py) + t.c[q];
                    const float es = s * e[q], as = s * t.a[q], bs = s * t.b[q];
                    // on an edge: owned by the triangle whose edge normal points right, or straight down
                    inside = inside && (es > 0.f || (es == 0.f && (as > 0.f || (as == 0.f && bs > 0.f))));
                }
                if (!inside) continue;
                if (!(s * ((e[0] + e[1]) + e[2]) > 0.f)) continue;  // 1 / w at the sample must be positive
                float depth = ((t.z[0] * e[0] + t.z[1] * e[1]) + t.z[2] * e[2]) / t.det;
                if (clamp_depth) depth = depth < 0.f ? 0.f : (depth > 1.f ? 1.f : depth);  // a NaN stays one and is dropped
                if (!(depth >= 0.f && depth <= 1.f)) continue;
                const unsigned long long word = ((unsigned long long)__float_as_uint(depth) << 32) | inst;
                atomicMin(a.target + (size_t)j * a.width + i, word);
            }
        }
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

void launch_raster(const RasterArgs& a, cudaStream_t st)
{
    if (a.visible) k_raster<<<(a.visible + kRasterThreads - 1) / kRasterThreads, kRasterThreads, 0, st>>>(a);
}

}  // namespace ss
