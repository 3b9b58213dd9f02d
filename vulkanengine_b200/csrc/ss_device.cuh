// Device-side building blocks of libscenesubmit (sm_100a).  Compiled with -fmad=false: every float
// expression below is evaluated exactly as written, which is what makes the masks bit-identical to
// the CPU oracle (SURVEY.md §7 hard part 2).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/scenesubmit.h"
#include "../../include/scenesubmit_dettrig.h"

namespace ss {

constexpr int kWarp = 32;

// Affine world matrix in TransformData layout (VC/Core/Components.h:95-98):
// row r = (M[0][r], M[1][r], M[2][r], M[3][r]); the fourth row (0,0,0,1) is implicit.
struct Aff {
    float4 r0, r1, r2;
};

// Submesh table row as the kernels read it.
struct SlotDev {
    float4 r0, r1, r2;       // Submesh::LocalTransform rows
    float mn[3], mx[3];      // local AABB
    uint32_t index_count, first_index;
    int32_t vertex_offset;
    uint32_t pad;
    uint64_t key_hi;         // (meshRank << (mb+sb)) | (submeshIndex << mb)
};
static_assert(sizeof(SlotDev) == 96, "SlotDev layout");

// Planes of one view, (a,b,c,d), inside when a x + b y + c z + d >= 0; order L R B T N F, i.e. three
// axes of two planes each.  flags bit k (k = 0..2): the two planes of axis k have exactly opposite
// normals (always true for the affine cascade matrices), so their dot products are shared;
// flags bit 3: skip the near plane (SS_VIEW_SKIP_NEAR).
struct ViewPlanes {
    float4 pl[6];
    uint32_t flags;
    uint32_t pad[3];
};
constexpr uint32_t kViewSkipNear = 8u;

struct CullParams {
    ViewPlanes v[SS_MAX_VIEWS];
    uint32_t n_views;
};

__device__ __forceinline__ float4 ld_nc_f4(const float4* p)
{
    float4 r;
    asm("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void st_na_f4(float4* p, float4 v)
{
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
                 "f"(v.w));
}

// T * R * S for TransformComponent::GetTransform (VC/Core/Components.h:22-29).  The generic glm
// products multiply by the structural 0/1 entries of translate()/scale()/mat4(mat3); for finite inputs
// those terms are exact (x*1 = x, x*0 = 0, x+0 = x), so the reduced form below yields the same values
// (up to the sign of a zero, which nothing downstream can observe).  Every surviving operation keeps
// glm's order: quat(euler) GLM/detail/type_quat.inl:208-217, mat3_cast GLM/gtc/quaternion.inl:47-72.
template <int TRIG>
__device__ __forceinline__ Aff trs_to_affine(const float* __restrict__ t9)
{
    float sx, cx, sy, cy, sz, cz;
    float hx = t9[3] * 0.5f, hy = t9[4] * 0.5f, hz = t9[5] * 0.5f;
    if (TRIG == SS_TRIG_DETERMINISTIC) {
        ss_det_sincosf(hx, &sx, &cx);
        ss_det_sincosf(hy, &sy, &cy);
        ss_det_sincosf(hz, &sz, &cz);
    } else {
        sx = sinf(hx); cx = cosf(hx);
        sy = sinf(hy); cy = cosf(hy);
        sz = sinf(hz); cz = cosf(hz);
    }
    float qw = cx * cy * cz + sx * sy * sz;
    float qx = sx * cy * cz - cx * sy * sz;
    float qy = cx * sy * cz + sx * cy * sz;
    float qz = cx * cy * sz - sx * sy * cz;

    float qxx = qx * qx, qyy = qy * qy, qzz = qz * qz;
    float qxz = qx * qz, qxy = qx * qy, qyz = qy * qz;
    float qwx = qw * qx, qwy = qw * qy, qwz = qw * qz;

    // R[c][r]
    float R00 = 1.0f - 2.0f * (qyy + qzz), R01 = 2.0f * (qxy + qwz), R02 = 2.0f * (qxz - qwy);
    float R10 = 2.0f * (qxy - qwz), R11 = 1.0f - 2.0f * (qxx + qzz), R12 = 2.0f * (qyz + qwx);
    float R20 = 2.0f * (qxz + qwy), R21 = 2.0f * (qyz - qwx), R22 = 1.0f - 2.0f * (qxx + qyy);

    float s0 = t9[6], s1 = t9[7], s2 = t9[8];
    Aff m;
    m.r0 = make_float4(R00 * s0, R10 * s1, R20 * s2, t9[0]);
    m.r1 = make_float4(R01 * s0, R11 * s1, R21 * s2, t9[1]);
    m.r2 = make_float4(R02 * s0, R12 * s1, R22 * s2, t9[2]);
    return m;
}

// One row of A*B for affine A, B: glm accumulates ((A0*b0 + A1*b1) + A2*b2) + A3*b3 per column
// (GLM/detail/type_mat4x4.inl:643-646); b3 is 0 for columns 0..2 and 1 for column 3.
__device__ __forceinline__ float4 aff_row_mul(float4 a, const Aff& b)
{
    float4 o;
    o.x = (a.x * b.r0.x + a.y * b.r1.x) + a.z * b.r2.x;
    o.y = (a.x * b.r0.y + a.y * b.r1.y) + a.z * b.r2.y;
    o.z = (a.x * b.r0.z + a.y * b.r1.z) + a.z * b.r2.z;
    o.w = ((a.x * b.r0.w + a.y * b.r1.w) + a.z * b.r2.w) + a.w;
    return o;
}

__device__ __forceinline__ Aff aff_mul(const Aff& a, const Aff& b)
{
    Aff o;
    o.r0 = aff_row_mul(a.r0, b);
    o.r1 = aff_row_mul(a.r1, b);
    o.r2 = aff_row_mul(a.r2, b);
    return o;
}

// N2 + N4: world AABB by centre/extent, then the plane tests, no FMA.
__device__ __forceinline__ uint32_t cull_box(const Aff& w, const float* __restrict__ mn, const float* __restrict__ mx,
                                             const CullParams& cp)
{
    float cl0 = (mn[0] + mx[0]) * 0.5f, cl1 = (mn[1] + mx[1]) * 0.5f, cl2 = (mn[2] + mx[2]) * 0.5f;
    float el0 = (mx[0] - mn[0]) * 0.5f, el1 = (mx[1] - mn[1]) * 0.5f, el2 = (mx[2] - mn[2]) * 0.5f;

    float cw0 = (w.r0.x * cl0 + w.r0.y * cl1) + (w.r0.z * cl2 + w.r0.w);
    float cw1 = (w.r1.x * cl0 + w.r1.y * cl1) + (w.r1.z * cl2 + w.r1.w);
    float cw2 = (w.r2.x * cl0 + w.r2.y * cl1) + (w.r2.z * cl2 + w.r2.w);
    float ew0 = (fabsf(w.r0.x) * el0 + fabsf(w.r0.y) * el1) + fabsf(w.r0.z) * el2;
    float ew1 = (fabsf(w.r1.x) * el0 + fabsf(w.r1.y) * el1) + fabsf(w.r1.z) * el2;
    float ew2 = (fabsf(w.r2.x) * el0 + fabsf(w.r2.y) * el1) + fabsf(w.r2.z) * el2;

    // Spec (oracle): visible iff for every plane  s + r >= 0  with  s = ((a*cx + b*cy) + c*cz) + d  and
    // r = (|a|*ex + |b|*ey) + |c|*ez.  For finite operands  s + r >= 0  <=>  s >= -r  (rounding is monotone
    // and a non-zero sum never rounds to zero), which saves the add.  When the second plane of an axis
    // has exactly the negated normal, its dot product is the exact negation of the first one's
    // (rounding is symmetric), so  s1 = (-t) + d1  with the shared t, and r is shared as well.
    uint32_t mask = 0;
#pragma unroll
    for (int v = 0; v < (int)SS_MAX_VIEWS; ++v) {
        if (v < (int)cp.n_views) {
            const uint32_t fl = cp.v[v].flags;
            bool inside = true;
#pragma unroll
            for (int ax = 0; ax < 3; ++ax) {
                const float4 p0 = cp.v[v].pl[2 * ax], p1 = cp.v[v].pl[2 * ax + 1];
                const bool skip0 = (ax == 2) && (fl & kViewSkipNear);
                const float t0 = (p0.x * cw0 + p0.y * cw1) + p0.z * cw2;
                const float r0 = (fabsf(p0.x) * ew0 + fabsf(p0.y) * ew1) + fabsf(p0.z) * ew2;
                const bool ok0 = skip0 || (t0 + p0.w >= -r0);
                bool ok1;
                if (fl & (1u << ax)) {
                    ok1 = (p1.w - t0 >= -r0);
                } else {
                    const float t1 = (p1.x * cw0 + p1.y * cw1) + p1.z * cw2;
                    const float r1 = (fabsf(p1.x) * ew0 + fabsf(p1.y) * ew1) + fabsf(p1.z) * ew2;
                    ok1 = (t1 + p1.w >= -r1);
                }
                inside = inside && ok0 && ok1;
            }
            if (inside) mask |= (1u << v);
        }
    }
    return mask;
}

// ---- decoupled look-back (single-pass prefix over tiles) -------------------------------------------
// One 32-bit status word per (tile, counter): value | flag.  A tile first publishes its own count
// (kLbAgg), then walks back over earlier tiles summing their counts until it meets one whose inclusive
// prefix is known (kLbPrefix), then publishes its own inclusive prefix.  Tile ids are handed out by an
// atomic ticket at CTA start, so every tile waited on is already running: the walk cannot deadlock.
constexpr uint32_t kLbAgg = 1u << 30;
constexpr uint32_t kLbPrefix = 1u << 31;
constexpr uint32_t kLbValue = kLbAgg - 1;

// Sum of the counters `words[t * stride]` of tiles t < tile.  The first wave of CTAs starts together,
// none has a prefix yet, and tile k must add up k aggregates: the loads are issued kLbBatch at a time so
// the walk costs one L2 round trip per batch instead of one per tile.
constexpr int kLbBatch = 8;
__device__ __forceinline__ uint32_t lookback_exclusive(const uint32_t* words, size_t stride, int64_t tile)
{
    uint32_t sum = 0;
    int64_t t = tile - 1;
    while (t >= 0) {
        uint32_t v[kLbBatch];
#pragma unroll
        for (int u = 0; u < kLbBatch; ++u)
            v[u] = (t - u >= 0) ? *(const volatile uint32_t*)(words + (size_t)(t - u) * stride) : kLbPrefix;
        int used = 0;
        bool done = false;
#pragma unroll
        for (int u = 0; u < kLbBatch; ++u) {
            if (!done && used == u) {
                if (v[u] & kLbPrefix) {
                    sum += v[u] & kLbValue;
                    done = true;
                } else if (v[u] & kLbAgg) {
                    sum += v[u] & kLbValue;
                    used = u + 1;
                }
            }
        }
        if (done) break;
        t -= used;  // used < kLbBatch: the next tile back has not published yet, poll it again
    }
    return sum;
}

}  // namespace ss
