// Transform -> cull -> key kernels, sm_100a.
//
// HBM-bound streaming work: every input stream is a dense array of small records
// (TransformComponent 36 B, AABB 24 B, ids 4+4 B).  A CTA copies its tile of each stream into shared
// memory with 128-bit coalesced loads, each thread then reads its own record at an odd word stride
// (bank-conflict free), computes, and the 48 B output rows leave through shared memory as 128-bit
// coalesced stores.  No tensor cores: a 4x4 affine product per instance is not a dense contraction.
#include <algorithm>
#include <cstdlib>

#include "ss_kernels.h"

namespace ss {

namespace {

constexpr int kTile = 256;  // instances per CTA == threads per CTA

// Shared-memory slot of float4 number k of the tile's world rows (k = instance*4 + row).  A thread
// writes its own 4 rows (stride 4 float4 = 16 words: every fourth lane would hit the same banks), so the
// row index is XOR-ed with bits of the instance number; 8 consecutive lanes then cover all 32 banks on
// both the per-thread writes and the linear copy-out.
__device__ __forceinline__ int row_slot(int k) { return (k & ~3) | ((k ^ (k >> 3)) & 3); }

// Cooperative copy of `count` floats from global to 16B-aligned shared memory.
// R = floats per record: the trip count of the vector loop is a compile-time constant, so the loads of a
// tile are all in flight before the first shared-memory store.
template <int R>
__device__ __forceinline__ void tile_load_f32(float* __restrict__ dst, const float* __restrict__ src, int n_tile)
{
    const int tid = threadIdx.x;
    const int count = n_tile * R;
    if ((reinterpret_cast<uintptr_t>(src) & 15u) == 0) {
        const int n4 = count >> 2;
        const float4* s4 = reinterpret_cast<const float4*>(src);
        float4* d4 = reinterpret_cast<float4*>(dst);
        constexpr int kTrips = (R * kTile / 4 + kTile - 1) / kTile;
        float4 v[kTrips];
#pragma unroll
        for (int u = 0; u < kTrips; ++u) {
            const int i = tid + u * kTile;
            if (i < n4) v[u] = ld_nc_f4(s4 + i);
        }
#pragma unroll
        for (int u = 0; u < kTrips; ++u) {
            const int i = tid + u * kTile;
            if (i < n4) d4[i] = v[u];
        }
        for (int i = (n4 << 2) + tid; i < count; i += kTile) dst[i] = __ldg(src + i);
    } else {
        for (int i = tid; i < count; i += kTile) dst[i] = __ldg(src + i);
    }
}

__device__ __forceinline__ Aff load_rows(const float4* __restrict__ p)
{
    Aff a;
    a.r0 = __ldg(p + 0);
    a.r1 = __ldg(p + 1);
    a.r2 = __ldg(p + 2);
    return a;
}

__device__ __forceinline__ float4 shfl_f4(float4 v, int src)
{
    v.x = __shfl_sync(0xffffffffu, v.x, src);
    v.y = __shfl_sync(0xffffffffu, v.y, src);
    v.z = __shfl_sync(0xffffffffu, v.z, src);
    v.w = __shfl_sync(0xffffffffu, v.w, src);
    return v;
}

// Parent rows for a whole warp.  Siblings are adjacent in a level-ordered stream, so most warps
// share one parent: lanes 0..2 fetch one row each and the warp gets them by shuffle; mixed warps
// fall back to per-lane loads.  Every lane of the warp must call this (p < 0 = no parent).
__device__ __forceinline__ bool warp_parent_rows(const float4* __restrict__ node_world, int p, Aff& out)
{
    const int lane = threadIdx.x & 31;
    int same;
    __match_all_sync(0xffffffffu, p, &same);
    if (same) {
        if (p < 0) return false;
        float4 mine = make_float4(0.f, 0.f, 0.f, 0.f);
        if (lane < 3) mine = __ldg(node_world + (size_t)p * 3 + lane);
        out.r0 = shfl_f4(mine, 0);
        out.r1 = shfl_f4(mine, 1);
        out.r2 = shfl_f4(mine, 2);
        return true;
    }
    if (p < 0) return false;
    out = load_rows(node_world + (size_t)p * 3);
    return true;
}

template <int TRIG>
__global__ void __launch_bounds__(kTile) k_node_level(NodeLevelArgs a)
{
    __shared__ __align__(16) float s_trs[kTile * 9];
    __shared__ __align__(16) float4 s_rows[kTile * 3];
    const uint32_t base = a.lo + blockIdx.x * kTile;
    const int n_tile = min((uint32_t)kTile, a.hi - base);
    const int tid = threadIdx.x;
    tile_load_f32<9>(s_trs, a.trs + (size_t)base * 9, n_tile);
    __syncthreads();
    const bool valid = tid < n_tile;
    int p = -1;
    if (valid && a.parent) p = __ldg(a.parent + base + tid);
    Aff parent;
    const bool has_parent = warp_parent_rows(a.node_world, p, parent);
    if (valid) {
        Aff m = trs_to_affine<TRIG>(s_trs + tid * 9);
        if (has_parent) m = aff_mul(parent, m);
        s_rows[tid * 3 + 0] = m.r0;
        s_rows[tid * 3 + 1] = m.r1;
        s_rows[tid * 3 + 2] = m.r2;
    }
    __syncthreads();
    float4* dst = a.node_world + (size_t)base * 3;
    for (int i = tid; i < n_tile * 3; i += kTile) dst[i] = s_rows[i];
}

template <int TRIG>
__global__ void __launch_bounds__(kTile) k_instances(InstanceArgs a, const __grid_constant__ CullParams cp)
{
    __shared__ __align__(16) float s_trs[kTile * 9];
    __shared__ __align__(16) float s_box[kTile * 6];
    __shared__ __align__(16) float4 s_rows[kTile * kWorldStride];

    const uint32_t base = blockIdx.x * kTile;
    const int n_tile = min((uint32_t)kTile, a.n - base);
    const int tid = threadIdx.x;
    const bool valid = tid < n_tile;
    const uint32_t i = base + tid;
    const bool from_trs = !a.world_given && a.node == nullptr;

    if (from_trs) tile_load_f32<9>(s_trs, a.trs + ((size_t)a.node_base + base) * 9, n_tile);
    if (a.inst_aabb) tile_load_f32<6>(s_box, a.inst_aabb + (size_t)base * 6, n_tile);
    if (a.world_given) {
        const float4* src = a.world + (size_t)base * kWorldStride;  // given rows always come at the 64 B stride
        for (int k = tid; k < n_tile * kWorldStride; k += kTile) s_rows[row_slot(k)] = ld_nc_f4(src + k);
    }
    uint32_t slot = SS_SLOT_NULL, mw = 0;
    if (valid) {
        slot = __ldg(a.slot + i);
        mw = __ldg(a.matword + i);
    }
    int p = -1;
    if (from_trs && valid && a.parent) p = __ldg(a.parent + a.node_base + i);
    __syncthreads();

    Aff parent;
    const bool has_parent = warp_parent_rows(a.node_world, p, parent);

    const bool live = valid && slot != SS_SLOT_NULL;
    Aff w;
    w.r0 = w.r1 = w.r2 = make_float4(0.f, 0.f, 0.f, 0.f);
    uint32_t mask = 0;
    uint64_t key = 2ull << a.tess_shift;  // dropped: the bit above tess, behind every real key
    if (live) {
        const SlotDev* sd = a.slots + slot;
        if (a.world_given) {
            w.r0 = s_rows[row_slot(tid * kWorldStride + 0)];
            w.r1 = s_rows[row_slot(tid * kWorldStride + 1)];
            w.r2 = s_rows[row_slot(tid * kWorldStride + 2)];
        } else {
            Aff m;
            if (from_trs) {
                m = trs_to_affine<TRIG>(s_trs + tid * 9);
                if (has_parent) m = aff_mul(parent, m);  // N1: world = world[parent] * local
            } else {
                m = load_rows(a.node_world + (size_t)__ldg(a.node + i) * 3);
            }
            Aff loc;
            loc.r0 = __ldg(&sd->r0);
            loc.r1 = __ldg(&sd->r1);
            loc.r2 = __ldg(&sd->r2);
            w = aff_mul(m, loc);  // VC/Scene/SceneRenderer.cpp:1150
        }
        float mn[3], mx[3];
        if (a.inst_aabb) {
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                mn[k] = s_box[tid * 6 + k];
                mx[k] = s_box[tid * 6 + 3 + k];
            }
        } else {
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                mn[k] = __ldg(&sd->mn[k]);
                mx[k] = __ldg(&sd->mx[k]);
            }
        }
        mask = cull_box(w, mn, mx, cp);
        // N5: [tess][meshRank][submeshIndex][materialRank]
        key = __ldg(&sd->key_hi) | (uint64_t)(mw & ~SS_MATERIAL_TESS_BIT) |
              ((uint64_t)(mw >> 31) << a.tess_shift);
    }
    if (!a.world_given) {
        __syncthreads();  // s_rows is free (world_given reads finished above / never written)
        if (valid) {
            s_rows[row_slot(tid * kWorldStride + 0)] = w.r0;
            s_rows[row_slot(tid * kWorldStride + 1)] = w.r1;
            s_rows[row_slot(tid * kWorldStride + 2)] = w.r2;
            s_rows[row_slot(tid * kWorldStride + 3)] = make_float4(0.f, 0.f, 0.f, 1.f);
        }
        __syncthreads();
        if (a.world_stride == 3) {  // packed TransformData: rows 0..2 of every instance, 48 B apart
            float4* dst = a.world + (size_t)base * 3;
            for (int k = tid; k < n_tile * 3; k += kTile) dst[k] = s_rows[row_slot((k / 3) * kWorldStride + k % 3)];
        } else {
            float4* dst = a.world + (size_t)base * kWorldStride;
#pragma unroll
            for (int u = 0; u < kWorldStride; ++u) {
                const int k = tid + u * kTile;
                if (k < n_tile * kWorldStride) dst[k] = s_rows[row_slot(k)];
            }
        }
    }
    if (valid) {
        a.mask[i] = (uint8_t)mask;
        if (a.keys == nullptr) {
            // SS_FRAME_REUSE_SORT: the sorted order of an earlier frame is kept, no keys are produced
        } else if (a.idx_bits) {  // key and instance index travel through the sort as one 64-bit word
            // [key | visibility byte (when it fits) | instance]: emit then needs no gather of the mask
            a.keys[i] = (key << (a.idx_bits + a.mask_bits)) | ((uint64_t)(a.mask_bits ? mask : 0u) << a.idx_bits) | i;
        } else {
            a.keys[i] = key;
            a.vals[i] = i;
        }
    }
}

}  // namespace

namespace {
// one thread per float of a record; the host has already dropped all but the last record of a node named twice
__global__ void k_trs_scatter(float* trs, const uint32_t* node_index, const float* records, uint32_t count)
{
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t k = g / 9u, f = g - k * 9u;
    if (k < count) trs[(size_t)node_index[k] * 9 + f] = records[(size_t)k * 9 + f];
}
// The physics write-back loop (Platform/JoltPhysics/JoltPhysicsWorld.cpp:112-127) on the device: body position and
// rotation quaternion -> TransformComponent::Translation and ::Rotation; Scale is left as it is.  The Euler angles are
// JPH::Quat::GetEulerAngles (vendor/JoltPhysics/Jolt/Math/Quat.inl:185-203) in its operation order; Jolt evaluates
// ATan2 / ASin with its own polynomials, here they are CUDA's atan2f / asinf (a few ulp apart: the angles agree to
// 1e-6, they are not bit-identical).
__global__ void k_pose_scatter(float* trs, const uint32_t* node_index, const float* poses, uint32_t count)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= count) return;
    const float* p = poses + (size_t)k * 7;
    const float x = p[3], y = p[4], z = p[5], w = p[6];
    const float y_sq = y * y;
    const float t0 = 2.0f * (w * x + y * z);
    const float t1 = 1.0f - 2.0f * (x * x + y_sq);
    float t2 = 2.0f * (w * y - z * x);
    t2 = t2 > 1.0f ? 1.0f : t2;
    t2 = t2 < -1.0f ? -1.0f : t2;
    const float t3 = 2.0f * (w * z + x * y);
    const float t4 = 1.0f - 2.0f * (y_sq + z * z);
    float* t = trs + (size_t)node_index[k] * 9;
    t[0] = p[0];
    t[1] = p[1];
    t[2] = p[2];
    t[3] = atan2f(t0, t1);
    t[4] = asinf(t2);
    t[5] = atan2f(t3, t4);
}
}  // namespace

namespace {
// Scene::UpdateLightsBuffer / OnUpdateLights (VC/Scene/Scene.cpp:167-264): one thread per light
__global__ void k_pack_lights(const float* trs, const ss_point_light* point, uint32_t n_point, const ss_spot_light* spot,
                              uint32_t n_spot, unsigned char* out)
{
    const uint32_t t = threadIdx.x;
    float* ubp = reinterpret_cast<float*>(out);
    float* ubs = reinterpret_cast<float*>(out + SS_UB_POINT_LIGHTS_BYTES);
    float* posv = reinterpret_cast<float*>(out + SS_UB_POINT_LIGHTS_BYTES + SS_UB_SPOT_LIGHTS_BYTES);
    if (t == 0) {
        reinterpret_cast<int*>(ubp)[0] = (int)n_point;
        reinterpret_cast<int*>(ubs)[0] = (int)n_spot;
    }
    if (t < n_point) {
        const ss_point_light L = point[t];
        const float* tr = trs + (size_t)L.node * 9;
        float* o = ubp + 4 + t * 12;
        o[0] = tr[0]; o[1] = tr[1]; o[2] = tr[2]; o[3] = 1.0f;
        o[4] = L.color[0]; o[5] = L.color[1]; o[6] = L.color[2]; o[7] = L.color[3];
        o[8] = L.radius; o[9] = L.falloff;
        float* q = posv + t * 4;
        q[0] = tr[0]; q[1] = tr[1]; q[2] = tr[2]; q[3] = L.radius;
    } else if (t >= 16 && t - 16 < n_spot) {
        const uint32_t k = t - 16;
        const ss_spot_light L = spot[k];
        const float* tr = trs + (size_t)L.node * 9;
        // glm::normalize(Rotation) = v * inversesqrt(dot(v, v)), inversesqrt(x) = 1 / sqrt(x) (GLM/detail/func_exponential.inl)
        const float rx = tr[3], ry = tr[4], rz = tr[5];
        const float inv = 1.0f / sqrtf(rx * rx + ry * ry + rz * rz);
        float* o = ubs + 4 + k * 16;
        o[0] = tr[0]; o[1] = tr[1]; o[2] = tr[2]; o[3] = 1.0f;
        o[4] = L.color[0]; o[5] = L.color[1]; o[6] = L.color[2]; o[7] = L.color[3];
        o[8] = rx * inv; o[9] = ry * inv; o[10] = rz * inv;
        o[11] = L.inner_cutoff; o[12] = L.outer_cutoff; o[13] = L.radius; o[14] = L.falloff;
        float* q = posv + (n_point + k) * 4;
        q[0] = tr[0]; q[1] = tr[1]; q[2] = tr[2]; q[3] = L.radius;
    }
}
}  // namespace

void launch_pack_lights(const float* trs, const ss_point_light* point, uint32_t n_point, const ss_spot_light* spot, uint32_t n_spot,
                        unsigned char* out, cudaStream_t st)
{
    k_pack_lights<<<1, 32, 0, st>>>(trs, point, n_point, spot, n_spot, out);
}

namespace {
// one thread per 32-bit word of the 52 B records: words 0..11 = the instance rows, word 12 = the entity id
__global__ void k_select_records(const uint32_t* __restrict__ rows, const uint32_t* __restrict__ idx, const int32_t* __restrict__ entity,
                                 uint32_t visible, uint32_t* __restrict__ out)
{
    const uint64_t total = (uint64_t)visible * 13u, step = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += step) {
        const uint32_t e = (uint32_t)(g / 13u), w = (uint32_t)(g - (uint64_t)e * 13u);
        out[g] = w < 12 ? rows[(size_t)e * 12 + w] : (entity ? (uint32_t)entity[idx[e]] : 0xFFFFFFFFu);
    }
}
}  // namespace

void launch_select_records(const float4* rows, const uint32_t* idx, const int32_t* entity, uint32_t visible, uint32_t* out13, cudaStream_t st)
{
    if (!visible) return;
    const uint64_t want = ((uint64_t)visible * 13 + 255) / 256;
    k_select_records<<<(unsigned)std::min<uint64_t>(want, 148 * 16), 256, 0, st>>>(reinterpret_cast<const uint32_t*>(rows), idx, entity, visible, out13);
}

void launch_pose_scatter(float* trs, const uint32_t* node_index, const float* poses, uint32_t count, cudaStream_t st)
{
    if (count) k_pose_scatter<<<(count + 255u) / 256u, 256, 0, st>>>(trs, node_index, poses, count);
}

void launch_trs_scatter(float* trs, const uint32_t* node_index, const float* records, uint32_t count, cudaStream_t st)
{
    if (count) k_trs_scatter<<<(count * 9u + 255u) / 256u, 256, 0, st>>>(trs, node_index, records, count);
}

void launch_node_level(const NodeLevelArgs& a, cudaStream_t st)
{
    if (a.hi <= a.lo) return;
    const uint32_t grid = (a.hi - a.lo + kTile - 1) / kTile;
    if (a.trig_mode == SS_TRIG_DETERMINISTIC)
        k_node_level<SS_TRIG_DETERMINISTIC><<<grid, kTile, 0, st>>>(a);
    else
        k_node_level<SS_TRIG_LIBM><<<grid, kTile, 0, st>>>(a);
}

bool launch_instances(const InstanceArgs& a, const CullParams& cp, cudaStream_t st)
{
    if (a.n == 0) return false;
    // SS_SCALAR_KERNEL=1 forces the one-instance-per-thread kernel (A/B measurements, parity tests of both)
    static const bool force_scalar = getenv("SS_SCALAR_KERNEL") && atoi(getenv("SS_SCALAR_KERNEL")) != 0;
    if (!force_scalar && instances_x2_eligible(a)) {
        launch_instances_x2(a, cp, st);
        return a.hist_passes > 0;
    }
    const uint32_t grid = (a.n + kTile - 1) / kTile;
    if (a.trig_mode == SS_TRIG_DETERMINISTIC)
        k_instances<SS_TRIG_DETERMINISTIC><<<grid, kTile, 0, st>>>(a, cp);
    else
        k_instances<SS_TRIG_LIBM><<<grid, kTile, 0, st>>>(a, cp);
    return false;  // the general kernel leaves the histograms to the sort
}

}  // namespace ss
