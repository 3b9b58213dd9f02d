// Host-side view matrices of libscenesubmit: the editor camera and the four shadow cascades.
// These stay on the host (5 x 64 B per frame) and are handed to ss_frame.
// Scalar binary32 in the operation order of the GLM 0.9.9.9 routines the reference calls, so the
// matrices equal the reference's bit for bit (compiled with -ffp-contract=off).
#include <cmath>
#include <cstring>

#include "../../include/scenesubmit.h"

#ifdef __CUDACC__
#include <cuda_runtime.h>
#define SS_HD __host__ __device__
#else
#define SS_HD
#endif

namespace {

struct V3 {
    float x, y, z;
};
struct M4 {
    float c[4][4];  // c[col][row], glm::mat4 memory
};

SS_HD inline V3 sub(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
SS_HD inline V3 add(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
SS_HD inline V3 mul(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
SS_HD inline float dot(V3 a, V3 b)
{
    const float px = a.x * b.x, py = a.y * b.y, pz = a.z * b.z;
    return px + py + pz;
}
SS_HD inline V3 cross(V3 a, V3 b) { return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; }
SS_HD inline V3 normalize(V3 v) { return mul(v, 1.0f / sqrtf(dot(v, v))); }

SS_HD M4 identity()
{
    M4 m;
    std::memset(&m, 0, sizeof m);
    m.c[0][0] = m.c[1][1] = m.c[2][2] = m.c[3][3] = 1.0f;
    return m;
}

// column j of a*b = ((a0*b[j][0] + a1*b[j][1]) + a2*b[j][2]) + a3*b[j][3]
SS_HD M4 matmul(const M4& a, const M4& b)
{
    M4 r;
    for (int j = 0; j < 4; ++j)
        for (int i = 0; i < 4; ++i) {
            float acc = a.c[0][i] * b.c[j][0] + a.c[1][i] * b.c[j][1];
            acc = acc + a.c[2][i] * b.c[j][2];
            acc = acc + a.c[3][i] * b.c[j][3];
            r.c[j][i] = acc;
        }
    return r;
}

SS_HD void mulvec(const M4& m, const float v[4], float out[4])
{
    for (int i = 0; i < 4; ++i) {
        const float lo = m.c[0][i] * v[0] + m.c[1][i] * v[1];
        const float hi = m.c[2][i] * v[2] + m.c[3][i] * v[3];
        out[i] = lo + hi;
    }
}

// cofactor expansion in glm's arrangement (six 2x2 sub-determinant triples, alternating signs)
SS_HD M4 inverse(const M4& m)
{
    const float (*a)[4] = m.c;
    float sub2[6][3];
    const int rows[6][2] = {{2, 3}, {1, 3}, {1, 2}, {0, 3}, {0, 2}, {0, 1}};
    for (int k = 0; k < 6; ++k) {
        const int p = rows[k][0], q = rows[k][1];
        sub2[k][0] = a[2][p] * a[3][q] - a[3][p] * a[2][q];
        sub2[k][1] = a[1][p] * a[3][q] - a[3][p] * a[1][q];
        sub2[k][2] = a[1][p] * a[2][q] - a[2][p] * a[1][q];
    }
    float fac[6][4];
    for (int k = 0; k < 6; ++k) {
        fac[k][0] = fac[k][1] = sub2[k][0];
        fac[k][2] = sub2[k][1];
        fac[k][3] = sub2[k][2];
    }
    float vec[4][4];
    for (int r = 0; r < 4; ++r) {
        vec[r][0] = a[1][r];
        vec[r][1] = vec[r][2] = vec[r][3] = a[0][r];
    }
    const float sign_a[4] = {1.f, -1.f, 1.f, -1.f}, sign_b[4] = {-1.f, 1.f, -1.f, 1.f};
    M4 inv;
    for (int k = 0; k < 4; ++k) {
        float c0 = vec[1][k] * fac[0][k] - vec[2][k] * fac[1][k];
        c0 = c0 + vec[3][k] * fac[2][k];
        float c1 = vec[0][k] * fac[0][k] - vec[2][k] * fac[3][k];
        c1 = c1 + vec[3][k] * fac[4][k];
        float c2 = vec[0][k] * fac[1][k] - vec[1][k] * fac[3][k];
        c2 = c2 + vec[3][k] * fac[5][k];
        float c3 = vec[0][k] * fac[2][k] - vec[1][k] * fac[4][k];
        c3 = c3 + vec[2][k] * fac[5][k];
        inv.c[0][k] = c0 * sign_a[k];
        inv.c[1][k] = c1 * sign_b[k];
        inv.c[2][k] = c2 * sign_a[k];
        inv.c[3][k] = c3 * sign_b[k];
    }
    const float t0 = a[0][0] * inv.c[0][0], t1 = a[0][1] * inv.c[1][0];
    const float t2 = a[0][2] * inv.c[2][0], t3 = a[0][3] * inv.c[3][0];
    const float inv_det = 1.0f / ((t0 + t1) + (t2 + t3));
    for (int j = 0; j < 4; ++j)
        for (int i = 0; i < 4; ++i) inv.c[j][i] = inv.c[j][i] * inv_det;
    return inv;
}

// rotation matrix of quat(vec3(ex, ey, ez)), host libm sin/cos like glm::sin/glm::cos
M4 euler_rotation(float ex, float ey, float ez)
{
    const float cx = std::cos(ex * 0.5f), cy = std::cos(ey * 0.5f), cz = std::cos(ez * 0.5f);
    const float sx = std::sin(ex * 0.5f), sy = std::sin(ey * 0.5f), sz = std::sin(ez * 0.5f);
    const float w = cx * cy * cz + sx * sy * sz;
    const float x = sx * cy * cz - cx * sy * sz;
    const float y = cx * sy * cz + sx * cy * sz;
    const float z = cx * cy * sz - sx * sy * cz;
    const float xx = x * x, yy = y * y, zz = z * z, xz = x * z, xy = x * y, yz = y * z;
    const float wx = w * x, wy = w * y, wz = w * z;
    M4 r = identity();
    r.c[0][0] = 1.0f - 2.0f * (yy + zz);
    r.c[0][1] = 2.0f * (xy + wz);
    r.c[0][2] = 2.0f * (xz - wy);
    r.c[1][0] = 2.0f * (xy - wz);
    r.c[1][1] = 1.0f - 2.0f * (xx + zz);
    r.c[1][2] = 2.0f * (yz + wx);
    r.c[2][0] = 2.0f * (xz + wy);
    r.c[2][1] = 2.0f * (yz - wx);
    r.c[2][2] = 1.0f - 2.0f * (xx + yy);
    return r;
}

M4 translation(V3 t)
{
    const M4 one = identity();
    M4 r = one;
    for (int i = 0; i < 4; ++i) {
        float acc = one.c[0][i] * t.x + one.c[1][i] * t.y;
        acc = acc + one.c[2][i] * t.z;
        r.c[3][i] = acc + one.c[3][i];
    }
    return r;
}

// The part of UpdateCascadeMap that depends on the camera (SceneRenderer.cpp:605-687), given the split fractions:
// slice the frustum, bound it with a sphere snapped to 1/16, look along the light, orthographic projection.
// Host and device compile this same text (-fmad=false / -ffp-contract=off, IEEE division and square root on both),
// so the batched device version gives the host version's bits.
// snap_size > 0: the CascadeOffset branch (SceneRenderer.cpp:662-679): the projection is shifted so that CascadeOrigin
// lands on a texel centre grid of a snap_size^2 map.
SS_HD void cascades_from_splits(const M4& vp, const float split[4], float z_near, float clip_range, float near_offset,
                                float far_offset, V3 light_in, float splits_out[4], float* light_matrices_out /* [4][16] */,
                                V3 snap_origin = {0.f, 0.f, 0.f}, float snap_size = 0.0f)
{
    const float ndc[8][3] = {{-1, 1, 0}, {1, 1, 0}, {1, -1, 0}, {-1, -1, 0}, {-1, 1, 1}, {1, 1, 1}, {1, -1, 1}, {-1, -1, 1}};
    const V3 light = normalize(light_in);
    float prev = 0.0f;
    for (unsigned i = 0; i < 4; ++i) {
        const M4 unproject = inverse(vp);
        V3 corner[8];
        for (int j = 0; j < 8; ++j) {
            const float h[4] = {ndc[j][0], ndc[j][1], ndc[j][2], 1.0f};
            float w[4];
            mulvec(unproject, h, w);
            corner[j] = {w[0] / w[3], w[1] / w[3], w[2] / w[3]};
        }
        for (int j = 0; j < 4; ++j) {  // cut the slice [prev, split[i]] out of each frustum edge
            const V3 edge = sub(corner[j + 4], corner[j]);
            corner[j + 4] = add(corner[j], mul(edge, split[i]));
            corner[j] = add(corner[j], mul(edge, prev));
        }
        V3 centre = {0.f, 0.f, 0.f};
        for (int j = 0; j < 8; ++j) centre = add(centre, corner[j]);
        centre = {centre.x / 8.0f, centre.y / 8.0f, centre.z / 8.0f};

        float radius = 0.0f;
        for (int j = 0; j < 8; ++j) {
            const V3 d = sub(corner[j], centre);
            const float len = sqrtf(dot(d, d));
            radius = radius < len ? len : radius;
        }
        radius = ceilf(radius * 16.0f) / 16.0f;
        const float lo = -radius, hi = radius;

        // lookAtRH(centre - light * radius, centre, +Y)
        const V3 eye = sub(centre, mul(light, -lo));
        const V3 f = normalize(sub(centre, eye));
        const V3 s = normalize(cross(f, {0.f, 1.f, 0.f}));
        const V3 u = cross(s, f);
        M4 lv = identity();
        lv.c[0][0] = s.x; lv.c[1][0] = s.y; lv.c[2][0] = s.z;
        lv.c[0][1] = u.x; lv.c[1][1] = u.y; lv.c[2][1] = u.z;
        lv.c[0][2] = -f.x; lv.c[1][2] = -f.y; lv.c[2][2] = -f.z;
        lv.c[3][0] = -dot(s, eye);
        lv.c[3][1] = -dot(u, eye);
        lv.c[3][2] = dot(f, eye);

        // orthoRH_ZO(lo, hi, lo, hi, near_offset, (hi - lo) + far_offset)
        const float zn = near_offset, zf = hi - lo + far_offset;
        M4 lo_m = identity();
        lo_m.c[0][0] = 2.0f / (hi - lo);
        lo_m.c[1][1] = 2.0f / (hi - lo);
        lo_m.c[2][2] = -1.0f / (zf - zn);
        lo_m.c[3][0] = -(hi + lo) / (hi - lo);
        lo_m.c[3][1] = -(hi + lo) / (hi - lo);
        lo_m.c[3][2] = -zn / (zf - zn);

        if (snap_size > 0.0f) {
            const M4 shadow = matmul(lo_m, lv);
            const float o[4] = {snap_origin.x, snap_origin.y, snap_origin.z, 1.0f};
            float so[4];
            mulvec(shadow, o, so);
            for (int r = 0; r < 2; ++r) {  // the z and w components of the offset are zeroed
                const float v = so[r] * snap_size / 2.0f;
                const float off = (roundf(v) - v) * 2.0f / snap_size;  // glm::round: half away from zero
                lo_m.c[3][r] = lo_m.c[3][r] + off;
            }
        }
        splits_out[i] = (z_near + split[i] * clip_range) * -1.0f;
        const M4 lm = matmul(lo_m, lv);
        for (int q = 0; q < 16; ++q) light_matrices_out[i * 16 + q] = (&lm.c[0][0])[q];
        prev = split[i];
    }
}

}  // namespace

extern "C" {

SS_API ss_status ss_camera_view_proj(float fov_y, float aspect, float z_near, float z_far, const float position[3],
                                     float pitch, float yaw, float view_out[16], float proj_out[16],
                                     float view_proj_out[16])
{
    if (!position) return SS_ERR_INVALID;
    M4 proj;
    std::memset(&proj, 0, sizeof proj);
    const float t = std::tan(fov_y / 2.0f);  // perspectiveRH_ZO
    proj.c[0][0] = 1.0f / (aspect * t);
    proj.c[1][1] = 1.0f / t;
    proj.c[2][2] = z_far / (z_near - z_far);
    proj.c[2][3] = -1.0f;
    proj.c[3][2] = -(z_far * z_near) / (z_far - z_near);

    const M4 cam_to_world = matmul(translation({position[0], position[1], position[2]}), euler_rotation(-pitch, -yaw, 0.0f));
    const M4 view = inverse(cam_to_world);
    const M4 vp = matmul(proj, view);
    if (view_out) std::memcpy(view_out, &view, 64);
    if (proj_out) std::memcpy(proj_out, &proj, 64);
    if (view_proj_out) std::memcpy(view_proj_out, &vp, 64);
    return SS_OK;
}

SS_API ss_status ss_cascade_matrices_snapped(const float view_proj[16], float z_near, float z_far, float split_lambda,
                                             float near_offset, float far_offset, const float light_dir[3],
                                             const float cascade_origin[3], float map_size, float splits_out[4],
                                             float light_matrices_out[4][16])
{
    if (!view_proj || !light_dir || !splits_out || !light_matrices_out || (cascade_origin && !(map_size > 0.0f))) return SS_ERR_INVALID;
    M4 vp;
    std::memcpy(&vp, view_proj, 64);
    const float clip_range = z_far - z_near;
    const float z_min = z_near, z_max = z_near + clip_range;
    const float range = z_max - z_min, ratio = z_max / z_min;

    float split[4];  // practical split scheme, fraction of the clip range
    for (unsigned i = 0; i < 4; ++i) {
        const float p = (float)(i + 1) / 4.0f;
        const float logarithmic = z_min * std::pow(ratio, p);
        const float uniform = z_min + range * p;
        const float d = split_lambda * (logarithmic - uniform) + uniform;
        split[i] = (d - z_near) / clip_range;
    }

    const V3 origin = cascade_origin ? V3{cascade_origin[0], cascade_origin[1], cascade_origin[2]} : V3{0.f, 0.f, 0.f};
    cascades_from_splits(vp, split, z_near, clip_range, near_offset, far_offset, {light_dir[0], light_dir[1], light_dir[2]}, splits_out,
                         &light_matrices_out[0][0], origin, cascade_origin ? map_size : 0.0f);
    return SS_OK;
}

SS_API ss_status ss_cascade_matrices(const float view_proj[16], float z_near, float z_far, float split_lambda,
                                     float near_offset, float far_offset, const float light_dir[3],
                                     float splits_out[4], float light_matrices_out[4][16])
{
    return ss_cascade_matrices_snapped(view_proj, z_near, z_far, split_lambda, near_offset, far_offset, light_dir, nullptr, 0.0f,
                                       splits_out, light_matrices_out);
}

}  // extern "C"

#ifdef __CUDACC__
namespace {
struct CascadeBatchArgs {
    const float* view_proj;  // [n][16] device
    float* out;              // [n][4][16] device
    uint32_t n;
    float split[4], z_near, clip_range, near_offset, far_offset;
    V3 light;
};
__global__ void k_cascades(const __grid_constant__ CascadeBatchArgs a)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    M4 vp;
    for (int q = 0; q < 16; ++q) (&vp.c[0][0])[q] = a.view_proj[(size_t)i * 16 + q];
    float splits[4];
    cascades_from_splits(vp, a.split, a.z_near, a.clip_range, a.near_offset, a.far_offset, a.light, splits, a.out + (size_t)i * 64);
}
}  // namespace

extern "C" SS_API ss_status ss_cascade_matrices_batch(int device, const float* view_proj, uint32_t n_cameras, float z_near, float z_far,
                                                      float split_lambda, float near_offset, float far_offset,
                                                      const float light_dir[3], float splits_out[4], float* light_matrices_out)
{
    if (!view_proj || !light_dir || !splits_out || !light_matrices_out) return SS_ERR_INVALID;
    if (n_cameras == 0) return SS_OK;
    CascadeBatchArgs a;
    const float clip_range = z_far - z_near, z_min = z_near, z_max = z_near + clip_range;
    const float range = z_max - z_min, ratio = z_max / z_min;
    for (unsigned i = 0; i < 4; ++i) {  // the camera-independent part stays on the host (libm pow, like the reference)
        const float p = (float)(i + 1) / 4.0f;
        const float logarithmic = z_min * std::pow(ratio, p);
        const float uniform = z_min + range * p;
        const float d = split_lambda * (logarithmic - uniform) + uniform;
        a.split[i] = (d - z_near) / clip_range;
        splits_out[i] = (z_near + a.split[i] * clip_range) * -1.0f;
    }
    a.n = n_cameras;
    a.z_near = z_near;
    a.clip_range = clip_range;
    a.near_offset = near_offset;
    a.far_offset = far_offset;
    a.light = {light_dir[0], light_dir[1], light_dir[2]};
    if (cudaSetDevice(device) != cudaSuccess) return SS_ERR_CUDA;
    float *d_in = nullptr, *d_out = nullptr;
    ss_status st = SS_OK;
    if (cudaMalloc(&d_in, (size_t)n_cameras * 64) != cudaSuccess || cudaMalloc(&d_out, (size_t)n_cameras * 256) != cudaSuccess) st = SS_ERR_CUDA;
    if (st == SS_OK && cudaMemcpy(d_in, view_proj, (size_t)n_cameras * 64, cudaMemcpyHostToDevice) != cudaSuccess) st = SS_ERR_CUDA;
    if (st == SS_OK) {
        a.view_proj = d_in;
        a.out = d_out;
        k_cascades<<<(n_cameras + 127) / 128, 128>>>(a);
        if (cudaMemcpy(light_matrices_out, d_out, (size_t)n_cameras * 256, cudaMemcpyDeviceToHost) != cudaSuccess) st = SS_ERR_CUDA;
    }
    if (d_in) cudaFree(d_in);
    if (d_out) cudaFree(d_out);
    return st;
}
#endif
