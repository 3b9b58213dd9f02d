/*
 * CPU oracle for libscenesubmit — TEST INFRASTRUCTURE, NOT PRODUCT (see scenesubmit_oracle.h).
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math -std=c11 -fPIC -shared -pthread  (oracle/Makefile).
 * Float op order follows the scalar GLM 0.9.9.9 the reference vendors; no FMA, no reassociation.
 * Citations: "VC/" = /root/reference/VulkanCore/src/VulkanCore/, "GLM/" = /root/reference/VulkanCore/vendor/glm/glm/.
 */
#define _GNU_SOURCE
#include "scenesubmit_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "../include/scenesubmit_dettrig.h" /* the deterministic sin/cos DEFINITION (spec header) */

/* m[c*4 + r] = glm m[c][r] */
#define M(m, c, r) ((m)[(c) * 4 + (r)])

void sso_det_sincos(float x, float* s, float* c) { ss_det_sincosf(x, s, c); }

/* GLM/detail/type_mat4x4.inl:630-648: Result[j] = A0*B[j][0] + A1*B[j][1] + A2*B[j][2] + A3*B[j][3],
   left to right. */
void sso_mat4_mul(const float a[16], const float b[16], float out[16])
{
    float r[16];
    for (int j = 0; j < 4; ++j) {
        for (int i = 0; i < 4; ++i) {
            float t0 = M(a, 0, i) * M(b, j, 0);
            float t1 = M(a, 1, i) * M(b, j, 1);
            float t2 = M(a, 2, i) * M(b, j, 2);
            float t3 = M(a, 3, i) * M(b, j, 3);
            float s = t0 + t1;
            s = s + t2;
            s = s + t3;
            M(r, j, i) = s;
        }
    }
    memcpy(out, r, sizeof r);
}

static void mat4_identity(float m[16])
{
    memset(m, 0, 16 * sizeof(float));
    M(m, 0, 0) = M(m, 1, 1) = M(m, 2, 2) = M(m, 3, 3) = 1.0f;
}

static void sincos_mode(float x, int trig_mode, float* s, float* c)
{
    if (trig_mode == SSO_TRIG_DETERMINISTIC) {
        ss_det_sincosf(x, s, c);
    } else { /* glm::sin / glm::cos -> std::sin(float) -> libm sinf */
        *s = sinf(x);
        *c = cosf(x);
    }
}

/* GLM/detail/type_quat.inl:208-217 quat(eulerAngle) and GLM/gtc/quaternion.inl:47-72 mat3_cast. */
static void euler_to_mat4(const float e[3], int trig_mode, float out[16])
{
    float cx, cy, cz, sx, sy, sz;
    sincos_mode(e[0] * 0.5f, trig_mode, &sx, &cx);
    sincos_mode(e[1] * 0.5f, trig_mode, &sy, &cy);
    sincos_mode(e[2] * 0.5f, trig_mode, &sz, &cz);

    float qw = cx * cy * cz + sx * sy * sz;
    float qx = sx * cy * cz - cx * sy * sz;
    float qy = cx * sy * cz + sx * cy * sz;
    float qz = cx * cy * sz - sx * sy * cz;

    float qxx = qx * qx, qyy = qy * qy, qzz = qz * qz;
    float qxz = qx * qz, qxy = qx * qy, qyz = qy * qz;
    float qwx = qw * qx, qwy = qw * qy, qwz = qw * qz;

    mat4_identity(out);
    M(out, 0, 0) = 1.0f - 2.0f * (qyy + qzz);
    M(out, 0, 1) = 2.0f * (qxy + qwz);
    M(out, 0, 2) = 2.0f * (qxz - qwy);
    M(out, 1, 0) = 2.0f * (qxy - qwz);
    M(out, 1, 1) = 1.0f - 2.0f * (qxx + qzz);
    M(out, 1, 2) = 2.0f * (qyz + qwx);
    M(out, 2, 0) = 2.0f * (qxz + qwy);
    M(out, 2, 1) = 2.0f * (qyz - qwx);
    M(out, 2, 2) = 1.0f - 2.0f * (qxx + qyy);
}

/* GLM/ext/matrix_transform.inl:10-15 with m = identity:
   Result[3] = m[0]*v[0] + m[1]*v[1] + m[2]*v[2] + m[3]. */
static void mat4_translate_identity(const float v[3], float out[16])
{
    float m[16];
    mat4_identity(m);
    mat4_identity(out);
    for (int r = 0; r < 4; ++r) {
        float s = M(m, 0, r) * v[0] + M(m, 1, r) * v[1];
        s = s + M(m, 2, r) * v[2];
        s = s + M(m, 3, r);
        M(out, 3, r) = s;
    }
}

/* GLM/ext/matrix_transform.inl:78-86 with m = identity. */
static void mat4_scale_identity(const float v[3], float out[16])
{
    float m[16];
    mat4_identity(m);
    for (int r = 0; r < 4; ++r) {
        M(out, 0, r) = M(m, 0, r) * v[0];
        M(out, 1, r) = M(m, 1, r) * v[1];
        M(out, 2, r) = M(m, 2, r) * v[2];
        M(out, 3, r) = M(m, 3, r);
    }
}

/* VC/Core/Components.h:22-29: translate(I,T) * toMat4(quat(R)) * scale(I,S), left-associated. */
void sso_trs_to_mat4(const float trs[9], int trig_mode, float out[16])
{
    float t[16], r[16], s[16], tr[16];
    mat4_translate_identity(trs + 0, t);
    euler_to_mat4(trs + 3, trig_mode, r);
    mat4_scale_identity(trs + 6, s);
    sso_mat4_mul(t, r, tr);
    sso_mat4_mul(tr, s, out);
}

/* VC/Scene/SceneRenderer.cpp:1151-1153: MRow[r] = (m[0][r], m[1][r], m[2][r], m[3][r]). */
void sso_pack_rows(const float m[16], float rows[12])
{
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 4; ++c) rows[r * 4 + c] = M(m, c, r);
}

/* N1: world[i] = world[parent[i]] * local[i]; roots: world = local. Nodes are level-ordered so a
   single forward sweep sees every parent first. */
void sso_node_worlds(const sso_scene* sc, float* node_world16)
{
    for (uint32_t i = 0; i < sc->n_nodes; ++i) {
        float local[16];
        sso_trs_to_mat4(sc->trs + (size_t)i * 9, (int)sc->trig_mode, local);
        int32_t p = sc->parent ? sc->parent[i] : -1;
        if (p < 0)
            memcpy(node_world16 + (size_t)i * 16, local, sizeof local);
        else
            sso_mat4_mul(node_world16 + (size_t)p * 16, local, node_world16 + (size_t)i * 16);
    }
}

static void instance_rows_range(const sso_scene* sc, const float* node_world16, float* rows12,
                                uint32_t lo, uint32_t hi)
{
    for (uint32_t i = lo; i < hi; ++i) {
        float* rows = rows12 + (size_t)i * 12;
        uint32_t slot = sc->slot[i];
        if (slot == SSO_SLOT_NULL) { /* VC/Scene/SceneRenderer.cpp:1128-1136: dropped */
            memset(rows, 0, 12 * sizeof(float));
            continue;
        }
        uint32_t n = sc->node ? sc->node[i] : sc->node_base + i;
        float w[16];
        /* :1150 submeshTransform = transform * submeshData[submeshIndex].LocalTransform */
        sso_mat4_mul(node_world16 + (size_t)n * 16, sc->slots[slot].local, w);
        sso_pack_rows(w, rows);
    }
}

void sso_instance_rows(const sso_scene* sc, const float* node_world16, float* rows12)
{
    instance_rows_range(sc, node_world16, rows12, 0, sc->n_instances);
}

/* N3: Gribb-Hartmann planes of a zero-to-one-depth clip matrix (GLM_FORCE_DEPTH_ZERO_TO_ONE,
   /root/reference/VulkanCore/premake5.lua:47).  Row r of VP = (m[0][r], m[1][r], m[2][r], m[3][r]).
   A point is inside when  -w <= x <= w, -w <= y <= w, 0 <= z <= w
   (the clip volume the vertex stages feed, VulkanCore/assets/shaders/CorePBR.vert:21-29). */
void sso_planes_from_view_proj(const float vp[16], float planes[6][4])
{
    for (int c = 0; c < 4; ++c) {
        float r0 = M(vp, c, 0), r1 = M(vp, c, 1), r2 = M(vp, c, 2), r3 = M(vp, c, 3);
        planes[0][c] = r3 + r0; /* left   */
        planes[1][c] = r3 - r0; /* right  */
        planes[2][c] = r3 + r1; /* bottom */
        planes[3][c] = r3 - r1; /* top    */
        planes[4][c] = r2;      /* near (z >= 0) */
        planes[5][c] = r3 - r2; /* far  (z <= w) */
    }
}

/* N2 + N4 for one instance; rows = TransformData rows (affine world matrix). */
static uint8_t cull_one(const float rows[12], const float box[6], float planes[][6][4],
                        const uint32_t* view_flags, uint32_t n_views)
{
    float cl[3], el[3];
    for (int k = 0; k < 3; ++k) {
        cl[k] = (box[k] + box[3 + k]) * 0.5f;
        el[k] = (box[3 + k] - box[k]) * 0.5f;
    }
    float cw[3], ew[3];
    for (int r = 0; r < 3; ++r) {
        const float* R = rows + r * 4; /* R[c] = m[c][r] */
        /* centre: GLM mat4*vec4 association (GLM/detail/type_mat4x4.inl:561-572) with w = 1 */
        float a0 = R[0] * cl[0] + R[1] * cl[1];
        float a1 = R[2] * cl[2] + R[3];
        cw[r] = a0 + a1;
        float e0 = fabsf(R[0]) * el[0] + fabsf(R[1]) * el[1];
        ew[r] = e0 + fabsf(R[2]) * el[2];
    }
    uint8_t mask = 0;
    for (uint32_t v = 0; v < n_views; ++v) {
        int inside = 1;
        for (int p = 0; p < 6; ++p) {
            if (p == 4 && (view_flags[v] & SSO_VIEW_SKIP_NEAR)) continue;
            const float* pl = planes[v][p];
            float s = pl[0] * cw[0] + pl[1] * cw[1];
            s = s + pl[2] * cw[2];
            s = s + pl[3];
            float r = fabsf(pl[0]) * ew[0] + fabsf(pl[1]) * ew[1];
            r = r + fabsf(pl[2]) * ew[2];
            if (!(s + r >= 0.0f)) inside = 0;
        }
        if (inside) mask |= (uint8_t)(1u << v);
    }
    return mask;
}

static void cull_range(const sso_scene* sc, const float* rows12, const sso_view* views,
                       uint32_t n_views, uint8_t* masks, uint32_t lo, uint32_t hi)
{
    float planes[8][6][4];
    uint32_t flags[8];
    for (uint32_t v = 0; v < n_views; ++v) {
        sso_planes_from_view_proj(views[v].view_proj, planes[v]);
        flags[v] = views[v].flags;
    }
    for (uint32_t i = lo; i < hi; ++i) {
        uint32_t slot = sc->slot[i];
        if (slot == SSO_SLOT_NULL) {
            masks[i] = 0;
            continue;
        }
        float box[6];
        if (sc->inst_aabb) {
            memcpy(box, sc->inst_aabb + (size_t)i * 6, sizeof box);
        } else {
            memcpy(box, sc->slots[slot].aabb_min, 3 * sizeof(float));
            memcpy(box + 3, sc->slots[slot].aabb_max, 3 * sizeof(float));
        }
        masks[i] = cull_one(rows12 + (size_t)i * 12, box, planes, flags, n_views);
    }
}

void sso_cull(const sso_scene* sc, const float* rows12, const sso_view* views, uint32_t n_views,
              uint8_t* masks)
{
    cull_range(sc, rows12, views, n_views, masks, 0, sc->n_instances);
}

/* N5: [tess][meshRank][submeshIndex][materialRank], right-aligned; isomorphic to
   MeshKey::operator< (VC/Scene/SceneRenderer.h:64-85), tessellated list after the static one
   (VC/Scene/SceneRenderer.cpp:1368-1382). */
static void keys_range(const sso_scene* sc, uint64_t* keys, uint32_t lo, uint32_t hi)
{
    const uint32_t mb = sc->material_bits, sb = sc->submesh_bits, hb = sc->mesh_bits;
    for (uint32_t i = lo; i < hi; ++i) {
        uint32_t slot = sc->slot[i];
        if (slot == SSO_SLOT_NULL) {
            keys[i] = ~0ull;
            continue;
        }
        uint32_t mw = sc->matword[i];
        uint64_t tess = (mw & SSO_TESS_BIT) ? 1u : 0u;
        uint64_t mat = mw & ~SSO_TESS_BIT;
        uint64_t k = mat;
        k |= (uint64_t)sc->slots[slot].submesh_index << mb;
        k |= (uint64_t)sc->slots[slot].mesh_rank << (mb + sb);
        k |= tess << (mb + sb + hb);
        keys[i] = k;
    }
}

void sso_keys(const sso_scene* sc, uint64_t* keys) { keys_range(sc, keys, 0, sc->n_instances); }

typedef struct {
    uint64_t key;
    uint32_t idx;
} key_idx;

static int cmp_key_idx(const void* a, const void* b)
{
    const key_idx* x = (const key_idx*)a;
    const key_idx* y = (const key_idx*)b;
    if (x->key != y->key) return x->key < y->key ? -1 : 1;
    if (x->idx != y->idx) return x->idx < y->idx ? -1 : 1;
    return 0;
}

/* N6 for one view: what iterating std::map<MeshKey, MeshTransform> yields
   (VC/Scene/SceneRenderer.cpp:1336-1337,1368-1369), restricted to the visible instances. */
static void view_list_one(const sso_scene* sc, const float* rows12, const uint8_t* masks,
                          const uint64_t* keys, uint32_t v, uint32_t* n_visible, uint32_t* n_draws,
                          uint32_t* out_idx, uint64_t* out_keys, float* out_rows12,
                          sso_draw_cmd* out_draws, uint64_t* out_draw_keys)
{
    uint32_t n = sc->n_instances, m = 0;
    key_idx* list = (key_idx*)malloc((size_t)(n ? n : 1) * sizeof(key_idx));
    for (uint32_t i = 0; i < n; ++i) {
        if (masks[i] & (1u << v)) {
            list[m].key = keys[i];
            list[m].idx = i;
            ++m;
        }
    }
    qsort(list, m, sizeof(key_idx), cmp_key_idx); /* (key, idx) is a total order: stable by construction */
    uint32_t d = 0;
    for (uint32_t j = 0; j < m; ++j) {
        out_idx[j] = list[j].idx;
        out_keys[j] = list[j].key;
        if (out_rows12) memcpy(out_rows12 + (size_t)j * 12, rows12 + (size_t)list[j].idx * 12, 48);
        if (j == 0 || list[j].key != list[j - 1].key) {
            const sso_submesh* sm = &sc->slots[sc->slot[list[j].idx]];
            /* VC/../Platform/Vulkan/VulkanRenderer.cpp:628
               vkCmdDrawIndexed(IndexCount, instanceCount, BaseIndex, BaseVertex, firstInstance) */
            out_draws[d].index_count = sm->index_count;
            out_draws[d].instance_count = 0;
            out_draws[d].first_index = sm->first_index;
            out_draws[d].vertex_offset = sm->vertex_offset;
            out_draws[d].first_instance = j;
            out_draw_keys[d] = list[j].key;
            ++d;
        }
        out_draws[d - 1].instance_count++;
    }
    *n_visible = m;
    *n_draws = d;
    free(list);
}

int sso_view_lists(const sso_scene* sc, const float* rows12, const uint8_t* masks,
                   const uint64_t* keys, uint32_t n_views, uint32_t* n_visible, uint32_t* n_draws,
                   uint32_t* out_idx, uint64_t* out_keys, float* out_rows12,
                   sso_draw_cmd* out_draws, uint64_t* out_draw_keys)
{
    size_t n = sc->n_instances;
    for (uint32_t v = 0; v < n_views; ++v)
        view_list_one(sc, rows12, masks, keys, v, &n_visible[v], &n_draws[v], out_idx + v * n,
                      out_keys + v * n, out_rows12 ? out_rows12 + v * n * 12 : NULL,
                      out_draws + v * n, out_draw_keys + v * n);
    return 0;
}

/* ---- threaded frame (CPU baseline) ---------------------------------------------------------- */

typedef struct {
    const sso_scene* sc;
    const sso_view* views;
    uint32_t n_views;
    const float* node_world;
    float* rows12;
    uint8_t* masks;
    uint64_t* keys;
    uint32_t lo, hi;
} range_job;

static void* range_worker(void* p)
{
    range_job* j = (range_job*)p;
    instance_rows_range(j->sc, j->node_world, j->rows12, j->lo, j->hi);
    cull_range(j->sc, j->rows12, j->views, j->n_views, j->masks, j->lo, j->hi);
    keys_range(j->sc, j->keys, j->lo, j->hi);
    return NULL;
}

typedef struct {
    const sso_scene* sc;
    const float* rows12;
    const uint8_t* masks;
    const uint64_t* keys;
    uint32_t v;
    uint32_t *n_visible, *n_draws, *out_idx;
    uint64_t* out_keys;
    float* out_rows12;
    sso_draw_cmd* out_draws;
    uint64_t* out_draw_keys;
} view_job;

static void* view_worker(void* p)
{
    view_job* j = (view_job*)p;
    view_list_one(j->sc, j->rows12, j->masks, j->keys, j->v, j->n_visible, j->n_draws, j->out_idx,
                  j->out_keys, j->out_rows12, j->out_draws, j->out_draw_keys);
    return NULL;
}

static double now_s(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

double sso_frame(const sso_scene* sc, const sso_view* views, uint32_t n_views, int threads,
                 float* rows12, uint8_t* masks, uint64_t* keys, uint32_t* n_visible,
                 uint32_t* n_draws, uint32_t* out_idx, uint64_t* out_keys, float* out_rows12,
                 sso_draw_cmd* out_draws, uint64_t* out_draw_keys)
{
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    float* node_world = (float*)malloc((size_t)(sc->n_nodes ? sc->n_nodes : 1) * 16 * sizeof(float));
    double t0 = now_s();
    /* the parent chain is a dependent sweep; interior nodes are few (levels above the leaves), so it
       stays on one thread like the reference's main-thread loop (VC/Scene/Scene.cpp:138-146) */
    sso_node_worlds(sc, node_world);
    size_t n = sc->n_instances;
    if (threads == 1) {
        range_job j = {sc, views, n_views, node_world, rows12, masks, keys, 0, (uint32_t)n};
        range_worker(&j);
        sso_view_lists(sc, rows12, masks, keys, n_views, n_visible, n_draws, out_idx, out_keys,
                       out_rows12, out_draws, out_draw_keys);
    } else {
        pthread_t th[256];
        range_job jobs[256];
        for (int t = 0; t < threads; ++t) {
            jobs[t] = (range_job){sc, views, n_views, node_world, rows12, masks, keys,
                                  (uint32_t)(n * (size_t)t / (size_t)threads),
                                  (uint32_t)(n * (size_t)(t + 1) / (size_t)threads)};
            pthread_create(&th[t], NULL, range_worker, &jobs[t]);
        }
        for (int t = 0; t < threads; ++t) pthread_join(th[t], NULL);
        view_job vj[8];
        for (uint32_t v = 0; v < n_views; ++v) {
            vj[v] = (view_job){sc, rows12, masks, keys, v, &n_visible[v], &n_draws[v],
                               out_idx + v * n, out_keys + v * n,
                               out_rows12 ? out_rows12 + v * n * 12 : NULL, out_draws + v * n,
                               out_draw_keys + v * n};
            pthread_create(&th[v], NULL, view_worker, &vj[v]);
        }
        for (uint32_t v = 0; v < n_views; ++v) pthread_join(th[v], NULL);
    }
    double t1 = now_s();
    free(node_world);
    return t1 - t0;
}

/* ---- view matrices ---------------------------------------------------------------------------- */

/* GLM/detail/func_matrix.inl:347-403 */
void sso_mat4_inverse(const float m[16], float out[16])
{
    float Coef00 = M(m, 2, 2) * M(m, 3, 3) - M(m, 3, 2) * M(m, 2, 3);
    float Coef02 = M(m, 1, 2) * M(m, 3, 3) - M(m, 3, 2) * M(m, 1, 3);
    float Coef03 = M(m, 1, 2) * M(m, 2, 3) - M(m, 2, 2) * M(m, 1, 3);
    float Coef04 = M(m, 2, 1) * M(m, 3, 3) - M(m, 3, 1) * M(m, 2, 3);
    float Coef06 = M(m, 1, 1) * M(m, 3, 3) - M(m, 3, 1) * M(m, 1, 3);
    float Coef07 = M(m, 1, 1) * M(m, 2, 3) - M(m, 2, 1) * M(m, 1, 3);
    float Coef08 = M(m, 2, 1) * M(m, 3, 2) - M(m, 3, 1) * M(m, 2, 2);
    float Coef10 = M(m, 1, 1) * M(m, 3, 2) - M(m, 3, 1) * M(m, 1, 2);
    float Coef11 = M(m, 1, 1) * M(m, 2, 2) - M(m, 2, 1) * M(m, 1, 2);
    float Coef12 = M(m, 2, 0) * M(m, 3, 3) - M(m, 3, 0) * M(m, 2, 3);
    float Coef14 = M(m, 1, 0) * M(m, 3, 3) - M(m, 3, 0) * M(m, 1, 3);
    float Coef15 = M(m, 1, 0) * M(m, 2, 3) - M(m, 2, 0) * M(m, 1, 3);
    float Coef16 = M(m, 2, 0) * M(m, 3, 2) - M(m, 3, 0) * M(m, 2, 2);
    float Coef18 = M(m, 1, 0) * M(m, 3, 2) - M(m, 3, 0) * M(m, 1, 2);
    float Coef19 = M(m, 1, 0) * M(m, 2, 2) - M(m, 2, 0) * M(m, 1, 2);
    float Coef20 = M(m, 2, 0) * M(m, 3, 1) - M(m, 3, 0) * M(m, 2, 1);
    float Coef22 = M(m, 1, 0) * M(m, 3, 1) - M(m, 3, 0) * M(m, 1, 1);
    float Coef23 = M(m, 1, 0) * M(m, 2, 1) - M(m, 2, 0) * M(m, 1, 1);

    float Fac0[4] = {Coef00, Coef00, Coef02, Coef03};
    float Fac1[4] = {Coef04, Coef04, Coef06, Coef07};
    float Fac2[4] = {Coef08, Coef08, Coef10, Coef11};
    float Fac3[4] = {Coef12, Coef12, Coef14, Coef15};
    float Fac4[4] = {Coef16, Coef16, Coef18, Coef19};
    float Fac5[4] = {Coef20, Coef20, Coef22, Coef23};
    float Vec0[4] = {M(m, 1, 0), M(m, 0, 0), M(m, 0, 0), M(m, 0, 0)};
    float Vec1[4] = {M(m, 1, 1), M(m, 0, 1), M(m, 0, 1), M(m, 0, 1)};
    float Vec2[4] = {M(m, 1, 2), M(m, 0, 2), M(m, 0, 2), M(m, 0, 2)};
    float Vec3[4] = {M(m, 1, 3), M(m, 0, 3), M(m, 0, 3), M(m, 0, 3)};
    static const float SignA[4] = {+1, -1, +1, -1};
    static const float SignB[4] = {-1, +1, -1, +1};
    float inv[16];
    for (int k = 0; k < 4; ++k) {
        float i0 = Vec1[k] * Fac0[k] - Vec2[k] * Fac1[k];
        i0 = i0 + Vec3[k] * Fac2[k];
        float i1 = Vec0[k] * Fac0[k] - Vec2[k] * Fac3[k];
        i1 = i1 + Vec3[k] * Fac4[k];
        float i2 = Vec0[k] * Fac1[k] - Vec1[k] * Fac3[k];
        i2 = i2 + Vec3[k] * Fac5[k];
        float i3 = Vec0[k] * Fac2[k] - Vec1[k] * Fac4[k];
        i3 = i3 + Vec2[k] * Fac5[k];
        M(inv, 0, k) = i0 * SignA[k];
        M(inv, 1, k) = i1 * SignB[k];
        M(inv, 2, k) = i2 * SignA[k];
        M(inv, 3, k) = i3 * SignB[k];
    }
    float d0 = M(m, 0, 0) * M(inv, 0, 0), d1 = M(m, 0, 1) * M(inv, 1, 0);
    float d2 = M(m, 0, 2) * M(inv, 2, 0), d3 = M(m, 0, 3) * M(inv, 3, 0);
    float det = (d0 + d1) + (d2 + d3);
    float ood = 1.0f / det;
    for (int k = 0; k < 16; ++k) out[k] = inv[k] * ood;
}

/* GLM mat4 * vec4 (GLM/detail/type_mat4x4.inl:561-572) */
static void mat4_mul_vec4(const float m[16], const float v[4], float out[4])
{
    for (int r = 0; r < 4; ++r) {
        float a0 = M(m, 0, r) * v[0] + M(m, 1, r) * v[1];
        float a1 = M(m, 2, r) * v[2] + M(m, 3, r) * v[3];
        out[r] = a0 + a1;
    }
}

static float dot3(const float a[3], const float b[3])
{ /* GLM/detail/func_geometric.inl:48-55 */
    float t0 = a[0] * b[0], t1 = a[1] * b[1], t2 = a[2] * b[2];
    return t0 + t1 + t2;
}

static void cross3(const float x[3], const float y[3], float out[3])
{ /* GLM/detail/func_geometric.inl:68-79 */
    out[0] = x[1] * y[2] - y[1] * x[2];
    out[1] = x[2] * y[0] - y[2] * x[0];
    out[2] = x[0] * y[1] - y[0] * x[1];
}

static void normalize3(const float v[3], float out[3])
{ /* v * inversesqrt(dot(v,v)); inversesqrt = 1/sqrt (GLM/detail/func_exponential.inl:134-139) */
    float inv = 1.0f / sqrtf(dot3(v, v));
    out[0] = v[0] * inv;
    out[1] = v[1] * inv;
    out[2] = v[2] * inv;
}

/* VC/Renderer/EditorCamera.cpp:15-33 + EditorCamera.h:45.  The orbit parameters
   (focal point, distance) only produce m_Position; here the position is given directly. */
void sso_camera_view_proj(float fov_y, float aspect, float z_near, float z_far,
                          const float position[3], float pitch, float yaw, float view[16],
                          float proj[16], float view_proj[16])
{
    /* GLM/ext/matrix_clip_space.inl:233-246 perspectiveRH_ZO */
    float tanHalfFovy = tanf(fov_y / 2.0f);
    memset(proj, 0, 16 * sizeof(float));
    M(proj, 0, 0) = 1.0f / (aspect * tanHalfFovy);
    M(proj, 1, 1) = 1.0f / tanHalfFovy;
    M(proj, 2, 2) = z_far / (z_near - z_far);
    M(proj, 2, 3) = -1.0f;
    M(proj, 3, 2) = -(z_far * z_near) / (z_far - z_near);

    /* GetOrientation: quat(vec3(-pitch, -yaw, 0)) (EditorCamera.cpp:172-175) */
    float e[3] = {-pitch, -yaw, 0.0f};
    float rot[16], tr[16], inv_view[16];
    euler_to_mat4(e, SSO_TRIG_LIBM, rot);
    mat4_translate_identity(position, tr);
    sso_mat4_mul(tr, rot, inv_view);
    sso_mat4_inverse(inv_view, view);
    sso_mat4_mul(proj, view, view_proj);
}

/* VC/Scene/SceneRenderer.cpp:579-687.  cascade_origin == NULL: CascadeOffset == 0; otherwise the texel-snap branch
   (:662-679) with CascadeOrigin and MapSize. */
void sso_cascade_matrices_snapped(const float view_proj[16], float z_near, float z_far, float lambda,
                                  float near_offset, float far_offset, const float light_dir[3],
                                  const float* cascade_origin, float map_size, float splits[4], float light_matrices[64])
{
    float cascadeSplits[4];
    float clipRange = z_far - z_near;
    float minZ = z_near;
    float maxZ = z_near + clipRange;
    float range = maxZ - minZ;
    float ratio = maxZ / minZ;
    for (uint32_t i = 0; i < 4; ++i) {
        float p = (float)(i + 1) / (float)4;
        float lg = minZ * powf(ratio, p);
        float uniform = minZ + range * p;
        float d = lambda * (lg - uniform) + uniform;
        cascadeSplits[i] = (d - z_near) / clipRange;
    }

    float lastSplitDist = 0.0f;
    for (uint32_t i = 0; i < 4; ++i) {
        float splitDist = cascadeSplits[i];
        float corners[8][3] = {{-1.0f, 1.0f, 0.0f}, {1.0f, 1.0f, 0.0f},  {1.0f, -1.0f, 0.0f},
                               {-1.0f, -1.0f, 0.0f}, {-1.0f, 1.0f, 1.0f}, {1.0f, 1.0f, 1.0f},
                               {1.0f, -1.0f, 1.0f},  {-1.0f, -1.0f, 1.0f}};
        float inv[16];
        sso_mat4_inverse(view_proj, inv);
        for (int j = 0; j < 8; ++j) {
            float v4[4] = {corners[j][0], corners[j][1], corners[j][2], 1.0f}, o[4];
            mat4_mul_vec4(inv, v4, o);
            corners[j][0] = o[0] / o[3];
            corners[j][1] = o[1] / o[3];
            corners[j][2] = o[2] / o[3];
        }
        for (int j = 0; j < 4; ++j) {
            for (int k = 0; k < 3; ++k) {
                float dist = corners[j + 4][k] - corners[j][k];
                corners[j + 4][k] = corners[j][k] + (dist * splitDist);
                corners[j][k] = corners[j][k] + (dist * lastSplitDist);
            }
        }
        float center[3] = {0.0f, 0.0f, 0.0f};
        for (int j = 0; j < 8; ++j)
            for (int k = 0; k < 3; ++k) center[k] += corners[j][k];
        for (int k = 0; k < 3; ++k) center[k] /= 8.0f;

        float radius = 0.0f;
        for (int j = 0; j < 8; ++j) {
            float dlt[3] = {corners[j][0] - center[0], corners[j][1] - center[1],
                            corners[j][2] - center[2]};
            float distance = sqrtf(dot3(dlt, dlt));
            radius = radius < distance ? distance : radius; /* glm::max(radius, distance) */
        }
        radius = ceilf(radius * 16.0f) / 16.0f;

        float maxE = radius, minE = -radius;
        float ld[3];
        normalize3(light_dir, ld);
        float eye[3];
        for (int k = 0; k < 3; ++k) eye[k] = center[k] - ld[k] * -minE;

        /* GLM/ext/matrix_transform.inl:99-119 lookAtRH(eye, center, up = (0,1,0)) */
        float up[3] = {0.0f, 1.0f, 0.0f};
        float fdir[3] = {center[0] - eye[0], center[1] - eye[1], center[2] - eye[2]};
        float f[3], s[3], u[3], sc[3];
        normalize3(fdir, f);
        cross3(f, up, sc);
        normalize3(sc, s);
        cross3(s, f, u);
        float lv[16];
        mat4_identity(lv);
        M(lv, 0, 0) = s[0]; M(lv, 1, 0) = s[1]; M(lv, 2, 0) = s[2];
        M(lv, 0, 1) = u[0]; M(lv, 1, 1) = u[1]; M(lv, 2, 1) = u[2];
        M(lv, 0, 2) = -f[0]; M(lv, 1, 2) = -f[1]; M(lv, 2, 2) = -f[2];
        M(lv, 3, 0) = -dot3(s, eye);
        M(lv, 3, 1) = -dot3(u, eye);
        M(lv, 3, 2) = dot3(f, eye);

        /* GLM/ext/matrix_clip_space.inl:42-52 orthoRH_ZO */
        float left = minE, right = maxE, bottom = minE, top = maxE;
        float zn = near_offset, zf = maxE - minE + far_offset;
        float lo[16];
        mat4_identity(lo);
        M(lo, 0, 0) = 2.0f / (right - left);
        M(lo, 1, 1) = 2.0f / (top - bottom);
        M(lo, 2, 2) = -1.0f / (zf - zn);
        M(lo, 3, 0) = -(right + left) / (right - left);
        M(lo, 3, 1) = -(top + bottom) / (top - bottom);
        M(lo, 3, 2) = -zn / (zf - zn);

        if (cascade_origin) {
            /* shadowOrigin = (lo * lv) * vec4(origin, 1); glm mat4 * vec4: (m0*x + m1*y) + (m2*z + m3*w) */
            float sm[16];
            sso_mat4_mul(lo, lv, sm);
            const float ov[4] = {cascade_origin[0], cascade_origin[1], cascade_origin[2], 1.0f};
            for (int r = 0; r < 2; ++r) { /* z and w of the offset are zeroed (:674-675) */
                float a = M(sm, 0, r) * ov[0] + M(sm, 1, r) * ov[1];
                float b = M(sm, 2, r) * ov[2] + M(sm, 3, r) * ov[3];
                float so = a + b;
                so = so * map_size / 2.0f;
                float off = roundf(so) - so; /* glm::round = round half away from zero */
                off = off * 2.0f / map_size;
                M(lo, 3, r) = M(lo, 3, r) + off;
            }
        }
        splits[i] = (z_near + splitDist * clipRange) * -1.0f;
        sso_mat4_mul(lo, lv, light_matrices + i * 16);
        lastSplitDist = cascadeSplits[i];
    }
}

void sso_cascade_matrices(const float view_proj[16], float z_near, float z_far, float lambda,
                          float near_offset, float far_offset, const float light_dir[3],
                          float splits[4], float light_matrices[64])
{
    sso_cascade_matrices_snapped(view_proj, z_near, z_far, lambda, near_offset, far_offset, light_dir, NULL, 0.0f, splits,
                                 light_matrices);
}

/* ---- depth + id rasterisation --------------------------------------------------------------------
   Per (instance, triangle), the fixed-function steps of the pipeline the reference's two passes run on:
     1. clip = view_proj * (transform * position)           (ShadowDepth.vert/.geom, CoreEditor.vert)
     2. clip the triangle against the near plane z >= 0 (with DepthClamp: against w >= 1e-5 instead, depth is then
        clamped per sample); new vertices are interpolated from the inside vertex towards the outside one
     3. perspective divide, viewport (y down), snap to 1/256 pixel; coverage by exact integer edge functions at pixel
        centres with a tie rule that gives a shared edge to exactly one triangle; depth = screen-linear z/w
     3b. a triangle that cannot be snapped (a vertex at w <= 0 after clipping, or farther than 2^20 pixels out) is
        rasterised in homogeneous float coordinates instead (edge functions = adjugate rows of the (X, Y, W) matrix)
   Target word = [depth float bits : 32 | instance : 32], smallest wins.  The float operations are written in one fixed
   order (this file is built with -ffp-contract=off) which raster.cu follows. */
#define SSO_SUBPIXEL 256.0f
#define SSO_SNAP_LIMIT 268435456.0f /* 2^28 sub-pixel units = 2^20 pixels */

static void sso_clip_of(const float vp[16], const float* r, const float* p, float out[4])
{
    const float wx = (r[0] * p[0] + r[1] * p[1]) + (r[2] * p[2] + r[3]);
    const float wy = (r[4] * p[0] + r[5] * p[1]) + (r[6] * p[2] + r[7]);
    const float wz = (r[8] * p[0] + r[9] * p[1]) + (r[10] * p[2] + r[11]);
    for (int j = 0; j < 4; ++j) out[j] = (vp[j] * wx + vp[4 + j] * wy) + (vp[8 + j] * wz + vp[12 + j]);
}

static void sso_put(uint64_t* target, uint32_t width, uint32_t i, uint32_t j, float depth, uint32_t flags, uint32_t inst)
{
    if (flags & 1u) depth = depth < 0.f ? 0.f : (depth > 1.f ? 1.f : depth);
    if (!(depth >= 0.f && depth <= 1.f)) return;
    uint32_t bits;
    memcpy(&bits, &depth, 4);
    const uint64_t word = ((uint64_t)bits << 32) | inst;
    uint64_t* dst = &target[(size_t)j * width + i];
    if (word < *dst) *dst = word;
}

/* 3b: homogeneous float rasterisation over every sample of the target */
static void sso_raster_2dh(const float c[3][4], uint32_t width, uint32_t height, uint32_t flags, uint32_t inst, uint64_t* target)
{
    const float half_w = 0.5f * (float)width, half_h = 0.5f * (float)height;
    float X[3], Y[3], W[3], a[3], b[3], cc[3];
    for (int v = 0; v < 3; ++v) {
        X[v] = (c[v][0] + c[v][3]) * half_w;
        Y[v] = (c[v][1] + c[v][3]) * half_h;
        W[v] = c[v][3];
    }
    a[0] = Y[1] * W[2] - Y[2] * W[1]; b[0] = X[2] * W[1] - X[1] * W[2]; cc[0] = X[1] * Y[2] - X[2] * Y[1];
    a[1] = Y[2] * W[0] - Y[0] * W[2]; b[1] = X[0] * W[2] - X[2] * W[0]; cc[1] = X[2] * Y[0] - X[0] * Y[2];
    a[2] = Y[0] * W[1] - Y[1] * W[0]; b[2] = X[1] * W[0] - X[0] * W[1]; cc[2] = X[0] * Y[1] - X[1] * Y[0];
    const float det = (X[0] * a[0] + X[1] * a[1]) + X[2] * a[2];
    if (!(det > 0.f || det < 0.f)) return;
    if (!(W[0] > 0.f || W[1] > 0.f || W[2] > 0.f)) return;
    const float s = det > 0.f ? 1.f : -1.f;
    for (uint32_t j = 0; j < height; ++j) {
        const float py = (float)j + 0.5f;
        for (uint32_t i = 0; i < width; ++i) {
            const float px = (float)i + 0.5f;
            float e[3];
            int inside = 1;
            for (int q = 0; q < 3; ++q) {
                e[q] = (a[q] * px + b[q] * py) + cc[q];
                const float es = s * e[q], as = s * a[q], bs = s * b[q];
                if (!(es > 0.f || (es == 0.f && (as > 0.f || (as == 0.f && bs > 0.f))))) inside = 0;
            }
            if (!inside) continue;
            if (!(s * ((e[0] + e[1]) + e[2]) > 0.f)) continue;
            sso_put(target, width, i, j, ((c[0][2] * e[0] + c[1][2] * e[1]) + c[2][2] * e[2]) / det, flags, inst);
        }
    }
}

static int64_t sso_floor_div256(int64_t v) { return v >= 0 ? v / 256 : -((-v + 255) / 256); }

/* 3: snapped, exact coverage */
static void sso_raster_tri(const float c[3][4], uint32_t width, uint32_t height, uint32_t flags, uint32_t inst, uint64_t* target)
{
    const float half_w = 0.5f * (float)width, half_h = 0.5f * (float)height;
    int64_t x[3], y[3];
    float zn[3];
    int snap = 1;
    for (int v = 0; v < 3; ++v) {
        if (!(c[v][3] > 0.f)) { snap = 0; break; }
        const float iw = 1.0f / c[v][3];
        const float sx = ((c[v][0] * iw) * half_w + half_w) * SSO_SUBPIXEL;
        const float sy = ((c[v][1] * iw) * half_h + half_h) * SSO_SUBPIXEL;
        if (!(sx > -SSO_SNAP_LIMIT && sx < SSO_SNAP_LIMIT && sy > -SSO_SNAP_LIMIT && sy < SSO_SNAP_LIMIT)) { snap = 0; break; }
        x[v] = llrintf(sx);
        y[v] = llrintf(sy);
        zn[v] = c[v][2] * iw;
    }
    if (!snap) {
        sso_raster_2dh(c, width, height, flags, inst, target);
        return;
    }
    /* edge q is opposite vertex q: from vertex q+1 to vertex q+2 */
    int64_t dx[3], dy[3];
    for (int q = 0; q < 3; ++q) {
        const int p0 = (q + 1) % 3, p1 = (q + 2) % 3;
        dx[q] = x[p1] - x[p0];
        dy[q] = y[p1] - y[p0];
    }
    const int64_t area = dx[2] * (y[2] - y[0]) - dy[2] * (x[2] - x[0]);  /* edge 2 (v0 -> v1) evaluated at v2 */
    if (area == 0) return;
    const int64_t s = area > 0 ? 1 : -1;
    int64_t minx = x[0], maxx = x[0], miny = y[0], maxy = y[0];
    for (int v = 1; v < 3; ++v) {
        if (x[v] < minx) minx = x[v];
        if (x[v] > maxx) maxx = x[v];
        if (y[v] < miny) miny = y[v];
        if (y[v] > maxy) maxy = y[v];
    }
    /* pixel i is sampled at 256 i + 128 */
    int64_t i0 = sso_floor_div256(minx - 128 + 255), i1 = sso_floor_div256(maxx - 128);
    int64_t j0 = sso_floor_div256(miny - 128 + 255), j1 = sso_floor_div256(maxy - 128);
    if (i0 < 0) i0 = 0;
    if (j0 < 0) j0 = 0;
    if (i1 > (int64_t)width - 1) i1 = (int64_t)width - 1;
    if (j1 > (int64_t)height - 1) j1 = (int64_t)height - 1;
    const float fa = (float)(s * area);
    for (int64_t j = j0; j <= j1; ++j) {
        const int64_t py = 256 * j + 128;
        for (int64_t i = i0; i <= i1; ++i) {
            const int64_t px = 256 * i + 128;
            int64_t e[3];
            int inside = 1;
            for (int q = 0; q < 3; ++q) {
                const int p0 = (q + 1) % 3;
                e[q] = s * (dx[q] * (py - y[p0]) - dy[q] * (px - x[p0]));
                /* a sample exactly on an edge belongs to the triangle for which that edge runs down, or level to the right */
                const int64_t sdx = s * dx[q], sdy = s * dy[q];
                if (!(e[q] > 0 || (e[q] == 0 && (sdy > 0 || (sdy == 0 && sdx > 0))))) inside = 0;
            }
            if (!inside) continue;
            const float l0 = (float)e[0] / fa, l1 = (float)e[1] / fa, l2 = (float)e[2] / fa;
            sso_put(target, width, (uint32_t)i, (uint32_t)j, (zn[0] * l0 + zn[1] * l1) + zn[2] * l2, flags, inst);
        }
    }
}

void sso_raster_view(const float* rows12, const uint32_t* instance_idx, uint32_t visible, const uint32_t* slot_of_instance,
                     const sso_submesh* slots, const float* positions_xyz, uint32_t n_vertices, const uint32_t* indices,
                     uint32_t n_indices, const float view_proj[16], uint32_t width, uint32_t height, uint32_t flags,
                     uint64_t* target)
{
    for (size_t i = 0; i < (size_t)width * height; ++i) target[i] = ~(uint64_t)0;
    for (uint32_t k = 0; k < visible; ++k) {
        const float* r = rows12 + (size_t)k * 12;
        const uint32_t inst = instance_idx[k];
        const sso_submesh* sm = &slots[slot_of_instance[inst]];
        for (uint32_t t = 0; t < sm->index_count / 3u; ++t) {
            float c[3][4];
            int in_range = 1;
            for (int v = 0; v < 3 && in_range; ++v) {
                const uint64_t ii = (uint64_t)sm->first_index + (uint64_t)t * 3u + (uint64_t)v;
                if (ii >= n_indices) { in_range = 0; break; }
                const int64_t vi = (int64_t)sm->vertex_offset + (int64_t)indices[ii];
                if (vi < 0 || vi >= (int64_t)n_vertices) { in_range = 0; break; }
                sso_clip_of(view_proj, r, positions_xyz + (size_t)vi * 3, c[v]);
            }
            if (!in_range) continue;
            /* 2: near plane.  d >= 0 is inside */
            float d[3], poly[4][4];
            int n_out = 0;
            for (int v = 0; v < 3; ++v) d[v] = (flags & 1u) ? c[v][3] - 1e-5f : c[v][2];
            for (int v = 0; v < 3; ++v) {
                const int nx = (v + 1) % 3;
                const int in_v = d[v] >= 0.f, in_n = d[nx] >= 0.f;
                if (in_v) memcpy(poly[n_out++], c[v], 16);
                if (in_v != in_n) {
                    const int a = in_v ? v : nx, b = in_v ? nx : v; /* from the inside vertex a towards the outside one b */
                    const float tt = d[a] / (d[a] - d[b]);
                    for (int q = 0; q < 4; ++q) poly[n_out][q] = c[a][q] + tt * (c[b][q] - c[a][q]);
                    ++n_out;
                }
            }
            if (n_out < 3) continue;
            {
                const float t0[3][4] = {{poly[0][0], poly[0][1], poly[0][2], poly[0][3]}, {poly[1][0], poly[1][1], poly[1][2], poly[1][3]},
                                        {poly[2][0], poly[2][1], poly[2][2], poly[2][3]}};
                sso_raster_tri(t0, width, height, flags, inst, target);
            }
            if (n_out == 4) {
                const float t1[3][4] = {{poly[0][0], poly[0][1], poly[0][2], poly[0][3]}, {poly[2][0], poly[2][1], poly[2][2], poly[2][3]},
                                        {poly[3][0], poly[3][1], poly[3][2], poly[3][3]}};
                sso_raster_tri(t1, width, height, flags, inst, target);
            }
        }
    }
}

/* JPH::Quat::GetEulerAngles (VulkanCore/vendor/JoltPhysics/Jolt/Math/Quat.inl:185-203), the conversion
   JoltPhysicsWorld::Update applies to every body (Platform/JoltPhysics/JoltPhysicsWorld.cpp:121-127).  Jolt's ATan2 / ASin
   are its own polynomial approximations; libm stands in for them here (agreement to ~1e-6, checked against the Jolt
   build in tests/test_oracle_vs_ref.py). */
void sso_quat_to_euler(const float* quat_xyzw, uint32_t n, float* euler_out)
{
    for (uint32_t i = 0; i < n; ++i) {
        const float x = quat_xyzw[i * 4 + 0], y = quat_xyzw[i * 4 + 1], z = quat_xyzw[i * 4 + 2], w = quat_xyzw[i * 4 + 3];
        float y_sq = y * y;
        float t0 = 2.0f * (w * x + y * z);
        float t1 = 1.0f - 2.0f * (x * x + y_sq);
        float t2 = 2.0f * (w * y - z * x);
        t2 = t2 > 1.0f ? 1.0f : t2;
        t2 = t2 < -1.0f ? -1.0f : t2;
        float t3 = 2.0f * (w * z + x * y);
        float t4 = 1.0f - 2.0f * (y_sq + z * z);
        euler_out[i * 3 + 0] = atan2f(t0, t1);
        euler_out[i * 3 + 1] = asinf(t2);
        euler_out[i * 3 + 2] = atan2f(t3, t4);
    }
}

/* Scene::UpdateLightsBuffer + OnUpdateLights (VC/Scene/Scene.cpp:167-264): uniform blocks UBPointLights (496 B) and
   UBSpotLights (656 B) of VC/Core/Components.h:229-241, and vec4(Translation, Radius) per light.  Offsets in floats:
   block = {count, pad x3, lights...}; point light 12 floats {pos4, color4, radius, falloff, pad x2}; spot light 16 floats
   {pos4, color4, dir3, inner, outer, radius, falloff, pad}.  glm::normalize(v) = v * (1 / sqrt(dot(v, v))). */
void sso_pack_lights(const float* trs, const uint32_t* point_node, const float* point_par, uint32_t n_point,
                     const uint32_t* spot_node, const float* spot_par, uint32_t n_spot, void* ub_point, void* ub_spot,
                     float* positions)
{
    float* p = (float*)ub_point;
    float* s = (float*)ub_spot;
    memset(p, 0, 496);
    memset(s, 0, 656);
    ((int*)p)[0] = (int)n_point;
    ((int*)s)[0] = (int)n_spot;
    for (uint32_t k = 0; k < n_point; ++k) {
        const float* t = trs + (size_t)point_node[k] * 9;
        const float* q = point_par + k * 6;
        float* o = p + 4 + k * 12;
        o[0] = t[0]; o[1] = t[1]; o[2] = t[2]; o[3] = 1.0f;
        for (int c = 0; c < 4; ++c) o[4 + c] = q[c];
        o[8] = q[4]; o[9] = q[5];
        positions[k * 4 + 0] = t[0]; positions[k * 4 + 1] = t[1]; positions[k * 4 + 2] = t[2]; positions[k * 4 + 3] = q[4];
    }
    for (uint32_t k = 0; k < n_spot; ++k) {
        const float* t = trs + (size_t)spot_node[k] * 9;
        const float* q = spot_par + k * 8;
        float* o = s + 4 + k * 16;
        const float inv = 1.0f / sqrtf(t[3] * t[3] + t[4] * t[4] + t[5] * t[5]);
        o[0] = t[0]; o[1] = t[1]; o[2] = t[2]; o[3] = 1.0f;
        for (int c = 0; c < 4; ++c) o[4 + c] = q[c];
        o[8] = t[3] * inv; o[9] = t[4] * inv; o[10] = t[5] * inv;
        o[11] = q[4]; o[12] = q[5]; o[13] = q[6]; o[14] = q[7];
        float* w = positions + (n_point + k) * 4;
        w[0] = t[0]; w[1] = t[1]; w[2] = t[2]; w[3] = q[6];
    }
}

