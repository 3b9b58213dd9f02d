// Reference-arithmetic build for the oracle — TEST INFRASTRUCTURE, NOT PRODUCT.
//
// This translation unit is compiled against the reference's own vendored headers WHERE THEY LIE
// (-I/root/reference/VulkanCore/vendor/glm -I/root/reference/VulkanCore/vendor/entt, see
// oracle/Makefile) into oracle/_ref/libss_ref.so.  Nothing from /root/reference is copied into this
// repository.  It exists to (1) pin oracle/scenesubmit_oracle.c bit-for-bit against real GLM 0.9.9.9
// and real entt 3.15.0 iteration order, and (2) serve as the timed CPU baseline "kind: reference":
// an entt registry walked in view order, glm T*R*S, glm mat4 product with the submesh transform, and
// std::map<MeshKey, ...> grouping, i.e. the shape of Scene::OnUpdateGeometry
// (VulkanCore/src/VulkanCore/Scene/Scene.cpp:135-147) + SceneRenderer::SubmitMesh
// (VulkanCore/src/VulkanCore/Scene/SceneRenderer.cpp:1124-1167).
//
// The reference's Scene.cpp / SceneRenderer.cpp themselves cannot be compiled here (Vulkan SDK,
// <Windows.h>, Optick, MSVC intrinsics), hence this shim around the header-only part.
#define GLM_FORCE_DEPTH_ZERO_TO_ONE // VulkanCore/premake5.lua:47
#define GLM_ENABLE_EXPERIMENTAL     // VulkanCore/src/VulkanCore/Core/Components.h:7
#include <glm/glm.hpp>
#include <glm/gtc/matrix_transform.hpp>
#include <glm/gtx/quaternion.hpp>

#include <entt.hpp>

#include <array>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <thread>
#include <vector>

#include "../include/scenesubmit_dettrig.h"

namespace {

// Same members, same GetTransform body shape as VulkanCore/src/VulkanCore/Core/Components.h:13-29.
struct TransformComponent {
    glm::vec3 Translation{0.0f}, Rotation{0.0f}, Scale{1.0f};

    glm::mat4 GetTransform() const
    {
        glm::mat4 rotation = glm::toMat4(glm::quat(Rotation));
        return glm::translate(glm::mat4(1.0f), Translation) * rotation * glm::scale(glm::mat4(1.0f), Scale);
    }

    // deterministic-trig variant: only the six sin/cos differ; the quaternion assembly is the body of
    // glm's qua(vec3) constructor (glm/detail/type_quat.inl:208-217), everything after is GLM.
    glm::mat4 GetTransformDet() const
    {
        glm::vec3 h = Rotation * 0.5f, c, s;
        ss_det_sincosf(h.x, &s.x, &c.x);
        ss_det_sincosf(h.y, &s.y, &c.y);
        ss_det_sincosf(h.z, &s.z, &c.z);
        glm::quat q;
        q.w = c.x * c.y * c.z + s.x * s.y * s.z;
        q.x = s.x * c.y * c.z - c.x * s.y * s.z;
        q.y = c.x * s.y * c.z + s.x * c.y * s.z;
        q.z = c.x * c.y * s.z - s.x * s.y * c.z;
        return glm::translate(glm::mat4(1.0f), Translation) * glm::toMat4(q) * glm::scale(glm::mat4(1.0f), Scale);
    }
};

struct MeshRef {
    int32_t mesh; // index into the mesh arrays, -1 = null asset
    int32_t table; // index into the material tables, -1 = null table
    uint32_t entity_index; // position in the caller's arrays
};

struct ParentRef {
    int32_t parent;
};

// VulkanCore/src/VulkanCore/Scene/SceneRenderer.h:51-86
struct MeshKey {
    uint64_t MeshHandle;
    uint64_t MaterialHandle;
    uint32_t SubmeshIndex;
    bool operator<(const MeshKey& o) const
    {
        if (MeshHandle != o.MeshHandle) return MeshHandle < o.MeshHandle;
        if (SubmeshIndex != o.SubmeshIndex) return SubmeshIndex < o.SubmeshIndex;
        return MaterialHandle < o.MaterialHandle;
    }
};

struct TransformData { // VulkanCore/src/VulkanCore/Core/Components.h:95-98
    glm::vec4 MRow[3];
};

struct Group {
    std::vector<TransformData> Transforms;
    std::vector<uint32_t> Instances; // bookkeeping for the checker: caller's instance numbering
};

glm::mat4 load_mat4(const float* p)
{
    glm::mat4 m;
    std::memcpy(&m[0][0], p, 64);
    return m;
}
void store_mat4(const glm::mat4& m, float* p) { std::memcpy(p, &m[0][0], 64); }

} // namespace

extern "C" {

struct ssr_scene {
    uint32_t n_entities;
    const float* trs;          // [n][9]
    const int32_t* parent;     // [n] or null (extension N1)
    const int32_t* mesh;       // [n]  -1 = entity has no valid mesh asset
    const int32_t* table;      // [n]  -1 = null material table
    const uint8_t* has_mesh_component; // [n] or null (= all): entity carries a MeshComponent
    // meshes
    uint32_t n_meshes;
    const uint64_t* mesh_handle;        // [n_meshes]
    const uint32_t* mesh_vertex_count;  // [n_meshes]
    const uint32_t* mesh_sub_begin;     // [n_meshes+1] range in the submesh arrays
    const uint64_t* base_material_handle; // [n_meshes]
    const uint8_t* base_material_tess;    // [n_meshes]
    // submeshes (MeshSource::GetSubmeshes() order; Mesh::GetSubmeshes() lists them all in order)
    const int32_t* sub_material_index;  // [n_sub]
    const float* sub_local;             // [n_sub][16]
    const float* sub_aabb;              // [n_sub][6]
    // material tables: table t owns entries [table_begin[t], table_begin[t+1])
    uint32_t n_tables;
    const uint32_t* table_begin;
    const uint32_t* entry_index;        // material index (map key)
    const uint64_t* entry_handle;       // 0 = nullptr entry (SceneSerializer.cpp:414-434 padding)
    const uint8_t* entry_tess;
    const float* inst_aabb;             // [n_instances][6] or null; instance numbering = submission order
    int trig_mode;
};

void ssr_trs_to_mat4(const float* trs9, int trig_mode, float* out16)
{
    TransformComponent tc;
    tc.Translation = {trs9[0], trs9[1], trs9[2]};
    tc.Rotation = {trs9[3], trs9[4], trs9[5]};
    tc.Scale = {trs9[6], trs9[7], trs9[8]};
    store_mat4(trig_mode == 0 ? tc.GetTransformDet() : tc.GetTransform(), out16);
}

void ssr_mat4_mul(const float* a, const float* b, float* out) { store_mat4(load_mat4(a) * load_mat4(b), out); }
void ssr_mat4_inverse(const float* m, float* out) { store_mat4(glm::inverse(load_mat4(m)), out); }

// EditorCamera (VulkanCore/src/VulkanCore/Renderer/EditorCamera.cpp:15-33, EditorCamera.h:45)
void ssr_camera_view_proj(float fov, float aspect, float zn, float zf, const float* pos, float pitch,
                          float yaw, float* view, float* proj, float* view_proj)
{
    glm::mat4 P = glm::perspective(fov, aspect, zn, zf);
    glm::quat orientation = glm::quat(glm::vec3(-pitch, -yaw, 0.0f));
    glm::mat4 invV = glm::translate(glm::mat4(1.0f), glm::vec3(pos[0], pos[1], pos[2])) * glm::toMat4(orientation);
    glm::mat4 V = glm::inverse(invV);
    store_mat4(V, view);
    store_mat4(P, proj);
    store_mat4(P * V, view_proj);
}

// SceneRenderer::UpdateCascadeMap (VulkanCore/src/VulkanCore/Scene/SceneRenderer.cpp:579-687), evaluated with the same
// glm calls in the same order.  cascade_origin == nullptr: CascadeOffset == 0; otherwise the texel-snap branch (:662-679).
void ssr_cascade_matrices_snapped(const float* view_proj, float zn, float zf, float lambda, float near_off,
                                  float far_off, const float* light_dir, const float* cascade_origin, float map_size,
                                  float* splits, float* light_mats)
{
    const glm::mat4 VP = load_mat4(view_proj);
    const float clipRange = zf - zn;
    const float minZ = zn, maxZ = zn + clipRange;
    const float range = maxZ - minZ, ratio = maxZ / minZ;
    std::array<float, 4> cs{};
    for (uint32_t i = 0; i < 4; ++i) {
        float p = (i + 1) / (float)4;
        float lg = minZ * std::pow(ratio, p);
        float uni = minZ + range * p;
        float d = lambda * (lg - uni) + uni;
        cs[i] = (d - zn) / clipRange;
    }
    float last = 0.0f;
    for (uint32_t i = 0; i < 4; ++i) {
        const float split = cs[i];
        glm::vec3 fc[8] = {{-1.0f, 1.0f, 0.0f}, {1.0f, 1.0f, 0.0f},  {1.0f, -1.0f, 0.0f}, {-1.0f, -1.0f, 0.0f},
                           {-1.0f, 1.0f, 1.0f}, {1.0f, 1.0f, 1.0f},  {1.0f, -1.0f, 1.0f}, {-1.0f, -1.0f, 1.0f}};
        glm::mat4 inv = glm::inverse(VP);
        for (auto& c : fc) {
            glm::vec4 ic = inv * glm::vec4(c, 1.0f);
            c = ic / ic.w;
        }
        for (uint32_t j = 0; j < 4; ++j) {
            glm::vec3 dist = fc[j + 4] - fc[j];
            fc[j + 4] = fc[j] + (dist * split);
            fc[j] = fc[j] + (dist * last);
        }
        glm::vec3 center(0.0f);
        for (auto& c : fc) center += c;
        center /= 8.0f;
        float radius = 0.0f;
        for (auto& c : fc) radius = glm::max(radius, glm::length(c - center));
        radius = std::ceil(radius * 16.0f) / 16.0f;
        glm::vec3 maxE(radius), minE = -maxE;
        glm::vec3 ld = glm::normalize(glm::vec3(light_dir[0], light_dir[1], light_dir[2]));
        glm::mat4 lv = glm::lookAt(center - ld * -minE.z, center, glm::vec3(0.0f, 1.0f, 0.0f));
        glm::mat4 lo = glm::ortho(minE.x, maxE.x, minE.y, maxE.y, near_off, maxE.z - minE.z + far_off);
        if (cascade_origin) {
            float shadowMapSize = map_size;
            glm::mat4 shadowMatrix = lo * lv;
            glm::vec4 shadowOrigin = glm::vec4(glm::vec3(cascade_origin[0], cascade_origin[1], cascade_origin[2]), 1.0f);
            shadowOrigin = shadowMatrix * shadowOrigin;
            shadowOrigin = shadowOrigin * shadowMapSize / 2.0f;
            glm::vec4 roundedOrigin = glm::round(shadowOrigin);
            glm::vec4 roundOffset = roundedOrigin - shadowOrigin;
            roundOffset = roundOffset * 2.0f / shadowMapSize;
            roundOffset.z = 0.0f;
            roundOffset.w = 0.0f;
            glm::mat4 shadowProj = lo;
            shadowProj[3] += roundOffset;
            lo = shadowProj;
        }
        splits[i] = (zn + split * clipRange) * -1.0f;
        store_mat4(lo * lv, light_mats + i * 16);
        last = cs[i];
    }
}

void ssr_cascade_matrices(const float* view_proj, float zn, float zf, float lambda, float near_off,
                          float far_off, const float* light_dir, float* splits, float* light_mats)
{
    ssr_cascade_matrices_snapped(view_proj, zn, zf, lambda, near_off, far_off, light_dir, nullptr, 0.0f, splits, light_mats);
}

// One pass of the reference's submission over an entt registry, then grouped output.
//   pass_view < 0 : no culling (what the reference does today)
//   pass_view >= 0: keep only instances whose visibility mask (given by the caller) has that bit.
// Outputs (caller allocated, capacity n_instances_cap): groups in draw order (static map first, then the
// tessellated map), per group key + count, instance rows concatenated, and the submission-order
// instance number of every row.  Returns the number of groups; *n_rows_out = total rows.
// world_rows_out (may be null): rows in submission order, [n_instances][12].
// n_instances_out: how many (entity, submesh) pairs were submitted.
uint32_t ssr_submit_pass(const ssr_scene* sc, const uint8_t* masks, int pass_view, uint64_t* g_mesh,
                         uint64_t* g_material, uint32_t* g_submesh, uint8_t* g_tess, uint32_t* g_count,
                         float* rows_out, uint32_t* row_instance_out, uint32_t* n_rows_out,
                         float* world_rows_out, uint32_t* n_instances_out, uint32_t* inst_entity_out,
                         uint32_t* inst_submesh_out, double* seconds_out)
{
    entt::registry reg;
    std::vector<entt::entity> ents(sc->n_entities);
    for (uint32_t i = 0; i < sc->n_entities; ++i) {
        entt::entity e = reg.create();
        ents[i] = e;
        auto& tc = reg.emplace<TransformComponent>(e);
        const float* t = sc->trs + (size_t)i * 9;
        tc.Translation = {t[0], t[1], t[2]};
        tc.Rotation = {t[3], t[4], t[5]};
        tc.Scale = {t[6], t[7], t[8]};
        reg.emplace<ParentRef>(e, sc->parent ? sc->parent[i] : -1);
        if (!sc->has_mesh_component || sc->has_mesh_component[i]) reg.emplace<MeshRef>(e, MeshRef{sc->mesh[i], sc->table[i], i});
    }

    auto t0 = std::chrono::steady_clock::now();

    // extension N1: parent chain, resolved top-down in index order (parent[i] < i)
    std::vector<glm::mat4> world(sc->n_entities);
    for (uint32_t i = 0; i < sc->n_entities; ++i) {
        const auto& tc = reg.get<TransformComponent>(ents[i]);
        glm::mat4 local = sc->trig_mode == 0 ? tc.GetTransformDet() : tc.GetTransform();
        int32_t p = reg.get<ParentRef>(ents[i]).parent;
        world[i] = p < 0 ? local : world[p] * local;
    }

    std::map<MeshKey, Group> staticMap, tessMap;
    uint32_t inst = 0;
    auto view = reg.view<TransformComponent, MeshRef>(); // Scene.cpp:138
    for (auto ent : view) {
        auto [tc, mr] = view.get<TransformComponent, MeshRef>(ent);
        (void)tc;
        if (mr.mesh < 0 || mr.table < 0) continue;                  // SceneRenderer.cpp:1128-1129
        if (sc->mesh_vertex_count[mr.mesh] < 1) continue;           // :1135-1136
        const glm::mat4& transform = world[mr.entity_index];
        for (uint32_t s = sc->mesh_sub_begin[mr.mesh]; s < sc->mesh_sub_begin[mr.mesh + 1]; ++s) {
            uint32_t submeshIndex = s - sc->mesh_sub_begin[mr.mesh];
            int materialIndex = sc->sub_material_index[s];
            // MaterialTable::HasMaterial / GetMaterial (Asset/MaterialAsset.cpp:43-51), base fallback (:1141)
            uint64_t materialHandle = sc->base_material_handle[mr.mesh];
            bool tess = sc->base_material_tess[mr.mesh] != 0;
            for (uint32_t k = sc->table_begin[mr.table]; k < sc->table_begin[mr.table + 1]; ++k) {
                if ((int)sc->entry_index[k] == materialIndex && sc->entry_handle[k] != 0) {
                    materialHandle = sc->entry_handle[k];
                    tess = sc->entry_tess[k] != 0;
                }
            }
            glm::mat4 submeshTransform = transform * load_mat4(sc->sub_local + (size_t)s * 16); // :1150
            TransformData td;
            td.MRow[0] = {submeshTransform[0][0], submeshTransform[1][0], submeshTransform[2][0], submeshTransform[3][0]};
            td.MRow[1] = {submeshTransform[0][1], submeshTransform[1][1], submeshTransform[2][1], submeshTransform[3][1]};
            td.MRow[2] = {submeshTransform[0][2], submeshTransform[1][2], submeshTransform[2][2], submeshTransform[3][2]};
            if (world_rows_out) std::memcpy(world_rows_out + (size_t)inst * 12, &td, 48);
            if (inst_entity_out) inst_entity_out[inst] = mr.entity_index;
            if (inst_submesh_out) inst_submesh_out[inst] = s;
            bool keep = pass_view < 0 || (masks[inst] >> pass_view) & 1;
            if (keep) {
                MeshKey key{sc->mesh_handle[mr.mesh], materialHandle, submeshIndex};
                Group& g = tess ? tessMap[key] : staticMap[key];
                g.Transforms.emplace_back(td);
                g.Instances.push_back(inst);
            }
            ++inst;
        }
    }

    auto t1 = std::chrono::steady_clock::now();
    if (seconds_out) *seconds_out = std::chrono::duration<double>(t1 - t0).count();

    uint32_t ng = 0, nr = 0;
    for (int pass = 0; pass < 2; ++pass) {
        for (auto& [mk, g] : pass == 0 ? staticMap : tessMap) {
            if (g.Transforms.empty()) continue;
            g_mesh[ng] = mk.MeshHandle;
            g_material[ng] = mk.MaterialHandle;
            g_submesh[ng] = mk.SubmeshIndex;
            g_tess[ng] = (uint8_t)pass;
            g_count[ng] = (uint32_t)g.Transforms.size();
            for (size_t k = 0; k < g.Transforms.size(); ++k) {
                if (rows_out) std::memcpy(rows_out + (size_t)nr * 12, &g.Transforms[k], 48);
                row_instance_out[nr] = g.Instances[k];
                ++nr;
            }
            ++ng;
        }
    }
    *n_rows_out = nr;
    *n_instances_out = inst;
    return ng;
}

// Timed CPU baseline at scale: flat arrays instead of the asset indirections (each entity = one
// instance), `threads` range partitions each owning one std::map per view, merged at the end.
// The per-entity work is the reference's: glm TRS, glm product with the submesh transform, row pack,
// std::map<MeshKey,...> insert; plus the extension steps (AABB, planes, mask) so CPU and GPU do the
// same job.  Returns seconds.
struct ssr_flat {
    uint32_t n;
    const float* trs;           // [n_nodes][9]
    const int32_t* parent;      // [n_nodes] or null
    uint32_t n_nodes;
    uint32_t node_base;         // instance i uses node node_base + i
    const uint32_t* slot;       // [n]
    const uint32_t* matword;    // [n]
    const float* inst_aabb;     // [n][6]
    const float* slot_local;    // [n_slots][16]
    const uint64_t* slot_mesh_handle; // [n_slots]
    const uint32_t* slot_submesh;     // [n_slots]
    const uint64_t* material_handle;  // [n_materials] rank -> handle
    int trig_mode;
};

static uint8_t cull_glm(const glm::mat4& W, const float* box, const glm::vec4 (*planes)[6], const uint32_t* vflags, uint32_t nv)
{
    glm::vec3 mn(box[0], box[1], box[2]), mx(box[3], box[4], box[5]);
    glm::vec3 cl = (mn + mx) * 0.5f, el = (mx - mn) * 0.5f;
    glm::vec4 cw4 = W * glm::vec4(cl, 1.0f);
    glm::vec3 ew;
    for (int r = 0; r < 3; ++r) ew[r] = (std::fabs(W[0][r]) * el.x + std::fabs(W[1][r]) * el.y) + std::fabs(W[2][r]) * el.z;
    uint8_t mask = 0;
    for (uint32_t v = 0; v < nv; ++v) {
        bool in = true;
        for (int p = 0; p < 6; ++p) {
            if (p == 4 && (vflags[v] & 1u)) continue;
            const glm::vec4& pl = planes[v][p];
            float s = ((pl.x * cw4.x + pl.y * cw4.y) + pl.z * cw4.z) + pl.w;
            float r = (std::fabs(pl.x) * ew.x + std::fabs(pl.y) * ew.y) + std::fabs(pl.z) * ew.z;
            if (!(s + r >= 0.0f)) in = false;
        }
        if (in) mask |= (uint8_t)(1u << v);
    }
    return mask;
}

double ssr_flat_frame(const ssr_flat* f, const float* view_projs /* [nv][16] */, const uint32_t* vflags,
                      uint32_t nv, int threads, uint32_t* visible_out /* [nv] */, uint32_t* draws_out /* [nv] */,
                      uint8_t* masks_out /* [n] or null */, float* rows_out /* [n][12] or null */)
{
    if (threads < 1) threads = 1;
    glm::vec4 planes[8][6];
    for (uint32_t v = 0; v < nv; ++v) {
        glm::mat4 m = load_mat4(view_projs + v * 16);
        glm::vec4 r0(m[0][0], m[1][0], m[2][0], m[3][0]), r1(m[0][1], m[1][1], m[2][1], m[3][1]);
        glm::vec4 r2(m[0][2], m[1][2], m[2][2], m[3][2]), r3(m[0][3], m[1][3], m[2][3], m[3][3]);
        planes[v][0] = r3 + r0; planes[v][1] = r3 - r0; planes[v][2] = r3 + r1;
        planes[v][3] = r3 - r1; planes[v][4] = r2;      planes[v][5] = r3 - r2;
    }
    auto t0 = std::chrono::steady_clock::now();
    // interior nodes (levels above the instances) on the calling thread
    std::vector<glm::mat4> upper(f->node_base);
    for (uint32_t i = 0; i < f->node_base; ++i) {
        TransformComponent tc;
        const float* t = f->trs + (size_t)i * 9;
        tc.Translation = {t[0], t[1], t[2]}; tc.Rotation = {t[3], t[4], t[5]}; tc.Scale = {t[6], t[7], t[8]};
        glm::mat4 local = f->trig_mode == 0 ? tc.GetTransformDet() : tc.GetTransform();
        int32_t p = f->parent ? f->parent[i] : -1;
        upper[i] = p < 0 ? local : upper[p] * local;
    }
    using Map = std::map<MeshKey, std::vector<TransformData>>;
    std::vector<std::array<Map, 8>> maps(threads), tmaps(threads);
    auto work = [&](int t) {
        uint32_t lo = (uint32_t)((uint64_t)f->n * t / threads), hi = (uint32_t)((uint64_t)f->n * (t + 1) / threads);
        for (uint32_t i = lo; i < hi; ++i) {
            uint32_t slot = f->slot[i];
            if (slot == 0xFFFFFFFFu) { if (masks_out) masks_out[i] = 0; continue; }
            uint32_t node = f->node_base + i;
            TransformComponent tc;
            const float* tr = f->trs + (size_t)node * 9;
            tc.Translation = {tr[0], tr[1], tr[2]}; tc.Rotation = {tr[3], tr[4], tr[5]}; tc.Scale = {tr[6], tr[7], tr[8]};
            glm::mat4 local = f->trig_mode == 0 ? tc.GetTransformDet() : tc.GetTransform();
            int32_t p = f->parent ? f->parent[node] : -1;
            glm::mat4 transform = p < 0 ? local : upper[p] * local;
            glm::mat4 W = transform * load_mat4(f->slot_local + (size_t)slot * 16);
            TransformData td;
            td.MRow[0] = {W[0][0], W[1][0], W[2][0], W[3][0]};
            td.MRow[1] = {W[0][1], W[1][1], W[2][1], W[3][1]};
            td.MRow[2] = {W[0][2], W[1][2], W[2][2], W[3][2]};
            if (rows_out) std::memcpy(rows_out + (size_t)i * 12, &td, 48);
            uint8_t mask = cull_glm(W, f->inst_aabb + (size_t)i * 6, planes, vflags, nv);
            if (masks_out) masks_out[i] = mask;
            if (!mask) continue;
            uint32_t mw = f->matword[i];
            MeshKey key{f->slot_mesh_handle[slot], f->material_handle[mw & 0x7FFFFFFFu], f->slot_submesh[slot]};
            bool tess = (mw >> 31) != 0;
            for (uint32_t v = 0; v < nv; ++v)
                if ((mask >> v) & 1) (tess ? tmaps[t][v] : maps[t][v])[key].emplace_back(td);
        }
    };
    if (threads == 1) {
        work(0);
    } else {
        std::vector<std::thread> th;
        for (int t = 0; t < threads; ++t) th.emplace_back(work, t);
        for (auto& x : th) x.join();
    }
    // merge: thread ranges are in input order, so appending thread by thread keeps entt/input order
    for (uint32_t v = 0; v < nv; ++v) {
        uint32_t vis = 0, draws = 0;
        for (int pass = 0; pass < 2; ++pass) {
            Map merged;
            for (int t = 0; t < threads; ++t) {
                Map& m = pass == 0 ? maps[t][v] : tmaps[t][v];
                for (auto& [k, vec] : m) {
                    auto& dst = merged[k];
                    if (threads == 1) dst.swap(vec); else dst.insert(dst.end(), vec.begin(), vec.end());
                }
            }
            for (auto& [k, vec] : merged) { vis += (uint32_t)vec.size(); draws += 1; }
        }
        visible_out[v] = vis;
        draws_out[v] = draws;
    }
    auto t1 = std::chrono::steady_clock::now();
    return std::chrono::duration<double>(t1 - t0).count();
}

} // extern "C"

// JPH::Quat::GetEulerAngles through the reference's own Jolt headers (vendor/JoltPhysics/Jolt/Math/Quat.inl:185-203):
// what JoltPhysicsWorld::Update (Platform/JoltPhysics/JoltPhysicsWorld.cpp:121-122) writes into TransformComponent::Rotation.
#ifdef SSR_WITH_JOLT
#include <Jolt/Jolt.h>
#include <Jolt/Math/Quat.h>
extern "C" void ssr_jolt_euler(const float* quat_xyzw, uint32_t n, float* euler_out)
{
    for (uint32_t i = 0; i < n; ++i) {
        const float* q4 = quat_xyzw + (size_t)i * 4;
        JPH::Quat q(q4[0], q4[1], q4[2], q4[3]);
        JPH::Vec3 e = q.GetEulerAngles();
        euler_out[i * 3 + 0] = e.GetX();
        euler_out[i * 3 + 1] = e.GetY();
        euler_out[i * 3 + 2] = e.GetZ();
    }
}
#endif

// Scene::UpdateLightsBuffer / OnUpdateLights (VulkanCore/src/VulkanCore/Scene/Scene.cpp:167-264) with the reference's own
// light structs restated field for field (VulkanCore/src/VulkanCore/Core/Components.h:106-139, 229-241; the header itself
// drags in the asset system) and its glm calls.  trs: [n_nodes][9] TransformComponent memory; lights in view order.
namespace ssr_lights {
struct alignas(16) PointLightComponent { glm::vec4 Position{0.0f}; glm::vec4 Color{0.0f}; float Radius = 0.1f; float Falloff = 1.0f; };
struct alignas(8) SpotLightComponent {
    glm::vec4 Position{0.0f}; glm::vec4 Color{1.0f}; glm::vec3 Direction{1.0f};
    float InnerCutoff = 0.174f; float OuterCutoff = 0.261f; float Radius = 0.1f; float Falloff = 1.0f;
};
struct UBPointLights { int LightCount; glm::vec3 Padding{}; PointLightComponent PointLights[10]; };
struct UBSpotLights { int LightCount; glm::vec3 Padding{}; SpotLightComponent SpotLights[10]; };
static_assert(sizeof(UBPointLights) == 496 && sizeof(UBSpotLights) == 656, "uniform block sizes");
}  // namespace ssr_lights
extern "C" void ssr_pack_lights(const float* trs, const uint32_t* point_node, const float* point_par /* [n][6] color4 radius falloff */,
                                uint32_t n_point, const uint32_t* spot_node, const float* spot_par /* [n][8] color4 inner outer radius falloff */,
                                uint32_t n_spot, void* ub_point, void* ub_spot, float* positions)
{
    using namespace ssr_lights;
    UBPointLights pl;
    UBSpotLights sl;
    memset(&pl, 0, sizeof pl);
    memset(&sl, 0, sizeof sl);
    int li = 0;
    for (uint32_t k = 0; k < n_point; ++k) {
        const float* t = trs + (size_t)point_node[k] * 9;
        const float* q = point_par + k * 6;
        glm::vec3 Translation(t[0], t[1], t[2]);
        PointLightComponent c;
        c.Position = glm::vec4(Translation, 1.0f);
        c.Color = glm::vec4(q[0], q[1], q[2], q[3]);
        c.Radius = q[4];
        c.Falloff = q[5];
        memcpy(&pl.PointLights[li++], &c, 40);  // the bytes the aggregate initialisation sets; padding stays zero
        glm::vec4 position = glm::vec4(Translation, q[4]);
        memcpy(positions + k * 4, &position, 16);
    }
    pl.LightCount = li;
    li = 0;
    for (uint32_t k = 0; k < n_spot; ++k) {
        const float* t = trs + (size_t)spot_node[k] * 9;
        const float* q = spot_par + k * 8;
        glm::vec3 Translation(t[0], t[1], t[2]), Rotation(t[3], t[4], t[5]);
        glm::vec3 direction = glm::normalize(Rotation);
        SpotLightComponent c;
        c.Position = glm::vec4(Translation, 1.0f);
        c.Color = glm::vec4(q[0], q[1], q[2], q[3]);
        c.Direction = direction;
        c.InnerCutoff = q[4];
        c.OuterCutoff = q[5];
        c.Radius = q[6];
        c.Falloff = q[7];
        memcpy(&sl.SpotLights[li++], &c, 60);
        glm::vec4 position = glm::vec4(Translation, q[6]);
        memcpy(positions + (n_point + k) * 4, &position, 16);
    }
    sl.LightCount = li;
    memcpy(ub_point, &pl, sizeof pl);
    memcpy(ub_spot, &sl, sizeof sl);
}

