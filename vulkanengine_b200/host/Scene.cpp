// Headless mirror of Scene / SceneRenderer over libscenesubmit (see Scene.h).
#include "Scene.h"

#include <algorithm>
#include <stdexcept>

namespace VulkanCore {

// ---- Scene ---------------------------------------------------------------------------------------

// Scene.cpp:83-89: a fresh UUID and a default TransformComponent.
Entity Scene::CreateEntity(const std::string& name)
{
    Entity e = CreateEntityWithUUID(m_NextUUID++, name);
    e.AddComponent<TransformComponent>();
    return e;
}

// Scene.cpp:91-124: ID + Tag only (the tag bookkeeping is editor-side and not mirrored).  No TransformComponent:
// the deserializer adds one only when the file has it (SceneSerializer.cpp:352-360), so an entity saved without
// one stays outside view<TransformComponent, MeshComponent>.
Entity Scene::CreateEntityWithUUID(UUID uuid, const std::string&)
{
    Entity e{m_Registry.create(), this};
    m_Registry.emplace<IDComponent>((entity)(uint32_t)e, IDComponent{uuid});
    ++m_StructureVersion;
    return e;
}

// Scene.cpp:135-147.  The reference walks view<TransformComponent, MeshComponent>, resolves the mesh and
// calls SubmitMesh with GetTransform() per entity.  Here the per-entity work happens on the GPU: the
// TransformComponent pool is uploaded as it lies and the (entity, submesh) walk was done once, in the same
// view order, when the static streams were built.
void Scene::OnUpdateGeometry(SceneRenderer* renderer) { renderer->UploadTransforms(*this); }

// Scene.cpp:149-160: same walk plus the entity id per instance; the id stream is instance_idx -> entityOfInstance.
void Scene::OnSelectGeometry(SceneRenderer* renderer) { renderer->UploadTransforms(*this); }

// ---- flattening ----------------------------------------------------------------------------------

namespace {

uint32_t bits_for(size_t n)
{
    uint32_t b = 1;
    while ((1ull << b) < n) ++b;
    return b;
}

struct Resolved {
    AssetHandle material;
    bool tess;
};

// SceneRenderer.cpp:1140-1144: per-index material if the table has a non-null one, else the mesh's base material
Resolved resolve_material(const Mesh& mesh, const MaterialTable* table, const MaterialAsset* forced, uint32_t submeshIndex)
{
    if (forced) return {forced->Handle, forced->HasDisplacementTexture()};
    const int materialIndex = mesh.GetMeshSource()->GetSubmeshes()[submeshIndex].MaterialIndex;
    std::shared_ptr<MaterialAsset> m = (materialIndex >= 0 && table->HasMaterial((uint32_t)materialIndex))
                                           ? table->GetMaterial((uint32_t)materialIndex)
                                           : mesh.GetMeshSource()->GetBaseMaterial();
    // neither a table entry nor a base material: the reference would dereference null here (:1146); the mirror
    // leaves that submesh out
    if (!m) return {0, false};
    return {m->Handle, m->HasDisplacementTexture()};
}

bool submits(const std::shared_ptr<Mesh>& mesh, bool hasMaterial)
{
    // SceneRenderer.cpp:1128-1136: null mesh / null table / no vertices -> silent skip
    return mesh && hasMaterial && mesh->GetMeshSource()->GetVertexCount() >= 1;
}

}  // namespace

FlatScene SceneRenderer::Flatten(Scene& scene, const std::vector<AdHocInstance>& adhoc)
{
    FlatScene f;
    Registry& reg = scene.m_Registry;
    Pool<TransformComponent>& tpool = reg.storage<TransformComponent>();
    const uint32_t nT = (uint32_t)tpool.size();

    // -- nodes: level order when a parent chain exists, pool order otherwise
    bool anyParent = false;
    for (entity e : tpool.packed)
        if (reg.all_of<RelationshipComponent>(e) && reg.get<RelationshipComponent>(e).Parent != null_entity) anyParent = true;
    f.nodeOfPoolIndex.resize(nT);
    if (anyParent) {
        std::vector<uint32_t> depth(nT, 0);
        auto depth_of = [&](uint32_t k) {
            uint32_t d = 0;
            entity e = tpool.packed[k];
            while (reg.all_of<RelationshipComponent>(e) && reg.get<RelationshipComponent>(e).Parent != null_entity) {
                e = reg.get<RelationshipComponent>(e).Parent;
                if (++d > SS_MAX_LEVELS) throw std::runtime_error("RelationshipComponent cycle or chain deeper than SS_MAX_LEVELS");
            }
            return d;
        };
        uint32_t maxDepth = 0;
        for (uint32_t k = 0; k < nT; ++k) maxDepth = std::max(maxDepth, depth[k] = depth_of(k));
        std::vector<uint32_t> order(nT);
        for (uint32_t k = 0; k < nT; ++k) order[k] = k;
        std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return depth[a] < depth[b]; });
        f.levelOffsets.assign(maxDepth + 2, 0);
        for (uint32_t k = 0; k < nT; ++k) f.levelOffsets[depth[k] + 1]++;
        for (uint32_t l = 0; l <= maxDepth; ++l) f.levelOffsets[l + 1] += f.levelOffsets[l];
        for (uint32_t n = 0; n < nT; ++n) f.nodeOfPoolIndex[order[n]] = n;
        f.parent.assign(nT, -1);
        for (uint32_t k = 0; k < nT; ++k) {
            const entity e = tpool.packed[k];
            if (reg.all_of<RelationshipComponent>(e)) {
                const entity p = reg.get<RelationshipComponent>(e).Parent;
                if (p != null_entity) f.parent[f.nodeOfPoolIndex[k]] = (int32_t)f.nodeOfPoolIndex[tpool.sparse[p]];
            }
        }
        f.identityNodeOrder = true;  // a pool that already lies in level order can be uploaded as it is
        for (uint32_t k = 0; k < nT; ++k)
            if (f.nodeOfPoolIndex[k] != k) f.identityNodeOrder = false;
    } else {
        for (uint32_t k = 0; k < nT; ++k) f.nodeOfPoolIndex[k] = k;
        f.levelOffsets = {0, nT};
    }
    f.externalNodes = (uint32_t)adhoc.size();
    f.nodeCount = nT + f.externalNodes;
    if (f.externalNodes && f.parent.empty()) f.parent.assign(nT, -1);
    f.parent.resize(f.parent.empty() ? 0 : f.nodeCount, -1);

    // -- handle tables (ranks = position among the sorted unique handles, MeshKey::operator< order)
    struct Pending {
        std::shared_ptr<Mesh> mesh;
        const MaterialTable* table;
        const MaterialAsset* forced;
        uint32_t node, entityID;
    };
    std::vector<Pending> pend;
    reg.view<TransformComponent, MeshComponent>([&](entity e, TransformComponent&, MeshComponent& mc) {
        std::shared_ptr<Mesh> mesh = AssetManager::GetMesh(mc.MeshHandle);
        if (!submits(mesh, mc.MaterialTableHandle != nullptr)) return;
        pend.push_back({mesh, mc.MaterialTableHandle.get(), nullptr, f.nodeOfPoolIndex[tpool.sparse[e]], e});
    });
    for (uint32_t k = 0; k < adhoc.size(); ++k) {
        const AdHocInstance& a = adhoc[k];
        if (!submits(a.mesh, a.table != nullptr || a.material != nullptr)) continue;
        pend.push_back({a.mesh, a.table.get(), a.material.get(), nT + k, a.entityID});
    }
    std::map<AssetHandle, std::shared_ptr<Mesh>> meshes;
    std::vector<AssetHandle> mats;
    for (const Pending& p : pend) {
        meshes[p.mesh->Handle] = p.mesh;
        for (uint32_t s : p.mesh->GetSubmeshes()) {
            const AssetHandle m = resolve_material(*p.mesh, p.table, p.forced, s).material;
            if (m) mats.push_back(m);
        }
    }
    std::sort(mats.begin(), mats.end());
    mats.erase(std::unique(mats.begin(), mats.end()), mats.end());
    f.materialHandles = mats;

    // -- submesh table: meshes in handle order, their source submeshes in source order
    std::map<AssetHandle, uint32_t> slotBase;
    uint32_t maxSub = 1;
    bool allGeometry = !meshes.empty();
    std::map<const MeshSource*, std::pair<uint32_t, uint32_t>> arenaBase;  // source -> (first vertex, first index)
    for (auto& [handle, mesh] : meshes) {
        slotBase[handle] = (uint32_t)f.submeshTable.size();
        const MeshSource& source = *mesh->GetMeshSource();
        const auto& subs = source.GetSubmeshes();
        if (source.GetIndices().empty()) allGeometry = false;
        if (!arenaBase.count(&source)) {  // meshes naming one source share its vertices
            arenaBase[&source] = {(uint32_t)(f.positions.size() / 3), (uint32_t)f.indices.size()};
            for (const vec3& p : source.GetPositions()) {
                f.positions.push_back(p.x);
                f.positions.push_back(p.y);
                f.positions.push_back(p.z);
            }
            f.indices.insert(f.indices.end(), source.GetIndices().begin(), source.GetIndices().end());
        }
        maxSub = std::max<uint32_t>(maxSub, (uint32_t)subs.size());
        for (uint32_t s = 0; s < subs.size(); ++s) {
            ss_submesh row{};
            std::memcpy(row.local, subs[s].LocalTransform.data(), 64);
            row.aabb_min[0] = subs[s].BoundsMin.x; row.aabb_min[1] = subs[s].BoundsMin.y; row.aabb_min[2] = subs[s].BoundsMin.z;
            row.aabb_max[0] = subs[s].BoundsMax.x; row.aabb_max[1] = subs[s].BoundsMax.y; row.aabb_max[2] = subs[s].BoundsMax.z;
            row.index_count = subs[s].IndexCount;
            row.first_index = subs[s].BaseIndex;
            row.vertex_offset = (int32_t)subs[s].BaseVertex;
            row.mesh_rank = (uint32_t)f.meshHandles.size();
            row.submesh_index = s;
            f.submeshTable.push_back(row);
            f.vertexBaseOfSlot.push_back(arenaBase[&source].first);
            f.indexBaseOfSlot.push_back(arenaBase[&source].second);
        }
        f.meshHandles.push_back(handle);
    }
    if (!allGeometry) {
        f.positions.clear();
        f.indices.clear();
    }
    f.meshBits = bits_for(f.meshHandles.size());
    f.submeshBits = bits_for(maxSub);
    f.materialBits = bits_for(f.materialHandles.size());

    // -- instances, in submission order
    for (const Pending& p : pend) {
        for (uint32_t s : p.mesh->GetSubmeshes()) {
            const Resolved r = resolve_material(*p.mesh, p.table, p.forced, s);
            if (!r.material) continue;
            const uint32_t rank = (uint32_t)(std::lower_bound(mats.begin(), mats.end(), r.material) - mats.begin());
            f.node.push_back(p.node);
            f.slot.push_back(slotBase[p.mesh->Handle] + s);
            f.material.push_back(rank | (r.tess ? SS_MATERIAL_TESS_BIT : 0u));
            f.entityOfInstance.push_back(p.entityID);
        }
    }
    return f;
}

// ---- SceneRenderer -------------------------------------------------------------------------------

SceneRenderer::SceneRenderer(int device, uint32_t trigMode) : m_Device(device), m_TrigMode(trigMode) {}

SceneRenderer::~SceneRenderer()
{
    if (m_Submit) ss_destroy(m_Submit);
}

const char* SceneRenderer::LastError() const { return m_Error.empty() ? ss_last_error(m_Submit) : m_Error.c_str(); }

#define SS_TRY(call)                                                                   \
    do {                                                                               \
        ss_status st__ = (call);                                                       \
        if (st__ != SS_OK) {                                                           \
            m_Error = std::string(#call " -> ") + ss_last_error(m_Submit);             \
            throw std::runtime_error(m_Error);                                         \
        }                                                                              \
    } while (0)

void SceneRenderer::SetActiveScene(std::shared_ptr<Scene> scene)
{
    // SceneRenderer.cpp:1043-1050: a new scene clears every draw list
    m_Scene = std::move(scene);
    m_BuiltVersion = ~0ull;
    m_AdHoc.clear();
    for (auto& l : m_DrawLists) l.clear();
}

void SceneRenderer::EnsureContext()
{
    const uint32_t need[3] = {std::max<uint32_t>(m_Flat.nodeCount, 1), std::max<uint32_t>((uint32_t)m_Flat.node.size(), 1),
                              std::max<uint32_t>((uint32_t)m_Flat.submeshTable.size(), 1)};
    if (m_Submit && need[0] <= m_Capacity[0] && need[1] <= m_Capacity[1] && need[2] <= m_Capacity[2]) return;
    if (m_Submit) ss_destroy(m_Submit);
    m_Submit = nullptr;
    for (int k = 0; k < 3; ++k) m_Capacity[k] = std::max(need[k] * 2, 64u);  // head-room: editing adds entities
    ss_config cfg{};
    cfg.struct_size = sizeof(ss_config);
    cfg.device = m_Device;
    cfg.max_nodes = m_Capacity[0];
    cfg.max_instances = m_Capacity[1];
    cfg.max_submesh_slots = m_Capacity[2];
    cfg.max_views = 5;  // camera + SHADOW_MAP_CASCADE_COUNT (SceneRenderer.h:97)
    cfg.trig_mode = m_TrigMode;
    if (ss_create(&cfg, &m_Submit) != SS_OK) {
        m_Error = std::string("ss_create -> ") + ss_last_error(nullptr);
        throw std::runtime_error(m_Error);
    }
}

void SceneRenderer::RebuildStaticStreams()
{
    m_Flat = Flatten(*m_Scene, m_AdHoc);
    m_AdHocBuilt = m_AdHoc;
    m_BuiltVersion = m_Scene->m_StructureVersion;
    m_SlotOfMeshRank.assign(m_Flat.meshHandles.size(), 0);
    for (size_t k = m_Flat.submeshTable.size(); k-- > 0;) m_SlotOfMeshRank[m_Flat.submeshTable[k].mesh_rank] = (uint32_t)k;  // first row of each mesh
    EnsureContext();
    if (m_Flat.submeshTable.empty() || m_Flat.nodeCount == 0) return;
    SS_TRY(ss_set_key_layout(m_Submit, m_Flat.meshBits, m_Flat.submeshBits, m_Flat.materialBits));
    if (HasGeometry()) {
        // one geometry arena for the consumer passes: the table the library sees addresses it
        std::vector<ss_submesh> table = m_Flat.submeshTable;
        for (size_t k = 0; k < table.size(); ++k) {
            table[k].first_index += m_Flat.indexBaseOfSlot[k];
            table[k].vertex_offset += (int32_t)m_Flat.vertexBaseOfSlot[k];
        }
        SS_TRY(ss_set_submesh_table(m_Submit, table.data(), (uint32_t)table.size()));
        SS_TRY(ss_set_geometry(m_Submit, m_Flat.positions.data(), (uint32_t)(m_Flat.positions.size() / 3), m_Flat.indices.data(),
                               (uint32_t)m_Flat.indices.size()));
    } else {
        SS_TRY(ss_set_submesh_table(m_Submit, m_Flat.submeshTable.data(), (uint32_t)m_Flat.submeshTable.size()));
    }
    SS_TRY(ss_set_nodes(m_Submit, m_Flat.nodeCount, m_Flat.parent.empty() ? nullptr : m_Flat.parent.data(),
                        m_Flat.parent.empty() ? nullptr : m_Flat.levelOffsets.data(),
                        m_Flat.parent.empty() ? 0u : (uint32_t)m_Flat.levelOffsets.size() - 1));
    SS_TRY(ss_set_instances(m_Submit, (uint32_t)m_Flat.node.size(), m_Flat.node.data(), 0, m_Flat.slot.data(),
                            m_Flat.material.data()));
}

void SceneRenderer::UploadTransforms(Scene& scene)
{
    if (!m_Submit || m_Flat.nodeCount == 0) return;
    Pool<TransformComponent>& tpool = scene.m_Registry.storage<TransformComponent>();
    const uint32_t nT = (uint32_t)tpool.size();
    if (nT == 0) return;
    if (m_Flat.identityNodeOrder) {
        // the pool's packed component array is the TRS stream: no per-entity work on the host at all
        SS_TRY(ss_upload_trs(m_Submit, 0, nT, reinterpret_cast<const float*>(tpool.data.data())));
    } else {
        m_TrsStaging.resize((size_t)nT * 9);
        for (uint32_t k = 0; k < nT; ++k)
            std::memcpy(&m_TrsStaging[(size_t)m_Flat.nodeOfPoolIndex[k] * 9], &tpool.data[k], 36);
        SS_TRY(ss_upload_trs(m_Submit, 0, nT, m_TrsStaging.data()));
    }
}

void SceneRenderer::SubmitMesh(const std::shared_ptr<Mesh>& mesh, const std::shared_ptr<MaterialTable>& materialTable, const mat4& transform)
{
    if (!mesh || !materialTable) return;  // SceneRenderer.cpp:1128-1129
    m_AdHoc.push_back({mesh, materialTable, nullptr, transform, 0xFFFFFFFFu});
}

void SceneRenderer::SubmitSelectedMesh(const std::shared_ptr<Mesh>& mesh, const std::shared_ptr<MaterialTable>& materialTable,
                                       const mat4& transform, uint32_t entityID)
{
    if (!mesh || !materialTable) return;  // :1173-1174
    m_AdHoc.push_back({mesh, materialTable, nullptr, transform, entityID});
}

void SceneRenderer::SubmitPhysicsMesh(const std::shared_ptr<Mesh>& mesh, const std::shared_ptr<MaterialAsset>& materialAsset, const mat4& transform)
{
    if (!mesh || !materialAsset) return;  // :1215-1216
    m_AdHoc.push_back({mesh, nullptr, materialAsset, transform, 0xFFFFFFFFu});
}

void SceneRenderer::UpdateMeshInstanceData(std::shared_ptr<Mesh>, std::shared_ptr<MaterialTable>)
{
    // SceneRenderer.cpp:1247-1263 erases the key of an edited mesh/material; here the static streams are rebuilt
    m_BuiltVersion = ~0ull;
}

void SceneRenderer::RenderScene()
{
    if (!m_Scene) return;
    // static streams are rebuilt only when the registry or the set of ad-hoc submitters changed
    bool sameAdHoc = m_AdHoc.size() == m_AdHocBuilt.size();
    for (size_t k = 0; sameAdHoc && k < m_AdHoc.size(); ++k)
        sameAdHoc = m_AdHoc[k].mesh == m_AdHocBuilt[k].mesh && m_AdHoc[k].table == m_AdHocBuilt[k].table &&
                    m_AdHoc[k].material == m_AdHocBuilt[k].material && m_AdHoc[k].entityID == m_AdHocBuilt[k].entityID;
    if (m_BuiltVersion != m_Scene->m_StructureVersion || !sameAdHoc) RebuildStaticStreams();
    for (auto& l : m_DrawLists) l.clear();
    if (m_Flat.node.empty()) {
        m_Stats = ss_frame_stats{};
        m_AdHoc.clear();
        return;
    }

    m_Scene->OnUpdateGeometry(this);  // ShadowPass, SceneRenderer.cpp:1321
    if (!m_AdHoc.empty()) {            // finished transforms of the ad-hoc submitters -> external nodes
        std::vector<float> rows(m_AdHoc.size() * 12);
        for (size_t k = 0; k < m_AdHoc.size(); ++k)
            for (int r = 0; r < 3; ++r)
                for (int col = 0; col < 4; ++col) rows[k * 12 + r * 4 + col] = m_AdHoc[k].transform.c[col][r];  // :1151-1153
        const uint32_t first = m_Flat.nodeCount - m_Flat.externalNodes;
        SS_TRY(ss_upload_node_world_rows(m_Submit, first, (uint32_t)m_AdHoc.size(), rows.data()));
        SS_TRY(ss_flush_uploads(m_Submit));
    }

    // views: camera P*V (EditorCamera.h:45) + UpdateCascadeMap (SceneRenderer.cpp:579-687, m_CSMSettings defaults)
    ss_view views[5];
    std::memset(views, 0, sizeof views);
    const float pos[3] = {m_Camera.Position.x, m_Camera.Position.y, m_Camera.Position.z};
    SS_TRY(ss_camera_view_proj(m_Camera.FOV, m_Camera.AspectRatio, m_Camera.NearClip, m_Camera.FarClip, pos, m_Camera.Pitch,
                               m_Camera.Yaw, nullptr, nullptr, views[0].view_proj));
    const vec3 ld = m_Scene->GetDirectionalLightDirection();
    const float light[3] = {ld.x, ld.y, ld.z};
    float cascades[4][16];
    SS_TRY(ss_cascade_matrices(views[0].view_proj, m_Camera.NearClip, m_Camera.FarClip, 0.92f, -75.0f, 75.0f, light,
                               m_CascadeSplits, cascades));
    for (int i = 0; i < 4; ++i) {
        std::memcpy(views[1 + i].view_proj, cascades[i], 64);
        views[1 + i].flags = SS_VIEW_SKIP_NEAR;  // shadow pipeline: DepthClamp (SceneRenderer.cpp:188)
    }

    SS_TRY(ss_frame(m_Submit, views, 5, SS_FRAME_DEFAULT, &m_Slot));
    SS_TRY(ss_frame_sync(m_Submit, m_Slot, &m_Stats));

    // the pass loops `for (auto& [mk, dc] : m_MeshDrawList)` iterate these instead
    const uint32_t mb = m_Flat.materialBits, sb = m_Flat.submeshBits, hb = m_Flat.meshBits;
    for (uint32_t v = 0; v < 5; ++v) {
        const uint32_t nd = m_Stats.draws[v];
        std::vector<ss_draw_cmd> cmds(nd);
        std::vector<uint64_t> keys(nd);
        SS_TRY(ss_download_view(m_Submit, m_Slot, v, nullptr, nullptr, nullptr, cmds.data(), keys.data()));
        m_DrawLists[v].resize(nd);
        for (uint32_t d = 0; d < nd; ++d) {
            const uint64_t k = keys[d];
            MeshKey mk;
            mk.MaterialHandle = m_Flat.materialHandles[k & ((1ull << mb) - 1)];
            mk.SubmeshIndex = (uint32_t)((k >> mb) & ((1ull << sb) - 1));
            mk.MeshHandle = m_Flat.meshHandles[(k >> (mb + sb)) & ((1ull << hb) - 1)];
            mk.Tessellated = ((k >> (mb + sb + hb)) & 1) != 0;
            if (HasGeometry()) {  // back to the per-source offsets the reference's draw would carry (VulkanRenderer.cpp:628)
                const uint32_t row = m_SlotOfMeshRank[(k >> (mb + sb)) & ((1ull << hb) - 1)] + mk.SubmeshIndex;
                cmds[d].first_index -= m_Flat.indexBaseOfSlot[row];
                cmds[d].vertex_offset -= (int32_t)m_Flat.vertexBaseOfSlot[row];
            }
            m_DrawLists[v][d] = {mk, cmds[d]};
        }
    }
    m_AdHoc.clear();  // ResetDrawCommands (SceneRenderer.cpp:1102, 1559-1591): per-frame submissions do not persist
}

// ---- consumers of the lists ------------------------------------------------------------------------
// SceneRenderer.cpp:1421-1457: render the entity id of every selected-pass instance, read the pixel under the mouse.
void SceneRenderer::SelectionPass()
{
    m_HoveredEntity = -1;
    if (!m_Submit || !HasGeometry() || m_Flat.node.empty()) return;
    SS_TRY(ss_raster_view(m_Submit, m_Slot, 0, m_ViewportWidth, m_ViewportHeight, SS_RASTER_DEFAULT));
    if (m_MouseX >= m_ViewportWidth || m_MouseY >= m_ViewportHeight) return;
    uint32_t instance = 0xFFFFFFFFu;
    SS_TRY(ss_read_pixel(m_Submit, 0, m_MouseX, m_MouseY, &instance, nullptr));
    if (instance != 0xFFFFFFFFu && instance < m_Flat.entityOfInstance.size()) m_HoveredEntity = (int)m_Flat.entityOfInstance[instance];
}

// SceneRenderer.cpp:1266-1310: depth of the four cascades, one layer each (the geometry shader's gl_Layer)
void SceneRenderer::ShadowPass()
{
    if (!m_Submit || !HasGeometry() || m_Flat.node.empty()) return;
    for (uint32_t c = 1; c <= 4; ++c) SS_TRY(ss_raster_view(m_Submit, m_Slot, c, m_ShadowMapSize, m_ShadowMapSize, SS_RASTER_DEPTH_CLAMP));
}

std::vector<uint64_t> SceneRenderer::DownloadRasterTarget(uint32_t view)
{
    std::vector<uint64_t> out((size_t)GetRasterWidth(view) * GetRasterHeight(view));
    SS_TRY(ss_download_raster(m_Submit, view, out.data()));
    return out;
}

ss_outputs SceneRenderer::GetDeviceOutputs(uint32_t view)
{
    ss_outputs o{};
    SS_TRY(ss_map_outputs(m_Submit, m_Slot, view, &o));
    return o;
}

std::vector<float> SceneRenderer::DownloadInstanceRows(uint32_t view)
{
    std::vector<float> rows((size_t)m_Stats.visible[view] * 12);
    SS_TRY(ss_download_view(m_Submit, m_Slot, view, rows.data(), nullptr, nullptr, nullptr, nullptr));
    return rows;
}

std::vector<uint32_t> SceneRenderer::DownloadInstanceEntities(uint32_t view)
{
    std::vector<uint32_t> idx(m_Stats.visible[view]);
    SS_TRY(ss_download_view(m_Submit, m_Slot, view, nullptr, nullptr, idx.data(), nullptr, nullptr));
    for (uint32_t& i : idx) i = m_Flat.entityOfInstance[i];
    return idx;
}

}  // namespace VulkanCore
