// Headless mirror of the reference's Scene / SceneRenderer submission API over libscenesubmit.
//
// Same names and signatures as the reference (VulkanCore/src/VulkanCore/Scene/Scene.h:34-38,
// SceneRenderer.h:27-36), so that the call sites — EditorLayer::OnUpdate -> RenderScene, ShadowPass ->
// OnUpdateGeometry, JoltDebugRenderer -> SubmitPhysicsMesh — would compile unchanged; the bodies call
// the C ABI (include/scenesubmit.h) instead of walking std::map<MeshKey, ...>.
#pragma once
#include <map>
#include <string>
#include <vector>

#include "../../include/scenesubmit.h"
#include "Core.h"

namespace VulkanCore {

// ---- components (VulkanCore/src/VulkanCore/Core/Components.h) ------------------------------------------
struct IDComponent {
    UUID ID = 0;
};

struct TransformComponent {  // :13-15 — 36 bytes, the record ss_upload_trs takes
    vec3 Translation{0.f, 0.f, 0.f}, Rotation{0.f, 0.f, 0.f}, Scale{1.f, 1.f, 1.f};
};
static_assert(sizeof(TransformComponent) == 36, "TransformComponent must match the TRS stream record");

// extension N1 (the reference has no parent component)
struct RelationshipComponent {
    entity Parent = null_entity;
};

// ---- assets (only what the submission path reads) --------------------------------------------------
struct MaterialAsset {  // VulkanCore/src/VulkanCore/Asset/MaterialAsset.h
    AssetHandle Handle = 0;
    bool DisplacementTexture = false;
    bool HasDisplacementTexture() const { return DisplacementTexture; }
};

class MaterialTable {  // MaterialAsset.h:28-43, MaterialAsset.cpp:43-51
public:
    void SetMaterial(uint32_t index, std::shared_ptr<MaterialAsset> material) { m_Materials[index] = std::move(material); }
    // the reference's HasMaterial calls .at() and so throws on a missing index; the mirror reports "no"
    // (a missing or null entry falls back to the base material either way, SceneRenderer.cpp:1141)
    bool HasMaterial(uint32_t index) const
    {
        auto it = m_Materials.find(index);
        return it != m_Materials.end() && it->second != nullptr;
    }
    std::shared_ptr<MaterialAsset> GetMaterial(uint32_t index) const { return m_Materials.at(index); }
    const std::map<uint32_t, std::shared_ptr<MaterialAsset>>& GetMaterialMap() const { return m_Materials; }

private:
    std::map<uint32_t, std::shared_ptr<MaterialAsset>> m_Materials;
};

struct Submesh {  // VulkanCore/src/VulkanCore/Mesh/Mesh.h:39-46 + the bounds the reference lacks
    uint32_t BaseVertex = 0, BaseIndex = 0, VertexCount = 0, IndexCount = 0;
    int MaterialIndex = 0;
    mat4 LocalTransform = mat4::identity();
    vec3 BoundsMin{-1.f, -1.f, -1.f}, BoundsMax{1.f, 1.f, 1.f};
};

class MeshSource {
public:
    std::vector<Submesh>& GetSubmeshes() { return m_Submeshes; }
    const std::vector<Submesh>& GetSubmeshes() const { return m_Submeshes; }
    uint32_t GetVertexCount() const
    {
        uint32_t n = 0;
        for (const Submesh& s : m_Submeshes) n += s.VertexCount;
        return n;
    }
    std::shared_ptr<MaterialAsset> GetBaseMaterial() const { return m_BaseMaterial; }
    void SetBaseMaterial(std::shared_ptr<MaterialAsset> m) { m_BaseMaterial = std::move(m); }
    // vertex positions and indices the submeshes' BaseVertex / BaseIndex refer to (Mesh.h: m_Vertices / m_Indices;
    // only the position attribute is kept). Empty when the source carried no geometry (packs written by the tests).
    std::vector<vec3>& GetPositions() { return m_Positions; }
    const std::vector<vec3>& GetPositions() const { return m_Positions; }
    std::vector<uint32_t>& GetIndices() { return m_Indices; }
    const std::vector<uint32_t>& GetIndices() const { return m_Indices; }

private:
    std::vector<Submesh> m_Submeshes;
    std::shared_ptr<MaterialAsset> m_BaseMaterial;
    std::vector<vec3> m_Positions;
    std::vector<uint32_t> m_Indices;
};

class Mesh {  // Mesh.h; GetSubmeshes() lists indices into the source's submesh array (Mesh.cpp:47-54)
public:
    AssetHandle Handle = 0;
    explicit Mesh(std::shared_ptr<MeshSource> source) : m_MeshSource(std::move(source))
    {
        for (uint32_t i = 0; i < m_MeshSource->GetSubmeshes().size(); ++i) m_Submeshes.push_back(i);
    }
    std::shared_ptr<MeshSource> GetMeshSource() const { return m_MeshSource; }
    const std::vector<uint32_t>& GetSubmeshes() const { return m_Submeshes; }

private:
    std::shared_ptr<MeshSource> m_MeshSource;
    std::vector<uint32_t> m_Submeshes;
};

struct MeshComponent {  // Components.h:86-93
    AssetHandle MeshHandle = 0;
    std::shared_ptr<MaterialTable> MaterialTableHandle;
};

enum class AssetType { None, Scene, Texture2D, TextureCube, Mesh, MeshAsset, Material };  // Asset/Asset.cpp:10-42

struct AssetMetadata {  // Asset/AssetMetadata.h
    std::string FilePath;
    AssetType Type = AssetType::None;
};

class AssetManager {  // the one lookup the hot loop performs (Scene.cpp:144) + the on-disk registry behind it
public:
    static void Register(const std::shared_ptr<Mesh>& mesh) { Meshes()[mesh->Handle] = mesh; }
    // a mesh that is registered, or named by the loaded registry (then imported on first use, Asset/MeshImporter.cpp)
    static std::shared_ptr<Mesh> GetMesh(AssetHandle h)
    {
        auto it = Meshes().find(h);
        if (it != Meshes().end()) return it->second;
        return RegistryMap().empty() ? nullptr : LoadMesh(h);
    }
    static void Clear();

    // AssetRegistry.vkr in `assetDirectory` (Asset/AssetManager.cpp:32-65); file paths inside are relative to it
    static bool LoadRegistryFromFile(const std::string& assetDirectory);
    static bool IsAssetHandleValid(AssetHandle h) { return h != 0 && RegistryMap().count(h) != 0; }
    static const std::map<AssetHandle, AssetMetadata>& GetAssetRegistry() { return RegistryMap(); }
    // .vkmat -> the one property the submission path reads (Asset/MaterialAssetImporter.cpp:144-181); nullptr when unknown
    static std::shared_ptr<MaterialAsset> GetMaterialAsset(AssetHandle h);
    static const std::string& LastImportError() { return ImportError(); }

private:
    static std::shared_ptr<Mesh> LoadMesh(AssetHandle h);
    static std::map<AssetHandle, AssetMetadata>& RegistryMap()
    {
        static std::map<AssetHandle, AssetMetadata> s;
        return s;
    }
    static std::map<AssetHandle, std::shared_ptr<MaterialAsset>>& Materials()
    {
        static std::map<AssetHandle, std::shared_ptr<MaterialAsset>> s;
        return s;
    }
    static std::map<AssetHandle, std::shared_ptr<MeshSource>>& Sources()
    {
        static std::map<AssetHandle, std::shared_ptr<MeshSource>> s;
        return s;
    }
    static std::string& Directory()
    {
        static std::string s;
        return s;
    }
    static std::string& ImportError()
    {
        static std::string s;
        return s;
    }
    static std::map<AssetHandle, std::shared_ptr<Mesh>>& Meshes()
    {
        static std::map<AssetHandle, std::shared_ptr<Mesh>> s;
        return s;
    }
};

// ---- camera data handed to the renderer (SceneEditorData, EditorCamera) --------------------------------
struct EditorCameraData {
    float FOV = 0.785398f, AspectRatio = 1.635005f, NearClip = 0.1f, FarClip = 1000.0f;  // EditorLayer.cpp:75
    vec3 Position{0.f, 0.f, 10.f};
    float Pitch = 0.f, Yaw = 0.f;
};

class SceneRenderer;

class Entity {
public:
    Entity() = default;
    Entity(entity handle, class Scene* scene) : m_Handle(handle), m_Scene(scene) {}
    template <class T, class... Args>
    T& AddComponent(Args&&... args);
    template <class T>
    T& GetComponent();
    template <class T>
    bool HasComponent() const;
    operator uint32_t() const { return m_Handle; }

private:
    entity m_Handle = null_entity;
    class Scene* m_Scene = nullptr;
};

class Scene {
public:
    Entity CreateEntity(const std::string& name = std::string());            // Scene.cpp: CreateEntity
    Entity CreateEntityWithUUID(UUID uuid, const std::string& name = std::string());

    void OnUpdateGeometry(SceneRenderer* renderer);  // Scene.h:34, Scene.cpp:135-147
    void OnSelectGeometry(SceneRenderer* renderer);  // Scene.h:35, Scene.cpp:149-160

    vec3 GetDirectionalLightDirection() const { return m_LightDirection; }  // Components.h:144 default
    void SetDirectionalLightDirection(vec3 d) { m_LightDirection = d; }

    Registry m_Registry;
    uint64_t m_StructureVersion = 0;  // bumped when entities / meshes / parents change

private:
    vec3 m_LightDirection{1.f, -1.f, 1.f};
    UUID m_NextUUID = 1;
    friend class Entity;
};

template <class T, class... Args>
T& Entity::AddComponent(Args&&... args)
{
    ++m_Scene->m_StructureVersion;
    return m_Scene->m_Registry.emplace<T>(m_Handle, std::forward<Args>(args)...);
}
template <class T>
T& Entity::GetComponent()
{
    return m_Scene->m_Registry.get<T>(m_Handle);
}
template <class T>
bool Entity::HasComponent() const
{
    return m_Scene->m_Registry.all_of<T>(m_Handle);
}

// What the registry flattens to: the streams of include/scenesubmit.h, in entt view order.
struct FlatScene {
    std::vector<int32_t> parent;          // per node (level order); empty = flat
    std::vector<uint32_t> levelOffsets;
    std::vector<uint32_t> nodeOfPoolIndex;  // TransformComponent pool position -> node number
    bool identityNodeOrder = true;          // node number == pool position: upload the pool as it lies
    std::vector<uint32_t> node, slot, material, entityOfInstance;
    std::vector<ss_submesh> submeshTable;
    std::vector<AssetHandle> meshHandles, materialHandles;  // rank -> handle
    uint32_t meshBits = 1, submeshBits = 1, materialBits = 1;
    uint32_t nodeCount = 0, externalNodes = 0;
    // geometry arena of the consumer passes: every mesh source's positions / indices back to back, in submesh-table
    // order; the reference binds each MeshSource's own buffers per draw, so its BaseVertex / BaseIndex are per source
    // and these bases are what has to be added for one arena.  Empty when any mesh came without geometry.
    std::vector<float> positions;          // xyz
    std::vector<uint32_t> indices;
    std::vector<uint32_t> vertexBaseOfSlot, indexBaseOfSlot;  // per submesh-table row
};

struct MeshKey {  // SceneRenderer.h:51-86, decoded from the 64-bit key
    uint64_t MeshHandle, MaterialHandle;
    uint32_t SubmeshIndex;
    bool Tessellated;
};

struct DrawCommand {  // what `for (auto& [mk, dc] : m_MeshDrawList)` used to yield (SceneRenderer.h:107-114)
    MeshKey Key;
    ss_draw_cmd Indirect;
};

class SceneRenderer {
public:
    explicit SceneRenderer(int device = 0, uint32_t trigMode = SS_TRIG_DETERMINISTIC);
    ~SceneRenderer();
    SceneRenderer(const SceneRenderer&) = delete;

    void SetActiveScene(std::shared_ptr<Scene> scene);                     // SceneRenderer.h:27
    void SetSceneEditorData(const EditorCameraData& cam) { m_Camera = cam; }  // SceneRenderer.h:49
    void RenderScene();                                                       // SceneRenderer.cpp:1060-1107 (submission part)

    // ad-hoc submitters with a finished transform (SceneRenderer.h:32-35)
    void SubmitMesh(const std::shared_ptr<Mesh>& mesh, const std::shared_ptr<MaterialTable>& materialTable, const mat4& transform);
    void SubmitSelectedMesh(const std::shared_ptr<Mesh>& mesh, const std::shared_ptr<MaterialTable>& materialTable,
                            const mat4& transform, uint32_t entityID);
    void SubmitPhysicsMesh(const std::shared_ptr<Mesh>& mesh, const std::shared_ptr<MaterialAsset>& materialAsset, const mat4& transform);
    void UpdateMeshInstanceData(std::shared_ptr<Mesh> mesh, std::shared_ptr<MaterialTable> materialTable);  // SceneRenderer.h:36

    // results of the last RenderScene, per view (0 = camera, 1..4 = cascades)
    uint32_t GetViewCount() const { return 5; }
    const std::vector<DrawCommand>& GetDrawList(uint32_t view) const { return m_DrawLists[view]; }
    ss_outputs GetDeviceOutputs(uint32_t view);
    std::vector<float> DownloadInstanceRows(uint32_t view);
    std::vector<uint32_t> DownloadInstanceEntities(uint32_t view);  // EntityID per row (SelectTransformData, Components.h:100-104)
    const ss_frame_stats& GetStats() const { return m_Stats; }       // RendererStats (PV/VulkanRenderer.h:13-17)
    const FlatScene& GetFlatScene() const { return m_Flat; }
    const float* GetCascadeSplits() const { return m_CascadeSplits; }

    // The consumers of the lists (only with geometry: scenes read from files).  SelectionPass (SceneRenderer.cpp:
    // 1421-1457): entity id per pixel of the camera view, then the read-back of the pixel under the mouse.
    // ShadowPass (:1266-1310): depth of each cascade, MapSize x MapSize (CSMSettings, SceneRenderer.h:179), DepthClamp.
    void SetViewportSize(uint32_t width, uint32_t height) { m_ViewportWidth = width; m_ViewportHeight = height; }
    void SetViewportMousePos(uint32_t x, uint32_t y) { m_MouseX = x; m_MouseY = y; }
    void SetShadowMapSize(uint32_t size) { m_ShadowMapSize = size; }
    bool HasGeometry() const { return !m_Flat.indices.empty(); }
    void SelectionPass();
    void ShadowPass();
    int GetHoveredEntity() const { return m_HoveredEntity; }             // -1 where nothing was drawn
    std::vector<uint64_t> DownloadRasterTarget(uint32_t view);            // [depth bits : 32 | instance : 32] words
    uint32_t GetRasterWidth(uint32_t view) const { return view == 0 ? m_ViewportWidth : m_ShadowMapSize; }
    uint32_t GetRasterHeight(uint32_t view) const { return view == 0 ? m_ViewportHeight : m_ShadowMapSize; }
    const char* LastError() const;

    // called by Scene::OnUpdateGeometry / OnSelectGeometry
    void UploadTransforms(Scene& scene);
    // registry -> streams; public so the flattening rule can be tested without a GPU
    static FlatScene Flatten(Scene& scene, const std::vector<struct AdHocInstance>& adhoc);

private:
    void RebuildStaticStreams();
    void EnsureContext();

    int m_Device;
    uint32_t m_TrigMode;
    ss_ctx* m_Submit = nullptr;
    uint32_t m_Capacity[3] = {0, 0, 0};  // nodes, instances, slots the context was created for
    std::shared_ptr<Scene> m_Scene;
    uint64_t m_BuiltVersion = ~0ull;
    FlatScene m_Flat;
    EditorCameraData m_Camera;
    std::vector<struct AdHocInstance> m_AdHoc, m_AdHocBuilt;
    uint32_t m_Slot = 0;
    ss_frame_stats m_Stats{};
    float m_CascadeSplits[4] = {0, 0, 0, 0};
    std::vector<DrawCommand> m_DrawLists[SS_MAX_VIEWS];
    std::vector<float> m_TrsStaging;
    uint32_t m_ViewportWidth = 1280, m_ViewportHeight = 720, m_MouseX = 0, m_MouseY = 0, m_ShadowMapSize = 4096;
    int m_HoveredEntity = -1;
    std::vector<uint32_t> m_SlotOfMeshRank;  // first submesh-table row of each mesh
    std::string m_Error;
};

struct AdHocInstance {
    std::shared_ptr<Mesh> mesh;
    std::shared_ptr<MaterialTable> table;      // SubmitMesh / SubmitSelectedMesh
    std::shared_ptr<MaterialAsset> material;   // SubmitPhysicsMesh
    mat4 transform;
    uint32_t entityID = 0xFFFFFFFFu;
};

}  // namespace VulkanCore
