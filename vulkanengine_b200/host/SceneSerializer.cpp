#include "SceneSerializer.h"

#include <fstream>
#include <sstream>

#include "MeshImporter.h"
#include "Yaml.h"

namespace VulkanCore {

namespace {

bool read_file(const std::string& path, std::string& out)
{
    std::ifstream stream(path);
    if (!stream) return false;
    std::stringstream ss;
    ss << stream.rdbuf();
    out = ss.str();
    return true;
}

AssetType AssetTypeFromString(const std::string& s)  // Asset/Asset.cpp:27-40
{
    if (s == "Scene") return AssetType::Scene;
    if (s == "Texture2D") return AssetType::Texture2D;
    if (s == "TextureCube") return AssetType::TextureCube;
    if (s == "Mesh") return AssetType::Mesh;
    if (s == "MeshAsset") return AssetType::MeshAsset;
    if (s == "Material") return AssetType::Material;
    return AssetType::None;
}

// Registry paths are relative to the assets directory (the editor's working directory); an absolute path recorded
// on the author's machine is re-rooted at its last "assets/" component.
std::string resolve(const std::string& dir, std::string path)
{
    if (path.find(':') != std::string::npos || (!path.empty() && path[0] == '/')) {
        const size_t k = path.rfind("assets/");
        if (k != std::string::npos) path = path.substr(k + 7);
    }
    return dir + "/" + path;
}

vec3 as_vec3(const Yaml::Node& n, vec3 fallback)
{
    if (!n || n.size() != 3) return fallback;
    return {n[0].f32(), n[1].f32(), n[2].f32()};
}

}  // namespace

// ---- AssetManager: registry, materials, meshes ---------------------------------------------------------
void AssetManager::Clear()
{
    Meshes().clear();
    RegistryMap().clear();
    Materials().clear();
    Sources().clear();
    Directory().clear();
    ImportError().clear();
}

bool AssetManager::LoadRegistryFromFile(const std::string& assetDirectory)
{
    std::string text;
    if (!read_file(assetDirectory + "/AssetRegistry.vkr", text)) return false;
    const Yaml::Node data = Yaml::Load(text);
    const Yaml::Node& reg = data["Asset Registry"];
    if (!reg) return false;
    Directory() = assetDirectory;
    for (const Yaml::Node& asset : reg.seq) {
        const AssetHandle handle = asset["Handle"].u64();
        AssetMetadata md;
        if (const Yaml::Node& m = asset["Metadata"]) {
            md.FilePath = m["Filepath"].str();
            md.Type = AssetTypeFromString(m["Type"].str());
        }
        RegistryMap()[handle] = md;
    }
    return true;
}

std::shared_ptr<MaterialAsset> AssetManager::GetMaterialAsset(AssetHandle h)
{
    auto it = Materials().find(h);
    if (it != Materials().end()) return it->second;
    auto rit = RegistryMap().find(h);
    if (rit == RegistryMap().end() || rit->second.Type != AssetType::Material) return nullptr;
    std::string text;
    if (!read_file(resolve(Directory(), rit->second.FilePath), text)) return Materials()[h] = nullptr;
    const Yaml::Node m = Yaml::Load(text)["Material"];
    if (!m) return Materials()[h] = nullptr;
    auto mat = std::make_shared<MaterialAsset>();
    mat->Handle = h;
    // tessellated = the file names a displacement texture AND that handle is a valid asset
    // (MaterialAssetImporter.cpp:144, 161-181)
    const Yaml::Node& d = m["DisplacementHandle"];
    mat->DisplacementTexture = d && IsAssetHandleValid(d.u64());
    return Materials()[h] = mat;
}

// .vkmesh -> MeshAsset handle -> mesh source file -> submesh table (MeshImporter.cpp:73-95)
std::shared_ptr<Mesh> AssetManager::LoadMesh(AssetHandle h)
{
    auto rit = RegistryMap().find(h);
    if (rit == RegistryMap().end() || rit->second.Type != AssetType::Mesh) return nullptr;
    std::string text;
    if (!read_file(resolve(Directory(), rit->second.FilePath), text)) return nullptr;
    const Yaml::Node root = Yaml::Load(text)["Mesh"];
    if (!root || !root["MeshAsset"]) return nullptr;
    const AssetHandle sourceHandle = root["MeshAsset"].u64();
    std::shared_ptr<MeshSource> source;
    auto sit = Sources().find(sourceHandle);
    if (sit != Sources().end()) {
        source = sit->second;
    } else {
        auto mit = RegistryMap().find(sourceHandle);
        if (mit != RegistryMap().end() && mit->second.Type == AssetType::MeshAsset)
            source = MeshImporter::ImportAssimpMesh(resolve(Directory(), mit->second.FilePath), &ImportError());
        Sources()[sourceHandle] = source;
    }
    if (!source) return nullptr;
    auto mesh = std::make_shared<Mesh>(source);
    mesh->Handle = h;
    Meshes()[h] = mesh;
    return mesh;
}

// ---- SceneSerializer::Deserialize (SceneSerializer.cpp:322-472) ------------------------------------------
bool SceneSerializer::Deserialize(const std::string& filepath)
{
    std::string text;
    if (!read_file(filepath, text)) return false;
    const Yaml::Node data = Yaml::Load(text);
    if (!data["Scene"]) return false;
    m_SceneName = data["Scene"].str();

    const Yaml::Node& entities = data["Entities"];
    if (!entities) return true;
    for (const Yaml::Node& entity : entities.seq) {
        const uint64_t uuid = entity["Entity"].u64();
        std::string name;
        if (const Yaml::Node& tag = entity["TagComponent"]) name = tag["Tag"] ? tag["Tag"].str() : std::string();
        Entity e = m_Scene->CreateEntityWithUUID(uuid, name);

        if (const Yaml::Node& t = entity["TransformComponent"]) {
            auto& tc = e.AddComponent<TransformComponent>();
            tc.Translation = as_vec3(t["Translation"], tc.Translation);
            tc.Rotation = as_vec3(t["Rotation"], tc.Rotation);
            tc.Scale = as_vec3(t["Scale"], tc.Scale);
        }
        if (const Yaml::Node& dl = entity["DirectionalLightComponent"])
            m_Scene->SetDirectionalLightDirection(as_vec3(dl["Direction"], m_Scene->GetDirectionalLightDirection()));

        if (const Yaml::Node& mc = entity["MeshComponent"]) {
            MeshComponent comp;
            comp.MeshHandle = mc["MeshHandle"].u64();
            auto table = std::make_shared<MaterialTable>();
            const Yaml::Node& mt = mc["MaterialTable"];
            if (mt && mt.kind == Yaml::Node::Kind::Map)
                for (const auto& kv : mt.map)
                    table->SetMaterial((uint32_t)strtoul(kv.first.c_str(), nullptr, 10), AssetManager::GetMaterialAsset(kv.second.u64()));
            // a null entry for every submesh material slot the table does not name (:427-433), and the table's
            // slot 0 becomes the mesh source's base material (:437-439)
            if (std::shared_ptr<Mesh> mesh = AssetManager::GetMesh(comp.MeshHandle)) {
                for (const Submesh& sm : mesh->GetMeshSource()->GetSubmeshes())
                    if (!table->GetMaterialMap().count((uint32_t)sm.MaterialIndex)) table->SetMaterial((uint32_t)sm.MaterialIndex, nullptr);
                auto it = table->GetMaterialMap().find(0);
                mesh->GetMeshSource()->SetBaseMaterial(it == table->GetMaterialMap().end() ? nullptr : it->second);
            }
            comp.MaterialTableHandle = table;
            e.AddComponent<MeshComponent>(comp);
        }
    }
    return true;
}

}  // namespace VulkanCore
