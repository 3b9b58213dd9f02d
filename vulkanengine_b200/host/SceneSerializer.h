// Headless mirror of the reference's scene ingest: `.vkscene` -> Scene (VulkanCore/src/VulkanCore/Scene/
// SceneSerializer.h, SceneSerializer.cpp:322-472), with AssetRegistry.vkr / .vkmesh / .vkmat resolved through
// AssetManager (Asset/AssetManager.cpp:32-65, Asset/MeshImporter.cpp:73-95, Asset/MaterialAssetImporter.cpp:135-185).
// Only the components the submission path reads are materialised (ID, Transform, Mesh, the directional light's
// direction); lights' colours, physics bodies and the skybox are parsed past.
#pragma once
#include <memory>
#include <string>

#include "Scene.h"

namespace VulkanCore {

class SceneSerializer {
public:
    explicit SceneSerializer(std::shared_ptr<Scene> scene) : m_Scene(std::move(scene)) {}
    // false when the file is missing or has no "Scene" key (SceneSerializer.cpp:329-330)
    bool Deserialize(const std::string& filepath);
    const std::string& GetSceneName() const { return m_SceneName; }

private:
    std::shared_ptr<Scene> m_Scene;
    std::string m_SceneName;
};

}  // namespace VulkanCore
