// host_cli — drives the headless Scene / SceneRenderer mirror from a scene pack written by the tests.
//
//   host_cli flatten <pack> <out>     build the scene through the mirror API, dump the flattened streams (no GPU)
//   host_cli render  <pack> <out>     additionally RenderScene() on cuda:0 and dump per-view draw lists + rows
//   host_cli load-flatten <assets_dir> <scene.vkscene> <out>   the same two dumps for a scene read from the reference's
//   host_cli load-render  <assets_dir> <scene.vkscene> <out>   own files (AssetRegistry.vkr, .vkscene, .vkmesh, .vkmat, FBX),
//                                                               preceded by what the loader read (entities, submesh tables)
//   host_cli load-pick    <assets_dir> <scene.vkscene> <out> <w> <h> <shadow> <mx> <my>
//                                     load-render + SelectionPass (w x h, mouse at mx,my) + ShadowPass (shadow x shadow):
//                                     appends the geometry arena, the five raster targets and the hovered entity
//
// Pack / dump format: a sequence of blocks, each  uint64 byte_count  followed by the raw bytes.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <string>
#include <vector>

#include "Scene.h"
#include "SceneSerializer.h"

using namespace VulkanCore;

namespace {

struct Reader {
    FILE* f;
    template <class T>
    std::vector<T> block()
    {
        uint64_t bytes = 0;
        if (fread(&bytes, 8, 1, f) != 1) throw std::runtime_error("pack truncated");
        std::vector<T> v(bytes / sizeof(T));
        if (bytes && fread(v.data(), 1, bytes, f) != bytes) throw std::runtime_error("pack truncated");
        return v;
    }
};

struct Writer {
    FILE* f;
    template <class T>
    void block(const T* p, size_t n)
    {
        const uint64_t bytes = n * sizeof(T);
        fwrite(&bytes, 8, 1, f);
        if (bytes) fwrite(p, 1, bytes, f);
    }
    template <class T>
    void block(const std::vector<T>& v) { block(v.data(), v.size()); }
};

// scene described by a pack written by the tests (tests/test_host_mirror.py)
std::shared_ptr<Scene> scene_from_pack(const char* path, std::vector<float>& camera)
{
        FILE* in = fopen(path, "rb");
        if (!in) throw std::runtime_error("cannot open pack");
        Reader rd{in};
        auto trs = rd.block<float>();             // [n][9]
        auto parent = rd.block<int32_t>();        // [n] or empty
        auto meshOf = rd.block<int32_t>();        // [n] -1 = null asset
        auto tableOf = rd.block<int32_t>();       // [n] -1 = null table
        auto hasMesh = rd.block<uint8_t>();       // [n]
        auto meshHandle = rd.block<uint64_t>();   // [m]
        auto meshVertexCount = rd.block<uint32_t>();
        auto meshSubBegin = rd.block<uint32_t>(); // [m+1]
        auto baseMatHandle = rd.block<uint64_t>();
        auto baseMatTess = rd.block<uint8_t>();
        auto subMatIndex = rd.block<int32_t>();   // [s]
        auto subLocal = rd.block<float>();        // [s][16]
        auto subAabb = rd.block<float>();         // [s][6]
        auto tableBegin = rd.block<uint32_t>();   // [t+1]
        auto entryIndex = rd.block<uint32_t>();
        auto entryHandle = rd.block<uint64_t>();
        auto entryTess = rd.block<uint8_t>();
        camera = rd.block<float>();          // fov aspect near far px py pz pitch yaw
        std::vector<uint32_t> subCounts;     // optional: [s][BaseVertex, BaseIndex, VertexCount, IndexCount]
        {
            const long at = ftell(in);
            uint64_t bytes = 0;
            if (fread(&bytes, 8, 1, in) == 1) {
                fseek(in, at, SEEK_SET);
                subCounts = rd.block<uint32_t>();
            }
        }
        fclose(in);

        const size_t n = trs.size() / 9, m = meshHandle.size(), t = tableBegin.size() - 1;
        AssetManager::Clear();
        std::map<uint64_t, std::shared_ptr<MaterialAsset>> materials;
        auto material = [&](uint64_t h, bool tess) {
            auto& p = materials[h];
            if (!p) {
                p = std::make_shared<MaterialAsset>();
                p->Handle = h;
                p->DisplacementTexture = tess;
            }
            return p;
        };
        std::vector<std::shared_ptr<Mesh>> meshes(m);
        for (size_t k = 0; k < m; ++k) {
            auto src = std::make_shared<MeshSource>();
            const uint32_t ns = meshSubBegin[k + 1] - meshSubBegin[k];
            for (uint32_t s = meshSubBegin[k]; s < meshSubBegin[k + 1]; ++s) {
                Submesh sm;
                sm.MaterialIndex = subMatIndex[s];
                std::memcpy(&sm.LocalTransform, &subLocal[(size_t)s * 16], 64);
                sm.BoundsMin = {subAabb[s * 6 + 0], subAabb[s * 6 + 1], subAabb[s * 6 + 2]};
                sm.BoundsMax = {subAabb[s * 6 + 3], subAabb[s * 6 + 4], subAabb[s * 6 + 5]};
                if (subCounts.size() >= (size_t)(s + 1) * 4) {  // the pack carries real counts
                    sm.BaseVertex = subCounts[s * 4 + 0];
                    sm.BaseIndex = subCounts[s * 4 + 1];
                    sm.VertexCount = subCounts[s * 4 + 2];
                    sm.IndexCount = subCounts[s * 4 + 3];
                } else {
                    sm.IndexCount = 3 * (s + 10);
                    sm.BaseIndex = 1000 * s;
                    sm.BaseVertex = 77 * s;
                    // spread the mesh's vertex count over its submeshes (0 stays 0: the zero-vertex early-out)
                    sm.VertexCount = meshVertexCount[k] / ns + (s == meshSubBegin[k] ? meshVertexCount[k] % ns : 0);
                }
                src->GetSubmeshes().push_back(sm);
            }
            src->SetBaseMaterial(baseMatHandle[k] ? material(baseMatHandle[k], baseMatTess[k] != 0) : nullptr);
            meshes[k] = std::make_shared<Mesh>(src);
            meshes[k]->Handle = meshHandle[k];
            AssetManager::Register(meshes[k]);
        }
        std::vector<std::shared_ptr<MaterialTable>> tables(t);
        for (size_t k = 0; k < t; ++k) {
            tables[k] = std::make_shared<MaterialTable>();
            for (uint32_t e = tableBegin[k]; e < tableBegin[k + 1]; ++e)
                tables[k]->SetMaterial(entryIndex[e], entryHandle[e] ? material(entryHandle[e], entryTess[e] != 0) : nullptr);
        }

        auto scene = std::make_shared<Scene>();
        std::vector<Entity> ents(n);
        for (size_t i = 0; i < n; ++i) {
            ents[i] = scene->CreateEntity();
            auto& tc = ents[i].GetComponent<TransformComponent>();
            tc.Translation = {trs[i * 9 + 0], trs[i * 9 + 1], trs[i * 9 + 2]};
            tc.Rotation = {trs[i * 9 + 3], trs[i * 9 + 4], trs[i * 9 + 5]};
            tc.Scale = {trs[i * 9 + 6], trs[i * 9 + 7], trs[i * 9 + 8]};
            if (!parent.empty() && parent[i] >= 0) ents[i].AddComponent<RelationshipComponent>(RelationshipComponent{(entity)parent[i]});
            if (hasMesh[i]) {
                MeshComponent mc;
                mc.MeshHandle = meshOf[i] >= 0 ? meshHandle[meshOf[i]] : 0xDEADull;  // unknown handle -> null asset
                mc.MaterialTableHandle = tableOf[i] >= 0 ? tables[tableOf[i]] : nullptr;
                ents[i].AddComponent<MeshComponent>(mc);
            }
        }

        return scene;
}

// what the loader read, ahead of the flatten / render dump
void dump_loaded(Writer& wr, Scene& scene)
{
    Registry& reg = scene.m_Registry;
    std::vector<uint64_t> uuid, meshHandle;
    std::vector<uint8_t> hasTransform;
    std::vector<float> trs;
    for (entity e = 0; e < reg.alive(); ++e) {
        uuid.push_back(reg.get<IDComponent>(e).ID);
        hasTransform.push_back(reg.all_of<TransformComponent>(e));
        meshHandle.push_back(reg.all_of<MeshComponent>(e) ? reg.get<MeshComponent>(e).MeshHandle : 0);
        if (hasTransform.back()) {
            const TransformComponent& t = reg.get<TransformComponent>(e);
            const float v[9] = {t.Translation.x, t.Translation.y, t.Translation.z, t.Rotation.x, t.Rotation.y, t.Rotation.z, t.Scale.x, t.Scale.y, t.Scale.z};
            trs.insert(trs.end(), v, v + 9);
        }
    }
    wr.block(uuid);
    wr.block(hasTransform);
    wr.block(meshHandle);
    wr.block(trs);
    const vec3 ld = scene.GetDirectionalLightDirection();
    const float l[3] = {ld.x, ld.y, ld.z};
    wr.block(l, 3);
    // submesh tables of the meshes the scene names, in order of first use
    std::vector<uint64_t> seen;
    std::vector<uint32_t> counts, ints;
    std::vector<float> floats;
    for (uint64_t h : meshHandle) {
        if (!h || std::find(seen.begin(), seen.end(), h) != seen.end()) continue;
        seen.push_back(h);
        std::shared_ptr<Mesh> mesh = AssetManager::GetMesh(h);
        counts.push_back(mesh ? (uint32_t)mesh->GetMeshSource()->GetSubmeshes().size() : 0u);
        if (!mesh) continue;
        for (const Submesh& sm : mesh->GetMeshSource()->GetSubmeshes()) {
            const uint32_t iv[5] = {sm.BaseVertex, sm.BaseIndex, sm.VertexCount, sm.IndexCount, (uint32_t)sm.MaterialIndex};
            ints.insert(ints.end(), iv, iv + 5);
            floats.insert(floats.end(), sm.LocalTransform.data(), sm.LocalTransform.data() + 16);
            const float bb[6] = {sm.BoundsMin.x, sm.BoundsMin.y, sm.BoundsMin.z, sm.BoundsMax.x, sm.BoundsMax.y, sm.BoundsMax.z};
            floats.insert(floats.end(), bb, bb + 6);
        }
    }
    wr.block(seen);
    wr.block(counts);
    wr.block(ints);
    wr.block(floats);
}

}  // namespace

int main(int argc, char** argv)
{
    const std::string mode = argc > 1 ? argv[1] : "";
    const bool pick = mode == "load-pick";
    const bool from_files = mode == "load-flatten" || mode == "load-render" || pick;
    if (argc < (pick ? 10 : from_files ? 5 : 4) || !(from_files || mode == "flatten" || mode == "render")) {
        fprintf(stderr, "usage: host_cli flatten|render <pack> <out>\n       host_cli load-flatten|load-render <assets_dir> <scene.vkscene> <out>\n");
        return 2;
    }
    try {
        std::vector<float> camera;
        std::shared_ptr<Scene> scene;
        if (from_files) {
            AssetManager::Clear();
            if (!AssetManager::LoadRegistryFromFile(argv[2])) throw std::runtime_error("cannot read AssetRegistry.vkr");
            scene = std::make_shared<Scene>();
            SceneSerializer serializer(scene);
            if (!serializer.Deserialize(argv[3])) throw std::runtime_error("cannot deserialize scene");
            if (!AssetManager::LastImportError().empty()) fprintf(stderr, "host_cli: %s\n", AssetManager::LastImportError().c_str());
        } else {
            scene = scene_from_pack(argv[2], camera);
        }
        const bool render = mode == "render" || mode == "load-render" || pick;

        FILE* out = fopen(argv[from_files ? 4 : 3], "wb");
        if (!out) throw std::runtime_error("cannot open output");
        Writer wr{out};
        if (from_files) dump_loaded(wr, *scene);
        if (!render) {
            FlatScene f = SceneRenderer::Flatten(*scene, {});
            wr.block(f.node);
            wr.block(f.slot);
            wr.block(f.material);
            wr.block(f.entityOfInstance);
            wr.block(f.parent);
            wr.block(f.levelOffsets);
            wr.block(f.submeshTable);
            wr.block(f.meshHandles);
            wr.block(f.materialHandles);
            const uint32_t bits[3] = {f.meshBits, f.submeshBits, f.materialBits};
            wr.block(bits, 3);
            wr.block(f.positions);  // geometry arena of the consumer passes (empty for packs)
            wr.block(f.indices);
            wr.block(f.vertexBaseOfSlot);
            wr.block(f.indexBaseOfSlot);
        } else {
            SceneRenderer renderer(0);
            renderer.SetActiveScene(scene);
            EditorCameraData cam;  // EditorLayer.cpp:75 defaults unless the pack names a camera
            if (camera.size() >= 9) {
            cam.FOV = camera[0]; cam.AspectRatio = camera[1]; cam.NearClip = camera[2]; cam.FarClip = camera[3];
            cam.Position = {camera[4], camera[5], camera[6]};
            cam.Pitch = camera[7]; cam.Yaw = camera[8];
            }
            renderer.SetSceneEditorData(cam);
            renderer.RenderScene();
            renderer.RenderScene();  // a second frame: slots rotate, static streams must not be rebuilt
            const ss_frame_stats& st = renderer.GetStats();
            wr.block(st.visible, 5);
            wr.block(st.draws, 5);
            for (uint32_t v = 0; v < 5; ++v) {
                wr.block(renderer.DownloadInstanceRows(v));
                wr.block(renderer.DownloadInstanceEntities(v));
                const auto& dl = renderer.GetDrawList(v);
                std::vector<uint64_t> flat;
                for (const DrawCommand& d : dl) {
                    flat.push_back(d.Key.MeshHandle);
                    flat.push_back(d.Key.MaterialHandle);
                    flat.push_back(d.Key.SubmeshIndex);
                    flat.push_back(d.Key.Tessellated);
                    flat.push_back(d.Indirect.index_count);
                    flat.push_back(d.Indirect.instance_count);
                    flat.push_back(d.Indirect.first_index);
                    flat.push_back((uint64_t)(int64_t)d.Indirect.vertex_offset);
                    flat.push_back(d.Indirect.first_instance);
                }
                wr.block(flat);
            }
            if (pick) {
                const uint32_t w = (uint32_t)atoi(argv[5]), h = (uint32_t)atoi(argv[6]), sm = (uint32_t)atoi(argv[7]);
                renderer.SetViewportSize(w, h);
                renderer.SetShadowMapSize(sm);
                renderer.SetViewportMousePos((uint32_t)atoi(argv[8]), (uint32_t)atoi(argv[9]));
                renderer.SelectionPass();
                renderer.ShadowPass();
                const FlatScene& f = renderer.GetFlatScene();
                wr.block(f.positions);
                wr.block(f.indices);
                wr.block(f.vertexBaseOfSlot);
                wr.block(f.indexBaseOfSlot);
                wr.block(f.slot);
                wr.block(f.submeshTable);
                for (uint32_t v = 0; v < 5; ++v) wr.block(renderer.DownloadRasterTarget(v));
                const int32_t hovered = renderer.GetHoveredEntity();
                wr.block(&hovered, 1);
            }
        }
        fclose(out);
    } catch (const std::exception& e) {
        fprintf(stderr, "host_cli: %s\n", e.what());
        return 1;
    }
    return 0;
}
