// host_cli — drives the headless Scene / SceneRenderer mirror from a scene pack written by the tests.
//
//   host_cli flatten <pack> <out>     build the scene through the mirror API, dump the flattened streams (no GPU)
//   host_cli render  <pack> <out>     additionally RenderScene() on cuda:0 and dump per-view draw lists + rows
//
// Pack / dump format: a sequence of blocks, each  uint64 byte_count  followed by the raw bytes.
#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <string>
#include <vector>

#include "Scene.h"

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

}  // namespace

int main(int argc, char** argv)
{
    if (argc < 4) {
        fprintf(stderr, "usage: host_cli flatten|render <pack> <out>\n");
        return 2;
    }
    const std::string mode = argv[1];
    try {
        FILE* in = fopen(argv[2], "rb");
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
        auto camera = rd.block<float>();          // fov aspect near far px py pz pitch yaw
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
                sm.IndexCount = 3 * (s + 10);
                sm.BaseIndex = 1000 * s;
                sm.BaseVertex = 77 * s;
                // spread the mesh's vertex count over its submeshes (0 stays 0: the zero-vertex early-out)
                sm.VertexCount = meshVertexCount[k] / ns + (s == meshSubBegin[k] ? meshVertexCount[k] % ns : 0);
                src->GetSubmeshes().push_back(sm);
            }
            src->SetBaseMaterial(material(baseMatHandle[k], baseMatTess[k] != 0));
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

        FILE* out = fopen(argv[3], "wb");
        if (!out) throw std::runtime_error("cannot open output");
        Writer wr{out};
        if (mode == "flatten") {
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
        } else {
            SceneRenderer renderer(0);
            renderer.SetActiveScene(scene);
            EditorCameraData cam;
            cam.FOV = camera[0]; cam.AspectRatio = camera[1]; cam.NearClip = camera[2]; cam.FarClip = camera[3];
            cam.Position = {camera[4], camera[5], camera[6]};
            cam.Pitch = camera[7]; cam.Yaw = camera[8];
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
        }
        fclose(out);
    } catch (const std::exception& e) {
        fprintf(stderr, "host_cli: %s\n", e.what());
        return 1;
    }
    return 0;
}
