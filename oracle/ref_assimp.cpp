// Submesh tables of mesh sources through the reference's own importer path: the vendored Assimp (compiled from
// /root/reference/VulkanCore/vendor/Assimp by oracle/Makefile, FBX importer only) with the reference's import flags
// (VulkanCore/src/VulkanCore/Mesh/Mesh.cpp:62-72), then the bookkeeping of MeshSource's constructor (Mesh.cpp:93-110:
// BaseVertex / BaseIndex / VertexCount / IndexCount / MaterialIndex per aiMesh) and of MeshImporter::TraverseNodes
// (VulkanCore/src/VulkanCore/Asset/MeshImporter.cpp:162-188: Submesh::LocalTransform = the node's OWN transformation).
// Test infrastructure: prints one JSON object per file; tests/golden/make_assimp_fixtures.py stores them.
//   ref_assimp <file.fbx> [...]
#include <assimp/postprocess.h>
#include <assimp/scene.h>

#include <assimp/Importer.hpp>
#include <cfloat>
#include <cstdio>
#include <string>
#include <vector>

static const unsigned kMeshImportFlags = aiProcess_Triangulate | aiProcess_CalcTangentSpace | aiProcess_JoinIdenticalVertices |
                                         aiProcess_GenUVCoords | aiProcess_GenNormals | aiProcess_GlobalScale |
                                         aiProcess_OptimizeMeshes | aiProcess_SortByPType | aiProcess_ValidateDataStructure;

struct Sub {
    unsigned base_vertex, base_index, vertex_count, index_count, material;
    std::string mesh_name, node_name;
    float local[16];  // glm::mat4 memory (column-major), Utils::Mat4FromAIMatrix4
    float mn[3], mx[3];
    bool placed = false;
};

static void traverse(const aiNode* n, std::vector<Sub>& subs)
{
    for (unsigned i = 0; i < n->mNumMeshes; ++i) {
        Sub& s = subs[n->mMeshes[i]];
        const aiMatrix4x4& m = n->mTransformation;  // row-major a1..d4; glm column c row r = m[r][c]
        const float rows[4][4] = {{m.a1, m.a2, m.a3, m.a4}, {m.b1, m.b2, m.b3, m.b4}, {m.c1, m.c2, m.c3, m.c4}, {m.d1, m.d2, m.d3, m.d4}};
        for (int c = 0; c < 4; ++c)
            for (int r = 0; r < 4; ++r) s.local[c * 4 + r] = rows[r][c];
        s.node_name = n->mName.C_Str();
        s.placed = true;
    }
    for (unsigned i = 0; i < n->mNumChildren; ++i) traverse(n->mChildren[i], subs);
}

int main(int argc, char** argv)
{
    Assimp::Importer importer;
    std::printf("{");
    for (int f = 1; f < argc; ++f) {
        const aiScene* scene = importer.ReadFile(argv[f], kMeshImportFlags);
        if (!scene || !scene->HasMeshes()) {
            std::fprintf(stderr, "ref_assimp: failed to load %s: %s\n", argv[f], importer.GetErrorString());
            return 1;
        }
        std::vector<Sub> subs(scene->mNumMeshes);
        unsigned vertex_count = 0, index_count = 0;
        for (unsigned m = 0; m < scene->mNumMeshes; ++m) {
            const aiMesh* mesh = scene->mMeshes[m];
            Sub& s = subs[m];
            s.base_vertex = vertex_count;
            s.base_index = index_count;
            s.vertex_count = mesh->mNumVertices;
            s.index_count = mesh->mNumFaces * 3;
            s.mesh_name = mesh->mName.C_Str();
            s.material = mesh->mMaterialIndex;
            for (int k = 0; k < 16; ++k) s.local[k] = (k % 5 == 0) ? 1.0f : 0.0f;  // Submesh::LocalTransform{1.0f} when no node names it
            for (int k = 0; k < 3; ++k) { s.mn[k] = FLT_MAX; s.mx[k] = -FLT_MAX; }
            for (unsigned i = 0; i < mesh->mNumVertices; ++i) {
                const float p[3] = {mesh->mVertices[i].x, mesh->mVertices[i].y, mesh->mVertices[i].z};
                for (int k = 0; k < 3; ++k) { s.mn[k] = p[k] < s.mn[k] ? p[k] : s.mn[k]; s.mx[k] = p[k] > s.mx[k] ? p[k] : s.mx[k]; }
            }
            vertex_count += mesh->mNumVertices;
            index_count += s.index_count;
        }
        traverse(scene->mRootNode, subs);
        std::printf("%s\n \"%s\": [", f > 1 ? "," : "", argv[f]);
        for (size_t m = 0; m < subs.size(); ++m) {
            const Sub& s = subs[m];
            std::printf("%s\n  {\"BaseVertex\": %u, \"BaseIndex\": %u, \"VertexCount\": %u, \"IndexCount\": %u, \"MaterialIndex\": %u, "
                        "\"MeshName\": \"%s\", \"NodeName\": \"%s\", \"LocalTransform\": [",
                        m ? "," : "", s.base_vertex, s.base_index, s.vertex_count, s.index_count, s.material, s.mesh_name.c_str(),
                        s.node_name.c_str());
            for (int k = 0; k < 16; ++k) std::printf("%s%.9g", k ? ", " : "", s.local[k]);
            std::printf("], \"BoundsMin\": [%.9g, %.9g, %.9g], \"BoundsMax\": [%.9g, %.9g, %.9g]}", s.mn[0], s.mn[1], s.mn[2], s.mx[0],
                        s.mx[1], s.mx[2]);
        }
        std::printf("\n ]");
    }
    std::printf("\n}\n");
    return 0;
}
