// Mesh source ingest for the headless host: what the submission path needs from a mesh file — the submesh table
// (counts, material slot, node-local transform) and, for the cull step the reference lacks, a local-space box.
//
// The reference gets these through Assimp (VulkanCore/src/VulkanCore/Mesh/Mesh.cpp:60-112 import flags and submesh
// loop, Asset/MeshImporter.cpp:160-188 TraverseNodes).  Assimp is vendored there as source behind a CMake build with
// generated headers: unbuildable here.  This reader opens the binary FBX (7.x) files themselves and restates the
// parts of Assimp's FBX path that decide those fields:
//   * one submesh per (Model of type Mesh, Geometry, material slot in use), in model connection order;
//   * IndexCount = 3 x triangles after fan triangulation (aiProcess_Triangulate), polygons with < 3 corners dropped
//     (aiProcess_SortByPType splits them off);
//   * vertices scaled by UnitScaleFactor x 0.01 and node translations likewise (FBXImporter.cpp:171-179 +
//     aiProcess_GlobalScale, PostProcessing/ScaleProcess.cpp:118-137);
//   * LocalTransform = the model's own T * Rz*Ry*Rx * S from Lcl Translation / Rotation / Scaling (the reference
//     keeps the node-LOCAL matrix, MeshImporter.cpp:165-171), valid for models without pivots / pre-rotation —
//     true of every bundled asset (all Blender exports).
// Not reproduced: VertexCount after aiProcess_JoinIdenticalVertices (depends on Assimp's generated normals and
// tangents); VertexCount here is the number of polygon corners, an upper bound, so BaseVertex of later submeshes
// can differ from the reference's.  Parity of this file's output is UNPINNED (no Assimp to compare with); the
// expected values in tests/golden/vkmeshes.json come from an independent Python restatement.
#pragma once
#include <memory>
#include <string>

#include "Scene.h"

namespace VulkanCore {

class MeshImporter {
public:
    // returns nullptr when the file is missing or not a binary FBX; `error` (optional) says why
    static std::shared_ptr<MeshSource> ImportAssimpMesh(const std::string& filepath, std::string* error = nullptr);
};

}  // namespace VulkanCore
