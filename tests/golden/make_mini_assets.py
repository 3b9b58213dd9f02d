"""Writes tests/golden/mini_assets/ — a small asset directory in the reference's on-disk formats (AssetRegistry.vkr,
.vkscene, .vkmesh, .vkmat, one binary FBX 7.4) authored for the loader tests, so that they run where /root/reference
does not exist (the GPU box) — and the two fixtures derived from it the same way the reference's files are
(mini_vkscenes.json, mini_vkmeshes.json).      python tests/golden/make_mini_assets.py

The scene exercises what the bundled scenes do not: an entity saved without a TransformComponent, a mesh whose
source file is missing, a mesh handle the registry does not list, an absolute registry path from another machine,
a material table naming an unknown handle, a displacement handle that is not a valid asset, quoted scalars,
a child model, polygons of 2/3/4/5 corners, per-polygon material slots, zlib and raw arrays."""
import os
import struct
import zlib

import make_vkmesh_fixtures as MM
import make_vkscene_fixtures as MS

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "mini_assets")


# ---- binary FBX writer (the record layout MeshImporter.cpp documents) ----------------------------------
def prop(v):
    if isinstance(v, bool):
        return b"C" + struct.pack("<b", int(v))
    if isinstance(v, int):
        return b"L" + struct.pack("<q", v) if abs(v) > 0x7FFFFFFF else b"I" + struct.pack("<i", v)
    if isinstance(v, float):
        return b"D" + struct.pack("<d", v)
    if isinstance(v, (bytes, str)):
        b = v.encode() if isinstance(v, str) else v
        return b"S" + struct.pack("<I", len(b)) + b
    kind, values, compress = v  # ("d" | "i", list, bool)
    raw = struct.pack("<%d%s" % (len(values), kind), *values)
    data = zlib.compress(raw) if compress else raw
    return kind.encode() + struct.pack("<III", len(values), 1 if compress else 0, len(data)) + data


def node(name, props=(), kids=()):
    return (name, list(props), list(kids))


def emit(n, offset):
    name, props, kids = n
    pb = b"".join(prop(p) for p in props)
    head = 12 + 1 + len(name)
    body, at = b"", offset + head + len(pb)
    for k in kids:
        kb = emit(k, at)
        body += kb
        at += len(kb)
    if kids:
        body += b"\0" * 13
        at += 13
    return struct.pack("<III", at, len(props), len(pb)) + bytes([len(name)]) + name.encode() + pb + body


def P(name, *vals):
    return node("P", [name, name, "", "A", *vals])


def model(mid, name, kind, t=None, r=None, s=None):
    ps = []
    if t:
        ps.append(P("Lcl Translation", *t))
    if r:
        ps.append(P("Lcl Rotation", *r))
    if s:
        ps.append(P("Lcl Scaling", *s))
    return node("Model", [mid, name + "\0\1Model", kind], [node("Version", [232]), node("Properties70", [], ps)])


def geometry(gid, name, verts, polys, mapping, mats, compress):
    pvi = []
    for poly in polys:
        pvi += list(poly[:-1]) + [~poly[-1]]
    return node("Geometry", [gid, name + "\0\1Geometry", "Mesh"], [
        node("Vertices", [("d", [float(x) for v in verts for x in v], compress)]),
        node("PolygonVertexIndex", [("i", pvi, not compress)]),
        node("LayerElementNormal", [0], [node("Normals", [("d", [0.0, 0.0, 1.0] * len(pvi), True)])]),  # skipped by the reader
        node("LayerElementMaterial", [0], [node("Version", [101]), node("Name", [""]),
                                           node("MappingInformationType", [mapping]),
                                           node("ReferenceInformationType", ["IndexToDirect"]),
                                           node("Materials", [("i", mats, False)])])])


def write_fbx(path):
    va = [(-1, -1, 0), (1, -1, 0), (1, 1, 0), (-1, 1, 0), (0, 2, 3)]
    vb = [(0, 0, 0), (2, 0, 0), (3, 1, 0.5), (2, 2, 0), (0, 2, -0.5), (9, 9, 9)]   # vertex 5 only used by the 2-corner "polygon"
    top = [
        node("FBXHeaderExtension", [], [node("FBXVersion", [7400])]),
        node("GlobalSettings", [], [node("Version", [1000]), node("Properties70", [], [
            node("P", ["UpAxis", "int", "Integer", "", 1]), node("P", ["UnitScaleFactor", "double", "Number", "", 2.0])])]),
        node("Objects", [], [
            geometry(1001, "GA", va, [(0, 1, 2, 3), (2, 3, 4)], "ByPolygon", [0, 1], True),
            geometry(1002, "GB", vb, [(0, 1, 2, 3, 4), (4, 5)], "AllSame", [0], False),
            model(2001, "A", "Mesh", t=(100.0, 0.0, 50.0), r=(0.0, 90.0, 0.0), s=(2.0, 2.0, 2.0)),
            model(2002, "B", "Mesh", t=(0.0, 10.0, 0.0), r=(-90.0, 0.0, 30.0)),
            model(2003, "Lamp", "Light", t=(1.0, 2.0, 3.0)),
            node("Material", [3001, "M1\0\1Material", ""]),
            node("Material", [3002, "M2\0\1Material", ""])]),
        node("Connections", [], [
            node("C", ["OO", 2001, 0]), node("C", ["OO", 2003, 0]), node("C", ["OO", 2002, 2001]),
            node("C", ["OO", 1001, 2001]), node("C", ["OO", 3002, 2001]), node("C", ["OO", 3001, 2001]),
            node("C", ["OO", 1002, 2002]), node("C", ["OO", 3001, 2002])]),
    ]
    data = b"Kaydara FBX Binary  \0\x1a\0" + struct.pack("<I", 7400)
    for n in top:
        data += emit(n, len(data))
    data += b"\0" * 13 + b"footer-bytes-the-reader-never-looks-at"
    open(path, "wb").write(data)


REGISTRY = """Asset Registry:
  - Handle: 101
    Metadata:
      Filepath: meshes/A.vkmesh
      Type: Mesh
  - Handle: 102
    Metadata:
      Filepath: meshes/Missing.vkmesh
      Type: Mesh
  - Handle: 103
    Metadata:
      Filepath: D:/Somewhere/Else/assets/meshes/Abs.vkmesh
      Type: Mesh
  - Handle: 201
    Metadata:
      Filepath: meshes/Source/mini.fbx
      Type: MeshAsset
  - Handle: 202
    Metadata:
      Filepath: meshes/Source/not_bundled.fbx
      Type: MeshAsset
  - Handle: 18446744073709551000
    Metadata:
      Filepath: materials/Plain.vkmat
      Type: Material
  - Handle: 302
    Metadata:
      Filepath: materials/Displaced.vkmat
      Type: Material
  - Handle: 303
    Metadata:
      Filepath: materials/DanglingDisplacement.vkmat
      Type: Material
  - Handle: 401
    Metadata:
      Filepath: textures/height.png
      Type: Texture2D
"""

MATERIAL = """Material:
  Albedo: [0.5, 0.25, 1, 1]
  Emission: 1
  Metallic: 0
  Roughness: 0.5
  UseNormalMap: 0
  AlbedoHandle: 0
  NormalHandle: 0
  ARMHandle: 0
"""

SCENE = """Scene: Mini
Entities:
  - Entity: 11
    TagComponent:
      Tag: Sky
    SkyLightComponent:
      TextureHandle: 401
  - Entity: 12
    TagComponent:
      Tag: "Lamp: big"
    TransformComponent:
      Translation: [1.5, -2.25, 3]
      Rotation: [0.100000001, 0.200000003, -0.300000012]
      Scale: [1, 2, 0.5]
    MeshComponent:
      MeshHandle: 101
      MaterialTable:
        0: 18446744073709551000
        1: 302
  - Entity: 13
    TagComponent:
      Tag: second
    TransformComponent:
      Translation: [-4, 0, 1e-3]
      Rotation: [3.14159274, -0.797807634, 3.14159274]
      Scale: [1, 1, 1]
    MeshComponent:
      MeshHandle: 101
      MaterialTable:
        0: 303
  - Entity: 14
    TagComponent:
      Tag: source not bundled
    TransformComponent:
      Translation: [0, 0, 0]
      Rotation: [0, 0, 0]
      Scale: [1, 1, 1]
    MeshComponent:
      MeshHandle: 102
      MaterialTable:
        0: 302
  - Entity: 15
    TagComponent:
      Tag: absolute registry path
    TransformComponent:
      Translation: [2, 2, -7.5]
      Rotation: [0, 1.57079637, 0]
      Scale: [3, 3, 3]
    MeshComponent:
      MeshHandle: 103
      MaterialTable:
        0: 302
        1: 777
  - Entity: 16
    TagComponent:
      Tag: Sun
    TransformComponent:
      Translation: [0, 10, 0]
      Rotation: [0, 0, 0]
      Scale: [1, 1, 1]
    DirectionalLightComponent:
      Direction: [0.5, -1, 0.25]
      Color: [1, 1, 1, 2]
      Falloff: 0
  - Entity: 17
    TagComponent:
      Tag: handle the registry does not list
    TransformComponent:
      Translation: [0, 0, 0]
      Rotation: [0, 0, 0]
      Scale: [1, 1, 1]
    MeshComponent:
      MeshHandle: 555
      MaterialTable:
        {}
"""


def main():
    for d in ("", "meshes", "meshes/Source", "materials", "scenes", "textures"):
        os.makedirs(os.path.join(OUT, d), exist_ok=True)
    write_fbx(os.path.join(OUT, "meshes/Source/mini.fbx"))
    open(os.path.join(OUT, "AssetRegistry.vkr"), "w").write(REGISTRY)
    open(os.path.join(OUT, "meshes/A.vkmesh"), "w").write("Mesh:\n  MeshAsset: 201")
    open(os.path.join(OUT, "meshes/Missing.vkmesh"), "w").write("Mesh:\n  MeshAsset: 202")
    open(os.path.join(OUT, "meshes/Abs.vkmesh"), "w").write("Mesh:\n  MeshAsset: 201\n")
    open(os.path.join(OUT, "materials/Plain.vkmat"), "w").write(MATERIAL)
    open(os.path.join(OUT, "materials/Displaced.vkmat"), "w").write(MATERIAL + "  DisplacementHandle: 401\n")
    open(os.path.join(OUT, "materials/DanglingDisplacement.vkmat"), "w").write(MATERIAL + "  DisplacementHandle: 999\n")
    open(os.path.join(OUT, "textures/height.png"), "wb").write(b"")
    open(os.path.join(OUT, "scenes/Mini.vkscene"), "w").write(SCENE)
    MS.main(OUT, "mini_vkscenes.json", "tests/golden/mini_assets (authored by make_mini_assets.py)")
    MM.main(OUT, "mini_vkmeshes.json")


if __name__ == "__main__":
    main()
