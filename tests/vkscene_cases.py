"""configs[0]: the reference's six bundled scenes (tests/golden/vkscenes.json, extracted from
VulkanCore/assets/scenes/*.vkscene by tests/golden/make_vkscene_fixtures.py) as cases in the format of
golden_cases.make_case.  Meshes carry one stand-in submesh with a unit box (Assimp import is unbuildable
here); entity order, transforms, mesh handles, material tables and tessellation flags are the reference's."""
import json
import os
from types import SimpleNamespace

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SCENES = ["ClosedSpace", "FirstScene", "LampDemo", "PhysicsDemo", "ShadowDemo", "TessellationDemo"]


def load(prefix=""):
    return json.load(open(os.path.join(HERE, "golden", prefix + "vkscenes.json")))


def load_meshes(prefix=""):
    return json.load(open(os.path.join(HERE, "golden", prefix + "vkmeshes.json")))


def vkscene_case(name, real_meshes=False, prefix=""):
    """real_meshes=False: every entity of the file, one stand-in unit-box submesh per mesh (the round-1 cases).
    real_meshes=True: what the headless loader (host/SceneSerializer.cpp) builds — only entities saved with a
    TransformComponent get one, submesh tables come from the FBX sources (tests/golden/vkmeshes.json), a mesh whose
    source file is not bundled resolves to a null asset, and a mesh source's base material is slot 0 of the last
    entity deserialized with it (SceneSerializer.cpp:437-439)."""
    doc = load(prefix)
    ents = doc["scenes"][name]
    if real_meshes:
        return _real_case(doc, [e for e in ents if e["has_transform"]], load_meshes(prefix))
    n = len(ents)
    trs = np.array([e["translation"] + e["rotation"] + e["scale"] for e in ents], np.float32)
    mesh_handles = sorted({int(e["mesh"]) for e in ents if e["mesh"]})
    mesh_index = {h: i for i, h in enumerate(mesh_handles)}
    mat_handles = sorted({int(h) for e in ents for h in e["materials"].values()} | {1})  # 1 = stand-in base material
    tess = {int(h): bool(m["tessellated"]) for h, m in doc["materials"].items()}
    has_mesh = np.array([1 if e["mesh"] else 0 for e in ents], np.uint8)
    mesh = np.array([mesh_index[int(e["mesh"])] if e["mesh"] else -1 for e in ents], np.int32)
    # one material table per mesh entity
    table = np.full(n, -1, np.int32)
    table_begin, entry_index, entry_handle, entry_tess = [0], [], [], []
    t = 0
    for i, e in enumerate(ents):
        if not e["mesh"]:
            continue
        table[i] = t
        for idx, h in sorted(e["materials"].items(), key=lambda kv: int(kv[0])):
            entry_index.append(int(idx))
            entry_handle.append(int(h))
            entry_tess.append(1 if tess.get(int(h), False) else 0)
        table_begin.append(len(entry_index))
        t += 1
    m = len(mesh_handles)
    ident = np.eye(4, dtype=np.float32).reshape(16)
    light = next((e["light_direction"] for e in ents if "light_direction" in e), [1.0, -1.0, 1.0])
    return SimpleNamespace(
        n_entities=n, trs=trs, parent=None, level_offsets=np.array([0, n], np.uint32), mesh=mesh, table=table,
        has_mesh=has_mesh, mesh_handle=np.array(mesh_handles, np.uint64),
        mesh_vertex_count=np.full(m, 24, np.uint32), mesh_sub_begin=np.arange(m + 1, dtype=np.uint32),
        base_material_handle=np.full(m, 1, np.uint64), base_material_tess=np.zeros(m, np.uint8),
        sub_material_index=np.zeros(m, np.int32), sub_local=np.tile(ident, (m, 1)),
        sub_aabb=np.tile(np.array([-1, -1, -1, 1, 1, 1], np.float32), (m, 1)),
        table_begin=np.array(table_begin, np.uint32), entry_index=np.array(entry_index, np.uint32),
        entry_handle=np.array(entry_handle, np.uint64), entry_tess=np.array(entry_tess, np.uint8),
        material_handle=np.array(mat_handles, np.uint64),
        material_tess=np.array([tess.get(h, False) for h in mat_handles], bool), light_direction=light)


def _real_case(doc, ents, md):
    n = len(ents)
    trs = np.array([e["translation"] + e["rotation"] + e["scale"] for e in ents], np.float32)
    tess = {int(h): bool(m["tessellated"]) for h, m in doc["materials"].items()}
    usable = sorted({int(e["mesh"]) for e in ents if e["mesh"] and md["meshes"].get(e["mesh"], {}).get("source")})
    mesh_index = {h: i for i, h in enumerate(usable)}
    has_mesh = np.array([1 if e["mesh"] else 0 for e in ents], np.uint8)
    mesh = np.array([mesh_index.get(int(e["mesh"]), -1) if e["mesh"] else -1 for e in ents], np.int32)
    subs = [md["sources"][md["meshes"][str(h)]["source"]] for h in usable]
    sub_begin = np.concatenate([[0], np.cumsum([len(x) for x in subs])]).astype(np.uint32)
    flat = [sm for x in subs for sm in x]
    base_handle, base_tess = np.zeros(len(usable), np.uint64), np.zeros(len(usable), np.uint8)
    table = np.full(n, -1, np.int32)
    table_begin, entry_index, entry_handle, entry_tess = [0], [], [], []
    mats = set()
    t = 0
    for i, e in enumerate(ents):
        if not e["mesh"]:
            continue
        table[i] = t
        for idx, h in sorted(e["materials"].items(), key=lambda kv: int(kv[0])):
            known = h in doc["materials"]          # a handle the registry does not list resolves to a null material
            entry_index.append(int(idx))
            entry_handle.append(int(h) if known else 0)
            entry_tess.append(1 if known and tess.get(int(h), False) else 0)
            if known:
                mats.add(int(h))
        table_begin.append(len(entry_index))
        t += 1
        if mesh[i] >= 0:
            # the base material lives on the mesh SOURCE, which meshes naming the same source file share
            h0 = e["materials"].get("0")
            ok = h0 is not None and h0 in doc["materials"]
            src = md["meshes"][e["mesh"]]["source"]
            for k, h in enumerate(usable):
                if md["meshes"][str(h)]["source"] == src:
                    base_handle[k] = int(h0) if ok else 0
                    base_tess[k] = 1 if ok and tess.get(int(h0), False) else 0
    mat_handles = sorted(mats)
    return SimpleNamespace(
        n_entities=n, trs=trs, parent=None, level_offsets=np.array([0, n], np.uint32), mesh=mesh, table=table,
        has_mesh=has_mesh, mesh_handle=np.array(usable, np.uint64),
        mesh_vertex_count=np.array([sum(sm["VertexCount"] for sm in x) for x in subs], np.uint32), mesh_sub_begin=sub_begin,
        base_material_handle=base_handle, base_material_tess=base_tess,
        sub_material_index=np.array([sm["MaterialIndex"] for sm in flat], np.int32),
        sub_local=np.array([sm["LocalTransform"] for sm in flat], np.float32).reshape(-1, 16),
        sub_aabb=np.array([sm["BoundsMin"] + sm["BoundsMax"] for sm in flat], np.float32).reshape(-1, 6),
        sub_counts=np.array([[sm["BaseVertex"], sm["BaseIndex"], sm["VertexCount"], sm["IndexCount"]] for sm in flat], np.uint32).reshape(-1, 4),
        table_begin=np.array(table_begin, np.uint32), entry_index=np.array(entry_index, np.uint32),
        entry_handle=np.array(entry_handle, np.uint64), entry_tess=np.array(entry_tess, np.uint8),
        material_handle=np.array(mat_handles, np.uint64),
        material_tess=np.array([tess.get(h, False) for h in mat_handles], bool),
        light_direction=next((e["light_direction"] for e in ents if "light_direction" in e), [1.0, -1.0, 1.0]),
        uuid=[int(e["uuid"]) for e in ents])
