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


def load():
    return json.load(open(os.path.join(HERE, "golden", "vkscenes.json")))


def vkscene_case(name):
    doc = load()
    ents = doc["scenes"][name]
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
