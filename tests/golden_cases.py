"""Scenes in the reference's own terms (entities with TransformComponent + MeshComponent, meshes with
submeshes, material tables with null padding) for pinning the oracle against real entt + GLM, plus
the Python statement of the host adapter's flattening rule: walk the view in entt
order, resolve materials like SubmitMesh, emit one instance per (entity, submesh)."""
import ctypes as C
from types import SimpleNamespace

import numpy as np

import oracle_lib as O
from vulkanengine_b200 import synth


def _rand_local(rng):
    t = rng.uniform(-1, 1, 3)
    e = rng.uniform(-np.pi, np.pi, 3)
    s = rng.uniform(0.7, 1.4, 3)
    return synth._trs_matrix(t, e, s)


def make_case(seed, n_entities, n_meshes, n_materials, n_tables, hierarchy=False, frac_no_mesh=0.1):
    rng = np.random.default_rng(seed)
    trs = np.concatenate([rng.uniform(-50, 50, (n_entities, 3)), rng.uniform(-np.pi, np.pi, (n_entities, 3)),
                          rng.uniform(0.5, 2, (n_entities, 3))], axis=1).astype(np.float32)
    parent = None
    if hierarchy:  # parent[i] < i, level-ordered by construction below
        depth = np.zeros(n_entities, np.int32)
        parent = np.full(n_entities, -1, np.int32)
        n_roots = max(1, n_entities // 6)
        for i in range(n_roots, n_entities):
            parent[i] = rng.integers(0, i)
            depth[i] = depth[parent[i]] + 1
        order = np.argsort(depth, kind="stable")  # level order, parents first
        inv = np.empty_like(order)
        inv[order] = np.arange(n_entities)
        trs = trs[order]
        parent = np.where(parent[order] >= 0, inv[np.maximum(parent[order], 0)], -1).astype(np.int32)
        depth = depth[order]
        level_offsets = np.concatenate([[0], np.cumsum(np.bincount(depth))]).astype(np.uint32)
    else:
        level_offsets = np.array([0, n_entities], np.uint32)

    mesh_handle = np.unique(rng.integers(1, 2**63, n_meshes * 2, dtype=np.uint64))[:n_meshes]
    rng.shuffle(mesh_handle)  # handle order != creation order, so the rank mapping is exercised
    material_handle = np.unique(rng.integers(1, 2**63, n_materials * 2, dtype=np.uint64))[:n_materials]
    rng.shuffle(material_handle)
    material_tess = rng.random(n_materials) < 0.2
    sub_count = rng.integers(1, 5, n_meshes)
    mesh_sub_begin = np.concatenate([[0], np.cumsum(sub_count)]).astype(np.uint32)
    n_sub = int(mesh_sub_begin[-1])
    sub_material_index = rng.integers(0, 4, n_sub).astype(np.int32)
    sub_local = np.stack([np.eye(4, dtype=np.float32).reshape(16) if rng.random() < 0.3 else _rand_local(rng)
                          for _ in range(n_sub)]).astype(np.float32)
    half = rng.uniform(0.25, 2, (n_sub, 3))
    ctr = rng.uniform(-0.5, 0.5, (n_sub, 3))
    sub_aabb = np.concatenate([ctr - half, ctr + half], axis=1).astype(np.float32)
    mesh_vertex_count = rng.integers(3, 5000, n_meshes).astype(np.uint32)
    mesh_vertex_count[rng.integers(0, n_meshes)] = 0  # one zero-vertex mesh: SceneRenderer.cpp:1135-1136
    base_mat = rng.integers(0, n_materials, n_meshes)

    # material tables: indices 0..3, some entries null (SceneSerializer.cpp:414-434 pads with nullptr)
    table_begin, entry_index, entry_handle, entry_tess = [0], [], [], []
    for _ in range(n_tables):
        for idx in range(4):
            if rng.random() < 0.8:
                entry_index.append(idx)
                if rng.random() < 0.25:
                    entry_handle.append(0)
                    entry_tess.append(0)
                else:
                    m = rng.integers(0, n_materials)
                    entry_handle.append(int(material_handle[m]))
                    entry_tess.append(int(material_tess[m]))
        table_begin.append(len(entry_index))

    mesh = rng.integers(0, n_meshes, n_entities).astype(np.int32)
    table = rng.integers(0, n_tables, n_entities).astype(np.int32)
    mesh[rng.random(n_entities) < 0.05] = -1   # null mesh asset
    table[rng.random(n_entities) < 0.05] = -1  # null material table
    has_mesh = (rng.random(n_entities) >= frac_no_mesh).astype(np.uint8)
    return SimpleNamespace(
        n_entities=n_entities, trs=trs, parent=parent, level_offsets=level_offsets, mesh=mesh, table=table,
        has_mesh=has_mesh, mesh_handle=mesh_handle, mesh_vertex_count=mesh_vertex_count, mesh_sub_begin=mesh_sub_begin,
        base_material_handle=material_handle[base_mat].copy(), base_material_tess=material_tess[base_mat].astype(np.uint8),
        sub_material_index=sub_material_index, sub_local=sub_local, sub_aabb=sub_aabb,
        table_begin=np.array(table_begin, np.uint32), entry_index=np.array(entry_index, np.uint32),
        entry_handle=np.array(entry_handle, np.uint64), entry_tess=np.array(entry_tess, np.uint8),
        material_handle=material_handle, material_tess=material_tess)


def entt_scene_cases():
    return {"small": make_case(1, 10, 3, 4, 2, frac_no_mesh=0.2),
            "medium": make_case(2, 400, 12, 9, 6),
            "hier": make_case(3, 300, 8, 6, 4, hierarchy=True)}


class _RefScene(C.Structure):
    _fields_ = [("n_entities", C.c_uint32), ("trs", C.c_void_p), ("parent", C.c_void_p), ("mesh", C.c_void_p),
                ("table", C.c_void_p), ("has_mesh_component", C.c_void_p), ("n_meshes", C.c_uint32),
                ("mesh_handle", C.c_void_p), ("mesh_vertex_count", C.c_void_p), ("mesh_sub_begin", C.c_void_p),
                ("base_material_handle", C.c_void_p), ("base_material_tess", C.c_void_p),
                ("sub_material_index", C.c_void_p), ("sub_local", C.c_void_p), ("sub_aabb", C.c_void_p),
                ("n_tables", C.c_uint32), ("table_begin", C.c_void_p), ("entry_index", C.c_void_p),
                ("entry_handle", C.c_void_p), ("entry_tess", C.c_void_p), ("inst_aabb", C.c_void_p),
                ("trig_mode", C.c_int)]


def run_ref_pass(case, trig_mode=0):
    """One submission pass through real entt + glm + std::map<MeshKey> (oracle/ref_glm_entt.cpp)."""
    p = lambda a: None if a is None else np.ascontiguousarray(a).ctypes.data_as(C.c_void_p)
    keep = [np.ascontiguousarray(x) if x is not None else None for x in
            (case.trs, case.parent, case.mesh, case.table, case.has_mesh, case.mesh_handle, case.mesh_vertex_count,
             case.mesh_sub_begin, case.base_material_handle, case.base_material_tess, case.sub_material_index,
             case.sub_local, case.sub_aabb, case.table_begin, case.entry_index, case.entry_handle, case.entry_tess)]
    sc = _RefScene(case.n_entities, p(keep[0]), p(keep[1]), p(keep[2]), p(keep[3]), p(keep[4]), len(case.mesh_handle),
                   p(keep[5]), p(keep[6]), p(keep[7]), p(keep[8]), p(keep[9]), p(keep[10]), p(keep[11]), p(keep[12]),
                   len(case.table_begin) - 1, p(keep[13]), p(keep[14]), p(keep[15]), p(keep[16]), None, trig_mode)
    cap = case.n_entities * 4 + 4
    g_mesh, g_mat = np.zeros(cap, np.uint64), np.zeros(cap, np.uint64)
    g_sub, g_cnt = np.zeros(cap, np.uint32), np.zeros(cap, np.uint32)
    g_tess = np.zeros(cap, np.uint8)
    rows = np.zeros((cap, 12), np.float32)
    row_inst = np.zeros(cap, np.uint32)
    world = np.zeros((cap, 12), np.float32)
    inst_ent, inst_sub = np.zeros(cap, np.uint32), np.zeros(cap, np.uint32)
    n_rows, n_inst = C.c_uint32(), C.c_uint32()
    secs = C.c_double()
    ng = O.ref().ssr_submit_pass(C.byref(sc), None, C.c_int(-1), p(g_mesh), p(g_mat), p(g_sub), p(g_tess), p(g_cnt),
                                 p(rows), p(row_inst), C.byref(n_rows), p(world), C.byref(n_inst), p(inst_ent),
                                 p(inst_sub), C.byref(secs))
    ni, nr = n_inst.value, n_rows.value
    return {"g_mesh": g_mesh[:ng], "g_mat": g_mat[:ng], "g_sub": g_sub[:ng], "g_tess": g_tess[:ng], "g_cnt": g_cnt[:ng],
            "rows": rows[:nr], "row_inst": row_inst[:nr], "world": world[:ni], "inst_ent": inst_ent[:ni],
            "inst_sub": inst_sub[:ni]}


def flatten_case(case):
    """The host adapter's rule, restated: entt view<Transform, Mesh> order = the smaller pool's packed
    array walked back to front (ENTT:22231-22241, 20264-20267) = reverse creation order here; per entity
    the early-outs of SubmitMesh (SceneRenderer.cpp:1128-1136) and its material resolution (:1140-1141)."""
    mat_sorted = np.sort(case.material_handle)
    mesh_sorted = np.sort(case.mesh_handle)
    tess_of = dict(zip(case.material_handle.tolist(), case.material_tess.tolist()))
    node, slot, matword = [], [], []
    for e in range(case.n_entities - 1, -1, -1):
        if not case.has_mesh[e]:
            continue
        m, t = int(case.mesh[e]), int(case.table[e])
        if m < 0 or t < 0 or case.mesh_vertex_count[m] < 1:
            continue
        for s in range(int(case.mesh_sub_begin[m]), int(case.mesh_sub_begin[m + 1])):
            handle, tess = int(case.base_material_handle[m]), int(case.base_material_tess[m])
            for k in range(int(case.table_begin[t]), int(case.table_begin[t + 1])):
                if int(case.entry_index[k]) == int(case.sub_material_index[s]) and int(case.entry_handle[k]) != 0:
                    handle, tess = int(case.entry_handle[k]), int(case.entry_tess[k])
            if handle == 0:  # no table entry and no base material: left out (host/Scene.cpp resolve_material)
                continue
            rank = int(np.searchsorted(mat_sorted, np.uint64(handle)))
            node.append(e)
            slot.append(s)
            matword.append(rank | (synth.TESS_BIT if tess else 0))
    n_sub = int(case.mesh_sub_begin[-1])
    slots = np.zeros(n_sub, synth.SUBMESH_DTYPE)
    mesh_of_sub = np.repeat(np.arange(len(case.mesh_handle)), np.diff(case.mesh_sub_begin.astype(np.int64)))
    slots["local"] = case.sub_local
    slots["aabb_min"] = case.sub_aabb[:, :3]
    slots["aabb_max"] = case.sub_aabb[:, 3:]
    counts = getattr(case, "sub_counts", None)  # [s][BaseVertex, BaseIndex, VertexCount, IndexCount] when the case has real tables
    if counts is None:
        slots["index_count"] = 3 * (np.arange(n_sub) + 10)
        slots["first_index"] = 1000 * np.arange(n_sub)
        slots["vertex_offset"] = 77 * np.arange(n_sub)
    else:
        slots["index_count"] = counts[:, 3]
        slots["first_index"] = counts[:, 1]
        slots["vertex_offset"] = counts[:, 0]
    slots["mesh_rank"] = np.searchsorted(mesh_sorted, case.mesh_handle[mesh_of_sub])
    slots["submesh_index"] = np.arange(n_sub) - case.mesh_sub_begin[mesh_of_sub]
    sc = synth.SceneArrays(trs=case.trs, parent=case.parent, level_offsets=case.level_offsets,
                           node=np.array(node, np.uint32), node_base=0, slot=np.array(slot, np.uint32),
                           matword=np.array(matword, np.uint32), inst_aabb=None, slots=slots,
                           mesh_bits=synth._bits(len(mesh_sorted)), submesh_bits=3,
                           material_bits=synth._bits(len(mat_sorted)), mesh_handles=mesh_sorted,
                           material_handles=mat_sorted, seed=0, name="case")
    return sc


def all_visible_view():
    """A clip matrix every box passes: all six planes reduce to (0,0,0,1) or (0,0,0,0)."""
    v = np.zeros(1, synth.VIEW_DTYPE)
    v["view_proj"][0][15] = 1.0
    return v


def decode_key(sc, key):
    """key -> (mesh handle, submesh index, material handle, tess) through the rank tables."""
    mb, sb, hb = sc.material_bits, sc.submesh_bits, sc.mesh_bits
    mat = int(key) & ((1 << mb) - 1)
    sub = (int(key) >> mb) & ((1 << sb) - 1)
    mesh = (int(key) >> (mb + sb)) & ((1 << hb) - 1)
    tess = (int(key) >> (mb + sb + hb)) & 1
    return int(sc.mesh_handles[mesh]), sub, int(sc.material_handles[mat]), tess


def check_lists_against_ref_pass(sc, ref, idx, keys, rows, draws, draw_keys):
    """The per-view list of an all-visible view must BE the reference's grouped output."""
    assert np.array_equal(idx, ref["row_inst"]), "instance order inside/among groups differs from std::map order"
    assert np.array_equal(rows, ref["rows"]), "instance rows differ from glm's"
    assert len(draws) == len(ref["g_cnt"])
    assert np.array_equal(draws["instance_count"], ref["g_cnt"])
    for d in range(len(draws)):
        mesh, sub, mat, tess = decode_key(sc, draw_keys[d])
        assert (mesh, sub, mat, tess) == (int(ref["g_mesh"][d]), int(ref["g_sub"][d]), int(ref["g_mat"][d]),
                                          int(ref["g_tess"][d])), f"draw {d}: key differs from MeshKey"
    assert np.array_equal(draws["first_instance"], np.concatenate([[0], np.cumsum(ref["g_cnt"])[:-1]]))
