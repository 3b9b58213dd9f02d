"""The C++ host mirror (vulkanengine_b200/host: Scene / SceneRenderer with the reference's method names)
above the C ABI.  CPU part: its registry walks a view in entt order and its flattening equals the rule
pinned against real entt (golden vectors).  GPU part: RenderScene() through the mirror equals the oracle."""
import os
import subprocess

import numpy as np
import pytest

import golden_cases as G
import oracle_lib as O
from vulkanengine_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "vulkanengine_b200", "host", "host_cli.bin")
CAMERA = np.array([synth.CAMERA_FOV, synth.CAMERA_ASPECT, 0.1, 1000.0, 0.0, 0.0, 120.0, 0.1, -0.2], np.float32)


def write_pack(case, path):
    blocks = [case.trs.astype(np.float32), np.zeros(0, np.int32) if case.parent is None else case.parent.astype(np.int32),
              case.mesh.astype(np.int32), case.table.astype(np.int32), case.has_mesh.astype(np.uint8),
              case.mesh_handle.astype(np.uint64), case.mesh_vertex_count.astype(np.uint32),
              case.mesh_sub_begin.astype(np.uint32), case.base_material_handle.astype(np.uint64),
              case.base_material_tess.astype(np.uint8), case.sub_material_index.astype(np.int32),
              case.sub_local.astype(np.float32), case.sub_aabb.astype(np.float32), case.table_begin.astype(np.uint32),
              case.entry_index.astype(np.uint32), case.entry_handle.astype(np.uint64), case.entry_tess.astype(np.uint8),
              CAMERA]
    if getattr(case, "sub_counts", None) is not None:
        blocks.append(case.sub_counts.astype(np.uint32))
    with open(path, "wb") as f:
        for b in blocks:
            b = np.ascontiguousarray(b)
            f.write(np.uint64(b.nbytes).tobytes())
            f.write(b.tobytes())


def read_blocks(path):
    raw = open(path, "rb").read()
    out, off = [], 0
    while off < len(raw):
        n = int(np.frombuffer(raw, np.uint64, 1, off)[0])
        out.append(raw[off + 8: off + 8 + n])
        off += 8 + n
    return out


@pytest.mark.parametrize("name", ["small", "medium", "hier"])
def test_mirror_flattens_in_entt_order(name, tmp_path):
    case = G.entt_scene_cases()[name]
    pack, out = str(tmp_path / "pack.bin"), str(tmp_path / "flat.bin")
    write_pack(case, pack)
    subprocess.check_call([CLI, "flatten", pack, out])
    b = read_blocks(out)
    node, slot, mat, ent = (np.frombuffer(x, np.uint32) for x in b[:4])
    want = G.flatten_case(case)
    g = np.load(os.path.join(ROOT, "tests", "golden", "entt_pass.npz"))
    assert np.array_equal(ent, g[f"{name}__inst_ent"]), "mirror's view order differs from real entt"
    assert np.array_equal(node, want.node)
    table = np.frombuffer(b[6], synth.SUBMESH_DTYPE)
    mesh_handles, material_handles = np.frombuffer(b[7], np.uint64), np.frombuffer(b[8], np.uint64)
    # the mirror ranks only the handles in use and numbers slots per mesh in handle order:
    # compare what a slot / rank MEANS, not its number
    mine, theirs = table[slot], want.slots[want.slot]
    for f in ("local", "aabb_min", "aabb_max", "index_count", "first_index", "vertex_offset", "submesh_index"):
        assert np.array_equal(mine[f], theirs[f]), f
    assert np.array_equal(mesh_handles[mine["mesh_rank"]], want.mesh_handles[theirs["mesh_rank"]])
    tess = synth.TESS_BIT
    assert np.array_equal(material_handles[mat & ~np.uint32(tess)], want.material_handles[want.matword & ~np.uint32(tess)])
    assert np.array_equal(mat & np.uint32(tess), want.matword & np.uint32(tess))
    assert np.all(np.diff(mesh_handles.astype(np.float64)) > 0) and np.all(np.diff(material_handles.astype(np.float64)) > 0)
    if case.parent is not None:
        assert np.array_equal(np.frombuffer(b[4], np.int32), case.parent)
        assert np.array_equal(np.frombuffer(b[5], np.uint32), case.level_offsets)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["small", "medium", "hier"])
def test_mirror_render_scene_matches_oracle(name, tmp_path):
    case = G.entt_scene_cases()[name]
    pack, out = str(tmp_path / "pack.bin"), str(tmp_path / "render.bin")
    write_pack(case, pack)
    subprocess.check_call([CLI, "render", pack, out])
    b = read_blocks(out)
    visible, draws = np.frombuffer(b[0], np.uint32), np.frombuffer(b[1], np.uint32)
    sc = G.flatten_case(case)
    cam = lambda *a: O.camera_view_proj(*a)
    views = synth.make_views(cam, lambda *a: O.cascade_matrices(*a), 5, position=CAMERA[4:7], pitch=float(CAMERA[7]),
                             yaw=float(CAMERA[8]))
    ref = O.OracleScene(sc, 0).frame(views)
    for v in range(5):
        r = ref["views"][v]
        rows = np.frombuffer(b[2 + 3 * v], np.float32).reshape(-1, 12)
        ents = np.frombuffer(b[3 + 3 * v], np.uint32)
        dl = np.frombuffer(b[4 + 3 * v], np.uint64).reshape(-1, 9)
        assert int(visible[v]) == len(r["idx"]) and int(draws[v]) == len(r["draws"])
        assert np.array_equal(rows, r["rows"])
        assert np.array_equal(ents, sc.node[r["idx"]])  # EntityID per row (SelectTransformData)
        for d in range(len(dl)):
            mesh, sub, mat, tess = G.decode_key(sc, r["draw_keys"][d])
            assert (int(dl[d, 0]), int(dl[d, 1]), int(dl[d, 2]), int(dl[d, 3])) == (mesh, mat, sub, tess)
            assert int(dl[d, 5]) == int(r["draws"]["instance_count"][d]) and int(dl[d, 8]) == int(r["draws"]["first_instance"][d])
            assert int(dl[d, 4]) == int(r["draws"]["index_count"][d]) and int(dl[d, 6]) == int(r["draws"]["first_index"][d])


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["", "resident"])
def test_merge_from_a_cpp_host(mode):
    """ss_merge_views called from C++ through the C ABI alone (vulkanengine_b200/host/merge_cli.cpp): the merged lists of
    three shards equal one context holding the scene."""
    exe = os.path.join(ROOT, "vulkanengine_b200", "host", "merge_cli.bin")
    r = subprocess.run([exe, "150000", "3"] + ([mode] if mode else []), capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "merge identical" in r.stdout
