"""SURVEY.md 8(f)1 — the data formats in front of the path: AssetRegistry.vkr, .vkscene, .vkmesh, .vkmat and the binary
FBX mesh sources, read by the headless C++ loader (vulkanengine_b200/host/SceneSerializer.cpp, MeshImporter.cpp)
and checked against fixtures extracted from the same files by independent Python code (PyYAML + the FBX
restatement in tests/golden/make_vkmesh_fixtures.py).  "mini" is an asset directory authored for these tests
(tests/golden/make_mini_assets.py) and travels with the repo; the six bundled scenes are read from /root/reference
where that exists (the build container) and skipped elsewhere."""
import os
import subprocess

import numpy as np
import pytest

import golden_cases as G
import oracle_lib as O
import vkscene_cases as V
from test_host_mirror import CLI, read_blocks
from test_vkscenes import default_views
from vulkanengine_b200 import synth

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ASSETS = "/root/reference/VulkanCore/assets"
MINI_ASSETS = os.path.join(HERE, "golden", "mini_assets")
CASES = [("mini_", "Mini")] + [("", s) for s in V.SCENES]
IDS = [p + s for p, s in CASES]


def assets_of(prefix):
    if prefix == "mini_":
        return MINI_ASSETS
    if not os.path.isdir(REF_ASSETS):
        pytest.skip("the reference's asset directory is not on this machine")
    return REF_ASSETS


def run_loader(mode, prefix, name, tmp_path):
    assets = assets_of(prefix)
    out = str(tmp_path / "dump.bin")
    subprocess.check_call([CLI, mode, assets, os.path.join(assets, "scenes", name + ".vkscene"), out], stderr=subprocess.DEVNULL)
    b = read_blocks(out)
    loaded = dict(uuid=np.frombuffer(b[0], np.uint64), has_transform=np.frombuffer(b[1], np.uint8),
                  mesh_handle=np.frombuffer(b[2], np.uint64), trs=np.frombuffer(b[3], np.float32).reshape(-1, 9),
                  light=np.frombuffer(b[4], np.float32), meshes=np.frombuffer(b[5], np.uint64),
                  sub_count=np.frombuffer(b[6], np.uint32), sub_ints=np.frombuffer(b[7], np.uint32).reshape(-1, 5),
                  sub_floats=np.frombuffer(b[8], np.float32).reshape(-1, 22))
    return loaded, b[9:]


@pytest.mark.parametrize("prefix,name", CASES, ids=IDS)
def test_loader_reads_what_the_files_say(prefix, name, tmp_path):
    got, _ = run_loader("load-flatten", prefix, name, tmp_path)
    ents = V.load(prefix)["scenes"][name]
    assert got["uuid"].tolist() == [int(e["uuid"]) for e in ents]                      # file order = creation order
    assert got["has_transform"].tolist() == [int(e["has_transform"]) for e in ents]
    assert got["mesh_handle"].tolist() == [int(e["mesh"]) if e["mesh"] else 0 for e in ents]
    want_trs = np.array([e["translation"] + e["rotation"] + e["scale"] for e in ents if e["has_transform"]], np.float32)
    assert np.array_equal(got["trs"], want_trs.reshape(-1, 9))                          # same float32 as PyYAML's doubles round to
    light = next((e["light_direction"] for e in ents if "light_direction" in e), [1.0, -1.0, 1.0])
    assert np.array_equal(got["light"], np.array(light, np.float32))
    # submesh tables of the meshes the scene names, in order of first use
    md = V.load_meshes(prefix)
    seen = []
    for e in ents:
        if e["mesh"] and int(e["mesh"]) not in seen:
            seen.append(int(e["mesh"]))
    assert got["meshes"].tolist() == seen
    row = 0
    for h, cnt in zip(seen, got["sub_count"].tolist()):
        src = md["meshes"].get(str(h), {}).get("source")
        want = md["sources"][src] if src else []
        assert cnt == len(want)
        for sm in want:
            assert got["sub_ints"][row].tolist() == [sm["BaseVertex"], sm["BaseIndex"], sm["VertexCount"], sm["IndexCount"], sm["MaterialIndex"]]
            assert np.allclose(got["sub_floats"][row, :16], np.array(sm["LocalTransform"], np.float32), rtol=1e-6, atol=1e-7)
            assert np.array_equal(got["sub_floats"][row, 16:], np.array(sm["BoundsMin"] + sm["BoundsMax"], np.float32))
            row += 1
    assert row == len(got["sub_ints"])


def test_mini_fbx_known_answers(tmp_path):
    """Hand-checkable numbers of the authored FBX: quad + triangle on two material slots, a pentagon on a child model,
    a 2-corner polygon dropped; UnitScaleFactor 2 -> vertices x 0.02; model A = T(100,0,50)*0.02 . Ry(90) . S(2)."""
    got, _ = run_loader("load-flatten", "mini_", "Mini", tmp_path)
    assert got["meshes"].tolist() == [101, 102, 103, 555] and got["sub_count"].tolist() == [3, 0, 3, 0]  # 101 and 103 share a source
    assert got["sub_ints"][:3].tolist() == [[0, 0, 4, 6, 0], [4, 6, 3, 3, 1], [7, 9, 5, 9, 1]]
    assert np.array_equal(got["sub_ints"][3:], got["sub_ints"][:3])
    assert np.allclose(got["sub_floats"][0, 16:], [-0.02, -0.02, 0, 0.02, 0.02, 0])
    assert np.allclose(got["sub_floats"][1, 16:], [-0.02, 0.02, 0, 0.02, 0.04, 0.06])
    assert np.allclose(got["sub_floats"][2, 16:], [0, 0, -0.01, 0.06, 0.04, 0.01])
    a = got["sub_floats"][0, :16].reshape(4, 4).T                                          # column-major -> rows
    assert np.allclose(a, [[0, 0, 2, 2], [0, 2, 0, 0], [-2, 0, 0, 1], [0, 0, 0, 1]], atol=1e-6)


def test_mini_fbx_geometry_matches_second_reader(tmp_path):
    """Vertex positions (one per polygon corner, scaled) and fan-triangulated indices of the arena the consumer passes
    draw from == the Python restatement's, submesh by submesh."""
    import sys
    sys.path.insert(0, os.path.join(HERE, "golden"))
    import make_vkmesh_fixtures as MM
    geo = {"positions": [], "indices": []}
    subs = MM.submeshes_of(os.path.join(MINI_ASSETS, "meshes", "Source", "mini.fbx"), geo)
    _, blocks = run_loader("load-flatten", "mini_", "Mini", tmp_path)
    pos = np.frombuffer(blocks[10], np.float32).reshape(-1, 3)
    ind = np.frombuffer(blocks[11], np.uint32)
    assert np.array_equal(pos, np.concatenate(geo["positions"]))
    assert np.array_equal(ind, np.concatenate(geo["indices"]))
    # indices are relative to each submesh's BaseVertex and stay inside its VertexCount
    for sm, idx in zip(subs, geo["indices"]):
        assert len(idx) == sm["IndexCount"] and int(idx.max()) < sm["VertexCount"]


def flat_of(blocks):
    return dict(node=np.frombuffer(blocks[0], np.uint32), slot=np.frombuffer(blocks[1], np.uint32),
                matword=np.frombuffer(blocks[2], np.uint32), entity=np.frombuffer(blocks[3], np.uint32),
                parent=np.frombuffer(blocks[4], np.int32), levels=np.frombuffer(blocks[5], np.uint32),
                slots=np.frombuffer(blocks[6], synth.SUBMESH_DTYPE), mesh_handles=np.frombuffer(blocks[7], np.uint64),
                material_handles=np.frombuffer(blocks[8], np.uint64), bits=np.frombuffer(blocks[9], np.uint32))


@pytest.mark.parametrize("prefix,name", CASES, ids=IDS)
def test_loader_flatten_matches_fixture_flatten(prefix, name, tmp_path):
    """registry -> streams from the files == the flattening rule applied to the fixture (entt view order, early-outs,
    material resolution with the deserializer's base-material rule)."""
    got, blocks = run_loader("load-flatten", prefix, name, tmp_path)
    f = flat_of(blocks)
    case = V.vkscene_case(name, real_meshes=True, prefix=prefix)
    want = G.flatten_case(case)
    assert np.array_equal(f["node"], want.node) and np.array_equal(f["matword"], want.matword)
    # the mirror's table holds only meshes that submit, in handle order; the fixture's every usable mesh: compare per instance
    for k in ("index_count", "first_index", "vertex_offset", "submesh_index", "aabb_min", "aabb_max"):
        assert np.array_equal(f["slots"][k][f["slot"]], want.slots[k][want.slot]), k
    assert np.allclose(f["slots"]["local"][f["slot"]], want.slots["local"][want.slot], rtol=1e-6, atol=1e-7)
    assert np.array_equal(f["mesh_handles"][f["slots"]["mesh_rank"][f["slot"]]], want.mesh_handles[want.slots["mesh_rank"][want.slot]])
    used = np.unique(want.matword & np.uint32(synth.TESS_BIT - 1))
    assert np.array_equal(f["material_handles"], want.material_handles[used] if len(used) else f["material_handles"])
    # entity of instance = registry entity number (entities without a transform are numbered too)
    with_t = np.flatnonzero(got["has_transform"])
    assert np.array_equal(f["entity"], with_t[want.node])


@pytest.mark.gpu
def test_loaded_mini_scene_renders_like_the_oracle(tmp_path):
    """files -> loader -> Scene / SceneRenderer mirror -> libscenesubmit on cuda:0, against the oracle on the streams
    the loader itself flattened (EditorLayer.cpp:75 default camera + 4 cascades, light direction from the file)."""
    got, blocks = run_loader("load-flatten", "mini_", "Mini", tmp_path)
    f = flat_of(blocks)
    n_nodes = int(got["has_transform"].sum())
    sc = synth.SceneArrays(trs=got["trs"].copy(), parent=None, level_offsets=np.array([0, n_nodes], np.uint32), node=f["node"].copy(),
                           node_base=0, slot=f["slot"].copy(), matword=f["matword"].copy(), inst_aabb=None, slots=f["slots"].copy(),
                           mesh_bits=int(f["bits"][0]), submesh_bits=int(f["bits"][1]), material_bits=int(f["bits"][2]),
                           mesh_handles=f["mesh_handles"].copy(), material_handles=f["material_handles"].copy(), seed=0, name="mini")
    views = default_views([float(x) for x in got["light"]])
    ref = O.OracleScene(sc, 0).frame(views)
    _, rb = run_loader("load-render", "mini_", "Mini", tmp_path)
    visible, draws = np.frombuffer(rb[0], np.uint32), np.frombuffer(rb[1], np.uint32)
    assert sc.n_instances == 9 and int(visible.max()) > 0
    with_t = np.flatnonzero(got["has_transform"])
    for v in range(5):
        rows = np.frombuffer(rb[2 + 3 * v], np.float32).reshape(-1, 12)
        ents = np.frombuffer(rb[3 + 3 * v], np.uint32)
        dl = np.frombuffer(rb[4 + 3 * v], np.uint64).reshape(-1, 9)
        want = ref["views"][v]
        assert visible[v] == len(want["idx"]) and draws[v] == len(want["draws"])
        assert np.array_equal(rows, want["rows"].reshape(-1, 12))
        assert np.array_equal(ents, with_t[sc.node[want["idx"]]])
        assert np.array_equal(dl[:, 5].astype(np.uint32), want["draws"]["instance_count"])
        assert np.array_equal(dl[:, 4].astype(np.uint32), want["draws"]["index_count"])


@pytest.mark.gpu
def test_loaded_mini_scene_selection_and_shadow_passes(tmp_path):
    """files -> loader (FBX vertices, fan indices) -> SceneRenderer mirror -> ss_frame -> SelectionPass / ShadowPass as
    CUDA rasterisation over the compacted lists: the five targets equal the CPU restatement on the same arena, and the
    hovered-entity read-back (SceneRenderer.cpp:1449-1456) names the entity whose instance owns that sample."""
    W, H, SM = 320, 192, 256

    def pick(mx, my):
        out = str(tmp_path / f"pick_{mx}_{my}.bin")
        subprocess.check_call([CLI, "load-pick", MINI_ASSETS, os.path.join(MINI_ASSETS, "scenes", "Mini.vkscene"), out,
                               str(W), str(H), str(SM), str(mx), str(my)], stderr=subprocess.DEVNULL)
        return read_blocks(out)

    b = pick(0, 0)
    loaded_trs = np.frombuffer(b[3], np.float32).reshape(-1, 9)
    has_t = np.frombuffer(b[1], np.uint8)
    light = [float(x) for x in np.frombuffer(b[4], np.float32)]
    rb = b[9:]                                   # visible, draws, 5 x (rows, entities, draw list), then the pick blocks
    pk = rb[2 + 15:]
    pos = np.frombuffer(pk[0], np.float32).reshape(-1, 3)
    ind = np.frombuffer(pk[1], np.uint32)
    vbase, ibase = np.frombuffer(pk[2], np.uint32), np.frombuffer(pk[3], np.uint32)
    slot_of = np.frombuffer(pk[4], np.uint32)
    table = np.frombuffer(pk[5], synth.SUBMESH_DTYPE).copy()
    # meshes 101 and 103 name the same source file: its vertices sit in the arena once
    assert len(ind) == 6 + 3 + 9 and len(pos) == 4 + 3 + 5
    table["first_index"] += ibase
    table["vertex_offset"] += vbase.astype(np.int32)
    # the frame the mirror ran, restated: flattened streams -> oracle lists -> oracle rasterisation
    got, blocks = run_loader("load-flatten", "mini_", "Mini", tmp_path)
    f = flat_of(blocks)
    n_nodes = int(has_t.sum())
    sc = synth.SceneArrays(trs=loaded_trs.copy(), parent=None, level_offsets=np.array([0, n_nodes], np.uint32), node=f["node"].copy(),
                           node_base=0, slot=f["slot"].copy(), matword=f["matword"].copy(), inst_aabb=None, slots=f["slots"].copy(),
                           mesh_bits=int(f["bits"][0]), submesh_bits=int(f["bits"][1]), material_bits=int(f["bits"][2]),
                           mesh_handles=f["mesh_handles"].copy(), material_handles=f["material_handles"].copy(), seed=0, name="mini")
    assert np.array_equal(slot_of, sc.slot)
    views = default_views(light)
    ref = O.OracleScene(sc, 0).frame(views)
    covered = None
    for v in range(5):
        w, h = (W, H) if v == 0 else (SM, SM)
        img = np.frombuffer(pk[6 + v], np.uint64).reshape(h, w)
        want = O.raster_view(ref["views"][v]["rows"], ref["views"][v]["idx"], sc.slot, table, pos, ind, views["view_proj"][v], w, h,
                             0 if v == 0 else 1)
        assert np.array_equal(img, want), f"view {v}: {np.count_nonzero(img != want)} samples differ"
        if v == 0:
            covered = img
    assert int(np.frombuffer(pk[11], np.int32)[0]) == -1 or covered[0, 0] != np.uint64(0xFFFFFFFFFFFFFFFF)
    ys, xs = np.nonzero(covered != np.uint64(0xFFFFFFFFFFFFFFFF))
    assert len(ys) > 0, "the mini scene must land on at least one sample of the selection target"
    # hover over a covered sample: the entity of the instance that owns it
    y, x = int(ys[len(ys) // 2]), int(xs[len(xs) // 2])
    b2 = pick(x, y)
    hovered = int(np.frombuffer(b2[9 + 17 + 11], np.int32)[0])
    inst = int(covered[y, x] & np.uint64(0xFFFFFFFF))
    assert hovered == int(f["entity"][inst])


def test_submesh_tables_against_the_reference_assimp_import():
    """tests/golden/vkmeshes_assimp.json is the submesh table the REFERENCE's importer path gives for the bundled FBX
    sources: its vendored Assimp compiled by oracle/Makefile (`make -C oracle assimp`) with Mesh.cpp:62-72's flags, then
    MeshSource / TraverseNodes bookkeeping (oracle/ref_assimp.cpp).  The headless reader (host/MeshImporter.cpp, pinned to
    tests/golden/vkmeshes.json by the tests above) must agree on everything the submission path reads:
      exact      number and order of submeshes, BaseIndex, IndexCount, MaterialIndex
      1e-6       LocalTransform (Assimp accumulates in float, the reader in double), local bounds
      different  VertexCount / BaseVertex: Assimp's JoinIdenticalVertices welds corners that share every attribute
                 (after CalcTangentSpace); the reader keeps one vertex per polygon corner, so its counts are >= Assimp's
                 and its vertex_offset refers to its own (unwelded) vertex arena, which is what its index buffer indexes."""
    import json
    here = os.path.dirname(os.path.abspath(__file__))
    ours = json.load(open(os.path.join(here, "golden", "vkmeshes.json")))["sources"]
    ref = json.load(open(os.path.join(here, "golden", "vkmeshes_assimp.json")))["sources"]
    assert len(ref) >= 11
    for src, table in ref.items():
        mine = ours[src]
        assert len(mine) == len(table), src
        base_vertex = 0
        for a, b in zip(table, mine):
            for f in ("BaseIndex", "IndexCount", "MaterialIndex"):
                assert a[f] == b[f], (src, f)
            la, lb = np.array(a["LocalTransform"]), np.array(b["LocalTransform"])
            assert np.abs(la - lb).max() <= 1e-6 * max(np.abs(la).max(), 1.0), src
            for f in ("BoundsMin", "BoundsMax"):
                xa, xb = np.array(a[f]), np.array(b[f])
                assert np.abs(xa - xb).max() <= 1e-6 * max(np.abs(xa).max(), 1e-30), (src, f)
            assert b["VertexCount"] >= a["VertexCount"] and b["BaseVertex"] >= a["BaseVertex"]
            assert b["VertexCount"] <= a["IndexCount"]          # never more than one vertex per corner
            assert b["BaseVertex"] == base_vertex                # the reader's own arena is dense
            base_vertex += b["VertexCount"]
