"""configs[0]: the reference's bundled scenes, headless, default editor camera + 4 cascades.
CPU: oracle == real entt/GLM/std::map (where oracle/_ref exists) and the C++ host mirror flattens them in entt
order.  GPU: the mirror's RenderScene() == oracle."""
import os
import subprocess

import numpy as np
import pytest

import golden_cases as G
import oracle_lib as O
import vkscene_cases as V
from test_host_mirror import CLI, read_blocks, write_pack
from vulkanengine_b200 import synth


def default_views(light):
    cam = lambda *a: O.camera_view_proj(*a)
    casc = lambda vp, zn, zf, lam, no, fo, _ld: O.cascade_matrices(vp, zn, zf, lam, no, fo, light)
    # EditorCamera(radians(45), 1.635005, 0.1, 1000), distance 10 (EditorLayer.cpp:75, EditorCamera.h:81)
    return synth.make_views(cam, casc, 5, position=(0.0, 0.0, 10.0), pitch=0.0, yaw=0.0)


@pytest.mark.parametrize("name", V.SCENES)
def test_bundled_scene_oracle_matches_reference_grouping(name):
    case = V.vkscene_case(name)
    sc = G.flatten_case(case)
    assert sc.n_instances == int(case.has_mesh.sum())  # every mesh entity submits (one stand-in submesh each)
    res = O.OracleScene(sc, 1).frame(G.all_visible_view())
    if O.ref_available():
        ref = G.run_ref_pass(case, trig_mode=1)
        assert np.array_equal(sc.node, ref["inst_ent"])
        v = res["views"][0]
        G.check_lists_against_ref_pass(sc, ref, v["idx"], v["keys"], v["rows"], v["draws"], v["draw_keys"])
    # tessellated materials (MedievalBrick, Rock) form the second list
    tess_draws = [G.decode_key(sc, k)[3] for k in res["views"][0]["draw_keys"]]
    assert tess_draws == sorted(tess_draws)


@pytest.mark.parametrize("name", V.SCENES)
def test_bundled_scene_mirror_flatten(name, tmp_path):
    case = V.vkscene_case(name)
    pack, out = str(tmp_path / "pack.bin"), str(tmp_path / "flat.bin")
    write_pack(case, pack)
    subprocess.check_call([CLI, "flatten", pack, out])
    b = read_blocks(out)
    want = G.flatten_case(case)
    assert np.array_equal(np.frombuffer(b[0], np.uint32), want.node)
    assert np.array_equal(np.frombuffer(b[3], np.uint32), want.node)  # entity of instance (one submesh each)


@pytest.mark.gpu
@pytest.mark.parametrize("name", V.SCENES)
def test_bundled_scene_gpu_frame(name):
    """Through the C ABI with the reference's default camera + cascades: bit-exact against the oracle."""
    from vulkanengine_b200 import scenesubmit as S
    case = V.vkscene_case(name)
    sc = G.flatten_case(case)
    views = default_views(case.light_direction)
    ref = O.OracleScene(sc, 0).frame(views)
    ctx = S.SceneSubmit(sc.n_nodes, sc.n_instances, len(sc.slots), max_views=5)
    ctx.load_scene(sc)
    slot = ctx.frame(views)
    st = ctx.frame_sync(slot)
    assert np.array_equal(ctx.download_world(), ref["rows"]) and np.array_equal(ctx.download_masks(), ref["masks"])
    for v in range(5):
        got = ctx.download_view(slot, v, st)
        assert np.array_equal(got["idx"], ref["views"][v]["idx"]) and np.array_equal(got["draws"], ref["views"][v]["draws"])
        assert np.array_equal(got["rows"], ref["views"][v]["rows"])
    assert int(st.visible[0]) <= sc.n_instances
    ctx.close()
