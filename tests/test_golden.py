"""Golden vectors produced by the reference-arithmetic build (tests/golden/make_golden.py) pin, without
/root/reference: the oracle, and the product's host-side view-matrix functions."""
import os

import numpy as np
import pytest

import golden_cases as G
import oracle_lib as O
from vulkanengine_b200 import scenesubmit as S

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_oracle_trs_to_mat4_golden():
    g = np.load(os.path.join(GOLD, "trs_to_mat4.npz"))
    det = np.stack([O.trs_to_mat4(t, 0) for t in g["trs"]])
    assert det.tobytes() == g["mat_det"].tobytes()
    # libm mode depends on the host's sinf/cosf: same glibc here and on the GPU box, but keep a tolerance
    libm = np.stack([O.trs_to_mat4(t, 1) for t in g["trs"]])
    assert np.allclose(libm, g["mat_libm"], rtol=0, atol=1e-6 * np.abs(g["mat_libm"]).max())


def test_det_sincos_close_to_libm():
    xs = np.random.default_rng(3).uniform(-20, 20, 20000).astype(np.float32)
    err = 0.0
    for x in xs:
        s, c = O.det_sincos(x)
        err = max(err, abs(s - np.sin(np.float64(x))), abs(c - np.cos(np.float64(x))))
    assert err < 2.0 * 2.0 ** -24  # <= 2 ulp of 1.0


@pytest.mark.parametrize("which", ["oracle", "product"])
def test_views_golden(which):
    g = np.load(os.path.join(GOLD, "views.npz"))
    for k, (cam, out) in enumerate(zip(g["cams"], g["outs"])):
        fov, aspect, zn, zf = (float(x) for x in cam[:4])
        pos, pitch, yaw, lam, light = cam[4:7], float(cam[7]), float(cam[8]), float(cam[9]), cam[10:13]
        if which == "oracle":
            v, p, vp = O.camera_view_proj(fov, aspect, zn, zf, pos, pitch, yaw)
            splits, lm = O.cascade_matrices(vp, zn, zf, lam, -75.0, 75.0, light)
        else:  # host functions inside libscenesubmit.so; no GPU needed
            v, p, vp = S.camera_view_proj(fov, aspect, zn, zf, pos, pitch, yaw)
            splits, lm = S.cascade_matrices(vp, zn, zf, lam, -75.0, 75.0, light)
        got = np.concatenate([v, p, vp, splits, lm.reshape(-1)]).astype(np.float32)
        assert got.tobytes() == out.tobytes()
        for snap in g["snapped"][2 * k: 2 * k + 2]:  # the CascadeOffset branch, two origins per camera
            fn = O.cascade_matrices if which == "oracle" else S.cascade_matrices
            s2, lm2 = fn(vp, zn, zf, lam, -75.0, 75.0, light, origin=snap[:3], map_size=4096.0)
            assert np.concatenate([s2, lm2.reshape(-1)]).astype(np.float32).tobytes() == snap[3:].tobytes()


@pytest.mark.parametrize("name", ["small", "medium", "hier"])
def test_oracle_against_entt_pass_golden(name):
    g = np.load(os.path.join(GOLD, "entt_pass.npz"))
    ref = {k.split("__", 1)[1]: g[k] for k in g.files if k.startswith(name + "__")}
    case = G.entt_scene_cases()[name]
    sc = G.flatten_case(case)
    assert np.array_equal(sc.node, ref["inst_ent"]) and np.array_equal(sc.slot, ref["inst_sub"])
    res = O.OracleScene(sc, 0).frame(G.all_visible_view())
    assert np.array_equal(res["rows"], ref["world"])
    v = res["views"][0]
    G.check_lists_against_ref_pass(sc, ref, v["idx"], v["keys"], v["rows"], v["draws"], v["draw_keys"])
