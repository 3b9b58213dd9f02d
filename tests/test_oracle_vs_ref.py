"""Pins the plain-C oracle against the reference's own arithmetic: oracle/_ref/libss_ref.so is GLM 0.9.9.9 +
entt 3.15.0 compiled from /root/reference.  Bit-for-bit.  Skipped where that build is absent."""
import numpy as np
import pytest

import golden_cases as G
import oracle_lib as O
from vulkanengine_b200 import synth

pytestmark = pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built (needs /root/reference)")


def test_trs_to_mat4_matches_glm_bitwise():
    rng = np.random.default_rng(11)
    for i in range(3000):
        trs = np.concatenate([rng.uniform(-500, 500, 3), rng.uniform(-7, 7, 3), rng.uniform(-2, 2, 3)]).astype(np.float32)
        for mode in (0, 1):
            assert O.trs_to_mat4(trs, mode, "oracle").tobytes() == O.trs_to_mat4(trs, mode, "ref").tobytes()


def test_camera_and_cascades_match_glm_bitwise():
    rng = np.random.default_rng(12)
    for _ in range(200):
        args = (rng.uniform(0.4, 1.6), rng.uniform(0.7, 2.3), 0.1, float(rng.choice([100.0, 1000.0])),
                rng.uniform(-600, 600, 3), rng.uniform(-1.5, 1.5), rng.uniform(-3, 3))
        a, b = O.camera_view_proj(*args, which="oracle"), O.camera_view_proj(*args, which="ref")
        for x, y in zip(a, b):
            assert x.tobytes() == y.tobytes()
        light = rng.uniform(-1, 1, 3)
        light[1] = -abs(light[1]) - 0.05
        ca = O.cascade_matrices(a[2], 0.1, args[3], 0.92, -75.0, 75.0, light, "oracle")
        cb = O.cascade_matrices(a[2], 0.1, args[3], 0.92, -75.0, 75.0, light, "ref")
        assert ca[0].tobytes() == cb[0].tobytes() and ca[1].tobytes() == cb[1].tobytes()
        org = rng.uniform(-300, 300, 3)  # the CascadeOffset branch
        for size in (1024.0, 4096.0):
            sa = O.cascade_matrices(a[2], 0.1, args[3], 0.92, -75.0, 75.0, light, "oracle", origin=org, map_size=size)
            sb = O.cascade_matrices(a[2], 0.1, args[3], 0.92, -75.0, 75.0, light, "ref", origin=org, map_size=size)
            assert sa[1].tobytes() == sb[1].tobytes()
            assert sa[1].tobytes() != ca[1].tobytes()  # the snap really moves the projection


@pytest.mark.parametrize("name", ["small", "medium", "hier"])
@pytest.mark.parametrize("trig", [0, 1])
def test_submission_pass_matches_entt_and_std_map(name, trig):
    """entt view order + SubmitMesh early-outs + material fallback + std::map<MeshKey> order, against the
    oracle's key/sort/compact formulation fed by the flattening rule of the host adapter."""
    case = G.entt_scene_cases()[name]
    ref = G.run_ref_pass(case, trig_mode=trig)
    sc = G.flatten_case(case)
    assert np.array_equal(sc.node, ref["inst_ent"]), "flattening order differs from entt view order"
    assert np.array_equal(sc.slot, ref["inst_sub"])
    res = O.OracleScene(sc, trig).frame(G.all_visible_view())
    assert np.array_equal(res["rows"], ref["world"])
    v = res["views"][0]
    G.check_lists_against_ref_pass(sc, ref, v["idx"], v["keys"], v["rows"], v["draws"], v["draw_keys"])


def test_flat_frame_of_ref_matches_oracle_masks_and_counts():
    """The timed CPU baseline (glm + std::map per view) computes the same masks/rows/counts as the oracle."""
    import ctypes as C
    sc = synth.flat_scene(20000, n_keys=300)
    views = synth.make_views(lambda *a: O.camera_view_proj(*a), lambda *a: O.cascade_matrices(*a), 5)
    res = O.OracleScene(sc, 1).frame(views)

    class Flat(C.Structure):
        _fields_ = [("n", C.c_uint32), ("trs", C.c_void_p), ("parent", C.c_void_p), ("n_nodes", C.c_uint32),
                    ("node_base", C.c_uint32), ("slot", C.c_void_p), ("matword", C.c_void_p), ("inst_aabb", C.c_void_p),
                    ("slot_local", C.c_void_p), ("slot_mesh_handle", C.c_void_p), ("slot_submesh", C.c_void_p),
                    ("material_handle", C.c_void_p), ("trig_mode", C.c_int)]
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    local = np.ascontiguousarray(sc.slots["local"])
    mh = np.ascontiguousarray(sc.mesh_handles[sc.slots["mesh_rank"]])
    sub = np.ascontiguousarray(sc.slots["submesh_index"])
    f = Flat(sc.n_instances, p(sc.trs), None, sc.n_nodes, 0, p(sc.slot), p(sc.matword), p(sc.inst_aabb), p(local), p(mh),
             p(sub), p(sc.material_handles), 1)
    vps, vfl = np.ascontiguousarray(views["view_proj"]), np.ascontiguousarray(views["flags"])
    vis, drw = np.zeros(5, np.uint32), np.zeros(5, np.uint32)
    masks, rows = np.zeros(sc.n_instances, np.uint8), np.zeros((sc.n_instances, 12), np.float32)
    for threads in (1, 4):
        O.ref().ssr_flat_frame(C.byref(f), p(vps), p(vfl), C.c_uint32(5), C.c_int(threads), p(vis), p(drw), p(masks), p(rows))
        assert np.array_equal(masks, res["masks"])
        live = sc.slot != synth.SLOT_NULL
        assert np.array_equal(rows[live], res["rows"][live])
        assert [int(x) for x in vis] == [len(v["idx"]) for v in res["views"]]
        assert [int(x) for x in drw] == [len(v["draws"]) for v in res["views"]]


def test_quat_to_euler_matches_jolt():
    """The oracle's restatement of JPH::Quat::GetEulerAngles against the reference's own Jolt headers: same formula,
    Jolt evaluates ATan2 / ASin with its own polynomials, libm here: agreement to 1e-6 (absolute, radians)."""
    import ctypes as C
    rng = np.random.default_rng(31)
    q = rng.normal(size=(20000, 4)).astype(np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True).astype(np.float32)
    q[0] = [0, 0, 0, 1]
    q[1] = [0, 0.70710678, 0, 0.70710678]  # pitch of 90 degrees: the clamped asin branch
    a, b = np.empty((len(q), 3), np.float32), np.empty((len(q), 3), np.float32)
    p = lambda x: x.ctypes.data_as(C.c_void_p)
    O.lib().sso_quat_to_euler(p(q), C.c_uint32(len(q)), p(a))
    O.ref().ssr_jolt_euler(p(q), C.c_uint32(len(q)), p(b))
    assert np.abs(a - b).max() <= 2e-6, np.abs(a - b).max()


def _light_case(seed, n_nodes=50):
    rng = np.random.default_rng(seed)
    trs = np.concatenate([rng.uniform(-100, 100, (n_nodes, 3)), rng.uniform(-3, 3, (n_nodes, 3)), rng.uniform(0.5, 2, (n_nodes, 3))],
                         axis=1).astype(np.float32)
    n_pt, n_sp = int(rng.integers(0, 11)), int(rng.integers(0, 11))
    pn, sn = rng.integers(0, n_nodes, n_pt).astype(np.uint32), rng.integers(0, n_nodes, n_sp).astype(np.uint32)
    pp, spp = rng.uniform(0, 5, (n_pt, 6)).astype(np.float32), rng.uniform(0, 5, (n_sp, 8)).astype(np.float32)
    return trs, pn, pp, sn, spp


def pack_lights_cpu(fn, trs, pn, pp, sn, spp):
    import ctypes as C
    p = lambda x: x.ctypes.data_as(C.c_void_p)
    ubp, ubs = np.zeros(496, np.uint8), np.zeros(656, np.uint8)
    pos = np.zeros((len(pn) + len(sn) + 1, 4), np.float32)
    fn(p(np.ascontiguousarray(trs)), p(pn), p(np.ascontiguousarray(pp)), C.c_uint32(len(pn)), p(sn), p(np.ascontiguousarray(spp)),
       C.c_uint32(len(sn)), p(ubp), p(ubs), p(pos))
    return ubp, ubs, pos[:len(pn) + len(sn)]


def test_pack_lights_matches_reference_structs_bytewise():
    """UBPointLights / UBSpotLights as Scene::UpdateLightsBuffer fills them (reference structs + glm::normalize in
    oracle/_ref) against the oracle's float-offset restatement: identical bytes, padding included."""
    for seed in range(40):
        case = _light_case(seed)
        a = pack_lights_cpu(O.lib().sso_pack_lights, *case)
        b = pack_lights_cpu(O.ref().ssr_pack_lights, *case)
        for x, y in zip(a, b):
            assert x.tobytes() == y.tobytes()
