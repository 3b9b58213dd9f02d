"""GPU parity: libscenesubmit (through the C ABI) against the CPU oracle on the same seeded inputs.
Bit-exact for masks, keys, instance order and draws; world rows bit-exact in deterministic-trig mode
(up to the sign of a zero, hence value comparison) and within 1e-6 relative in libm mode."""
import os

import numpy as np
import pytest

import oracle_lib as O
from vulkanengine_b200 import scenesubmit as S
from vulkanengine_b200 import synth

pytestmark = pytest.mark.gpu


def oracle_views(n_views=5, **kw):
    return synth.make_views(lambda *a: O.camera_view_proj(*a), lambda *a: O.cascade_matrices(*a), n_views, **kw)


# the two pipelines a frame can take (ss_config.flags): the general one (transform -> sort -> gather, every frame) and the
# draw-order-resident layout (sorted when a static stream changes; a frame streams)
PIPELINES = [pytest.param(0, id="general"), pytest.param(S.CONFIG_DRAW_ORDER_RESIDENT, id="resident")]
EXPECT_PIPELINE = {0: S.PIPELINE_GENERAL, S.CONFIG_DRAW_ORDER_RESIDENT: S.PIPELINE_RESIDENT}


def run_gpu(sc, views, trig_mode=S.TRIG_DETERMINISTIC, flags=0, world=None, pipe=None, **cfg):
    """One frame through the C ABI.  pipe: ss_config.flags choosing the pipeline (None: the library's default, or
    SS_CONFIG_FLAGS from the environment); where given, the frame must really have taken that pipeline."""
    expect_pipeline = None if pipe is None else EXPECT_PIPELINE[pipe]
    ctx = S.SceneSubmit(sc.n_nodes, max(sc.n_instances, 1), len(sc.slots), max_views=len(views), trig_mode=trig_mode,
                        flags=pipe, **cfg)
    ctx.load_scene(sc)
    if world is not None:
        ctx.upload_world_rows(world)
    slot = ctx.frame(views, flags)
    st = ctx.frame_sync(slot)
    if expect_pipeline is not None and sc.n_instances:
        assert int(st.pipeline) == expect_pipeline, f"frame took pipeline {int(st.pipeline)}, expected {expect_pipeline}"
    out = {"stats": st, "world": ctx.download_world(), "masks": ctx.download_masks(),
           "views": [ctx.download_view(slot, v, st) for v in range(len(views))]}
    ctx.close()
    return out


def check_against_oracle(sc, views, got, ref, exact_rows=True):
    live = sc.slot != synth.SLOT_NULL
    if exact_rows:
        assert np.array_equal(got["world"], ref["rows"]), "world rows differ"
    else:
        scale = np.maximum(np.abs(ref["rows"]).max(axis=1, keepdims=True), 1e-30)
        assert (np.abs(got["world"] - ref["rows"]) <= 1e-6 * scale)[live].all(), "world rows beyond 1e-6 relative"
    assert np.array_equal(got["masks"], ref["masks"]), "visibility masks differ"
    for v in range(len(views)):
        g, r = got["views"][v], ref["views"][v]
        assert int(got["stats"].visible[v]) == len(r["idx"]), f"view {v}: visible count"
        assert np.array_equal(g["idx"], r["idx"]), f"view {v}: draw order (instance ids)"
        assert np.array_equal(g["keys"], r["keys"]), f"view {v}: sort keys"
        assert np.array_equal(g["rows"], got["world"][g["idx"]]), f"view {v}: gathered rows"
        assert int(got["stats"].draws[v]) == len(r["draws"]), f"view {v}: draw count"
        assert np.array_equal(g["draws"], r["draws"]), f"view {v}: indirect commands"
        assert np.array_equal(g["draw_keys"], r["draw_keys"]), f"view {v}: draw keys"


@pytest.mark.parametrize("pipe", PIPELINES)
@pytest.mark.parametrize("n", [1, 31, 256, 257, 4096, 4097, 100_003])
def test_flat_bit_exact(n, pipe):
    sc = synth.flat_scene(n, n_keys=512)
    views = oracle_views(5)
    ref = O.OracleScene(sc, 0).frame(views)
    got = run_gpu(sc, views, pipe=pipe)
    check_against_oracle(sc, views, got, ref)


def test_flat_libm_mode_tolerance():
    sc = synth.flat_scene(50_000, n_keys=512)
    views = oracle_views(5)
    ref = O.OracleScene(sc, 1).frame(views)
    got = run_gpu(sc, views, trig_mode=S.TRIG_LIBM)
    live = sc.slot != synth.SLOT_NULL
    scale = np.maximum(np.abs(ref["rows"]).max(axis=1, keepdims=True), 1e-30)
    assert (np.abs(got["world"] - ref["rows"]) <= 1e-6 * scale)[live].all()
    # masks may differ only on borderline boxes: well under 0.1 % of instances
    assert (got["masks"] != ref["masks"]).mean() < 1e-3


def test_stagewise_cull_bit_exact_on_oracle_matrices():
    """SURVEY §7 hard part 1 (i): feed the oracle's libm-trig matrices, everything after must be bit-exact."""
    sc = synth.flat_scene(30_000, n_keys=300)
    views = oracle_views(5)
    osc = O.OracleScene(sc, 1)
    ref = osc.frame(views)
    got = run_gpu(sc, views, trig_mode=S.TRIG_LIBM, flags=S.FRAME_WORLD_ROWS_GIVEN, world=ref["rows"])
    check_against_oracle(sc, views, got, ref)


@pytest.mark.parametrize("pipe", PIPELINES)
def test_hierarchy_three_levels(pipe):
    sc = synth.hierarchy_scene(n_roots=16, fan1=8, fan2=32, n_keys=200)
    views = oracle_views(1)
    ref = O.OracleScene(sc, 0).frame(views)
    got = run_gpu(sc, views, pipe=pipe)
    check_against_oracle(sc, views, got, ref)


@pytest.mark.parametrize("pipe", PIPELINES)
def test_hierarchy_mixed_parents_in_warp(pipe):
    sc = synth.hierarchy_scene(n_roots=7, fan1=3, fan2=5, n_keys=64)
    views = oracle_views(5)
    ref = O.OracleScene(sc, 0).frame(views)
    got = run_gpu(sc, views, pipe=pipe)
    check_against_oracle(sc, views, got, ref)


def test_explicit_node_map():
    sc = synth.with_explicit_nodes(synth.hierarchy_scene(n_roots=8, fan1=4, fan2=16, n_keys=100))
    views = oracle_views(5)
    ref = O.OracleScene(sc, 0).frame(views)
    got = run_gpu(sc, views)
    check_against_oracle(sc, views, got, ref)


@pytest.mark.parametrize("pipe", PIPELINES)
def test_submesh_table_boxes(pipe):
    sc = synth.flat_scene(5000, n_keys=128)
    sc.inst_aabb = None
    views = oracle_views(5)
    ref = O.OracleScene(sc, 0).frame(views)
    got = run_gpu(sc, views, pipe=pipe)
    check_against_oracle(sc, views, got, ref)


@pytest.mark.parametrize("pipe", PIPELINES)
def test_all_dropped_and_single_key(pipe):
    sc = synth.flat_scene(3000, n_keys=1)
    views = oracle_views(2)
    ref = O.OracleScene(sc, 0).frame(views)
    got = run_gpu(sc, views, pipe=pipe)
    check_against_oracle(sc, views, got, ref)
    sc.slot[:] = synth.SLOT_NULL
    ref = O.OracleScene(sc, 0).frame(views)
    got = run_gpu(sc, views, pipe=pipe)
    assert got["masks"].sum() == 0 and all(int(got["stats"].visible[v]) == 0 for v in range(2))
    assert all(int(got["stats"].draws[v]) == 0 for v in range(2))


@pytest.mark.parametrize("pipe", PIPELINES)
def test_every_key_distinct(pipe):
    """Worst case for run bookkeeping: as many draws as instances."""
    n = 3000
    sc = synth.flat_scene(n, n_keys=4096)
    sc.material_bits = 12
    sc.matword = (np.arange(n, dtype=np.uint32) % 4096).astype(np.uint32)
    sc.slot = (np.arange(n, dtype=np.uint32) // 4096 % len(sc.slots)).astype(np.uint32)
    views = oracle_views(5)
    ref = O.OracleScene(sc, 0).frame(views)
    got = run_gpu(sc, views, pipe=pipe)
    check_against_oracle(sc, views, got, ref)


def test_frames_in_flight_slots_and_reuse():
    sc = synth.flat_scene(20_000, n_keys=256)
    views = oracle_views(5)
    ref = O.OracleScene(sc, 0).frame(views)
    ctx = S.SceneSubmit(sc.n_nodes, sc.n_instances, len(sc.slots), max_views=5)
    ctx.load_scene(sc)
    slots = [ctx.frame(views) for _ in range(5)]
    assert slots == [0, 1, 2, 0, 1]
    for slot in (2, 0, 1):
        st = ctx.frame_sync(slot)
        for v in range(5):
            assert np.array_equal(ctx.download_view(slot, v, st)["idx"], ref["views"][v]["idx"])
    # move everything far away: nothing visible in the camera view
    trs = sc.trs.copy()
    trs[:, 2] = 5000.0
    ctx.upload_trs(trs)
    st = ctx.frame_sync(ctx.frame(views))
    assert int(st.visible[0]) == 0 and int(st.draws[0]) == 0
    ctx.close()


def test_capacity_overflow_reports_error():
    sc = synth.flat_scene(20_000, n_keys=256)
    views = oracle_views(5)
    ctx = S.SceneSubmit(sc.n_nodes, sc.n_instances, len(sc.slots), max_views=5, max_visible_per_view=100)
    ctx.load_scene(sc)
    slot = ctx.frame(views)
    with pytest.raises(S.SceneSubmitError) as e:
        ctx.frame_sync(slot)
    assert e.value.status == S.SS_ERR_CAPACITY
    ctx.close()


def test_invalid_arguments():
    with pytest.raises(S.SceneSubmitError):
        S.SceneSubmit(0, 0, 0)
    sc = synth.flat_scene(100, n_keys=16)
    ctx = S.SceneSubmit(sc.n_nodes, sc.n_instances, len(sc.slots), max_views=2)
    with pytest.raises(S.SceneSubmitError):  # frame before streams
        ctx.frame(oracle_views(2))
    ctx.load_scene(sc)
    with pytest.raises(S.SceneSubmitError):  # too many views
        ctx.frame(oracle_views(5))
    bad = sc.slots.copy()
    bad["local"][0][3] = 1.0  # not affine
    with pytest.raises(S.SceneSubmitError):
        ctx.set_submesh_table(bad)
    ctx.close()


# ---- the sort on its own ---------------------------------------------------------------------------

@pytest.mark.parametrize("n", [0, 1, 2, 255, 4096, 4097, 70_001, 1 << 20])
def test_radix_sort_random_64bit(n):
    rng = np.random.default_rng(n)
    keys = rng.integers(0, 2**64, size=n, dtype=np.uint64)
    vals = np.arange(n, dtype=np.uint32)
    ko, vo = S.sort_pairs_host(keys, vals)
    order = np.argsort(keys, kind="stable")
    assert np.array_equal(ko, keys[order]) and np.array_equal(vo, vals[order])


def test_radix_sort_stability_and_digit_skip():
    rng = np.random.default_rng(5)
    n = 300_000
    keys = (rng.integers(0, 37, size=n, dtype=np.uint64) << np.uint64(20)) | np.uint64(0xABC)
    vals = np.arange(n, dtype=np.uint32)
    ko, vo = S.sort_pairs_host(keys, vals)
    order = np.argsort(keys, kind="stable")
    assert np.array_equal(ko, keys[order]) and np.array_equal(vo, vals[order])


def test_radix_sort_bit_range():
    rng = np.random.default_rng(6)
    n = 50_000
    keys = rng.integers(0, 2**64, size=n, dtype=np.uint64)
    vals = np.arange(n, dtype=np.uint32)
    ko, vo = S.sort_pairs_host(keys, vals, begin_bit=8, end_bit=29)
    sub = (keys >> np.uint64(8)) & np.uint64((1 << 21) - 1)
    order = np.argsort(sub, kind="stable")
    assert np.array_equal(vo, vals[order]) and np.array_equal(ko, keys[order])


# ---- against the reference's grouped output (golden vectors from real entt + glm + std::map) ---------------

@pytest.mark.parametrize("name", ["small", "medium", "hier"])
def test_gpu_matches_reference_submission_pass(name):
    import os

    import golden_cases as G
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "entt_pass.npz"))
    ref = {k.split("__", 1)[1]: g[k] for k in g.files if k.startswith(name + "__")}
    sc = G.flatten_case(G.entt_scene_cases()[name])
    got = run_gpu(sc, G.all_visible_view())
    assert np.array_equal(got["world"], ref["world"]), "world rows differ from glm's"
    v = got["views"][0]
    G.check_lists_against_ref_pass(sc, ref, v["idx"], v["keys"], v["rows"], v["draws"], v["draw_keys"])


_FULL_REF = {}


@pytest.mark.parametrize("pipe", PIPELINES)
def test_full_size_16m_properties_and_mask_parity(pipe):
    """configs[2] at full size: masks + world rows bit-exact against the oracle (threaded), and the
    size-independent properties of the lists: sorted by (key, instance), exactly the visible set,
    rows gathered from the world rows, draws partition the list."""
    n = 1 << 24
    sc = synth.flat_scene(n)
    views = oracle_views(5)
    ctx = S.SceneSubmit(sc.n_nodes, n, len(sc.slots), max_views=5, max_draws_per_view=1 << 16, flags=pipe)
    ctx.load_scene(sc)
    slot = ctx.frame(views)
    st = ctx.frame_sync(slot)
    assert int(st.pipeline) == EXPECT_PIPELINE[pipe]
    masks = ctx.download_masks()
    if "rows" not in _FULL_REF:  # one oracle run serves both pipelines
        osc = O.OracleScene(sc, 0)
        import ctypes as C
        rows = np.empty((n, 12), np.float32)
        # oracle per-instance stages only (the per-view lists at 16M x 5 would need ~7 GB of host memory)
        nw = osc.node_worlds()
        O.lib().sso_instance_rows(C.byref(osc.s), nw.ctypes.data_as(C.c_void_p), rows.ctypes.data_as(C.c_void_p))
        del nw
        _FULL_REF.update(rows=rows, omask=osc.cull(rows, views), okeys=osc.keys())
    rows, omask, okeys = _FULL_REF["rows"], _FULL_REF["omask"], _FULL_REF["okeys"]
    assert np.array_equal(masks, omask)
    world = ctx.download_world()
    assert np.array_equal(world, rows)
    for v in range(5):
        got = ctx.download_view(slot, v, st)
        vis = np.flatnonzero((masks >> v) & 1).astype(np.uint32)
        assert len(got["idx"]) == len(vis)
        assert np.array_equal(np.sort(got["idx"]), vis), f"view {v}: not exactly the visible set"
        assert np.array_equal(got["keys"], okeys[got["idx"]]), f"view {v}: keys"
        k, i = got["keys"], got["idx"].astype(np.int64)
        assert np.all((k[1:] > k[:-1]) | ((k[1:] == k[:-1]) & (i[1:] > i[:-1]))), f"view {v}: not sorted by (key, instance)"
        assert np.array_equal(got["rows"], world[got["idx"]]), f"view {v}: gathered rows"
        d = got["draws"]
        assert int(d["instance_count"].sum()) == len(vis)
        assert np.array_equal(d["first_instance"], np.concatenate([[0], np.cumsum(d["instance_count"])[:-1]]))
        assert np.array_equal(got["draw_keys"], k[d["first_instance"]]) and np.all(np.diff(got["draw_keys"].astype(np.int64)) > 0)
        sl = sc.slots[sc.slot[got["idx"][d["first_instance"]]]]
        assert np.array_equal(d["index_count"], sl["index_count"]) and np.array_equal(d["vertex_offset"], sl["vertex_offset"])
    ctx.close()


def test_external_nodes_with_given_world_rows():
    """Ad-hoc submitters (SubmitPhysicsMesh etc.) hand over a finished matrix: nodes past the last level take
    their world rows from ss_upload_node_world_rows instead of TRS."""
    base = synth.with_explicit_nodes(synth.hierarchy_scene(n_roots=4, fan1=4, fan2=8, n_keys=40))
    n_ext = 16
    n_nodes = base.n_nodes
    rng = np.random.default_rng(9)
    node = base.node.copy()
    ext_inst = rng.choice(base.n_instances, 40, replace=False)
    node[ext_inst] = n_nodes + rng.integers(0, n_ext, 40)
    # oracle side: give the external nodes an ordinary TRS, so their world matrix is well defined
    trs_ext = synth._trs(77, 90, 0, n_ext, 100.0, (0.5, 2.0))
    full = synth.SceneArrays(**{**base.__dict__, "trs": np.concatenate([base.trs, trs_ext]),
                                "parent": np.concatenate([base.parent, np.full(n_ext, -1, np.int32)]), "node": node})
    views = oracle_views(5)
    osc = O.OracleScene(full, 0)
    ref = osc.frame(views)
    ext_rows = np.stack([O.trs_to_mat4(t, 0).reshape(4, 4).T[:3].reshape(12) for t in trs_ext])

    ctx = S.SceneSubmit(n_nodes + n_ext, base.n_instances, len(base.slots), max_views=5)
    ctx.set_key_layout(base.mesh_bits, base.submesh_bits, base.material_bits)
    ctx.set_submesh_table(base.slots)
    ctx.set_nodes(n_nodes + n_ext, full.parent, base.level_offsets)  # levels end at n_nodes: the rest is external
    ctx.set_instances(base.slot, base.matword, node, 0)
    ctx.set_instance_aabbs(base.inst_aabb)
    ctx.upload_trs(base.trs)
    ctx.upload_node_world_rows(ext_rows, n_nodes)
    slot = ctx.frame(views)
    st = ctx.frame_sync(slot)
    got = {"stats": st, "world": ctx.download_world(), "masks": ctx.download_masks(),
           "views": [ctx.download_view(slot, v, st) for v in range(5)]}
    check_against_oracle(full, views, got, ref)
    with pytest.raises(S.SceneSubmitError):  # computed nodes cannot be overwritten this way
        ctx.upload_node_world_rows(ext_rows[:1], 0)
    ctx.close()


def test_batched_views_reuse_world_and_sort():
    """configs[4] shape in small: many cameras over one scene, 8 views per frame; after the first full frame
    the later batches reuse the world rows and the sorted order (SS_FRAME_REUSE_WORLD | SS_FRAME_REUSE_SORT)
    and must still equal the oracle view by view."""
    sc = synth.hierarchy_scene(n_roots=16, fan1=8, fan2=32, n_keys=200)
    n_cams = 24
    cams = synth.fibonacci_cameras(lambda *a: O.camera_view_proj(*a), n_cams, radius=800.0)
    osc = O.OracleScene(sc, 0)
    ctx = S.SceneSubmit(sc.n_nodes, sc.n_instances, len(sc.slots), max_views=8)
    ctx.load_scene(sc)
    seen_any = 0
    for b in range(0, n_cams, 8):
        batch = cams[b:b + 8]
        flags = 0 if b == 0 else (S.FRAME_REUSE_WORLD | S.FRAME_REUSE_SORT)
        slot = ctx.frame(batch, flags)
        st = ctx.frame_sync(slot)
        ref = osc.frame(batch)
        assert np.array_equal(ctx.download_masks(), ref["masks"])
        assert np.array_equal(ctx.download_world(), ref["rows"])
        for v in range(len(batch)):
            got = ctx.download_view(slot, v, st)
            assert np.array_equal(got["idx"], ref["views"][v]["idx"]) and np.array_equal(got["keys"], ref["views"][v]["keys"])
            assert np.array_equal(got["draws"], ref["views"][v]["draws"]) and np.array_equal(got["rows"], ref["views"][v]["rows"])
            seen_any += len(got["idx"])
        if b > 0:
            assert int(st.kernel_launches) < 6  # no hierarchy, no sort: cull + emit + draws
    assert seen_any > 0
    # a TRS upload invalidates the kept world rows: the flag is then ignored and the transform runs again
    trs = sc.trs.copy()
    trs[:, 0] += 1.0
    ctx.upload_trs(trs)
    sc2 = synth.SceneArrays(**{**sc.__dict__, "trs": trs})
    slot = ctx.frame(cams[:8], S.FRAME_REUSE_WORLD | S.FRAME_REUSE_SORT)
    st = ctx.frame_sync(slot)
    assert np.array_equal(ctx.download_world(), O.OracleScene(sc2, 0).frame(cams[:8])["rows"])
    ctx.close()


_CFG2_REF = {}


def _cfg2_reference(sc, views):
    if "ref" not in _CFG2_REF:  # one oracle run serves both pipelines
        _CFG2_REF["ref"] = O.OracleScene(sc, 0).frame(views, threads=8)
    return _CFG2_REF["ref"]


@pytest.mark.parametrize("pipe", PIPELINES)
def test_config2_full_size_hierarchy_1m(pipe):
    """configs[1]: 2^20 leaves under a 3-level hierarchy, one camera: bit-exact against the oracle."""
    sc = synth.hierarchy_scene()
    views = oracle_views(1)
    ref = _cfg2_reference(sc, views)
    got = run_gpu(sc, views, pipe=pipe, max_draws_per_view=1 << 16)
    check_against_oracle(sc, views, got, ref)


def test_wide_key_layout_pairs_mode_and_digit_skip():
    """Default-width layout (22/16/24 bits): key + instance index no longer fit one 64-bit word, so the sort
    runs on (key, value) pairs with 8-bit digits and skips the constant ones."""
    sc = synth.flat_scene(40_000, n_keys=300)
    wide = synth.SceneArrays(**{**sc.__dict__, "mesh_bits": 22, "submesh_bits": 16, "material_bits": 24})
    views = oracle_views(5)
    ref = O.OracleScene(wide, 0).frame(views)
    got = run_gpu(wide, views)
    check_against_oracle(wide, views, got, ref)
    assert 1 <= int(got["stats"].sort_passes_run) <= 5  # 8 digit positions, most of them constant


def test_eight_views():
    sc = synth.flat_scene(30_000, n_keys=200)
    cams = synth.fibonacci_cameras(lambda *a: O.camera_view_proj(*a), 8, radius=900.0)
    ref = O.OracleScene(sc, 0).frame(cams)
    got = run_gpu(sc, cams)
    check_against_oracle(sc, cams, got, ref)
    assert any(int(got["stats"].visible[v]) > 0 for v in range(8))


def test_item_major_output_ragged_tile_and_capacity():
    """Eight views that all see (nearly) everything: every k_emit tile takes the item-major output order (each row
    fetched once, stored to every view); 5000 instances leave a ragged last tile.  Then the same frame into a
    context whose per-view capacity is too small: reported as SS_ERR_CAPACITY, not written out of bounds."""
    sc = synth.flat_scene(5000, n_keys=37)
    cam = synth.fibonacci_cameras(lambda *a: O.camera_view_proj(*a), 1, radius=1400.0)
    cams = np.repeat(cam, 8)
    ref = O.OracleScene(sc, 0).frame(cams)
    got = run_gpu(sc, cams)
    check_against_oracle(sc, cams, got, ref)
    assert int(got["stats"].visible[0]) > 3 * 5000 // 4
    ctx = S.SceneSubmit(sc.n_nodes, sc.n_instances, len(sc.slots), max_views=8, max_visible_per_view=1000)
    ctx.load_scene(sc)
    with pytest.raises(S.SceneSubmitError):
        ctx.frame_sync(ctx.frame(cams))
    ctx.close()


def test_whole_stream_trs_uploads_overlap_frames_and_stay_ordered():
    """Whole-stream TRS uploads of >= 2^16 nodes go to a second device buffer on their own stream (so the copy of the
    next frame's transforms overlaps this frame's kernels); ss_frame adopts the newest one.  Frames must see exactly
    the transforms uploaded before them: several frames in flight, two whole uploads in a row, a dirty range on top of
    a whole upload, a frame with no upload in between."""
    n = 70_000
    sc = synth.flat_scene(n, n_keys=64)
    views = oracle_views(5)
    ctx = S.SceneSubmit(sc.n_nodes, n, len(sc.slots), max_views=5)
    ctx.load_scene(sc)
    rng = np.random.default_rng(5)

    def variant(k):
        t = sc.trs.copy()
        t[:, 0] += np.float32(0.25 * k)
        t[:, 4] += np.float32(0.01 * k)
        return t

    def expect(trs):
        return O.OracleScene(synth.SceneArrays(**{**sc.__dict__, "trs": trs}), 0).frame(views)

    # three frames in flight, each preceded by its own whole upload; results read afterwards
    pinned = [np.ascontiguousarray(variant(k)) for k in range(1, 4)]
    slots = []
    for t in pinned:
        ctx.upload_trs(t)
        slots.append(ctx.frame(views))
    for t, slot in zip(pinned, slots):
        st = ctx.frame_sync(slot)
        ref = expect(t)
        for v in range(5):
            got = ctx.download_view(slot, v, st)
            assert np.array_equal(got["idx"], ref["views"][v]["idx"]) and np.array_equal(got["rows"], ref["views"][v]["rows"])
    assert np.array_equal(ctx.download_world(), expect(pinned[-1])["rows"])
    # two whole uploads in a row: the later one wins; then a dirty range on top of it
    a, b = variant(7), variant(8)
    ctx.upload_trs(a)
    ctx.upload_trs(b)
    lo, cnt = 1234, 4321
    b2 = b.copy()
    b2[lo:lo + cnt, :3] = rng.uniform(-100, 100, (cnt, 3)).astype(np.float32)
    ctx.upload_trs(np.ascontiguousarray(b2[lo:lo + cnt]), first=lo)
    st = ctx.frame_sync(ctx.frame(views))
    ref = expect(b2)
    assert np.array_equal(ctx.download_world(), ref["rows"]) and np.array_equal(ctx.download_masks(), ref["masks"])
    # a dirty range alone, then a frame with no upload at all
    b3 = b2.copy()
    b3[:100, 2] -= np.float32(3.0)
    ctx.upload_trs(np.ascontiguousarray(b3[:100]), first=0)
    ctx.frame_sync(ctx.frame(views))
    assert np.array_equal(ctx.download_world(), expect(b3)["rows"])
    ctx.frame_sync(ctx.frame(views))
    assert np.array_equal(ctx.download_world(), expect(b3)["rows"])
    ctx.close()


@pytest.mark.parametrize("case", range(int(os.environ.get("SS_SWEEP_CASES", "24"))))
def test_randomised_configurations(case):
    """Seeded sweep over what a host can combine: scene shape (flat / 3-level / explicit node map), size, number of
    views and where they look from, per-instance boxes or the submesh table's, key layout, frames queued back to back,
    reuse flags on an unchanged scene.  Everything must equal the oracle, every time."""
    rng = np.random.default_rng(1000 + case)
    shape = case % 3
    n_keys = int(rng.integers(1, 400))
    if shape == 0:
        sc = synth.flat_scene(int(rng.integers(1, 9000)), n_keys=n_keys, seed=synth.SEED_BASE + 100 + case)
    else:
        sc = synth.hierarchy_scene(n_roots=int(rng.integers(1, 9)), fan1=int(rng.integers(1, 9)), fan2=int(rng.integers(1, 70)),
                                   n_keys=n_keys, seed=synth.SEED_BASE + 100 + case)
        if shape == 2:
            sc = synth.with_explicit_nodes(sc, seed=case)
    if rng.random() < 0.4:
        sc = synth.SceneArrays(**{**sc.__dict__, "inst_aabb": None})       # cull with the submesh table's boxes
    n_views = int(rng.integers(1, 9))
    if n_views == 5 and rng.random() < 0.5:
        views = oracle_views(5, position=tuple(float(x) for x in rng.uniform(-300, 300, 3)), yaw=float(rng.uniform(-3, 3)),
                             pitch=float(rng.uniform(-1, 1)))
    else:
        views = synth.fibonacci_cameras(lambda *a: O.camera_view_proj(*a), max(n_views, 1), radius=float(rng.uniform(100, 1500)),
                                        seed=synth.SEED_BASE + case)[:n_views]
        if rng.random() < 0.5:
            views = views.copy()
            views["flags"][::2] = synth.VIEW_SKIP_NEAR
    osc = O.OracleScene(sc, 0)
    ref = osc.frame(views)
    pipe = [0, S.CONFIG_DRAW_ORDER_RESIDENT][(case // 3) % 2]
    ctx = S.SceneSubmit(sc.n_nodes, sc.n_instances, len(sc.slots), max_views=n_views, frames_in_flight=int(rng.integers(1, 4)),
                        flags=pipe)
    ctx.load_scene(sc)
    slots = [ctx.frame(views) for _ in range(int(rng.integers(1, 4)))]
    if rng.random() < 0.5:
        slots.append(ctx.frame(views, S.FRAME_REUSE_WORLD | S.FRAME_REUSE_SORT))
    for slot in slots[-min(len(slots), 1):]:
        st = ctx.frame_sync(slot)
        got = {"stats": st, "world": ctx.download_world(), "masks": ctx.download_masks(),
               "views": [ctx.download_view(slot, v, st) for v in range(n_views)]}
        check_against_oracle(sc, views, got, ref)
    ctx.close()


def test_cascade_matrices_batch_equals_host_per_camera():
    """4096 cameras' cascade sets fitted on the GPU == ss_cascade_matrices on the host, bit for bit
    (and == the oracle, which is pinned against the reference's UpdateCascadeMap built with its own GLM)."""
    cams = synth.fibonacci_cameras(S.camera_view_proj, 4096)
    light = np.array(synth.LIGHT_DIR, np.float32)
    splits, mats = S.cascade_matrices_batch(cams["view_proj"], synth.CAMERA_NEAR, 2000.0, synth.CSM_LAMBDA, synth.CSM_NEAR_OFFSET,
                                            synth.CSM_FAR_OFFSET, light)
    for i in list(range(0, 4096, 97)) + [4095]:
        hs, hm = S.cascade_matrices(cams["view_proj"][i], synth.CAMERA_NEAR, 2000.0, synth.CSM_LAMBDA, synth.CSM_NEAR_OFFSET,
                                    synth.CSM_FAR_OFFSET, light)
        assert np.array_equal(np.asarray(hs, np.float32), splits)
        assert np.array_equal(np.asarray(hm, np.float32).reshape(4, 16), mats[i]), i
        om = O.cascade_matrices(cams["view_proj"][i], synth.CAMERA_NEAR, 2000.0, synth.CSM_LAMBDA, synth.CSM_NEAR_OFFSET,
                                synth.CSM_FAR_OFFSET, light)
        assert np.array_equal(np.asarray(om[1], np.float32).reshape(4, 16), mats[i])


def test_sparse_trs_upload_scatter():
    """a producer that moves few entities (the physics write-back loop) uploads (node, TRS) records; a
    node named twice keeps the later record; records land on top of earlier whole-stream and range uploads."""
    n = 70_000                                   # above the threshold of the double-buffered whole-stream path
    sc = synth.flat_scene(n, n_keys=50)
    views = oracle_views(5)
    ctx = S.SceneSubmit(sc.n_nodes, n, len(sc.slots), max_views=5)
    ctx.load_scene(sc)
    rng = np.random.default_rng(11)
    trs = sc.trs.copy()
    for round_ in range(3):
        k = [5000, 17, 30_000][round_]
        nodes = rng.integers(0, n, k).astype(np.uint32)
        nodes[-1] = nodes[0]                     # a duplicate: the later record must win
        rec = trs[nodes].copy()
        rec[:, :3] += rng.uniform(-50, 50, (k, 3)).astype(np.float32)
        rec[:, 3:6] += rng.uniform(-1, 1, (k, 3)).astype(np.float32)
        if round_ == 1:
            trs[:, 1] += np.float32(2.0)
            ctx.upload_trs(trs)                  # a whole-stream upload first, the sparse records on top of it
        ctx.upload_trs_sparse(nodes, rec)
        for j in range(k):
            trs[nodes[j]] = rec[j]
        st = ctx.frame_sync(ctx.frame(views))
        ref = O.OracleScene(synth.SceneArrays(**{**sc.__dict__, "trs": trs}), 0).frame(views)
        assert np.array_equal(ctx.download_world(), ref["rows"]) and np.array_equal(ctx.download_masks(), ref["masks"])
        assert [int(st.visible[v]) for v in range(5)] == [len(ref["views"][v]["idx"]) for v in range(5)]
    with pytest.raises(S.SceneSubmitError):
        ctx.upload_trs_sparse(np.array([n], np.uint32), trs[:1])
    ctx.close()


def test_zero_instances():
    sc = synth.flat_scene(64, n_keys=8)
    empty = synth.SceneArrays(**{**sc.__dict__, "slot": sc.slot[:0], "matword": sc.matword[:0], "inst_aabb": sc.inst_aabb[:0]})
    views = oracle_views(5)
    ctx = S.SceneSubmit(sc.n_nodes, 16, len(sc.slots), max_views=5)
    ctx.set_key_layout(sc.mesh_bits, sc.submesh_bits, sc.material_bits)
    ctx.set_submesh_table(sc.slots)
    ctx.set_nodes(sc.n_nodes)
    ctx.set_instances(empty.slot, empty.matword, None, 0)
    ctx.upload_trs(sc.trs)
    st = ctx.frame_sync(ctx.frame(views))
    assert all(int(st.visible[v]) == 0 and int(st.draws[v]) == 0 for v in range(5))
    ctx.close()


def test_scalar_and_packed_instance_kernels_agree(monkeypatch):
    """The general one-instance-per-thread kernel and the packed-FP32 fast path must give identical bits.
    node_base % 4 != 0 makes the fast path ineligible (TMA alignment), so the same scene shifted by one
    interior node runs on the general kernel."""
    base = synth.hierarchy_scene(n_roots=5, fan1=4, fan2=64, n_keys=100)      # node_base = 25: general kernel
    views = oracle_views(5)
    ref = O.OracleScene(base, 0).frame(views)
    got = run_gpu(base, views)
    check_against_oracle(base, views, got, ref)
    aligned = synth.hierarchy_scene(n_roots=4, fan1=7, fan2=64, n_keys=100)   # node_base = 32: fast path
    ref = O.OracleScene(aligned, 0).frame(views)
    got = run_gpu(aligned, views)
    check_against_oracle(aligned, views, got, ref)


def test_view_count_changes_between_frames():
    """ss_frame takes any n_views up to max_views, frame by frame.  The per-view run counters are double-buffered by
    frame and cleaned one frame ahead, so 5 views, then 1, then 5 again reuses a buffer the first frame dirtied in
    rows 1-4: every frame must still equal the oracle."""
    sc = synth.flat_scene(20_000, n_keys=256)
    v5 = oracle_views(5)
    v1 = v5[:1]
    osc = O.OracleScene(sc, 0)
    ref5, ref1 = osc.frame(v5), osc.frame(v1)
    ctx = S.SceneSubmit(sc.n_nodes, sc.n_instances, len(sc.slots), max_views=5, frames_in_flight=2)
    ctx.load_scene(sc)
    for views, ref in ((v5, ref5), (v1, ref1), (v5, ref5), (v5, ref5), (v1, ref1), (v1, ref1), (v5, ref5)):
        slot = ctx.frame(views)
        st = ctx.frame_sync(slot)
        got = {"stats": st, "world": ctx.download_world(), "masks": ctx.download_masks(),
               "views": [ctx.download_view(slot, v, st) for v in range(len(views))]}
        check_against_oracle(sc, views, got, ref)
    ctx.close()


def test_implicit_node_map_must_stay_in_last_level():
    """With node == NULL only the levels above node_base are composed before the instance kernel: a range that
    spans levels (or starts above the last one) is refused instead of reading rows nobody computed."""
    sc = synth.hierarchy_scene(n_roots=4, fan1=4, fan2=8, n_keys=40)
    ctx = S.SceneSubmit(sc.n_nodes, sc.n_nodes, len(sc.slots), max_views=1)
    ctx.set_key_layout(sc.mesh_bits, sc.submesh_bits, sc.material_bits)
    ctx.set_submesh_table(sc.slots)
    ctx.set_nodes(sc.n_nodes, sc.parent, sc.level_offsets)
    n_all = sc.n_nodes
    slot = np.zeros(n_all, np.uint32)
    mw = np.zeros(n_all, np.uint32)
    with pytest.raises(S.SceneSubmitError) as e:      # every node as an instance: spans three levels
        ctx.set_instances(slot, mw, None, 0)
    assert e.value.status == S.SS_ERR_INVALID
    with pytest.raises(S.SceneSubmitError):           # starts in the middle level
        ctx.set_instances(slot[:8], mw[:8], None, int(sc.level_offsets[1]))
    ctx.set_instances(sc.slot, sc.matword, None, sc.node_base)  # the leaves: fine
    # narrowing the key layout below what the streams hold is refused
    with pytest.raises(S.SceneSubmitError):
        ctx.set_key_layout(1, 1, 1)
    ctx.close()


def _ref_flat_frame(sc, views, trig_mode):
    """oracle/_ref (the reference's own GLM): per-instance masks and world rows of a flat or hierarchical scene."""
    import ctypes as C

    class Flat(C.Structure):
        _fields_ = [("n", C.c_uint32), ("trs", C.c_void_p), ("parent", C.c_void_p), ("n_nodes", C.c_uint32),
                    ("node_base", C.c_uint32), ("slot", C.c_void_p), ("matword", C.c_void_p), ("inst_aabb", C.c_void_p),
                    ("slot_local", C.c_void_p), ("slot_mesh_handle", C.c_void_p), ("slot_submesh", C.c_void_p),
                    ("material_handle", C.c_void_p), ("trig_mode", C.c_int)]
    n = sc.n_instances
    keep = [np.ascontiguousarray(x) for x in (sc.trs, sc.slot, sc.matword, sc.inst_aabb, sc.slots["local"],
                                              sc.mesh_handles[sc.slots["mesh_rank"]], sc.slots["submesh_index"], sc.material_handles)]
    parent = None if sc.parent is None else np.ascontiguousarray(sc.parent)
    p = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
    f = Flat(n, p(keep[0]), p(parent), sc.n_nodes, sc.node_base, p(keep[1]), p(keep[2]), p(keep[3]), p(keep[4]), p(keep[5]),
             p(keep[6]), p(keep[7]), trig_mode)
    vps, vfl = np.ascontiguousarray(views["view_proj"]), np.ascontiguousarray(views["flags"])
    vis, drw = np.zeros(len(views), np.uint32), np.zeros(len(views), np.uint32)
    masks, rows = np.zeros(n, np.uint8), np.zeros((n, 12), np.float32)
    O.ref().ssr_flat_frame(C.byref(f), p(vps), p(vfl), C.c_uint32(len(views)), C.c_int(os.cpu_count() or 1), p(vis), p(drw),
                           p(masks), p(rows))
    return masks, rows, vis


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref/libss_ref.so (the reference's GLM build) is not here")
def test_default_mode_against_the_reference_libm_arithmetic_at_full_size():
    """SURVEY §7 hard part 1 (iii), configs[2] at full size: the library's DEFAULT mode (deterministic trig) against the
    reference as it really computes (glm -> libm sinf/cosf, oracle/_ref).  World rows within 1e-6 relative; every
    visibility bit that differs is listed with the signed distance of the box from the deciding plane, which must be a
    rounding-sized sliver (|d| <= 1e-5 of the magnitudes that were added up)."""
    n = 1 << 24
    sc = synth.flat_scene(n)
    views = oracle_views(5)
    ref_masks, ref_rows, ref_vis = _ref_flat_frame(sc, views, trig_mode=1)
    ctx = S.SceneSubmit(sc.n_nodes, n, len(sc.slots), max_views=5, max_draws_per_view=1 << 16)
    ctx.load_scene(sc)
    st = ctx.frame_sync(ctx.frame(views))
    rows, masks = ctx.download_world(), ctx.download_masks()
    ctx.close()
    live = sc.slot != synth.SLOT_NULL
    scale = np.maximum(np.abs(ref_rows).max(axis=1, keepdims=True), 1e-30)
    rel = (np.abs(rows - ref_rows) / scale)[live]
    assert rel.max() <= 1e-6, f"world rows: {rel.max():.3g} relative"
    diff = np.flatnonzero(masks != ref_masks)
    assert len(diff) < n * 1e-4, f"{len(diff)} instances differ in visibility"
    worst = 0.0
    lines = []
    for i in diff:
        R = ref_rows[i].astype(np.float64).reshape(3, 4)
        mn, mx = sc.inst_aabb[i, :3].astype(np.float64), sc.inst_aabb[i, 3:].astype(np.float64)
        cl, el = (mn + mx) * 0.5, (mx - mn) * 0.5
        cw, ew = R[:, :3] @ cl + R[:, 3], np.abs(R[:, :3]) @ el
        for v in range(5):
            if ((int(masks[i]) ^ int(ref_masks[i])) >> v) & 1:
                planes = O.planes_from_view_proj(views["view_proj"][v]).astype(np.float64).reshape(6, 4)
                if views["flags"][v] & synth.VIEW_SKIP_NEAR:
                    planes = np.delete(planes, 4, axis=0)
                s = planes[:, :3] @ cw + planes[:, 3]
                r = np.abs(planes[:, :3]) @ ew
                mag = np.abs(planes[:, :3]) @ np.abs(cw) + np.abs(planes[:, 3]) + r
                k = int(np.argmin(np.abs(s + r) / mag))      # the plane that decides: the one nearest to the box's far corner
                d = float((s + r)[k] / mag[k])
                worst = max(worst, abs(d))
                lines.append(f"instance {i} view {v}: gpu {(int(masks[i]) >> v) & 1} reference {(int(ref_masks[i]) >> v) & 1} "
                             f"plane {k} signed distance {float((s + r)[k]):+.3e} ({d:+.2e} of the summed magnitudes)")
    print(f"{len(diff)} of {n} instances differ in a visibility bit (deterministic trig vs glm/libm); worst relative distance {worst:.2e}")
    for ln in lines[:40]:
        print("  " + ln)
    assert worst <= 1e-5, "a visibility bit differs on a box that is not borderline"
    # the lists' sizes move by exactly the listed bits
    for v in range(5):
        flips = int((((masks[diff] >> v) & 1).astype(np.int64) - ((ref_masks[diff] >> v) & 1).astype(np.int64)).sum())
        assert int(st.visible[v]) - int(ref_vis[v]) == flips


_HIER_REF = {}


@pytest.mark.parametrize("pipe", PIPELINES)
def test_full_size_hierarchy_variant_16m(pipe):
    """The 3-level variant of configs[2] at full size (2^24 leaves under 2^14 roots x 32 x 32, 17.3 M nodes, 5 views):
    world rows, masks bit-exact against the oracle; the lists sorted by (key, instance) and exactly the visible set."""
    sc = synth.hierarchy_scene(n_roots=1 << 14, seed=synth.SEED_BASE + 3)
    n = sc.n_instances
    views = oracle_views(5)
    ctx = S.SceneSubmit(sc.n_nodes, n, len(sc.slots), max_views=5, max_draws_per_view=1 << 16, flags=pipe)
    ctx.load_scene(sc)
    slot = ctx.frame(views)
    st = ctx.frame_sync(slot)
    assert int(st.pipeline) == EXPECT_PIPELINE[pipe]
    if "rows" not in _HIER_REF:
        import ctypes as C
        osc = O.OracleScene(sc, 0)
        rows = np.empty((n, 12), np.float32)
        nw = osc.node_worlds()
        O.lib().sso_instance_rows(C.byref(osc.s), nw.ctypes.data_as(C.c_void_p), rows.ctypes.data_as(C.c_void_p))
        del nw
        _HIER_REF.update(rows=rows, omask=osc.cull(rows, views), okeys=osc.keys())
    rows, omask, okeys = _HIER_REF["rows"], _HIER_REF["omask"], _HIER_REF["okeys"]
    masks = ctx.download_masks()
    assert np.array_equal(masks, omask)
    world = ctx.download_world()
    assert np.array_equal(world, rows)
    for v in range(5):
        got = ctx.download_view(slot, v, st)
        vis = np.flatnonzero((masks >> v) & 1).astype(np.uint32)
        assert np.array_equal(np.sort(got["idx"]), vis), f"view {v}: not exactly the visible set"
        assert np.array_equal(got["keys"], okeys[got["idx"]]), f"view {v}: keys"
        k, i = got["keys"], got["idx"].astype(np.int64)
        assert np.all((k[1:] > k[:-1]) | ((k[1:] == k[:-1]) & (i[1:] > i[:-1]))), f"view {v}: not sorted by (key, instance)"
        assert np.array_equal(got["rows"], world[got["idx"]]), f"view {v}: rows"
        d = got["draws"]
        assert int(d["instance_count"].sum()) == len(vis)
    ctx.close()


def test_lists_belong_to_their_slot_world_rows_to_the_last_frame():
    """The contract of ss_outputs with frames in flight: the per-view lists of slot f stay as they are while frames
    f+1, f+2 are issued (FramesInFlight, VC/Renderer/Renderer.h:28); the per-instance world rows and masks are ONE
    array that always holds the most recent frame."""
    sc = synth.flat_scene(30_000, n_keys=128)
    views = oracle_views(5)
    ctx = S.SceneSubmit(sc.n_nodes, sc.n_instances, len(sc.slots), max_views=5)
    ctx.load_scene(sc)
    variants, slots = [], []
    for k in range(3):  # three frames issued back to back on different transforms, nothing read in between
        t = sc.trs.copy()
        t[:, 0] += np.float32(40.0 * k)
        variants.append(t)
        ctx.upload_trs(t)
        slots.append(ctx.frame(views))
    refs = [O.OracleScene(synth.SceneArrays(**{**sc.__dict__, "trs": t}), 0).frame(views) for t in variants]
    for k in (0, 1, 2):
        st = ctx.frame_sync(slots[k])
        for v in range(5):
            got = ctx.download_view(slots[k], v, st)
            assert np.array_equal(got["idx"], refs[k]["views"][v]["idx"]) and np.array_equal(got["rows"], refs[k]["views"][v]["rows"])
    assert np.array_equal(ctx.download_world(), refs[2]["rows"]) and np.array_equal(ctx.download_masks(), refs[2]["masks"])
    ctx.close()


def test_pose_upload_physics_write_back():
    """SURVEY 8f-2, the producer's arithmetic on the device: (position, quaternion) of moved bodies ->
    TransformComponent translation + Euler rotation (JPH::Quat::GetEulerAngles), scale kept.  The frame that follows
    must equal the oracle's frame on transforms converted by the oracle's restatement, to the tolerance of the two
    libms (angles to 1e-6, rows to 1e-5 relative)."""
    import ctypes as C
    n = 70_000
    sc = synth.flat_scene(n, n_keys=50)
    views = oracle_views(5)
    ctx = S.SceneSubmit(sc.n_nodes, n, len(sc.slots), max_views=5)
    ctx.load_scene(sc)
    rng = np.random.default_rng(21)
    k = 20_000
    nodes = np.sort(rng.choice(n, k, replace=False)).astype(np.uint32)
    quat = rng.normal(size=(k, 4)).astype(np.float32)
    quat /= np.linalg.norm(quat, axis=1, keepdims=True).astype(np.float32)
    pos = rng.uniform(-400, 400, (k, 3)).astype(np.float32)
    ctx.upload_pose_sparse(nodes, np.concatenate([pos, quat], axis=1))
    st = ctx.frame_sync(ctx.frame(views))
    euler = np.empty((k, 3), np.float32)
    O.lib().sso_quat_to_euler(quat.ctypes.data_as(C.c_void_p), C.c_uint32(k), euler.ctypes.data_as(C.c_void_p))
    trs = sc.trs.copy()
    trs[nodes, 0:3] = pos
    trs[nodes, 3:6] = euler
    ref = O.OracleScene(synth.SceneArrays(**{**sc.__dict__, "trs": trs}), 0).frame(views)
    world = ctx.download_world()
    live = sc.slot != synth.SLOT_NULL
    scale = np.maximum(np.abs(ref["rows"]).max(axis=1, keepdims=True), 1e-30)
    assert (np.abs(world - ref["rows"]) <= 1e-5 * scale)[live].all()
    untouched = np.ones(n, bool)
    untouched[nodes] = False
    assert np.array_equal(world[untouched], ref["rows"][untouched])  # the others are bit-identical
    assert (ctx.download_masks() != ref["masks"]).mean() < 1e-3
    assert all(abs(int(st.visible[v]) - len(ref["views"][v]["idx"])) <= 20 for v in range(5))
    with pytest.raises(S.SceneSubmitError):
        ctx.upload_pose_sparse(np.array([n], np.uint32), np.zeros((1, 7), np.float32))
    ctx.close()


def test_pack_lights_from_device_transforms():
    """ss_pack_lights: the uniform blocks of Scene::UpdateLightsBuffer from the device-resident TRS stream, after a pose
    upload moved two of the lights (the host never saw their Euler angles): identical bytes to the oracle on the same
    transforms; more than ten lights of a kind are refused (the reference's arrays hold ten)."""
    import ctypes as C

    from test_oracle_vs_ref import _light_case, pack_lights_cpu
    sc = synth.flat_scene(2000, n_keys=16)
    ctx = S.SceneSubmit(sc.n_nodes, sc.n_instances, len(sc.slots), max_views=1)
    ctx.load_scene(sc)
    _, pn, pp, sn, spp = _light_case(3, n_nodes=sc.n_nodes)
    trs = sc.trs.copy()
    moved = np.unique(np.concatenate([pn[:1], sn[:1]])).astype(np.uint32)
    rec = trs[moved].copy()
    rec[:, :6] += np.float32(0.5)
    ctx.upload_trs_sparse(moved, rec)
    trs[moved] = rec
    got = ctx.pack_lights(pn, pp, sn, spp)
    ref = pack_lights_cpu(O.lib().sso_pack_lights, trs, pn, pp, sn, spp)
    for g, r in zip(got, ref):
        assert g.tobytes() == r.tobytes()
    with pytest.raises(S.SceneSubmitError):
        ctx.pack_lights(np.zeros(11, np.uint32), np.zeros((11, 6), np.float32), sn, spp)
    ctx.close()


@pytest.mark.parametrize("pipe", PIPELINES)
def test_select_view_records_carry_entity_ids(pipe):
    """SubmitSelectedMesh's instance stream (SelectTransformData, 52 B: rows + EntityID, Components.h:100-104) built on
    the device from a view's compacted list and the per-instance entity stream."""
    sc = synth.flat_scene(40_000, n_keys=100)
    views = oracle_views(5)
    ref = O.OracleScene(sc, 0).frame(views)
    ctx = S.SceneSubmit(sc.n_nodes, sc.n_instances, len(sc.slots), max_views=5, flags=pipe)
    ctx.load_scene(sc)
    ent = (np.arange(sc.n_instances, dtype=np.int64) * 7 + 3).astype(np.int32)  # any host numbering
    slot = ctx.frame(views)
    st = ctx.frame_sync(slot)
    rec = ctx.select_view(slot, 0, st)
    assert (rec["entity"] == -1).all() and np.array_equal(rec["rows"], ref["views"][0]["rows"])  # no entity stream yet
    ctx.set_instance_entities(ent)
    for v in (0, 4):
        rec = ctx.select_view(slot, v, st)
        assert rec.dtype.itemsize == 52
        assert np.array_equal(rec["rows"], ref["views"][v]["rows"])
        assert np.array_equal(rec["entity"], ent[ref["views"][v]["idx"]])
    ctx.close()
