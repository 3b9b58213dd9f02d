"""SURVEY.md 8f-4 — the consumer of the lists: depth + id rasterisation of a view's compacted instances
(vulkanengine_b200/csrc/raster.cu, through the C ABI) against the CPU restatement sso_raster_view, bit for bit, plus
hand-checkable answers (the SelectionPass read-back of one pixel).  The reference's own rasteriser is the Vulkan
device: unavailable on this box, so parity with it stays unpinned; what is pinned is CUDA == restatement."""
import numpy as np
import pytest

import oracle_lib as O
from vulkanengine_b200 import synth

pytestmark = pytest.mark.gpu
S = None


def setup_module(_m):
    global S
    from vulkanengine_b200 import scenesubmit
    S = scenesubmit


def close_up_scene(n, spread=30.0, seed_shift=0):
    """synthetic flat scene pulled in around the origin so that boxes cover pixels of a small target"""
    sc = synth.flat_scene(n, n_keys=40, seed=synth.SEED_BASE + 9 + seed_shift)
    trs = sc.trs.copy()
    trs[:, :3] *= np.float32(spread / 500.0)
    pos, ind, slots = synth.box_geometry(sc.slots)
    return synth.SceneArrays(**{**sc.__dict__, "trs": trs, "slots": slots}), pos, ind


def views_for(position, n_views=5):
    return synth.make_views(lambda *a: O.camera_view_proj(*a), lambda *a: O.cascade_matrices(*a), n_views, position=position)


def run(sc, pos, ind, views, W, H, flags_of_view):
    ctx = S.SceneSubmit(sc.n_nodes, sc.n_instances, len(sc.slots), max_views=len(views))
    ctx.load_scene(sc)
    ctx.set_geometry(pos, ind)
    slot = ctx.frame(views)
    st = ctx.frame_sync(slot)
    out = []
    for v in range(len(views)):
        got_view = ctx.download_view(slot, v, st)
        ctx.raster_view(slot, v, W, H, flags_of_view(v))
        img = ctx.download_raster(v)
        want = O.raster_view(got_view["rows"], got_view["idx"], sc.slot, sc.slots, pos, ind, views["view_proj"][v], W, H, flags_of_view(v))
        out.append((img, want, got_view))
    return ctx, slot, out


def test_raster_matches_restatement_camera_and_cascades():
    sc, pos, ind = close_up_scene(400)
    views = views_for((0.0, 0.0, 70.0))
    ctx, slot, out = run(sc, pos, ind, views, 96, 64, lambda v: 0 if v == 0 else S.RASTER_DEPTH_CLAMP)
    covered = 0
    for v, (img, want, _) in enumerate(out):
        assert np.array_equal(img, want), f"view {v}: {np.count_nonzero(img != want)} samples differ"
        covered += int(np.count_nonzero(img != np.uint64(S.RASTER_EMPTY)))
    assert covered > 2000
    # the read-back of one sample = the word in the image
    img = out[0][0]
    ys, xs = np.nonzero(img != np.uint64(S.RASTER_EMPTY))
    inst, depth = ctx.read_pixel(0, int(xs[0]), int(ys[0]))
    assert inst == int(img[ys[0], xs[0]] & np.uint64(0xFFFFFFFF))
    assert np.float32(depth).view(np.uint32) == np.uint32(img[ys[0], xs[0]] >> np.uint64(32))
    assert ctx.read_pixel(0, 0, 0)[0] in (0xFFFFFFFF, int(img[0, 0] & np.uint64(0xFFFFFFFF)))
    ctx.close()


def test_single_cube_known_answer():
    """A unit cube at the origin seen from (0, 0, 5) with the editor camera (fov 45 deg, near 0.1, far 1000): the centre
    sample hits the +z face at view distance 4; Vulkan ZO depth = f (d - n) / (d (f - n))."""
    sc = synth.flat_scene(1, n_keys=1)
    trs = np.array([[0, 0, 0, 0, 0, 0, 1, 1, 1]], np.float32)
    slots = sc.slots.copy()
    slots["local"] = np.eye(4, dtype=np.float32).reshape(16)
    slots["aabb_min"], slots["aabb_max"] = -1.0, 1.0
    pos, ind, slots = synth.box_geometry(slots)
    sc = synth.SceneArrays(**{**sc.__dict__, "trs": trs, "slots": slots, "slot": np.zeros(1, np.uint32),
                              "inst_aabb": None, "matword": sc.matword[:1] & np.uint32(synth.TESS_BIT - 1)})
    views = views_for((0.0, 0.0, 5.0), 1)
    W = H = 65
    ctx, slot, out = run(sc, pos, ind, views, W, H, lambda v: 0)
    img, want, _ = out[0]
    assert np.array_equal(img, want)
    inst, depth = ctx.read_pixel(0, W // 2, H // 2)
    n, f, d = 0.1, 1000.0, 4.0
    assert inst == 0 and abs(depth - f * (d - n) / (d * (f - n))) < 1e-6
    assert ctx.read_pixel(0, 0, 0)[0] == 0xFFFFFFFF and ctx.read_pixel(0, W - 1, H - 1)[0] == 0xFFFFFFFF
    # the front face spans tan(22.5 deg) * 4 = 1.657 half-height: 1 / 1.657 of the half-target
    ys, xs = np.nonzero(img != np.uint64(S.RASTER_EMPTY))
    half = (ys.max() - ys.min() + 1) / 2
    assert abs(half - (H / 2) / (np.tan(np.radians(22.5)) * 4.0)) <= 1.0
    ctx.close()


def test_camera_inside_a_box_needs_no_clipper():
    """Triangles crossing w = 0 (the camera sits inside a large box): every sample is covered, and identically."""
    sc = synth.flat_scene(3, n_keys=2)
    trs = np.zeros((3, 9), np.float32)
    trs[:, 6:] = 1.0
    trs[1, :3] = (3.0, 0.5, -6.0)
    trs[2, :3] = (-2.0, -1.0, -9.0)
    slots = sc.slots.copy()
    slots["local"] = np.eye(4, dtype=np.float32).reshape(16)
    slots["aabb_min"], slots["aabb_max"] = -1.0, 1.0
    slots["aabb_min"][0], slots["aabb_max"][0] = -40.0, 40.0   # slot 0: the room
    pos, ind, slots = synth.box_geometry(slots)
    slot_of = np.array([0, 1 % len(slots), 1 % len(slots)], np.uint32)
    sc = synth.SceneArrays(**{**sc.__dict__, "trs": trs, "slots": slots, "slot": slot_of, "inst_aabb": None,
                              "matword": sc.matword[:3] & np.uint32(synth.TESS_BIT - 1)})
    views = views_for((0.0, 0.0, 5.0), 1)
    ctx, slot, out = run(sc, pos, ind, views, 80, 48, lambda v: 0)
    img, want, got_view = out[0]
    assert len(got_view["idx"]) == 3
    assert np.array_equal(img, want)
    assert np.all(img != np.uint64(S.RASTER_EMPTY))                 # the room's walls are everywhere behind the cubes
    ids = np.unique(img & np.uint64(0xFFFFFFFF))
    assert set(ids.tolist()) == {0, 1, 2}
    ctx.close()


def test_far_small_triangles_at_scene_scale():
    """The shape of configs[2]: boxes of a few units seen from 600 away land on a handful of samples each; coverage is
    decided by exact integer edge functions on snapped vertices, so it cannot depend on float cancellation."""
    sc = synth.flat_scene(120_000, n_keys=300)
    pos, ind, slots = synth.box_geometry(sc.slots)
    sc = synth.SceneArrays(**{**sc.__dict__, "slots": slots})
    views = views_for((0.0, 0.0, 600.0))
    ctx, slot, out = run(sc, pos, ind, views, 512, 320, lambda v: 0 if v == 0 else S.RASTER_DEPTH_CLAMP)
    for v, (img, want, got_view) in enumerate(out):
        assert np.array_equal(img, want), f"view {v}: {np.count_nonzero(img != want)} samples differ"
    assert np.count_nonzero(out[0][0] != np.uint64(S.RASTER_EMPTY)) > 10_000
    ctx.close()


def test_near_plane_cuts_triangles_and_clamp_keeps_them():
    """Boxes straddling the camera's near plane: without DepthClamp the part in front of it is cut away (new vertices
    interpolated from the inside vertex), with DepthClamp the triangle is cut at w = 1e-5 instead and what is left is
    rasterised in homogeneous coordinates when its projection cannot be snapped."""
    sc = synth.flat_scene(40, n_keys=6)
    trs = sc.trs.copy()
    rng = np.random.default_rng(3)
    trs[:, :3] = rng.uniform(-3, 3, (40, 3)).astype(np.float32)
    trs[:, 2] = rng.uniform(2.0, 8.0, 40).astype(np.float32)       # camera at z = 5 looking down -z: some in front, some around it
    pos, ind, slots = synth.box_geometry(sc.slots)
    sc = synth.SceneArrays(**{**sc.__dict__, "trs": trs, "slots": slots, "inst_aabb": None})
    views = views_for((0.0, 0.0, 5.0), 1)
    for flags in (0, S.RASTER_DEPTH_CLAMP):
        ctx, slot, out = run(sc, pos, ind, views, 72, 40, lambda v: flags)
        img, want, got_view = out[0]
        assert len(got_view["idx"]) > 5
        assert np.array_equal(img, want), f"flags {flags}: {np.count_nonzero(img != want)} samples differ"
        assert np.count_nonzero(img != np.uint64(S.RASTER_EMPTY)) > 500
        ctx.close()


def test_geometry_out_of_range_draws_nothing_and_errors_are_reported():
    sc, pos, ind = close_up_scene(50, seed_shift=1)
    views = views_for((0.0, 0.0, 70.0), 1)
    ctx = S.SceneSubmit(sc.n_nodes, sc.n_instances, len(sc.slots), max_views=1)
    ctx.load_scene(sc)
    ctx.set_geometry(pos[:4], ind[:6])          # far too short for the table
    slot = ctx.frame(views)
    ctx.frame_sync(slot)
    ctx.raster_view(slot, 0, 32, 32)
    img = ctx.download_raster(0)
    want = O.raster_view(*[ctx.download_view(slot, 0, ctx.frame_sync(slot))[k] for k in ("rows", "idx")], sc.slot, sc.slots,
                         pos[:4], ind[:6], views["view_proj"][0], 32, 32)
    assert np.array_equal(img, want)
    with pytest.raises(S.SceneSubmitError):
        ctx.raster_view(slot, 0, 0, 32)
    with pytest.raises(S.SceneSubmitError):
        ctx.read_pixel(0, 32, 0)
    ctx.close()


@pytest.mark.parametrize("case", range(int(__import__("os").environ.get("SS_SWEEP_CASES", "10"))))
def test_randomised_raster(case):
    """Seeded sweep: scene spread, camera distance and direction, target size, clamp flag, number of instances — the
    image must equal the CPU restatement every time (small triangles, large ones, cut ones, queued ones)."""
    rng = np.random.default_rng(500 + case)
    n = int(rng.integers(1, 1500))
    spread = float(rng.choice([3.0, 30.0, 300.0]))
    sc, pos, ind = close_up_scene(n, spread=spread, seed_shift=case)
    dist = float(rng.choice([2.0, 10.0, 80.0, 600.0]))
    views = synth.fibonacci_cameras(lambda *a: O.camera_view_proj(*a), 3, radius=dist, seed=synth.SEED_BASE + case)
    W, H = int(rng.integers(8, 200)), int(rng.integers(8, 120))
    flags = [0, S.RASTER_DEPTH_CLAMP][int(rng.integers(0, 2))]
    ctx, slot, out = run(sc, pos, ind, views, W, H, lambda v: flags)
    for v, (img, want, _) in enumerate(out):
        assert np.array_equal(img, want), f"case {case} view {v}: {np.count_nonzero(img != want)} of {W * H} samples differ"
    ctx.close()


@pytest.mark.parametrize("flags_name", ["default", "depth_clamp"])
def test_heavy_overdraw_is_drawn_in_phases(flags_name):
    """Thousands of boxes around the camera, each covering a good part of a small target: the pass runs front to back in
    phases and skips what the coarse depth already hides (csrc/raster.cu, launch_raster).  What is skipped depends on
    the order; the image must not — and with DepthClamp most of it sits at depth 0, where the lowest instance wins."""
    flags = S.RASTER_DEPTH_CLAMP if flags_name == "depth_clamp" else 0
    sc, pos, ind = close_up_scene(3000, spread=30.0, seed_shift=77)
    views = synth.fibonacci_cameras(lambda *a: O.camera_view_proj(*a), 2, radius=10.0, seed=synth.SEED_BASE + 77)
    ctx, slot, out = run(sc, pos, ind, views, 160, 120, lambda v: flags)
    for v, (img, want, got_view) in enumerate(out):
        assert len(got_view["idx"]) >= 256, "too few instances for more than one phase"
        assert np.array_equal(img, want), f"view {v}: {np.count_nonzero(img != want)} samples differ"
        assert np.count_nonzero(img != np.uint64(S.RASTER_EMPTY)) > 160 * 120 // 2
    ctx.close()


@pytest.mark.parametrize("flags_name", ["default", "depth_clamp"])
def test_large_target_exercises_tiles_and_the_queue(flags_name):
    """A 512 x 384 target seen from inside the cloud: triangles of tens of thousands of samples go through the tile queue,
    whose entries are filtered against the 64 x 64 level of the coarse depth as they are queued, the middle sizes are
    sampled cell by cell; several phases.  Bit-equal to the restatement."""
    flags = S.RASTER_DEPTH_CLAMP if flags_name == "depth_clamp" else 0
    sc, pos, ind = close_up_scene(2400, spread=30.0, seed_shift=91)
    views = synth.fibonacci_cameras(lambda *a: O.camera_view_proj(*a), 1, radius=8.0, seed=synth.SEED_BASE + 91)
    ctx, slot, out = run(sc, pos, ind, views, 512, 384, lambda v: flags)
    img, want, got_view = out[0]
    assert len(got_view["idx"]) >= 256
    assert np.array_equal(img, want), f"{np.count_nonzero(img != want)} samples differ"
    assert np.count_nonzero(img != np.uint64(S.RASTER_EMPTY)) > 512 * 384 // 2
    ctx.close()
