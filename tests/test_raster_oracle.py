"""CPU checks of the rasterisation restatement itself (oracle/scenesubmit_oracle.c: sso_raster_view), which is the
checker of csrc/raster.cu: hand-computable answers, watertightness, order independence, near-plane behaviour."""
import numpy as np

import oracle_lib as O
from vulkanengine_b200 import synth

EMPTY = np.uint64(0xFFFFFFFFFFFFFFFF)


def unit_slots(n=1):
    slots = np.zeros(n, synth.SUBMESH_DTYPE)
    slots["local"] = np.eye(4, dtype=np.float32).reshape(16)
    slots["aabb_min"], slots["aabb_max"] = -1.0, 1.0
    return synth.box_geometry(slots)


def rows_of(translations, scale=1.0):
    rows = np.zeros((len(translations), 12), np.float32)
    for k, t in enumerate(translations):
        rows[k] = [scale, 0, 0, t[0], 0, scale, 0, t[1], 0, 0, scale, t[2]]
    return rows


def camera(pos=(0.0, 0.0, 5.0)):
    return O.camera_view_proj(synth.CAMERA_FOV, synth.CAMERA_ASPECT, synth.CAMERA_NEAR, synth.CAMERA_FAR, np.array(pos, np.float32), 0.0, 0.0)[2]


def depth_of(img):
    return (img >> np.uint64(32)).astype(np.uint32).view(np.float32)


def test_single_cube_centre_depth_and_silhouette():
    pos, ind, slots = unit_slots()
    W = H = 65
    img = O.raster_view(rows_of([(0, 0, 0)]), [0], [0], slots, pos, ind, camera(), W, H)
    n, f, d = 0.1, 1000.0, 4.0                     # the +z face is 4 in front of the camera
    assert abs(depth_of(img)[H // 2, W // 2] - f * (d - n) / (d * (f - n))) < 1e-6
    assert img[0, 0] == EMPTY and img[H - 1, W - 1] == EMPTY
    ys, xs = np.nonzero(img != EMPTY)
    half_h = (ys.max() - ys.min() + 1) / 2          # front face: half height 1 at distance 4, tan(22.5 deg) per unit
    assert abs(half_h - (H / 2) / (np.tan(np.radians(22.5)) * 4.0)) <= 1.0
    half_w = (xs.max() - xs.min() + 1) / 2
    assert abs(half_w - (W / 2) / (np.tan(np.radians(22.5)) * synth.CAMERA_ASPECT * 4.0)) <= 1.0


def test_shared_edges_are_owned_once_and_order_does_not_matter():
    """Two cubes side by side sharing a face plane: every covered sample has exactly one nearest owner whichever order
    the instances come in; inside one cube's front face the two triangles of the quad leave no hole along the diagonal."""
    pos, ind, slots = unit_slots()
    rows = rows_of([(-1, 0, 0), (1, 0, 0)])
    vp = camera((0.0, 0.0, 6.0))
    a = O.raster_view(rows, [0, 1], [0, 0], slots, pos, ind, vp, 96, 64)
    b = O.raster_view(rows[::-1].copy(), [1, 0], [0, 0], slots, pos, ind, vp, 96, 64)
    assert np.array_equal(a, b)
    covered = a != EMPTY
    ys, xs = np.nonzero(covered)
    box = covered[ys.min():ys.max() + 1, xs.min():xs.max() + 1]
    assert box.all()                                # the two front faces form one filled rectangle: no cracks
    ids = (a & np.uint64(0xFFFFFFFF))[covered]
    assert set(ids.tolist()) == {0, 1}


def test_near_plane_cut_and_depth_clamp():
    """A cube the camera looks at from inside its near half: without clamp nothing closer than the near plane is
    drawn (all depths in [0, 1]); with clamp the cut happens at w = 1e-5 and the nearest depths are clamped to 0."""
    pos, ind, slots = unit_slots()
    rows = rows_of([(0, 0, 4.5)])                   # cube spans z in [3.5, 5.5], the camera sits at z = 5 inside it
    vp = camera((0.0, 0.0, 5.0))
    plain = O.raster_view(rows, [0], [0], slots, pos, ind, vp, 48, 32, 0)
    clamp = O.raster_view(rows, [0], [0], slots, pos, ind, vp, 48, 32, 1)
    assert (plain != EMPTY).all() and (clamp != EMPTY).all()      # the far wall fills the view either way
    dp, dc = depth_of(plain), depth_of(clamp)
    assert dp.min() >= 0.0 and dp.max() <= 1.0 and dc.min() >= 0.0 and dc.max() <= 1.0
    # the far wall (z = 3.5, distance 1.5) is what the centre sees in both
    n, f, d = 0.1, 1000.0, 1.5
    want = f * (d - n) / (d * (f - n))
    assert abs(dp[16, 24] - want) < 1e-5 and abs(dc[16, 24] - want) < 1e-5


def test_empty_inputs_and_bad_tables():
    pos, ind, slots = unit_slots()
    img = O.raster_view(np.zeros((0, 12), np.float32), np.zeros(0, np.uint32), [0], slots, pos, ind, camera(), 16, 16)
    assert (img == EMPTY).all()
    bad = slots.copy()
    bad["first_index"] = 30                          # runs past the 36 indices: those triangles draw nothing
    img = O.raster_view(rows_of([(0, 0, 0)]), [0], [0], bad, pos, ind, camera(), 33, 33)
    whole = O.raster_view(rows_of([(0, 0, 0)]), [0], [0], slots, pos, ind, camera(), 33, 33)
    assert np.count_nonzero(img != EMPTY) < np.count_nonzero(whole != EMPTY)
