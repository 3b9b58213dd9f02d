"""ctypes bindings of the CPU oracle (oracle/liboracle.so) and of the reference-arithmetic build
(oracle/_ref/libss_ref.so).  Test infrastructure only: imported by tests/, smoke() and the
cpu_baseline / --impl reference legs of bench.py, never by the product package."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

SLOT_NULL = 0xFFFFFFFF
TESS_BIT = 0x80000000

f32p = C.POINTER(C.c_float)
u8p = C.POINTER(C.c_uint8)
u32p = C.POINTER(C.c_uint32)
i32p = C.POINTER(C.c_int32)
u64p = C.POINTER(C.c_uint64)


class Submesh(C.Structure):
    _fields_ = [("local", C.c_float * 16), ("aabb_min", C.c_float * 3), ("aabb_max", C.c_float * 3),
                ("index_count", C.c_uint32), ("first_index", C.c_uint32), ("vertex_offset", C.c_int32),
                ("mesh_rank", C.c_uint32), ("submesh_index", C.c_uint32), ("reserved", C.c_uint32)]


SUBMESH_DTYPE = np.dtype([("local", "<f4", (16,)), ("aabb_min", "<f4", (3,)), ("aabb_max", "<f4", (3,)),
                          ("index_count", "<u4"), ("first_index", "<u4"), ("vertex_offset", "<i4"),
                          ("mesh_rank", "<u4"), ("submesh_index", "<u4"), ("reserved", "<u4")])
VIEW_DTYPE = np.dtype([("view_proj", "<f4", (16,)), ("flags", "<u4")])
DRAW_DTYPE = np.dtype([("index_count", "<u4"), ("instance_count", "<u4"), ("first_index", "<u4"),
                       ("vertex_offset", "<i4"), ("first_instance", "<u4")])
assert SUBMESH_DTYPE.itemsize == C.sizeof(Submesh) == 112
assert VIEW_DTYPE.itemsize == 68 and DRAW_DTYPE.itemsize == 20


class Scene(C.Structure):
    _fields_ = [("n_nodes", C.c_uint32), ("trs", C.c_void_p), ("parent", C.c_void_p),
                ("n_instances", C.c_uint32), ("node", C.c_void_p), ("node_base", C.c_uint32),
                ("slot", C.c_void_p), ("matword", C.c_void_p), ("inst_aabb", C.c_void_p),
                ("n_slots", C.c_uint32), ("slots", C.c_void_p),
                ("mesh_bits", C.c_uint32), ("submesh_bits", C.c_uint32), ("material_bits", C.c_uint32),
                ("trig_mode", C.c_uint32)]


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def build(force=False):
    so = os.path.join(ORACLE_DIR, "liboracle.so")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(
            os.path.join(ORACLE_DIR, "scenesubmit_oracle.c")):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "all"], stdout=subprocess.DEVNULL)
    return so


_lib = None
_ref = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.sso_frame.restype = C.c_double
    return _lib


def ref_available():
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libss_ref.so"))


def ref():
    global _ref
    if _ref is None:
        _ref = C.CDLL(os.path.join(ORACLE_DIR, "_ref", "libss_ref.so"))
        _ref.ssr_flat_frame.restype = C.c_double
        _ref.ssr_submit_pass.restype = C.c_uint32
    return _ref


class OracleScene:
    """Holds numpy arrays alive and exposes the sso_scene struct. `sc` is a
    vulkanengine_b200.synth.SceneArrays-like object (duck-typed attributes)."""

    def __init__(self, sc, trig_mode=0):
        self.keep = sc
        self.trs = np.ascontiguousarray(sc.trs, dtype=np.float32)
        self.parent = None if sc.parent is None else np.ascontiguousarray(sc.parent, dtype=np.int32)
        self.node = None if sc.node is None else np.ascontiguousarray(sc.node, dtype=np.uint32)
        self.slot = np.ascontiguousarray(sc.slot, dtype=np.uint32)
        self.matword = np.ascontiguousarray(sc.matword, dtype=np.uint32)
        self.inst_aabb = None if sc.inst_aabb is None else np.ascontiguousarray(sc.inst_aabb, dtype=np.float32)
        self.slots = np.ascontiguousarray(sc.slots, dtype=SUBMESH_DTYPE)
        self.n = int(self.slot.shape[0])
        self.s = Scene(int(self.trs.shape[0]), _ptr(self.trs), _ptr(self.parent), self.n, _ptr(self.node),
                       int(sc.node_base), _ptr(self.slot), _ptr(self.matword), _ptr(self.inst_aabb),
                       int(self.slots.shape[0]), _ptr(self.slots), int(sc.mesh_bits), int(sc.submesh_bits),
                       int(sc.material_bits), int(trig_mode))

    def node_worlds(self):
        out = np.empty((self.s.n_nodes, 16), np.float32)
        lib().sso_node_worlds(C.byref(self.s), _ptr(out))
        return out

    def instance_rows(self, node_world=None):
        if node_world is None:
            node_world = self.node_worlds()
        rows = np.empty((self.n, 12), np.float32)
        lib().sso_instance_rows(C.byref(self.s), _ptr(node_world), _ptr(rows))
        return rows

    def cull(self, rows, views):
        views = np.ascontiguousarray(views, dtype=VIEW_DTYPE)
        masks = np.empty(self.n, np.uint8)
        lib().sso_cull(C.byref(self.s), _ptr(np.ascontiguousarray(rows, np.float32)), _ptr(views),
                       C.c_uint32(len(views)), _ptr(masks))
        return masks

    def keys(self):
        keys = np.empty(self.n, np.uint64)
        lib().sso_keys(C.byref(self.s), _ptr(keys))
        return keys

    def frame(self, views, threads=1, want_rows=True):
        """Whole frame. Returns dict with rows, masks, keys and per-view lists."""
        views = np.ascontiguousarray(views, dtype=VIEW_DTYPE)
        nv, n = len(views), self.n
        rows = np.empty((n, 12), np.float32)
        masks = np.empty(n, np.uint8)
        keys = np.empty(n, np.uint64)
        n_vis = np.zeros(nv, np.uint32)
        n_draws = np.zeros(nv, np.uint32)
        out_idx = np.empty((nv, n), np.uint32)
        out_keys = np.empty((nv, n), np.uint64)
        out_rows = np.empty((nv, n, 12), np.float32) if want_rows else None
        out_draws = np.empty((nv, n), DRAW_DTYPE)
        out_dkeys = np.empty((nv, n), np.uint64)
        secs = lib().sso_frame(C.byref(self.s), _ptr(views), C.c_uint32(nv), C.c_int(threads), _ptr(rows),
                               _ptr(masks), _ptr(keys), _ptr(n_vis), _ptr(n_draws), _ptr(out_idx),
                               _ptr(out_keys), _ptr(out_rows), _ptr(out_draws), _ptr(out_dkeys))
        res = {"rows": rows, "masks": masks, "keys": keys, "seconds": secs, "views": []}
        for v in range(nv):
            m, d = int(n_vis[v]), int(n_draws[v])
            res["views"].append({"idx": out_idx[v, :m], "keys": out_keys[v, :m],
                                 "rows": None if out_rows is None else out_rows[v, :m],
                                 "draws": out_draws[v, :d], "draw_keys": out_dkeys[v, :d]})
        return res


def planes_from_view_proj(vp):
    vp = np.ascontiguousarray(vp, np.float32)
    out = np.empty((6, 4), np.float32)
    lib().sso_planes_from_view_proj(_ptr(vp), _ptr(out))
    return out


def camera_view_proj(fov, aspect, zn, zf, pos, pitch, yaw, which="oracle"):
    pos = np.asarray(pos, np.float32)
    v, p, vp = (np.empty(16, np.float32) for _ in range(3))
    fn = lib().sso_camera_view_proj if which == "oracle" else ref().ssr_camera_view_proj
    fn(C.c_float(fov), C.c_float(aspect), C.c_float(zn), C.c_float(zf), _ptr(pos), C.c_float(pitch),
       C.c_float(yaw), _ptr(v), _ptr(p), _ptr(vp))
    return v, p, vp


def cascade_matrices(vp, zn, zf, lam, near_off, far_off, light_dir, which="oracle", origin=None, map_size=0.0):
    """origin given: the CascadeOffset branch (texel snap of CascadeOrigin on a map_size^2 shadow map)."""
    vp = np.ascontiguousarray(vp, np.float32)
    ld = np.asarray(light_dir, np.float32)
    org = None if origin is None else np.ascontiguousarray(origin, np.float32)
    splits = np.empty(4, np.float32)
    mats = np.empty((4, 16), np.float32)
    fn = lib().sso_cascade_matrices_snapped if which == "oracle" else ref().ssr_cascade_matrices_snapped
    fn(_ptr(vp), C.c_float(zn), C.c_float(zf), C.c_float(lam), C.c_float(near_off), C.c_float(far_off),
       _ptr(ld), _ptr(org), C.c_float(map_size), _ptr(splits), _ptr(mats))
    return splits, mats


def trs_to_mat4(trs9, trig_mode, which="oracle"):
    trs9 = np.ascontiguousarray(trs9, np.float32)
    out = np.empty(16, np.float32)
    (lib().sso_trs_to_mat4 if which == "oracle" else ref().ssr_trs_to_mat4)(_ptr(trs9), C.c_int(trig_mode), _ptr(out))
    return out


def det_sincos(x):
    s, c = C.c_float(), C.c_float()
    lib().sso_det_sincos(C.c_float(float(x)), C.byref(s), C.byref(c))
    return s.value, c.value


def raster_view(rows12, instance_idx, slot_of_instance, slots, positions, indices, view_proj, width, height, flags=0):
    """sso_raster_view: the CPU restatement of the consumer pass (every sample of the target is tested against every
    triangle, so keep targets and scenes small)."""
    rows12 = np.ascontiguousarray(rows12, np.float32).reshape(-1, 12)
    instance_idx = np.ascontiguousarray(instance_idx, np.uint32)
    slot_of_instance = np.ascontiguousarray(slot_of_instance, np.uint32)
    slots = np.ascontiguousarray(slots)
    positions = np.ascontiguousarray(positions, np.float32).reshape(-1, 3)
    indices = np.ascontiguousarray(indices, np.uint32)
    vp = np.ascontiguousarray(view_proj, np.float32).reshape(16)
    out = np.empty((height, width), np.uint64)
    lib().sso_raster_view(_ptr(rows12), _ptr(instance_idx), C.c_uint32(len(instance_idx)), _ptr(slot_of_instance), _ptr(slots),
                          _ptr(positions), C.c_uint32(len(positions)), _ptr(indices), C.c_uint32(len(indices)), _ptr(vp),
                          C.c_uint32(width), C.c_uint32(height), C.c_uint32(flags), _ptr(out))
    return out
