"""ctypes binding of libscenesubmit.so (include/scenesubmit.h) — the same entry points the C++ host
mirror and a reference-side maintainer would bind.  There is no CPU fallback: if the CUDA library
cannot be loaded, importing the product path fails loudly."""
import ctypes as C
import os

import numpy as np

from .synth import DRAW_DTYPE, SUBMESH_DTYPE, VIEW_DTYPE

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libscenesubmit.so")

SS_OK, SS_ERR_INVALID, SS_ERR_CUDA, SS_ERR_CAPACITY, SS_ERR_STATE, SS_ERR_NOMEM = range(6)
TRIG_DETERMINISTIC, TRIG_LIBM = 0, 1
FRAME_WORLD_ROWS_GIVEN = 1
FRAME_REUSE_WORLD = 4
FRAME_REUSE_SORT = 8
CONFIG_DRAW_ORDER_RESIDENT = 2
PIPELINE_GENERAL, PIPELINE_RESIDENT = 0, 2
PIPELINE_NAMES = {0: "general", 2: "resident"}
MAX_VIEWS = 8
SLOT_NULL = 0xFFFFFFFF

# every symbol include/scenesubmit.h declares (tests check the library exports exactly these)
EXPORTS = ["ss_create", "ss_destroy", "ss_last_error", "ss_version", "ss_set_key_layout", "ss_set_submesh_table",
           "ss_set_nodes", "ss_set_instances", "ss_set_instance_aabbs", "ss_upload_trs", "ss_upload_node_world_rows", "ss_flush_uploads",
           "ss_upload_world_rows", "ss_frame", "ss_frame_sync", "ss_frame_stage_ms", "ss_map_outputs", "ss_download_view",
           "ss_download_world", "ss_download_masks", "ss_stream", "ss_sort_scratch_bytes", "ss_sort_pairs",
           "ss_sort_pairs_host", "ss_sort_pairs_static", "ss_camera_view_proj", "ss_cascade_matrices",
           "ss_set_geometry", "ss_raster_view", "ss_map_raster", "ss_download_raster", "ss_read_pixel",
           "ss_cascade_matrices_batch", "ss_upload_trs_sparse", "ss_merge_place",
           "ss_export_idx_lists", "ss_ipc_open", "ss_ipc_close", "ss_merge_place_peers",
           "ss_merge_create", "ss_merge_destroy", "ss_merge_views", "ss_merge_stage_ms", "ss_merge_map", "ss_merge_last_error", "ss_merge_download",
           "ss_export_slot_buffers", "ss_cascade_matrices_snapped", "ss_upload_pose_sparse", "ss_pack_lights", "ss_set_instance_entities", "ss_select_view"]
RASTER_DEPTH_CLAMP = 1
RASTER_EMPTY = 0xFFFFFFFFFFFFFFFF


class SceneSubmitError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__(f"libscenesubmit status {status}: {msg}")
        self.status = status


class Config(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("device", C.c_int32), ("max_nodes", C.c_uint32),
                ("max_instances", C.c_uint32), ("max_submesh_slots", C.c_uint32), ("max_views", C.c_uint32),
                ("max_draws_per_view", C.c_uint32), ("frames_in_flight", C.c_uint32),
                ("max_visible_per_view", C.c_uint64), ("trig_mode", C.c_uint32), ("flags", C.c_uint32)]


class FrameStats(C.Structure):
    _fields_ = [("n_instances", C.c_uint32), ("n_views", C.c_uint32), ("visible", C.c_uint32 * MAX_VIEWS),
                ("draws", C.c_uint32 * MAX_VIEWS), ("sort_passes_run", C.c_uint32), ("overflow", C.c_uint32),
                ("gpu_ms", C.c_float), ("kernel_launches", C.c_uint32), ("pipeline", C.c_uint32)]


class Outputs(C.Structure):
    _fields_ = [("instance_rows", C.c_void_p), ("keys", C.c_void_p), ("instance_idx", C.c_void_p),
                ("draws", C.c_void_p), ("draw_keys", C.c_void_p), ("counts", C.c_void_p), ("world_rows", C.c_void_p),
                ("masks", C.c_void_p), ("visible", C.c_uint32), ("n_draws", C.c_uint32), ("world_words", C.c_void_p),
                ("world_idx_bits", C.c_uint32), ("world_stride", C.c_uint32)]


class MergePeer(C.Structure):
    _fields_ = [("idx", C.c_void_p), ("draws", C.c_void_p), ("draw_keys", C.c_void_p), ("counts", C.c_void_p),
                ("instance_base", C.c_uint64)]


class MergedView(C.Structure):
    _fields_ = [("instance_idx", C.c_void_p), ("keys", C.c_void_p), ("draws", C.c_void_p), ("draw_keys", C.c_void_p),
                ("visible", C.c_uint32), ("n_draws", C.c_uint32)]


_lib = None


def load_library():
    """Loads libscenesubmit.so or raises: the product has no other implementation to fall back to."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: run `make` (nvcc, sm_100a). libscenesubmit has no CPU fallback.")
        lib = C.CDLL(LIB_PATH)
        lib.ss_last_error.restype = C.c_char_p
        lib.ss_last_error.argtypes = [C.c_void_p]
        lib.ss_version.restype = C.c_uint32
        lib.ss_stream.restype = C.c_void_p
        lib.ss_stream.argtypes = [C.c_void_p]
        lib.ss_sort_scratch_bytes.restype = C.c_size_t
        lib.ss_sort_scratch_bytes.argtypes = [C.c_uint64]
        lib.ss_destroy.argtypes = [C.c_void_p]
        lib.ss_destroy.restype = None
        lib.ss_merge_destroy.argtypes = [C.c_void_p]
        lib.ss_merge_destroy.restype = None
        lib.ss_merge_last_error.argtypes = [C.c_void_p]
        lib.ss_merge_last_error.restype = C.c_char_p
        _lib = lib
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def camera_view_proj(fov, aspect, zn, zf, pos, pitch, yaw):
    lib = load_library()
    pos = np.asarray(pos, np.float32)
    v, p, vp = (np.empty(16, np.float32) for _ in range(3))
    st = lib.ss_camera_view_proj(C.c_float(fov), C.c_float(aspect), C.c_float(zn), C.c_float(zf), _p(pos),
                                 C.c_float(pitch), C.c_float(yaw), _p(v), _p(p), _p(vp))
    if st != SS_OK:
        raise SceneSubmitError(st, "ss_camera_view_proj")
    return v, p, vp


def cascade_matrices(vp, zn, zf, lam, near_off, far_off, light_dir, origin=None, map_size=0.0):
    lib = load_library()
    vp = np.ascontiguousarray(vp, np.float32)
    ld = np.asarray(light_dir, np.float32)
    org = None if origin is None else np.ascontiguousarray(origin, np.float32)
    splits = np.empty(4, np.float32)
    mats = np.empty((4, 16), np.float32)
    st = lib.ss_cascade_matrices_snapped(_p(vp), C.c_float(zn), C.c_float(zf), C.c_float(lam), C.c_float(near_off),
                                         C.c_float(far_off), _p(ld), _p(org), C.c_float(map_size), _p(splits), _p(mats))
    if st != SS_OK:
        raise SceneSubmitError(st, "ss_cascade_matrices")
    return splits, mats


def cascade_matrices_batch(view_projs, zn, zf, lam, near_off, far_off, light_dir, device=0):
    """[n][16] camera view-projections -> ([4] split distances, [n][4][16] light matrices), computed on the GPU."""
    lib = load_library()
    vps = np.ascontiguousarray(view_projs, np.float32).reshape(-1, 16)
    out = np.empty((len(vps), 4, 16), np.float32)
    splits = np.empty(4, np.float32)
    ld = np.ascontiguousarray(light_dir, np.float32)
    st = lib.ss_cascade_matrices_batch(C.c_int(device), _p(vps), C.c_uint32(len(vps)), C.c_float(zn), C.c_float(zf), C.c_float(lam),
                                       C.c_float(near_off), C.c_float(far_off), _p(ld), _p(splits), _p(out))
    if st != SS_OK:
        raise SceneSubmitError(st, "ss_cascade_matrices_batch")
    return splits, out


def ipc_open(device, handle):
    lib = load_library()
    buf = (C.c_ubyte * 64).from_buffer_copy(handle)
    ptr = C.c_void_p()
    st = lib.ss_ipc_open(C.c_int(device), buf, C.byref(ptr))
    if st != SS_OK:
        raise SceneSubmitError(st, "ss_ipc_open: " + lib.ss_last_error(None).decode())
    return int(ptr.value)


def ipc_close(device, ptr):
    load_library().ss_ipc_close(C.c_int(device), C.c_void_p(ptr))


def sort_pairs_host(keys, vals, begin_bit=0, end_bit=64, device=0):
    lib = load_library()
    keys = np.ascontiguousarray(keys, np.uint64)
    vals = np.ascontiguousarray(vals, np.uint32)
    ko, vo = np.empty_like(keys), np.empty_like(vals)
    st = lib.ss_sort_pairs_host(C.c_int(device), _p(keys), _p(vals), _p(ko), _p(vo), C.c_uint64(len(keys)),
                                C.c_int(begin_bit), C.c_int(end_bit))
    if st != SS_OK:
        raise SceneSubmitError(st, lib.ss_last_error(None).decode())
    return ko, vo


class Merger:
    """ss_merge_*: merged draw lists of several shards (ranks), every view in one call."""

    def __init__(self, n_ranks, n_views, max_draws_per_rank, max_visible_per_view, device=0):
        self.lib = load_library()
        self.n_ranks, self.n_views = n_ranks, n_views
        h = C.c_void_p()
        st = self.lib.ss_merge_create(C.c_int(device), C.c_uint32(n_ranks), C.c_uint32(n_views), C.c_uint32(max_draws_per_rank),
                                      C.c_uint64(max_visible_per_view), C.byref(h))
        if st != SS_OK:
            raise SceneSubmitError(st, self.lib.ss_last_error(None).decode())
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.lib.ss_merge_destroy(self.h)
            self.h = None

    __del__ = close

    def merge(self, peers, stream=None):
        """peers[v][r] = (idx_ptr, draws_ptr, draw_keys_ptr, counts_ptr, instance_base); asynchronous on `stream`."""
        arr = (MergePeer * (self.n_views * self.n_ranks))()
        for v in range(self.n_views):
            for r in range(self.n_ranks):
                arr[v * self.n_ranks + r] = MergePeer(*peers[v][r])
        st = self.lib.ss_merge_views(self.h, C.c_void_p(stream), arr)
        if st != SS_OK:
            raise SceneSubmitError(st, self.lib.ss_merge_last_error(self.h).decode())

    def stage_ms(self):
        """CUDA-event times of the last merge's kernels: fetch, rank, segments, placement."""
        ms = (C.c_float * 4)()
        st = self.lib.ss_merge_stage_ms(self.h, ms)
        if st != SS_OK:
            raise SceneSubmitError(st, self.lib.ss_merge_last_error(self.h).decode())
        return dict(zip(("fetch", "rank", "segments", "place"), (float(x) for x in ms)))

    def map(self, view, stream=None):
        out = MergedView()
        st = self.lib.ss_merge_map(self.h, C.c_void_p(stream), C.c_uint32(view), C.byref(out))
        if st != SS_OK:
            raise SceneSubmitError(st, self.lib.ss_merge_last_error(self.h).decode())
        return out


def peer_of_outputs(o, instance_base):
    """(idx, draws, draw_keys, counts, base) of an Outputs struct from SceneSubmit.map_outputs."""
    return (o.instance_idx, o.draws, o.draw_keys, o.counts, instance_base)


class SceneSubmit:
    """One libscenesubmit context (one GPU). Mirrors the C ABI one to one."""

    def __init__(self, max_nodes, max_instances, max_submesh_slots, max_views=5, device=0, trig_mode=TRIG_DETERMINISTIC,
                 frames_in_flight=3, max_draws_per_view=0, max_visible_per_view=0, flags=None):
        self.lib = load_library()
        if flags is None:  # SS_CONFIG_FLAGS=1 / 2 / 3 in the environment: run a whole test session on one pipeline
            flags = int(os.environ.get("SS_CONFIG_FLAGS", "0"))
        cfg = Config(C.sizeof(Config), device, max_nodes, max_instances, max_submesh_slots, max_views,
                     max_draws_per_view, frames_in_flight, max_visible_per_view, trig_mode, flags)
        h = C.c_void_p()
        st = self.lib.ss_create(C.byref(cfg), C.byref(h))
        if st != SS_OK:
            raise SceneSubmitError(st, self.lib.ss_last_error(None).decode())
        self.h = h
        self.n_instances = 0
        self.n_views = 0
        self.max_views = max_views

    def close(self):
        if getattr(self, "h", None):
            self.lib.ss_destroy(self.h)
            self.h = None

    __del__ = close

    def _ck(self, st):
        if st != SS_OK:
            raise SceneSubmitError(st, self.lib.ss_last_error(self.h).decode())

    # ---- static streams
    def set_key_layout(self, mesh_bits, submesh_bits, material_bits):
        self._ck(self.lib.ss_set_key_layout(self.h, C.c_uint32(mesh_bits), C.c_uint32(submesh_bits),
                                            C.c_uint32(material_bits)))

    def set_submesh_table(self, slots):
        slots = np.ascontiguousarray(slots, SUBMESH_DTYPE)
        self._ck(self.lib.ss_set_submesh_table(self.h, _p(slots), C.c_uint32(len(slots))))

    def set_nodes(self, n_nodes, parent=None, level_offsets=None):
        parent = None if parent is None else np.ascontiguousarray(parent, np.int32)
        lo = None if level_offsets is None else np.ascontiguousarray(level_offsets, np.uint32)
        self._ck(self.lib.ss_set_nodes(self.h, C.c_uint32(n_nodes), _p(parent), _p(lo),
                                       C.c_uint32(0 if lo is None else len(lo) - 1)))

    def set_instances(self, slot, matword, node=None, node_base=0):
        slot = np.ascontiguousarray(slot, np.uint32)
        matword = np.ascontiguousarray(matword, np.uint32)
        node = None if node is None else np.ascontiguousarray(node, np.uint32)
        self._ck(self.lib.ss_set_instances(self.h, C.c_uint32(len(slot)), _p(node), C.c_uint32(node_base), _p(slot),
                                           _p(matword)))
        self.n_instances = len(slot)

    def set_instance_aabbs(self, aabb, first=0):
        aabb = np.ascontiguousarray(aabb, np.float32)
        self._ck(self.lib.ss_set_instance_aabbs(self.h, C.c_uint32(first), C.c_uint32(len(aabb)), _p(aabb)))

    def load_scene(self, sc):
        """Convenience: every static stream + the TRS of a synth.SceneArrays-like object."""
        self.set_key_layout(sc.mesh_bits, sc.submesh_bits, sc.material_bits)
        self.set_submesh_table(sc.slots)
        self.set_nodes(sc.n_nodes, sc.parent, sc.level_offsets if sc.parent is not None else None)
        self.set_instances(sc.slot, sc.matword, sc.node, sc.node_base)
        if sc.inst_aabb is not None:
            self.set_instance_aabbs(sc.inst_aabb)
        self.upload_trs(sc.trs)
        self.flush_uploads()

    # ---- per frame
    def upload_trs(self, trs, first=0):
        trs = np.ascontiguousarray(trs, np.float32)
        self._ck(self.lib.ss_upload_trs(self.h, C.c_uint32(first), C.c_uint32(trs.shape[0]), _p(trs)))

    def upload_trs_sparse(self, node_index, trs):
        node_index = np.ascontiguousarray(node_index, np.uint32)
        trs = np.ascontiguousarray(trs, np.float32).reshape(-1, 9)
        assert len(node_index) == len(trs)
        self._ck(self.lib.ss_upload_trs_sparse(self.h, C.c_uint32(len(trs)), _p(node_index), _p(trs)))

    def upload_pose_sparse(self, node_index, pose7):
        """(position xyz, quaternion xyzw) per node: the physics write-back, converted to Euler angles on the device."""
        node_index = np.ascontiguousarray(node_index, np.uint32)
        pose7 = np.ascontiguousarray(pose7, np.float32).reshape(-1, 7)
        assert len(node_index) == len(pose7)
        self._ck(self.lib.ss_upload_pose_sparse(self.h, C.c_uint32(len(pose7)), _p(node_index), _p(pose7)))

    def set_instance_entities(self, entity_ids, first=0):
        e = np.ascontiguousarray(entity_ids, np.int32)
        self._ck(self.lib.ss_set_instance_entities(self.h, C.c_uint32(first), C.c_uint32(len(e)), _p(e)))

    def select_view(self, slot, view, stats):
        """SelectTransformData records (rows [12] + entity id) of the view, draw order."""
        m = int(stats.visible[view])
        rec = np.empty(m, np.dtype([("rows", "<f4", (12,)), ("entity", "<i4")]))
        ptr = C.c_void_p()
        self._ck(self.lib.ss_select_view(self.h, C.c_uint32(slot), C.c_uint32(view), C.byref(ptr), _p(rec)))
        return rec

    def pack_lights(self, point_nodes, point_par, spot_nodes, spot_par):
        """point_par [n][6] = color rgba, radius, falloff; spot_par [n][8] = color rgba, inner, outer, radius, falloff.
        Returns (UBPointLights bytes, UBSpotLights bytes, positions [n_point + n_spot][4])."""
        pt = np.zeros(len(point_nodes), np.dtype([("node", "<u4"), ("par", "<f4", (6,))]))
        pt["node"], pt["par"] = point_nodes, np.asarray(point_par, np.float32).reshape(-1, 6)
        sp = np.zeros(len(spot_nodes), np.dtype([("node", "<u4"), ("par", "<f4", (8,))]))
        sp["node"], sp["par"] = spot_nodes, np.asarray(spot_par, np.float32).reshape(-1, 8)
        ubp, ubs = np.zeros(496, np.uint8), np.zeros(656, np.uint8)
        pos = np.zeros((len(pt) + len(sp), 4), np.float32)
        self._ck(self.lib.ss_pack_lights(self.h, _p(pt), C.c_uint32(len(pt)), _p(sp), C.c_uint32(len(sp)), _p(ubp), _p(ubs), _p(pos)))
        return ubp, ubs, pos

    def upload_trs_sparse_ptr(self, node_index_ptr, trs_ptr, count):
        """Raw host pointers (pinned tensors): ascending node lists are copied straight from them."""
        self._ck(self.lib.ss_upload_trs_sparse(self.h, C.c_uint32(count), C.c_void_p(node_index_ptr), C.c_void_p(trs_ptr)))

    def upload_trs_ptr(self, host_ptr, count, first=0):
        """Raw host pointer (e.g. a pinned torch tensor's data_ptr())."""
        self._ck(self.lib.ss_upload_trs(self.h, C.c_uint32(first), C.c_uint32(count), C.c_void_p(host_ptr)))

    def upload_node_world_rows(self, rows, first):
        rows = np.ascontiguousarray(rows, np.float32)
        self._ck(self.lib.ss_upload_node_world_rows(self.h, C.c_uint32(first), C.c_uint32(rows.shape[0]), _p(rows)))

    def flush_uploads(self):
        self._ck(self.lib.ss_flush_uploads(self.h))

    def upload_world_rows(self, rows, first=0):
        rows = np.ascontiguousarray(rows, np.float32)
        self._ck(self.lib.ss_upload_world_rows(self.h, C.c_uint32(first), C.c_uint32(rows.shape[0]), _p(rows)))

    def frame(self, views, flags=0):
        views = np.ascontiguousarray(views, VIEW_DTYPE)
        slot = C.c_uint32()
        self._ck(self.lib.ss_frame(self.h, _p(views), C.c_uint32(len(views)), C.c_uint32(flags), C.byref(slot)))
        self.n_views = len(views)
        return slot.value

    def frame_sync(self, slot):
        st = FrameStats()
        self._ck(self.lib.ss_frame_sync(self.h, C.c_uint32(slot), C.byref(st)))
        return st

    def frame_stage_ms(self, slot):
        ms = (C.c_float * 5)()
        self._ck(self.lib.ss_frame_stage_ms(self.h, C.c_uint32(slot), ms))
        return dict(zip(("hierarchy", "instances", "sort", "emit", "draws"), (float(x) for x in ms)))

    # ---- outputs
    def map_outputs(self, slot, view):
        o = Outputs()
        self._ck(self.lib.ss_map_outputs(self.h, C.c_uint32(slot), C.c_uint32(view), C.byref(o)))
        return o

    def download_view(self, slot, view, stats, want_rows=True):
        m, d = int(stats.visible[view]), int(stats.draws[view])
        rows = np.empty((m, 12), np.float32) if want_rows else None
        keys = np.empty(m, np.uint64)
        idx = np.empty(m, np.uint32)
        draws = np.empty(d, DRAW_DTYPE)
        dkeys = np.empty(d, np.uint64)
        self._ck(self.lib.ss_download_view(self.h, C.c_uint32(slot), C.c_uint32(view), _p(rows), _p(keys), _p(idx),
                                           _p(draws), _p(dkeys)))
        return {"rows": rows, "keys": keys, "idx": idx, "draws": draws, "draw_keys": dkeys}

    def download_draws(self, slot, view, stats):
        d = int(stats.draws[view])
        draws = np.empty(d, DRAW_DTYPE)
        dkeys = np.empty(d, np.uint64)
        self._ck(self.lib.ss_download_view(self.h, C.c_uint32(slot), C.c_uint32(view), None, None, None, _p(draws),
                                           _p(dkeys)))
        return draws, dkeys

    def download_world(self, first=0, count=None):
        count = self.n_instances - first if count is None else count
        rows = np.empty((count, 12), np.float32)
        self._ck(self.lib.ss_download_world(self.h, C.c_uint32(first), C.c_uint32(count), _p(rows)))
        return rows

    def download_masks(self, first=0, count=None):
        count = self.n_instances - first if count is None else count
        m = np.empty(count, np.uint8)
        self._ck(self.lib.ss_download_masks(self.h, C.c_uint32(first), C.c_uint32(count), _p(m)))
        return m

    def export_idx_lists(self):
        """([frames_in_flight * max_views] 64-byte CUDA IPC handles, the same buffers' addresses in this process) of the
        per-view instance-index outputs: what a peer rank needs to read this rank's lists over NVLink."""
        cap = 64
        handles = (C.c_ubyte * (64 * cap))()
        ptrs = (C.c_void_p * cap)()
        count = C.c_uint32()
        self._ck(self.lib.ss_export_idx_lists(self.h, handles, ptrs, C.c_uint32(cap), C.byref(count)))
        raw = bytes(handles)
        return [raw[64 * k: 64 * (k + 1)] for k in range(count.value)], [int(ptrs[k] or 0) for k in range(count.value)]

    def export_slot_buffers(self):
        """CUDA IPC handles (64 bytes each) of every buffer a peer needs for ss_merge_views:
        [frames_in_flight][max_views][idx, draws, draw_keys], then [frames_in_flight] counts blocks."""
        cap = 256
        handles = (C.c_ubyte * (64 * cap))()
        count = C.c_uint32()
        self._ck(self.lib.ss_export_slot_buffers(self.h, handles, C.c_uint32(cap), C.byref(count)))
        raw = bytes(handles)
        return [raw[64 * k: 64 * (k + 1)] for k in range(count.value)]

    # ---- consumer pass: depth + id rasterisation of a view's lists ----
    def set_geometry(self, positions, indices):
        positions = np.ascontiguousarray(positions, np.float32).reshape(-1, 3)
        indices = np.ascontiguousarray(indices, np.uint32)
        self._ck(self.lib.ss_set_geometry(self.h, _p(positions), C.c_uint32(len(positions)), _p(indices), C.c_uint32(len(indices))))

    def raster_view(self, slot, view, width, height, flags=0):
        self._ck(self.lib.ss_raster_view(self.h, C.c_uint32(slot), C.c_uint32(view), C.c_uint32(width), C.c_uint32(height),
                                         C.c_uint32(flags)))
        if not hasattr(self, "_raster_shape"):
            self._raster_shape = {}
        self._raster_shape[view] = (height, width)

    def download_raster(self, view):
        out = np.empty(self._raster_shape[view], np.uint64)
        self._ck(self.lib.ss_download_raster(self.h, C.c_uint32(view), _p(out)))
        return out

    def read_pixel(self, view, x, y):
        inst, depth = C.c_uint32(), C.c_float()
        self._ck(self.lib.ss_read_pixel(self.h, C.c_uint32(view), C.c_uint32(x), C.c_uint32(y), C.byref(inst), C.byref(depth)))
        return inst.value, depth.value

    @property
    def stream(self):
        return self.lib.ss_stream(self.h)
