"""Seeded synthetic scenes of the shapes BASELINE.json names.

One counter-based SplitMix64 generator drives every stream, so a value depends only on
(seed, stream id, element index): shards of a scene can be generated independently on every rank
and the oracle and the GPU path see identical bytes.  Pure numpy; no oracle import.
"""
from dataclasses import dataclass
from typing import Optional

import numpy as np

SEED_BASE = 0x5CE7E5B1
SLOT_NULL = 0xFFFFFFFF
TESS_BIT = 0x80000000

SUBMESH_DTYPE = np.dtype([("local", "<f4", (16,)), ("aabb_min", "<f4", (3,)), ("aabb_max", "<f4", (3,)),
                          ("index_count", "<u4"), ("first_index", "<u4"), ("vertex_offset", "<i4"),
                          ("mesh_rank", "<u4"), ("submesh_index", "<u4"), ("reserved", "<u4")])
VIEW_DTYPE = np.dtype([("view_proj", "<f4", (16,)), ("flags", "<u4")])
DRAW_DTYPE = np.dtype([("index_count", "<u4"), ("instance_count", "<u4"), ("first_index", "<u4"),
                       ("vertex_offset", "<i4"), ("first_instance", "<u4")])

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x):
    """Vectorised SplitMix64 finaliser over uint64 counters."""
    with np.errstate(over="ignore"):
        z = (x + np.uint64(0x9E3779B97F4A7C15)) & _M64
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
        return z ^ (z >> np.uint64(31))


def _counters(seed, stream, first, count):
    with np.errstate(over="ignore"):
        base = splitmix64(np.array([np.uint64(seed) ^ (np.uint64(stream) << np.uint64(40))], dtype=np.uint64))[0]
        return base + np.arange(first, first + count, dtype=np.uint64)


def u01(seed, stream, first, count):
    """float32 uniform in [0,1): top 24 bits of the hash."""
    h = splitmix64(_counters(seed, stream, first, count))
    return (h >> np.uint64(40)).astype(np.float32) * np.float32(2.0 ** -24)


def uniform(seed, stream, first, count, lo, hi):
    lo, hi = np.float32(lo), np.float32(hi)
    return lo + (hi - lo) * u01(seed, stream, first, count)


def u64(seed, stream, first, count):
    return splitmix64(_counters(seed, stream, first, count))


@dataclass
class SceneArrays:
    trs: np.ndarray                  # [n_nodes, 9] float32  T, R(euler), S
    parent: Optional[np.ndarray]     # [n_nodes] int32 or None (flat)
    level_offsets: np.ndarray        # [n_levels+1] uint32
    node: Optional[np.ndarray]       # [n_inst] uint32 or None -> node_base + i
    node_base: int
    slot: np.ndarray                 # [n_inst] uint32
    matword: np.ndarray              # [n_inst] uint32
    inst_aabb: Optional[np.ndarray]  # [n_inst, 6] float32
    slots: np.ndarray                # [n_slots] SUBMESH_DTYPE
    mesh_bits: int
    submesh_bits: int
    material_bits: int
    mesh_handles: np.ndarray         # [n_meshes] uint64 sorted (rank -> handle)
    material_handles: np.ndarray     # [n_materials] uint64 sorted
    seed: int
    name: str = ""

    @property
    def n_instances(self):
        return int(self.slot.shape[0])

    @property
    def n_nodes(self):
        return int(self.trs.shape[0])


def _trs_matrix(t, e, s):
    """float64 helper for slot-local matrices only (inputs, not checked arithmetic)."""
    cx, cy, cz = np.cos(e * 0.5)
    sx, sy, sz = np.sin(e * 0.5)
    w = cx * cy * cz + sx * sy * sz
    x = sx * cy * cz - cx * sy * sz
    y = cx * sy * cz + sx * cy * sz
    z = cx * cy * sz - sx * sy * cz
    r = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                  [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                  [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
    m = np.eye(4)
    m[:3, :3] = r * s[None, :]
    m[:3, 3] = t
    return m.T.reshape(16).astype(np.float32)  # column-major like glm::mat4


def make_tables(seed, n_meshes=256, n_materials=64, max_submeshes=4, n_keys=4096, identity_locals=False):
    """Submesh table + the K (slot, material) combinations instances draw from."""
    sub_count = 1 + (u64(seed, 1, 0, n_meshes) % np.uint64(max_submeshes)).astype(np.int64)
    n_slots = int(sub_count.sum())
    slots = np.zeros(n_slots, SUBMESH_DTYPE)
    mesh_of_slot = np.repeat(np.arange(n_meshes), sub_count)
    first = np.concatenate([[0], np.cumsum(sub_count)[:-1]])
    sub_idx = np.arange(n_slots) - first[mesh_of_slot]
    slots["mesh_rank"] = mesh_of_slot
    slots["submesh_index"] = sub_idx
    ic = (3 * (12 + (u64(seed, 2, 0, n_slots) % np.uint64(5000)))).astype(np.uint32)
    slots["index_count"] = ic
    slots["first_index"] = np.concatenate([[0], np.cumsum(ic)[:-1]]).astype(np.uint32)
    vc = (8 + (u64(seed, 3, 0, n_slots) % np.uint64(3000))).astype(np.int64)
    slots["vertex_offset"] = np.concatenate([[0], np.cumsum(vc)[:-1]]).astype(np.int32)
    t = uniform(seed, 4, 0, n_slots * 3, -1, 1).reshape(n_slots, 3).astype(np.float64)
    e = uniform(seed, 5, 0, n_slots * 3, -np.pi, np.pi).reshape(n_slots, 3).astype(np.float64)
    s = uniform(seed, 6, 0, n_slots * 3, 0.8, 1.25).reshape(n_slots, 3).astype(np.float64)
    ident = np.eye(4, dtype=np.float32).reshape(16)
    for k in range(n_slots):
        # every 4th slot keeps an identity LocalTransform (common for single-node meshes)
        slots["local"][k] = ident if (identity_locals or k % 4 == 0) else _trs_matrix(t[k], e[k], s[k])
    half = uniform(seed, 7, 0, n_slots * 3, 0.25, 2).reshape(n_slots, 3)
    ctr = uniform(seed, 8, 0, n_slots * 3, -0.5, 0.5).reshape(n_slots, 3)
    slots["aabb_min"] = ctr - half
    slots["aabb_max"] = ctr + half

    # handles: random uint64 like VulkanCore/src/VulkanCore/Core/UUID.cpp:8-16, ranks = sorted position
    mesh_handles = np.sort(u64(seed, 9, 0, n_meshes) | np.uint64(1))
    material_handles = np.sort(u64(seed, 10, 0, n_materials) | np.uint64(1))
    tess_material = (u64(seed, 11, 0, n_materials) % np.uint64(20)) == 0  # ~5 %

    # K distinct (slot, material) pairs: a seeded permutation of the full cross product
    total = n_slots * n_materials
    n_keys = min(n_keys, total)
    order = np.argsort(u64(seed, 12, 0, total), kind="stable")[:n_keys]
    key_slot = (order // n_materials).astype(np.uint32)
    key_mat = (order % n_materials).astype(np.uint32)
    return slots, mesh_handles, material_handles, tess_material, key_slot, key_mat


def _bits(n):
    return max(1, int(n - 1).bit_length())


def _zipf_ranks(seed, stream, first, count, n_keys, a=1.1):
    w = 1.0 / np.power(np.arange(1, n_keys + 1, dtype=np.float64), a)
    cdf = np.cumsum(w)
    cdf /= cdf[-1]
    u = u01(seed, stream, first, count).astype(np.float64)
    return np.minimum(np.searchsorted(cdf, u, side="right"), n_keys - 1)


def _instance_streams(seed, first, count, tables):
    slots, mesh_handles, material_handles, tess_material, key_slot, key_mat = tables
    n_materials = len(material_handles)
    r = _zipf_ranks(seed, 20, first, count, len(key_slot))
    slot = key_slot[r].copy()
    mat = key_mat[r].copy()
    # 1 %: no per-index material -> base material of the mesh (SceneRenderer.cpp:1141), not dropped
    sel = u64(seed, 21, first, count) % np.uint64(100)
    fallback = sel == 1
    mat[fallback] = (slots["mesh_rank"][slot[fallback]] % n_materials).astype(np.uint32)
    matword = mat | np.where(tess_material[mat], np.uint32(TESS_BIT), np.uint32(0)).astype(np.uint32)
    # 1 %: null mesh / null material table / zero-vertex mesh -> dropped (SceneRenderer.cpp:1128-1136)
    slot[sel == 0] = SLOT_NULL
    half = uniform(seed, 22, first * 3, count * 3, 0.25, 2).reshape(count, 3)
    ctr = uniform(seed, 23, first * 3, count * 3, -0.5, 0.5).reshape(count, 3)
    aabb = np.concatenate([ctr - half, ctr + half], axis=1).astype(np.float32)
    return slot.astype(np.uint32), matword.astype(np.uint32), aabb


def _trs(seed, stream, first, count, t_range, scale_range):
    t = uniform(seed, stream, first * 3, count * 3, -t_range, t_range).reshape(count, 3)
    e = uniform(seed, stream + 1, first * 3, count * 3, -np.pi, np.pi).reshape(count, 3)
    if scale_range is None:
        s = np.ones((count, 3), np.float32)
    else:
        s = uniform(seed, stream + 2, first * 3, count * 3, scale_range[0], scale_range[1]).reshape(count, 3)
    return np.concatenate([t, e, s], axis=1).astype(np.float32)


def flat_scene(n, seed=SEED_BASE + 3, first=0, n_keys=4096, name="flat"):
    """cfg 3/4 flat: every node is a root carrying one instance. `first` = global index of the shard's
    first instance (entity-wise sharding)."""
    tables = make_tables(seed, n_keys=n_keys)
    slots, mesh_handles, material_handles = tables[0], tables[1], tables[2]
    slot, matword, aabb = _instance_streams(seed, first, n, tables)
    trs = _trs(seed, 30, first, n, 500.0, (0.5, 2.0))
    return SceneArrays(trs=trs, parent=None, level_offsets=np.array([0, n], np.uint32), node=None, node_base=0,
                       slot=slot, matword=matword, inst_aabb=aabb, slots=slots,
                       mesh_bits=_bits(len(mesh_handles)), submesh_bits=_bits(4),
                       material_bits=_bits(len(material_handles)), mesh_handles=mesh_handles,
                       material_handles=material_handles, seed=seed, name=name)


def hierarchy_scene(n_roots=1 << 10, fan1=1 << 5, fan2=1 << 5, seed=SEED_BASE + 2, first_root=0, n_keys=4096,
                    name="hier3"):
    """cfg 2: three levels, level-ordered, subtrees contiguous inside each level; only the leaves carry
    meshes.  `first_root` shifts the RNG counters so that shards by root range are slices of one scene."""
    tables = make_tables(seed, n_keys=n_keys)
    slots, mesh_handles, material_handles = tables[0], tables[1], tables[2]
    n1, n2 = n_roots * fan1, n_roots * fan1 * fan2
    trs0 = _trs(seed, 40, first_root, n_roots, 500.0, None)
    trs1 = _trs(seed, 50, first_root * fan1, n1, 20.0, None)
    trs2 = _trs(seed, 60, first_root * fan1 * fan2, n2, 2.0, (0.5, 2.0))
    trs = np.concatenate([trs0, trs1, trs2], axis=0)
    parent = np.empty(n_roots + n1 + n2, np.int32)
    parent[:n_roots] = -1
    parent[n_roots:n_roots + n1] = np.arange(n1) // fan1
    parent[n_roots + n1:] = n_roots + np.arange(n2) // fan2
    slot, matword, aabb = _instance_streams(seed, first_root * fan1 * fan2, n2, tables)
    return SceneArrays(trs=trs, parent=parent, level_offsets=np.array([0, n_roots, n_roots + n1, n_roots + n1 + n2],
                                                                      np.uint32),
                       node=None, node_base=n_roots + n1, slot=slot, matword=matword, inst_aabb=aabb, slots=slots,
                       mesh_bits=_bits(len(mesh_handles)), submesh_bits=_bits(4),
                       material_bits=_bits(len(material_handles)), mesh_handles=mesh_handles,
                       material_handles=material_handles, seed=seed, name=name)


def with_explicit_nodes(sc: SceneArrays, seed=7):
    """General form: an explicit (shuffled) instance -> node map, several instances per node."""
    n = sc.n_instances
    leaves = sc.n_nodes - sc.node_base
    node = (sc.node_base + (u64(seed, 70, 0, n) % np.uint64(leaves))).astype(np.uint32)
    return SceneArrays(**{**sc.__dict__, "node": node, "name": sc.name + "+nodes"})


# camera / cascade defaults of the reference -------------------------------------------------------
CAMERA_FOV = float(np.radians(np.float32(45.0)))  # EditorLayer.cpp:75
CAMERA_ASPECT = 1.635005
CAMERA_NEAR, CAMERA_FAR = 0.1, 1000.0
CSM_LAMBDA, CSM_NEAR_OFFSET, CSM_FAR_OFFSET = 0.92, -75.0, 75.0  # SceneRenderer.h:173-181
LIGHT_DIR = (1.0, -1.0, 1.0)  # Components.h:144
CAMERA_POS = (0.0, 0.0, 600.0)
VIEW_SKIP_NEAR = 1


def make_views(camera_fn, cascade_fn, n_views=5, position=CAMERA_POS, pitch=0.0, yaw=0.0):
    """Camera (+ 4 cascades when n_views == 5) through the given host functions
    (product: SceneSubmit.camera_view_proj / cascade_matrices; tests may pass the oracle's)."""
    _, _, vp = camera_fn(CAMERA_FOV, CAMERA_ASPECT, CAMERA_NEAR, CAMERA_FAR, position, pitch, yaw)
    views = np.zeros(n_views, VIEW_DTYPE)
    views["view_proj"][0] = vp
    if n_views > 1:
        _, mats = cascade_fn(vp, CAMERA_NEAR, CAMERA_FAR, CSM_LAMBDA, CSM_NEAR_OFFSET, CSM_FAR_OFFSET, LIGHT_DIR)
        for i in range(min(4, n_views - 1)):
            views["view_proj"][1 + i] = mats[i]
            views["flags"][1 + i] = VIEW_SKIP_NEAR
    return views


def fibonacci_cameras(camera_fn, n_views, radius=800.0, seed=SEED_BASE + 5):
    """cfg 5: cameras on a Fibonacci sphere looking at the origin, fov U(30deg, 90deg)."""
    views = np.zeros(n_views, VIEW_DTYPE)
    fov = uniform(seed, 80, 0, n_views, np.radians(30.0), np.radians(90.0))
    golden = np.pi * (3.0 - np.sqrt(5.0))
    for i in range(n_views):
        y = 1.0 - 2.0 * (i + 0.5) / n_views
        r = np.sqrt(max(0.0, 1.0 - y * y))
        th = golden * i
        pos = np.array([np.cos(th) * r, y, np.sin(th) * r]) * radius
        # orientation quat(vec3(-pitch, -yaw, 0)) looks down -Z rotated by yaw about Y then pitch about X
        yaw = -np.arctan2(pos[0], pos[2])
        pitch = np.arcsin(np.clip(pos[1] / radius, -1, 1))
        _, _, vp = camera_fn(float(fov[i]), CAMERA_ASPECT, CAMERA_NEAR, 2000.0, pos.astype(np.float32), float(pitch),
                             float(yaw))
        views["view_proj"][i] = vp
    return views


def box_geometry(slots):
    """One box per submesh slot, spanning the slot's local AABB: 8 vertices, 12 triangles.  Returns (positions [8S][3],
    indices [36S], slots') where slots' is a copy whose first_index / index_count / vertex_offset address that geometry
    (synthetic scenes carry made-up counts; the consumer pass needs real ones)."""
    corners = np.array([[x, y, z] for z in (0, 1) for y in (0, 1) for x in (0, 1)], np.float32)       # bit k of the corner id = axis k
    quads = [(0, 2, 3, 1), (4, 5, 7, 6), (0, 1, 5, 4), (2, 6, 7, 3), (0, 4, 6, 2), (1, 3, 7, 5)]      # -z +z -y +y -x +x
    tri = np.array([[q[0], q[1], q[2], q[0], q[2], q[3]] for q in quads], np.uint32).reshape(-1)
    s2 = slots.copy()
    n = len(slots)
    lo, hi = slots["aabb_min"].astype(np.float32), slots["aabb_max"].astype(np.float32)
    pos = lo[:, None, :] + corners[None, :, :] * (hi - lo)[:, None, :]
    s2["index_count"] = 36
    s2["first_index"] = 36 * np.arange(n, dtype=np.uint32)
    s2["vertex_offset"] = 8 * np.arange(n, dtype=np.int32)
    return pos.reshape(-1, 3).astype(np.float32), np.tile(tri, n), s2
