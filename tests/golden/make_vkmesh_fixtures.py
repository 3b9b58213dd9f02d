"""Submesh tables of the reference's bundled mesh sources -> tests/golden/vkmeshes.json.
Run in the build container, where /root/reference exists:  python tests/golden/make_vkmesh_fixtures.py

The reference imports these FBX files through Assimp (Mesh.cpp:60-112), which cannot be built here; this script is
an independent Python restatement of the same rules vulkanengine_b200/host/MeshImporter.cpp states in its header
(one submesh per mesh model x geometry x material slot, fan-triangulated index counts, vertices and node
translations scaled by UnitScaleFactor x 0.01, node-local T*Rz*Ry*Rx*S).  It pins the C++ reader against a
second implementation, not against Assimp: parity of the mesh tables with the reference stays UNPINNED.
VertexCount is the number of polygon corners (Assimp's JoinIdenticalVertices result is not reproduced)."""
import json
import math
import os
import struct
import zlib

import numpy as np
import yaml

REF = "/root/reference/VulkanCore/assets"
HERE = os.path.dirname(os.path.abspath(__file__))


def read_node(b, pos, ver):
    if ver >= 7500:
        end, nprops, _ = struct.unpack_from("<QQQ", b, pos)
        pos += 24
    else:
        end, nprops, _ = struct.unpack_from("<III", b, pos)
        pos += 12
    nlen = b[pos]
    name = b[pos + 1:pos + 1 + nlen].decode()
    pos += 1 + nlen
    if end == 0:
        return None, pos
    props = []
    for _ in range(nprops):
        t = chr(b[pos])
        pos += 1
        if t in "YCIFDL":
            fmt = {"Y": "<h", "C": "<b", "I": "<i", "F": "<f", "D": "<d", "L": "<q"}[t]
            props.append(struct.unpack_from(fmt, b, pos)[0])
            pos += struct.calcsize(fmt)
        elif t in "fdlib":
            n, enc, clen = struct.unpack_from("<III", b, pos)
            pos += 12
            raw = b[pos:pos + clen]
            pos += clen
            if enc:
                raw = zlib.decompress(raw)
            props.append(np.frombuffer(raw, {"f": "<f4", "d": "<f8", "l": "<i8", "i": "<i4", "b": "i1"}[t], n))
        elif t in "SR":
            (n,) = struct.unpack_from("<I", b, pos)
            pos += 4
            props.append(b[pos:pos + n])
            pos += n
        else:
            raise ValueError(t)
    kids = []
    while pos < end:
        k, pos = read_node(b, pos, ver)
        if k is None:
            break
        kids.append(k)
    return (name, props, kids), end


def load_fbx(path):
    b = open(path, "rb").read()
    assert b.startswith(b"Kaydara FBX Binary  ")
    ver = struct.unpack_from("<I", b, 23)[0]
    pos, top = 27, {}
    while pos + 13 <= len(b):
        n, pos = read_node(b, pos, ver)
        if n is None:
            break
        top[n[0]] = n
    return top


def child(node, name):
    return next((k for k in node[2] if k[0] == name), None)


def rot(axis, deg):
    a = math.radians(deg)
    c, s = math.cos(a), math.sin(a)
    m = np.eye(4)
    u, v = (axis + 1) % 3, (axis + 2) % 3
    m[u, u] = m[v, v] = c
    m[v, u] = s   # row v, column u
    m[u, v] = -s
    return m


def submeshes_of(path, geometry=None):
    """geometry: optional dict {"positions": [], "indices": []} that receives one array per submesh"""
    top = load_fbx(path)
    unit = 1.0
    p70 = child(top["GlobalSettings"], "Properties70") if "GlobalSettings" in top else None
    for p in (p70[2] if p70 else []):
        if p[1] and p[1][0] == b"UnitScaleFactor":
            unit = float(p[1][4])
    scale = unit * 0.01
    models, geoms, mats = {}, {}, set()
    for o in top["Objects"][2]:
        if len(o[1]) < 3:
            continue
        oid = o[1][0]
        if o[0] == "Model":
            t, r, s = [0, 0, 0], [0, 0, 0], [1, 1, 1]
            pp = child(o, "Properties70")
            for p in (pp[2] if pp else []):
                if p[1][0] == b"Lcl Translation":
                    t = [float(x) for x in p[1][4:7]]
                if p[1][0] == b"Lcl Rotation":
                    r = [float(x) for x in p[1][4:7]]
                if p[1][0] == b"Lcl Scaling":
                    s = [float(x) for x in p[1][4:7]]
            models[oid] = (o[1][2], t, r, s)
        elif o[0] == "Geometry" and o[1][2] == b"Mesh":
            geoms[oid] = o
        elif o[0] == "Material":
            mats.add(oid)
    links = [(c[1][1], c[1][2]) for c in top["Connections"][2] if c[0] == "C" and c[1][0] == b"OO"]
    order, stack = [], [0]
    while stack:
        parent = stack.pop()
        if parent:
            order.append(parent)
        stack.extend(reversed([ch for ch, pa in links if pa == parent and ch in models]))
    out, mat_index, base_v, base_i = [], {}, 0, 0
    for mid in order:
        subtype, t, r, s = models[mid]
        if subtype != b"Mesh":
            continue
        slots = [ch for ch, pa in links if pa == mid and ch in mats]
        T, S = np.eye(4), np.diag([s[0], s[1], s[2], 1.0])
        T[:3, 3] = np.array(t) * scale
        local = (T @ ((rot(2, r[2]) @ (rot(1, r[1]) @ rot(0, r[0]))) @ S)).astype(np.float32)
        for gid in [ch for ch, pa in links if pa == mid and ch in geoms]:
            g = geoms[gid]
            verts = child(g, "Vertices")[1][0].reshape(-1, 3)
            pvi = child(g, "PolygonVertexIndex")[1][0]
            per_poly, all_same = None, 0
            lm = child(g, "LayerElementMaterial")
            if lm is not None:
                mapping, mv = child(lm, "MappingInformationType"), child(lm, "Materials")
                if mv is not None and len(mv[1][0]):
                    if mapping is not None and mapping[1][0] == b"ByPolygon":
                        per_poly = mv[1][0]
                    else:
                        all_same = int(mv[1][0][0])
            acc, start, poly = {}, 0, 0
            for k, v in enumerate(pvi):
                if v >= 0:
                    continue
                corners = k - start + 1
                slot = all_same if per_poly is None else int(per_poly[poly]) if poly < len(per_poly) else 0
                if corners >= 3:
                    a = acc.setdefault(slot, {"corners": 0, "indices": 0, "ids": [], "fan": []})
                    for f in range(1, corners - 1):
                        a["fan"] += [a["corners"], a["corners"] + f, a["corners"] + f + 1]
                    a["corners"] += corners
                    a["indices"] += (corners - 2) * 3
                    a["ids"] += [int(x) if x >= 0 else int(~x) for x in pvi[start:k + 1]]
                start, poly = k + 1, poly + 1
            for slot in sorted(acc):
                a = acc[slot]
                pts = verts[a["ids"]].astype(np.float32) * np.float32(scale)
                if geometry is not None:  # one vertex per polygon corner, fans relative to the submesh's first vertex
                    geometry["positions"].append(pts)
                    geometry["indices"].append(np.array(a["fan"], np.uint32))
                mat = 0
                if 0 <= slot < len(slots):
                    mat = mat_index.setdefault(slots[slot], len(mat_index))
                out.append({"BaseVertex": base_v, "BaseIndex": base_i, "VertexCount": a["corners"], "IndexCount": a["indices"],
                            "MaterialIndex": mat, "LocalTransform": [float(x) for x in local.T.reshape(16)],  # column-major
                            "BoundsMin": [float(x) for x in pts.min(0)], "BoundsMax": [float(x) for x in pts.max(0)]})
                base_v += a["corners"]
                base_i += a["indices"]
    return out


def resolve(path):
    """Registry paths are relative to the assets directory; absolute paths from the author's machine are re-rooted."""
    if ":" in path or path.startswith("/"):
        k = path.rfind("assets/")
        path = path[k + len("assets/"):] if k >= 0 else path
    return path


def main(REF=REF, out_name="vkmeshes.json"):
    reg = yaml.safe_load(open(os.path.join(REF, "AssetRegistry.vkr")))["Asset Registry"]
    by_handle = {int(e["Handle"]): e["Metadata"] for e in reg}
    meshes, sources = {}, {}
    for h, md in by_handle.items():
        if md["Type"] != "Mesh":
            continue
        p = os.path.join(REF, resolve(md["Filepath"]))
        src = None
        if os.path.exists(p):
            sh = int(yaml.safe_load(open(p))["Mesh"]["MeshAsset"])
            smd = by_handle.get(sh)
            if smd and smd["Type"] == "MeshAsset":
                rel = resolve(smd["Filepath"])
                sp = os.path.join(REF, rel)
                if os.path.exists(sp) and open(sp, "rb").read(20) == b"Kaydara FBX Binary  ":
                    src = rel
                    if rel not in sources:
                        sources[rel] = submeshes_of(sp)
        meshes[str(h)] = {"file": md["Filepath"], "source": src}
    json.dump({"note": "generated by tests/golden/make_vkmesh_fixtures.py (Python restatement; not Assimp output)",
               "meshes": meshes, "sources": sources}, open(os.path.join(HERE, out_name), "w"), indent=1)
    print({k: len(v) for k, v in sources.items()})


if __name__ == "__main__":
    main()
