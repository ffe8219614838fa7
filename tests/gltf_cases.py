"""A hand-made glTF that exercises what the procedural writer (gltf_io) does not: every index component type,
an interleaved bufferView with byteStride, accessors at unaligned byte offsets, a primitive without TEXCOORD_0,
one without a whole triangle, nested node transforms with rotations. Shared by the CPU loader test and the GPU
assembly test."""
import base64
import json

import numpy as np


def handmade(tmp_path, rng, bad_index=False):
    """Four primitives under rotated / scaled nodes: u8 indices + interleaved POSITION/NORMAL (byteStride 24);
    u32 indices without TEXCOORD_0; u16 indices whose accessors start at byte offsets that are not multiples of 4
    (odd for the floats); an index accessor with fewer than 3 entries (no triangle)."""
    blob = bytearray()
    views, accs = [], []

    def view(data, stride=None, pad_to=4, lead=0):
        while len(blob) % pad_to:
            blob.append(0)
        blob.extend(b"\xAA" * lead)
        v = {"buffer": 0, "byteOffset": len(blob), "byteLength": len(data)}
        if stride:
            v["byteStride"] = stride
        blob.extend(data)
        views.append(v)
        return len(views) - 1

    def acc(view_idx, ctype, count, typ, off=0):
        accs.append({"bufferView": view_idx, "componentType": ctype, "count": count, "type": typ, "byteOffset": off})
        return len(accs) - 1

    def verts(n):
        p = rng.uniform(-1, 1, (n, 3)).astype(np.float32)
        nr = rng.normal(size=(n, 3)).astype(np.float32)
        uv = rng.uniform(0, 1, (n, 2)).astype(np.float32)
        return p, nr, uv

    prims, want = [], []
    # 0: u8 indices, interleaved position/normal
    p, nr, uv = verts(40)
    inter = np.concatenate([p, nr], axis=1).astype(np.float32).tobytes()
    vi = view(inter, stride=24)
    idx = rng.integers(0, 40, 3 * 50).astype(np.uint8)
    prims.append({"attributes": {"POSITION": acc(vi, 5126, 40, "VEC3"), "NORMAL": acc(vi, 5126, 40, "VEC3", off=12),
                                 "TEXCOORD_0": acc(view(uv.tobytes()), 5126, 40, "VEC2")},
                  "indices": acc(view(idx.tobytes()), 5121, len(idx), "SCALAR"), "material": 0})
    want.append((p, nr, uv, idx))
    # 1: u32 indices, no texcoords
    p, nr, uv = verts(70000)
    idx = rng.integers(0, 70000, 3 * 900).astype(np.uint32)
    if bad_index:
        idx[1234] = 70000
    prims.append({"attributes": {"POSITION": acc(view(p.tobytes()), 5126, 70000, "VEC3"), "NORMAL": acc(view(nr.tobytes()), 5126, 70000, "VEC3")},
                  "indices": acc(view(idx.tobytes()), 5125, len(idx), "SCALAR"), "material": 1})
    want.append((p, nr, None, idx))
    # 2: u16 indices, every accessor misaligned
    p, nr, uv = verts(300)
    idx = rng.integers(0, 300, 3 * 211 + 2).astype(np.uint16)        # two trailing indices that form no triangle
    prims.append({"attributes": {"POSITION": acc(view(p.tobytes(), lead=1), 5126, 300, "VEC3"),
                                 "NORMAL": acc(view(nr.tobytes(), lead=2), 5126, 300, "VEC3"),
                                 "TEXCOORD_0": acc(view(uv.tobytes(), lead=3), 5126, 300, "VEC2")},
                  "indices": acc(view(idx.tobytes(), lead=1), 5123, len(idx), "SCALAR"), "material": 0})
    want.append((p, nr, uv, idx))
    # 3: nothing to draw
    prims.append({"attributes": {"POSITION": prims[2]["attributes"]["POSITION"], "NORMAL": prims[2]["attributes"]["NORMAL"]},
                  "indices": acc(view(np.uint16([0, 1]).tobytes()), 5123, 2, "SCALAR"), "material": 1})
    want.append(None)
    q = np.float64([0.3, -0.2, 0.5, 0.7]); q /= np.linalg.norm(q)
    doc = {"asset": {"version": "2.0"}, "scene": 0, "scenes": [{"nodes": [0]}],
           "nodes": [{"children": [1, 2, 3], "translation": [0.5, -1.0, 2.0], "rotation": [float(x) for x in q]},
                     {"mesh": 0, "scale": [1.5, 0.5, 2.0]},
                     {"mesh": 1, "translation": [3.0, 0.0, -1.0], "rotation": [0.0, 0.38268343, 0.0, 0.92387953]},
                     {"camera": 0, "translation": [0.0, 0.0, 9.0]}],
           "meshes": [{"primitives": prims[:2]}, {"primitives": prims[2:]}],
           "materials": [{"pbrMetallicRoughness": {"baseColorFactor": [0.8, 0.7, 0.6, 1.0]}}, {"emissiveFactor": [2.0, 2.0, 2.0]}],
           "cameras": [{"type": "perspective", "perspective": {"yfov": 0.7, "aspectRatio": 1.5, "znear": 0.1}}],
           "accessors": accs, "bufferViews": views,
           "buffers": [{"byteLength": len(blob), "uri": "data:application/octet-stream;base64," + base64.b64encode(bytes(blob)).decode()}]}
    path = str(tmp_path / ("bad.gltf" if bad_index else "hand.gltf"))
    json.dump(doc, open(path, "w"))
    return path, doc, want


def oracle_soup(orc, doc, want):
    """The soup loadTriangles (src/stage1.zig:217-259) produces for handmade()'s file, by the oracle: per
    primitive the vertices are gathered through the index list first (the oracle's entry point reads u16
    indices), which leaves the per-vertex arithmetic untouched."""
    n0 = doc["nodes"][0]
    parent = orc.mat4_recompose(n0.get("translation", (0, 0, 0)), n0.get("rotation", (0, 0, 0, 1)), n0.get("scale", (1, 1, 1)))
    ps, ds = [], []
    k = 0
    for node in doc["nodes"]:
        if "mesh" not in node:
            continue
        m = orc.mat4_mul(parent, orc.mat4_recompose(node.get("translation", (0, 0, 0)), node.get("rotation", (0, 0, 0, 1)),
                                                    node.get("scale", (1, 1, 1))))
        for prim in doc["meshes"][node["mesh"]]["primitives"]:
            w = want[k]; k += 1
            if w is None:
                continue
            p, nr, uv, idx = w
            idx = idx[:3 * (len(idx) // 3)].astype(np.int64)
            seq = np.arange(len(idx), dtype=np.uint16)
            assert len(idx) < 65536
            op, od = orc.assemble_triangles(m, p[idx], nr[idx], uv[idx] if uv is not None else None, seq, prim["material"])
            ps.append(op); ds.append(od)
    p, d = np.concatenate(ps), np.concatenate(ds)
    d["src_tri"] = np.arange(len(d), dtype=np.uint32)
    return p, d
