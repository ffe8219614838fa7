"""glTF 2.0 writer for the generated scenes (the reference bundles no renderable asset) under
the constraints the reference's loader imposes (src/stage1.zig): u16 indices (<= 65535 vertices per
primitive), tightly packed f32 POSITION / NORMAL / TEXCOORD_0, every primitive with a material, a
perspective camera with aspectRatio on a node, PNG images. Textures are written as 8-bit sRGB-ish PNGs
(value = lin^(1/2.2)), which the loader linearises again with pow(c/255, 2.2) like stb.
"""
import base64
import json
import os
import struct
import zlib

import numpy as np

MAX_TRIS_PER_PRIMITIVE = 65535 // 3


def png_encode(img):
    """[h,w,C] uint8 (C = 1..4) -> PNG bytes."""
    img = np.ascontiguousarray(img, np.uint8)
    if img.ndim == 2:
        img = img[:, :, None]
    h, w, c = img.shape
    ctype = {1: 0, 2: 4, 3: 2, 4: 6}[c]
    raw = np.zeros((h, 1 + w * c), np.uint8)
    raw[:, 1:] = img.reshape(h, w * c)

    def chunk(tag, data):
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)
    return (b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, ctype, 0, 0, 0)) +
            chunk(b"IDAT", zlib.compress(raw.tobytes(), 6)) + chunk(b"IEND", b""))


def _to8(lin):
    return np.clip(np.rint(255.0 * np.power(np.clip(lin, 0.0, 1.0), 1.0 / 2.2)), 0, 255).astype(np.uint8)


def _is_const(tex):
    return tex.w == 1 and tex.h == 1


def write_gltf(spec, path, node_translation=(0.0, 0.0, 0.0), node_scale=1.0, glb=False, embed=False, camera_name="Camera 1"):
    """Writes spec (scenes.SceneSpec) to `path` (.gltf + .bin + .png files, or one .glb). Mesh nodes carry
    a TRS (translation, uniform scale) under a parent node so that global transforms are exercised; the
    vertices are stored in the node's local frame."""
    out_dir = os.path.dirname(os.path.abspath(path))
    stem = os.path.splitext(os.path.basename(path))[0]
    doc = {"asset": {"version": "2.0", "generator": "zig_raytracing_contest_b200.gltf_io"},
           "scene": 0, "scenes": [{"nodes": [0]}], "nodes": [], "meshes": [], "accessors": [], "bufferViews": [],
           "buffers": [], "materials": [], "cameras": [], "textures": [], "images": [], "samplers": []}
    blob = bytearray()

    def view(data):
        while len(blob) % 4:
            blob.append(0)
        doc["bufferViews"].append({"buffer": 0, "byteOffset": len(blob), "byteLength": len(data)})
        blob.extend(data)
        return len(doc["bufferViews"]) - 1

    def add_image(name, arr8):
        png = png_encode(arr8)
        if glb:
            doc["images"].append({"bufferView": view(png), "mimeType": "image/png"})
        elif embed:
            doc["images"].append({"uri": "data:image/png;base64," + base64.b64encode(png).decode()})
        else:
            fn = "%s_%s.png" % (stem, name)
            open(os.path.join(out_dir, fn), "wb").write(png)
            doc["images"].append({"uri": fn})
        return len(doc["images"]) - 1

    def add_texture(image, tex):
        s = {"wrapS": 33071 if tex.u_min == 0 else 10497, "wrapT": 33071 if tex.v_min == 0 else 10497}
        if s not in doc["samplers"]:
            doc["samplers"].append(s)
        doc["textures"].append({"source": image, "sampler": doc["samplers"].index(s)})
        return len(doc["textures"]) - 1

    # ---- materials ----
    for i, m in enumerate(spec.materials):
        mat = {"name": "material_%d" % i, "pbrMetallicRoughness": {}}
        base, alpha, emis = m.base_color, m.transparency, m.emissive
        alpha_const = _is_const(alpha)
        if _is_const(base) and alpha_const and float(alpha.texels.reshape(-1)[0]) == 1.0:
            mat["pbrMetallicRoughness"]["baseColorFactor"] = [float(x) for x in base.texels.reshape(-1)] + [1.0]
        else:
            ref = base if not _is_const(base) else alpha
            h, w = ref.h, ref.w
            rgb = np.broadcast_to(base.texels, (h, w, 3)) if _is_const(base) else base.texels
            a = np.broadcast_to(alpha.texels, (h, w, 1)) if alpha_const else alpha.texels
            if rgb.shape[:2] != a.shape[:2]:          # glTF keeps alpha in the base-colour image: nearest-resample it
                ys = (np.arange(h) * a.shape[0]) // h
                xs = (np.arange(w) * a.shape[1]) // w
                a = a[ys][:, xs]
            has_alpha = not (alpha_const and float(alpha.texels.reshape(-1)[0]) == 1.0)
            img = _to8(rgb)
            if has_alpha:
                a8 = np.clip(np.rint(255.0 * a), 0, 255).astype(np.uint8)
                img = np.concatenate([img, a8], axis=2)
                binary = bool(np.all((a == 0.0) | (a == 1.0)))
                mat["alphaMode"] = "MASK" if binary else "BLEND"
                if binary:
                    mat["alphaCutoff"] = 0.5
            mat["pbrMetallicRoughness"]["baseColorTexture"] = {"index": add_texture(add_image("base%d" % i, img), ref)}
        if _is_const(emis):
            mat["emissiveFactor"] = [float(x) for x in emis.texels.reshape(-1)]
        else:
            mat["emissiveFactor"] = [1.0, 1.0, 1.0]
            mat["emissiveTexture"] = {"index": add_texture(add_image("emis%d" % i, _to8(emis.texels)), emis)}
        doc["materials"].append(mat)

    # ---- geometry: node 0 = parent (translation), children = one mesh node per primitive chunk ----
    t = np.asarray(node_translation, np.float32)
    doc["nodes"].append({"name": "root", "translation": [float(x) for x in t], "children": []})
    pos = np.asarray(spec.positions, np.float32).reshape(-1, 3, 3)
    local = (pos - t) / np.float32(node_scale)
    td = spec.tri_data
    order = np.argsort(td["material"], kind="stable")
    start = 0
    while start < len(order):
        mat = int(td["material"][order[start]])
        end = start
        while end < len(order) and end - start < MAX_TRIS_PER_PRIMITIVE and int(td["material"][order[end]]) == mat:
            end += 1
        idx = order[start:end]
        n = len(idx) * 3
        p = np.ascontiguousarray(local[idx].reshape(n, 3), np.float32)
        accs = []
        for data, typ, ct in ((p, "VEC3", 5126), (np.ascontiguousarray(td["normal"][idx].reshape(n, 3), np.float32), "VEC3", 5126),
                              (np.ascontiguousarray(td["texcoord"][idx].reshape(n, 2), np.float32), "VEC2", 5126),
                              (np.arange(n, dtype=np.uint16), "SCALAR", 5123)):
            acc = {"bufferView": view(data.tobytes()), "componentType": ct, "count": int(data.shape[0]), "type": typ}
            if data is p:
                acc["min"], acc["max"] = [float(x) for x in p.min(0)], [float(x) for x in p.max(0)]
            doc["accessors"].append(acc)
            accs.append(len(doc["accessors"]) - 1)
        doc["meshes"].append({"primitives": [{"attributes": {"POSITION": accs[0], "NORMAL": accs[1], "TEXCOORD_0": accs[2]},
                                              "indices": accs[3], "material": mat, "mode": 4}]})
        doc["nodes"].append({"mesh": len(doc["meshes"]) - 1, "scale": [float(node_scale)] * 3})
        doc["nodes"][0]["children"].append(len(doc["nodes"]) - 1)
        start = end

    # ---- camera: a second root node holding the camera matrix (column-major, like glTF) ----
    doc["cameras"].append({"name": camera_name, "type": "perspective",
                           "perspective": {"yfov": float(spec.yfov), "aspectRatio": float(spec.aspect), "znear": 0.01, "zfar": 1000.0}})
    cm = np.asarray(spec.cam_matrix, np.float32).reshape(4, 4)
    doc["nodes"].append({"name": "camera_node", "camera": 0, "matrix": [float(x) for x in cm.reshape(-1)]})
    doc["scenes"][0]["nodes"].append(len(doc["nodes"]) - 1)
    for key in ("textures", "images", "samplers"):
        if not doc[key]:
            del doc[key]

    if glb:
        doc["buffers"].append({"byteLength": len(blob)})
        js = json.dumps(doc).encode()
        js += b" " * (-len(js) % 4)
        blob.extend(b"\0" * (-len(blob) % 4))
        total = 12 + 8 + len(js) + 8 + len(blob)
        with open(path, "wb") as f:
            f.write(struct.pack("<4sII", b"glTF", 2, total) + struct.pack("<II", len(js), 0x4E4F534A) + js +
                    struct.pack("<II", len(blob), 0x004E4942) + bytes(blob))
    else:
        if embed:
            doc["buffers"].append({"byteLength": len(blob), "uri": "data:application/octet-stream;base64," + base64.b64encode(bytes(blob)).decode()})
        else:
            open(os.path.join(out_dir, stem + ".bin"), "wb").write(bytes(blob))
            doc["buffers"].append({"byteLength": len(blob), "uri": stem + ".bin"})
        with open(path, "w") as f:
            json.dump(doc, f)
    return doc
