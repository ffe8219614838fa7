"""The host-side image code against THE REFERENCE'S OWN C UNIT: oracle/_ref/libc_libs.so is src/c_impl.c of the
reference (the stb_image / stb_image_write instantiations build.zig:29-39 links as `c_libs`), compiled from where it
lies by oracle/Makefile's `ref` target. What stage1 asks of it (src/stage1.zig:56-63:
stbi_loadf_from_memory(ptr, len, &w, &h, &c, 4)) and what main asks of it (src/main.zig:130-135: stbi_write_png) is
run here and compared bit for bit with the loader's texels and with the PNG reader / writer of librtc:
  - colour channels linearised as pow(c/255, 2.2), alpha as c/255, for every PNG colour type and bit depth,
    row filters, palette + tRNS, colour-key tRNS;
  - a frame written by rtc_png_write decodes in stb to the same pixels; a frame written by stbi_write_png decodes in
    rtc_png_read to the same pixels.
Skipped when the reference's unit has not been built (no /root/reference on this machine and no prebuilt copy)."""
import ctypes as C
import json
import os
import struct
import zlib

import numpy as np
import pytest

from zig_raytracing_contest_b200 import gltf_io, rtc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libc_libs.so")


@pytest.fixture(scope="module")
def stb():
    if not os.path.exists(REF_SO):
        pytest.skip("oracle/_ref/libc_libs.so (the reference's stb unit) is not built on this machine")
    L = C.CDLL(REF_SO)
    L.stbi_loadf_from_memory.restype = C.POINTER(C.c_float)
    L.stbi_loadf_from_memory.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_int]
    L.stbi_load_from_memory.restype = C.POINTER(C.c_uint8)
    L.stbi_load_from_memory.argtypes = L.stbi_loadf_from_memory.argtypes
    L.stbi_image_free.argtypes = [C.c_void_p]
    L.stbi_image_free.restype = None
    L.stbi_write_png.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
    return L


def _stb_load(L, data, req, as_float):
    w, h, c = C.c_int(), C.c_int(), C.c_int()
    fn = L.stbi_loadf_from_memory if as_float else L.stbi_load_from_memory
    p = fn(data, len(data), C.byref(w), C.byref(h), C.byref(c), req)
    assert p, "stb could not decode the image"
    out = np.ctypeslib.as_array(p, (h.value, w.value, req)).copy()
    L.stbi_image_free(p)
    return out, c.value


def _png(w, h, depth, ctype, samples, filters=None, plte=None, trns=None, adam7=False):
    """A PNG of any colour type / bit depth from per-pixel sample values ([h, w, channels], already in range), with a
    chosen filter type per row; adam7 = the seven-pass interlace (every pass filtered as an image of its own)."""
    ch = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[ctype]
    samples = np.asarray(samples).reshape(h, w, ch)
    bpp = max(1, ch * depth // 8)

    def rows_of(sub):
        hh, ww, _ = sub.shape
        flat = sub.reshape(hh, ww * ch)
        if depth == 16:
            return flat.astype(">u2").view(np.uint8).reshape(hh, -1)
        if depth == 8:
            return flat.astype(np.uint8)
        per = 8 // depth
        pad = (-flat.shape[1]) % per
        s = np.pad(flat, ((0, 0), (0, pad))).astype(np.uint8).reshape(hh, -1, per)
        rows = np.zeros(s.shape[:2], np.uint8)
        for k in range(per):
            rows |= s[:, :, k] << (depth * (per - 1 - k))
        return rows

    def filtered(rows):
        raw = bytearray()
        prev = np.zeros(rows.shape[1], np.int32)
        for y in range(rows.shape[0]):
            cur = rows[y].astype(np.int32)
            f = 0 if filters is None else filters[y % len(filters)]
            left = np.concatenate([np.zeros(bpp, np.int32), cur[:-bpp]])[:len(cur)]
            ul = np.concatenate([np.zeros(bpp, np.int32), prev[:-bpp]])[:len(cur)]
            if f == 0:
                enc = cur
            elif f == 1:
                enc = cur - left
            elif f == 2:
                enc = cur - prev
            elif f == 3:
                enc = cur - ((left + prev) >> 1)
            else:
                p = left + prev - ul
                pa, pb, pc = np.abs(p - left), np.abs(p - prev), np.abs(p - ul)
                pred = np.where((pa <= pb) & (pa <= pc), left, np.where(pb <= pc, prev, ul))
                enc = cur - pred
            raw.append(f)
            raw.extend((enc & 255).astype(np.uint8).tobytes())
            prev = cur
        return bytes(raw)

    if adam7:
        raw = b""
        for x0, y0, dx, dy in ((0, 0, 8, 8), (4, 0, 8, 8), (0, 4, 4, 8), (2, 0, 4, 4), (0, 2, 2, 4), (1, 0, 2, 2), (0, 1, 1, 2)):
            sub = samples[y0::dy, x0::dx]
            if sub.shape[0] and sub.shape[1]:
                raw += filtered(rows_of(sub))
    else:
        raw = filtered(rows_of(samples))

    def chunk(tag, data):
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)
    out = b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, ctype, 0, 0, 1 if adam7 else 0))
    if plte is not None:
        out += chunk(b"PLTE", bytes(plte))
    if trns is not None:
        out += chunk(b"tRNS", bytes(trns))
    return out + chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b"")


def _cases():
    rng = np.random.RandomState(11)
    w, h = 13, 9
    allf = [0, 1, 2, 3, 4]
    yield "rgb8", _png(w, h, 8, 2, rng.randint(0, 256, (h, w, 3)), allf)
    yield "rgba8", _png(w, h, 8, 6, rng.randint(0, 256, (h, w, 4)), allf)
    yield "grey8", _png(w, h, 8, 0, rng.randint(0, 256, (h, w, 1)), allf)
    yield "greya8", _png(w, h, 8, 4, rng.randint(0, 256, (h, w, 2)), allf)
    yield "rgb16", _png(w, h, 16, 2, rng.randint(0, 65536, (h, w, 3)), allf)
    yield "rgba16", _png(w, h, 16, 6, rng.randint(0, 65536, (h, w, 4)), allf)
    yield "grey16", _png(w, h, 16, 0, rng.randint(0, 65536, (h, w, 1)), allf)
    yield "greya16", _png(w, h, 16, 4, rng.randint(0, 65536, (h, w, 2)), allf)
    for d in (1, 2, 4):
        yield "grey%d" % d, _png(w, h, d, 0, rng.randint(0, 1 << d, (h, w, 1)), [0, 2])
    pal = rng.randint(0, 256, 3 * 16)
    yield "pal8", _png(w, h, 8, 3, rng.randint(0, 16, (h, w, 1)), allf, plte=pal)
    yield "pal8_trns", _png(w, h, 8, 3, rng.randint(0, 16, (h, w, 1)), allf, plte=pal, trns=rng.randint(0, 256, 9))
    for d in (1, 2, 4):
        yield "pal%d" % d, _png(w, h, d, 3, rng.randint(0, min(16, 1 << d), (h, w, 1)), [0, 2], plte=pal)
    key = rng.randint(0, 256, (h, w, 3))
    key[2, 3] = key[5, 5] = (7, 200, 31)
    yield "rgb8_colourkey", _png(w, h, 8, 2, key, allf, trns=struct.pack(">HHH", 7, 200, 31))
    gk = rng.randint(0, 256, (h, w, 1))
    gk[1, 1] = gk[4, 2] = 99
    yield "grey8_colourkey", _png(w, h, 8, 0, gk, allf, trns=struct.pack(">H", 99))
    k16 = rng.randint(0, 65536, (h, w, 3))
    k16[3, 3] = k16[0, 12] = (513, 40000, 77)
    k16[6, 1] = (513 + 1, 40000, 77)                                   # same high bytes, different low byte: not the key
    yield "rgb16_colourkey", _png(w, h, 16, 2, k16, allf, trns=struct.pack(">HHH", 513, 40000, 77))
    g16 = rng.randint(0, 65536, (h, w, 1))
    g16[2, 2] = 0x1234
    g16[2, 3] = 0x1235
    yield "grey16_colourkey", _png(w, h, 16, 0, g16, allf, trns=struct.pack(">H", 0x1234))
    g4 = rng.randint(0, 16, (h, w, 1))
    yield "grey4_colourkey", _png(w, h, 4, 0, g4, [0, 1, 4], trns=struct.pack(">H", 5))
    g2 = rng.randint(0, 4, (h, w, 1))
    yield "grey2_colourkey", _png(w, h, 2, 0, g2, [0, 2], trns=struct.pack(">H", 2))
    for (hh, ww) in ((1, 1), (2, 3), (5, 4), (9, 13), (16, 16), (33, 21)):
        yield "adam7_rgb8_%dx%d" % (ww, hh), _png(ww, hh, 8, 2, rng.randint(0, 256, (hh, ww, 3)), allf, adam7=True)
        yield "adam7_rgba16_%dx%d" % (ww, hh), _png(ww, hh, 16, 6, rng.randint(0, 65536, (hh, ww, 4)), allf, adam7=True)
        yield "adam7_grey1_%dx%d" % (ww, hh), _png(ww, hh, 1, 0, rng.randint(0, 2, (hh, ww, 1)), [0, 2], adam7=True)
        yield "adam7_pal4_%dx%d" % (ww, hh), _png(ww, hh, 4, 3, rng.randint(0, 16, (hh, ww, 1)), [0, 1, 2], plte=pal, trns=rng.randint(0, 256, 5), adam7=True)


CASES = list(_cases())


def _scene_with_images(tmp_path, names):
    """One triangle, one camera, one BLEND material per image (base colour texture, factor 1)."""
    pos = np.float32([[0, 0, 0], [1, 0, 0], [0, 1, 0]])
    nrm = np.float32([[0, 0, 1]] * 3)
    idx = np.uint16([0, 1, 2, 0])
    blob = pos.tobytes() + nrm.tobytes() + idx.tobytes()
    open(str(tmp_path / "s.bin"), "wb").write(blob)
    doc = {"asset": {"version": "2.0"}, "scene": 0, "scenes": [{"nodes": [0, 1]}],
           "nodes": [{"mesh": 0}, {"camera": 0, "translation": [0, 0, 3]}],
           "meshes": [{"primitives": [{"attributes": {"POSITION": 0, "NORMAL": 1}, "indices": 2, "material": 0}]}],
           "cameras": [{"type": "perspective", "perspective": {"yfov": 0.8, "aspectRatio": 1.0, "znear": 0.1}}],
           "buffers": [{"uri": "s.bin", "byteLength": len(blob)}],
           "bufferViews": [{"buffer": 0, "byteOffset": 0, "byteLength": 36}, {"buffer": 0, "byteOffset": 36, "byteLength": 36},
                           {"buffer": 0, "byteOffset": 72, "byteLength": 6}],
           "accessors": [{"bufferView": 0, "componentType": 5126, "count": 3, "type": "VEC3"},
                         {"bufferView": 1, "componentType": 5126, "count": 3, "type": "VEC3"},
                         {"bufferView": 2, "componentType": 5123, "count": 3, "type": "SCALAR"}],
           "images": [{"uri": n + ".png"} for n in names],
           "textures": [{"source": i} for i in range(len(names))],
           "materials": [{"alphaMode": "BLEND", "pbrMetallicRoughness": {"baseColorTexture": {"index": i}}} for i in range(len(names))]}
    path = str(tmp_path / "s.gltf")
    json.dump(doc, open(path, "w"))
    return path


def test_loader_texels_equal_stbi_loadf(stb, tmp_path):
    for name, data in CASES:
        open(str(tmp_path / (name + ".png")), "wb").write(data)
    hs = rtc.HostSceneData(_scene_with_images(tmp_path, [n for n, _ in CASES]), height=8, num_threads=3)
    for i, (name, data) in enumerate(CASES):
        want, actual_c = _stb_load(stb, data, 4, True)                # what src/stage1.zig:56-63 gets
        m = hs.materials[i]
        assert m.base_color.texels.shape == want[:, :, :3].shape, name
        assert np.array_equal(m.base_color.texels.view(np.uint32), want[:, :, :3].view(np.uint32)), name
        if actual_c in (2, 4):                                        # src/stage1.zig:455: only images that carry alpha
            assert np.array_equal(m.transparency.texels[:, :, 0].view(np.uint32), want[:, :, 3].view(np.uint32)), name
        else:
            assert m.transparency.texels.shape == (1, 1, 1) and m.transparency.texels[0, 0, 0] == 1.0, name


def test_png_reader_equals_stbi_load(stb, tmp_path):
    for name, data in CASES:
        path = str(tmp_path / (name + ".png"))
        open(path, "wb").write(data)
        want, _ = _stb_load(stb, data, 3, False)
        assert np.array_equal(rtc.png_read(path), want), name


def test_png_writer_and_stbi_write_png_are_interchangeable(stb, tmp_path):
    rng = np.random.RandomState(5)
    for (h, w) in ((1, 1), (7, 5), (270, 480), (700, 1100)):           # the last one spans several deflate slices
        img = rng.randint(0, 256, (h, w, 3)).astype(np.uint8)
        img[h // 2:] = (np.linspace(0, 255, w)[None, :, None]).astype(np.uint8)
        mine = str(tmp_path / "mine.png")
        rtc.png_write(mine, img)
        got, c = _stb_load(stb, open(mine, "rb").read(), 3, False)
        assert c == 3 and np.array_equal(got, img)
        theirs = str(tmp_path / "theirs.png")
        assert stb.stbi_write_png(theirs.encode(), w, h, 3, img.ctypes.data, 3 * w) == 1     # src/main.zig:130-139
        assert np.array_equal(rtc.png_read(theirs), img)


def test_texture_images_of_the_test_scenes_equal_stb(stb, tmp_path):
    """The PNGs the procedural scenes' writer stores (what every other loader test reads)."""
    rng = np.random.RandomState(2)
    for c in (1, 2, 3, 4):
        data = gltf_io.png_encode(rng.randint(0, 256, (17, 31, c)).astype(np.uint8))
        path = str(tmp_path / ("c%d.png" % c))
        open(path, "wb").write(data)
        want, actual = _stb_load(stb, data, 3, False)
        assert actual == c and np.array_equal(rtc.png_read(path), want)


def _jpeg_cases():
    PIL = pytest.importorskip("PIL.Image")
    import io
    rng = np.random.RandomState(21)

    def picture(h, w):
        y, x = np.mgrid[0:h, 0:w]
        img = np.stack([128 + 100 * np.sin(x / 7.0) * np.cos(y / 5.0), (x * 255.0 / max(1, w - 1)), (y * 255.0 / max(1, h - 1))], axis=2)
        img += rng.normal(0, 12, img.shape)
        img[h // 3:h // 2, w // 4:w // 2] = (250, 10, 10)                 # hard edges: chroma upsampling shows
        return np.clip(img, 0, 255).astype(np.uint8)

    def enc(img, mode="RGB", **kw):
        b = io.BytesIO()
        PIL.fromarray(img if mode == "RGB" else img[:, :, 0], mode).save(b, "JPEG", **kw)
        return b.getvalue()
    for (h, w) in ((1, 1), (9, 13), (33, 17), (48, 64), (67, 131)):
        img = picture(h, w)
        for sub in (0, 1, 2):
            yield "baseline_%dx%d_sub%d" % (w, h, sub), enc(img, quality=80, subsampling=sub)
            yield "progressive_%dx%d_sub%d" % (w, h, sub), enc(img, quality=70, subsampling=sub, progressive=True)
        yield "grey_%dx%d" % (w, h), enc(img, "L", quality=85)
        yield "grey_progressive_%dx%d" % (w, h), enc(img, "L", quality=60, progressive=True)
    img = picture(120, 200)
    yield "optimized_q30", enc(img, quality=30, optimize=True)
    yield "q100", enc(img, quality=100, subsampling=0)
    yield "q5", enc(img, quality=5, subsampling=2)
    try:
        yield "restart_blocks", enc(img, quality=75, subsampling=2, restart_marker_blocks=3)
        yield "restart_rows_progressive", enc(img, quality=75, subsampling=1, restart_marker_rows=1, progressive=True)
    except TypeError:
        pass
    b = io.BytesIO()
    cmyk = np.concatenate([img, 255 - img[:, :, :1] // 2], axis=2)
    PIL.fromarray(cmyk, "CMYK").save(b, "JPEG", quality=80)
    yield "cmyk", b.getvalue()
    b = io.BytesIO()
    PIL.fromarray(cmyk, "CMYK").save(b, "JPEG", quality=60, progressive=True, subsampling=2)
    yield "cmyk_progressive", b.getvalue()
    try:
        import cv2
        bgr = np.ascontiguousarray(img[:, :, ::-1])
        for name, sf in (("411", cv2.IMWRITE_JPEG_SAMPLING_FACTOR_411), ("440", cv2.IMWRITE_JPEG_SAMPLING_FACTOR_440)):
            ok, buf = cv2.imencode(".jpg", bgr, [cv2.IMWRITE_JPEG_QUALITY, 85, cv2.IMWRITE_JPEG_SAMPLING_FACTOR, sf, cv2.IMWRITE_JPEG_RST_INTERVAL, 5])
            if ok:
                yield "cv2_" + name, buf.tobytes()
    except (ImportError, AttributeError):
        pass


def test_jpeg_textures_equal_stbi_loadf(stb, tmp_path):
    """JPEG textures through the loader against the reference's stb: baseline and progressive, 4:4:4 / 4:2:2 / 4:2:0 /
    4:1:1 / 4:4:0, greyscale, custom Huffman tables, restart intervals, sizes that are not multiples of the MCU."""
    cases = list(_jpeg_cases())
    assert len(cases) > 40
    for name, data in cases:
        open(str(tmp_path / (name + ".png")), "wb").write(data)          # the loader goes by content, not by extension
    hs = rtc.HostSceneData(_scene_with_images(tmp_path, [n for n, _ in cases]), height=8, num_threads=4)
    for i, (name, data) in enumerate(cases):
        want, actual_c = _stb_load(stb, data, 4, True)
        m = hs.materials[i]
        assert m.base_color.texels.shape == want[:, :, :3].shape, name
        diff = m.base_color.texels.view(np.uint32) != want[:, :, :3].view(np.uint32)
        assert not diff.any(), (name, int(diff.sum()), diff.shape)
        assert m.transparency.texels.shape == (1, 1, 1), name             # JPEG has no alpha (actual_c 3 or 1)


def test_broken_jpegs_are_reported(tmp_path):
    PIL = pytest.importorskip("PIL.Image")
    import io
    good = io.BytesIO()
    PIL.fromarray(np.full((16, 16, 3), 90, np.uint8), "RGB").save(good, "JPEG")
    open(str(tmp_path / "cut.png"), "wb").write(good.getvalue()[:40])               # cut inside the tables
    with pytest.raises(rtc.RtcError):
        rtc.HostSceneData(_scene_with_images(tmp_path, ["cut"]), height=8)
    data = bytearray(good.getvalue())
    data[data.index(b"\xff\xc0") + 1] = 0xC9                                       # SOF9: arithmetic coding
    open(str(tmp_path / "arith.png"), "wb").write(bytes(data))
    with pytest.raises(rtc.RtcError, match="arithmetic"):
        rtc.HostSceneData(_scene_with_images(tmp_path, ["arith"]), height=8)


def test_stb_compatible_writer_is_byte_identical(stb, tmp_path, monkeypatch):
    """RTC_PNG_STB=1: the file is the one stbi_write_png (src/main.zig:130-135) writes, byte for byte — filter choice per
    row, stb's own deflate (hash buckets, lazy matching, fixed Huffman block, stored-block fallback), Adler, CRCs."""
    rng = np.random.RandomState(9)
    monkeypatch.setenv("RTC_PNG_STB", "1")
    frames = []
    for (h, w) in ((1, 1), (2, 3), (5, 7), (64, 64), (200, 333)):
        frames.append(rng.randint(0, 256, (h, w, 3)).astype(np.uint8))                           # noise: stored blocks
        y, x = np.mgrid[0:h, 0:w]
        g = np.stack([x * 255 // max(1, w - 1), y * 255 // max(1, h - 1), (x + y) % 256], axis=2).astype(np.uint8)
        frames.append(g)                                                                          # gradients: filters matter
        frames.append(np.full((h, w, 3), 77, np.uint8))                                           # flat: long matches
        frames.append(np.where(rng.rand(h, w, 1) < 0.02, rng.randint(0, 256, (h, w, 3)), g).astype(np.uint8))
    tile = rng.randint(0, 256, (16, 16, 3)).astype(np.uint8)
    frames.append(np.tile(tile, (40, 50, 1)))                                                     # repeats beyond the bucket depth and the window
    frames.append((np.clip(rng.normal(128, 6, (300, 400, 3)), 0, 255)).astype(np.uint8))         # what a rendered frame looks like
    for k, img in enumerate(frames):
        img = np.ascontiguousarray(img)
        h, w, _ = img.shape
        mine, theirs = str(tmp_path / "mine.png"), str(tmp_path / "theirs.png")
        rtc.png_write(mine, img)
        assert stb.stbi_write_png(theirs.encode(), w, h, 3, img.ctypes.data, 3 * w) == 1
        a, b = open(mine, "rb").read(), open(theirs, "rb").read()
        assert a == b, (k, img.shape, len(a), len(b))
        assert np.array_equal(rtc.png_read(mine), img)


def test_committed_stb_vectors(tmp_path, monkeypatch):
    """tests/golden/stb_golden.npz holds outputs of the reference's own unit (made by tests/golden/make_stb_golden.py where
    /root/reference exists): stbi_loadf_from_memory(..., 4) of 20 small PNG / JPEG files and the files stbi_write_png
    wrote for 3 frames. Checked on every machine, with or without the built unit."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "stb_golden.npz"))
    names = [str(n) for n in g["names"]]
    assert len(names) >= 20
    for n in names:
        open(str(tmp_path / (n + ".png")), "wb").write(g["file_" + n].tobytes())
    hs = rtc.HostSceneData(_scene_with_images(tmp_path, names), height=8, num_threads=2)
    for i, n in enumerate(names):
        want, m = g["loadf_" + n], hs.materials[i]
        assert np.array_equal(m.base_color.texels.view(np.uint32), np.ascontiguousarray(want[:, :, :3]).view(np.uint32)), n
        if int(g["comp_" + n]) in (2, 4):
            assert np.array_equal(m.transparency.texels[:, :, 0].view(np.uint32), np.ascontiguousarray(want[:, :, 3]).view(np.uint32)), n
        else:
            assert m.transparency.texels.shape == (1, 1, 1), n
    monkeypatch.setenv("RTC_PNG_STB", "1")
    for f in [str(x) for x in g["frames"]]:
        path = str(tmp_path / (f + "_out.png"))
        rtc.png_write(path, g["frame_" + f])
        assert open(path, "rb").read() == g["png_" + f].tobytes(), f
        theirs = str(tmp_path / (f + "_stb.png"))
        open(theirs, "wb").write(g["png_" + f].tobytes())
        assert np.array_equal(rtc.png_read(theirs), g["frame_" + f]), f
