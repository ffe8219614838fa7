"""Regenerates tests/golden/stb_golden.npz by RUNNING THE REFERENCE'S OWN C UNIT (oracle/_ref/libc_libs.so =
/root/reference/src/c_impl.c, built by `make -C oracle ref`): for a set of small PNG and JPEG files, what
stbi_loadf_from_memory(..., 4) returns (src/stage1.zig:56-63), and for a few frames the file stbi_write_png writes
(src/main.zig:130-135). tests/test_ref_stb.py::test_committed_stb_vectors checks the loader and the PNG writer against
these vectors on machines that have neither /root/reference nor the built unit.
    python tests/golden/make_stb_golden.py        (needs oracle/_ref/libc_libs.so, Pillow, OpenCV)"""
import ctypes as C
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import test_ref_stb as T  # noqa: E402


def small_jpegs():
    """Restart intervals, 4:1:1 / 4:4:0 and CMYK at a size that keeps the fixture small."""
    import io
    import cv2
    from PIL import Image
    rng = np.random.RandomState(33)
    y, x = np.mgrid[0:24, 0:40]
    img = np.clip(np.stack([128 + 100 * np.sin(x / 5.0), x * 6.0, y * 10.0], axis=2) + rng.normal(0, 10, (24, 40, 3)), 0, 255).astype(np.uint8)
    b = io.BytesIO(); Image.fromarray(img, "RGB").save(b, "JPEG", quality=75, subsampling=2, restart_marker_blocks=2); yield "small_restart", b.getvalue()
    b = io.BytesIO(); Image.fromarray(np.concatenate([img, 255 - img[:, :, :1] // 2], axis=2), "CMYK").save(b, "JPEG", quality=80); yield "small_cmyk", b.getvalue()
    bgr = np.ascontiguousarray(img[:, :, ::-1])
    for name, sf in (("small_411", cv2.IMWRITE_JPEG_SAMPLING_FACTOR_411), ("small_440", cv2.IMWRITE_JPEG_SAMPLING_FACTOR_440)):
        ok, buf = cv2.imencode(".jpg", bgr, [cv2.IMWRITE_JPEG_QUALITY, 85, cv2.IMWRITE_JPEG_SAMPLING_FACTOR, sf, cv2.IMWRITE_JPEG_RST_INTERVAL, 3])
        assert ok
        yield name, buf.tobytes()


def main():
    L = C.CDLL(T.REF_SO)
    L.stbi_loadf_from_memory.restype = C.POINTER(C.c_float)
    L.stbi_loadf_from_memory.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_int]
    L.stbi_load_from_memory.restype = C.POINTER(C.c_uint8)
    L.stbi_load_from_memory.argtypes = L.stbi_loadf_from_memory.argtypes
    L.stbi_image_free.argtypes = [C.c_void_p]
    L.stbi_write_png.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
    out = {}
    keep_png = ("rgb8", "rgba8", "greya8", "rgb16", "grey2", "pal8_trns", "pal4", "rgb8_colourkey", "grey16_colourkey",
                "adam7_rgb8_9x13", "adam7_rgba16_5x4", "adam7_pal4_33x21", "adam7_grey1_16x16")
    keep_jpg = ("baseline_13x9_sub0", "baseline_17x33_sub2", "progressive_17x33_sub1", "progressive_64x48_sub2", "grey_17x33",
                "grey_progressive_13x9")
    cases = [(n, d) for n, d in T.CASES if n in keep_png] + [(n, d) for n, d in T._jpeg_cases() if n in keep_jpg] + list(small_jpegs())
    names = []
    for name, data in cases:
        texels, actual_c = T._stb_load(L, data, 4, True)
        out["file_" + name] = np.frombuffer(data, np.uint8)
        out["loadf_" + name] = texels
        out["comp_" + name] = np.int32(actual_c)
        names.append(name)
    rng = np.random.RandomState(9)
    frames = {"noise_7x5": rng.randint(0, 256, (5, 7, 3)).astype(np.uint8),
              "gradient_40x24": np.stack(np.broadcast_arrays(np.arange(40)[None, :] * 6, np.arange(24)[:, None] * 10, 99), axis=2).astype(np.uint8),
              "render_like_96x64": np.clip(rng.normal(128, 6, (64, 96, 3)), 0, 255).astype(np.uint8)}
    tmp = tempfile.mkdtemp()
    for name, img in frames.items():
        img = np.ascontiguousarray(img)
        path = os.path.join(tmp, name + ".png")
        assert L.stbi_write_png(path.encode(), img.shape[1], img.shape[0], 3, img.ctypes.data, 3 * img.shape[1]) == 1
        out["frame_" + name] = img
        out["png_" + name] = np.frombuffer(open(path, "rb").read(), np.uint8)
        os.remove(path)
    os.rmdir(tmp)
    out["names"] = np.array(names)
    out["frames"] = np.array(list(frames))
    np.savez_compressed(os.path.join(HERE, "stb_golden.npz"), **out)
    print("wrote", len(names), "image cases and", len(frames), "frames:", os.path.getsize(os.path.join(HERE, "stb_golden.npz")), "bytes")


if __name__ == "__main__":
    main()
