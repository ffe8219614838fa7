"""The reference's file contracts as the path's callers: config.json,
glTF in (rtc_gltf_load = stage1), PNG out, and the `rtc` command line (src/main.zig:73-143).
CPU tests check the loader bit-for-bit against the oracle's restatement of stage1 and the CLI's
contract up to the device boundary; the GPU test renders through the CLI and compares the PNG with the
same frame rendered through the Python binding."""
import json
import os
import subprocess

import numpy as np
import pytest

from gltf_cases import handmade, oracle_soup
from zig_raytracing_contest_b200 import abi, gltf_io, rtc, scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "zig_raytracing_contest_b200", "bin", "rtc")
CONFIG = {"grid_resolution": [16, 16, 16], "num_threads": None, "num_samples": 3, "max_bounce": 4}


def _read_accessor(doc, blob, idx):
    acc = doc["accessors"][idx]
    view = doc["bufferViews"][acc["bufferView"]]
    dt, n = {5126: (np.float32, None), 5123: (np.uint16, 1)}[acc["componentType"]]
    ncomp = {"VEC3": 3, "VEC2": 2, "SCALAR": 1}[acc["type"]]
    off = view.get("byteOffset", 0) + acc.get("byteOffset", 0)
    return np.frombuffer(blob, dt, acc["count"] * ncomp, off).reshape(acc["count"], ncomp)


@pytest.mark.parametrize("flavour", ["gltf", "glb", "embed"])
def test_gltf_loader_matches_oracle_stage1(orc, tmp_path, flavour):
    spec = scenes.cornell(subdiv=2)
    path = str(tmp_path / ("scene.glb" if flavour == "glb" else "scene.gltf"))
    doc = gltf_io.write_gltf(spec, path, node_translation=(0.25, -0.5, 1.0), node_scale=2.0,
                             glb=flavour == "glb", embed=flavour == "embed")
    hs = rtc.HostSceneData(path, camera="Camera 1", height=72, num_threads=3)
    # ---- triangles: every mesh node in file order, parent * child transform (zgltf) ----
    if flavour == "gltf":
        blob = open(str(tmp_path / "scene.bin"), "rb").read()
        parent = orc.mat4_recompose(doc["nodes"][0]["translation"], (0, 0, 0, 1), (1, 1, 1))
        want_p, want_d = [], []
        for node in doc["nodes"]:
            if "mesh" not in node:
                continue
            prim = doc["meshes"][node["mesh"]]["primitives"][0]
            m = orc.mat4_mul(parent, orc.mat4_recompose((0, 0, 0), (0, 0, 0, 1), node["scale"]))
            a = prim["attributes"]
            p, d = orc.assemble_triangles(m, _read_accessor(doc, blob, a["POSITION"]), _read_accessor(doc, blob, a["NORMAL"]),
                                          _read_accessor(doc, blob, a["TEXCOORD_0"]), _read_accessor(doc, blob, prim["indices"]),
                                          prim["material"])
            want_p.append(p); want_d.append(d)
        want_p, want_d = np.concatenate(want_p), np.concatenate(want_d)
        assert np.array_equal(hs.positions.view(np.uint32), want_p.view(np.uint32))
        for f in ("normal", "texcoord"):
            assert np.array_equal(hs.tri_data[f].view(np.uint32), want_d[f].view(np.uint32)), f
        assert np.array_equal(hs.tri_data["material"], want_d["material"])
        assert np.array_equal(hs.tri_data["src_tri"], np.arange(len(want_d), dtype=np.uint32))
    assert hs.positions.shape == spec.positions.shape
    # ---- camera ----
    want = orc.load_camera(spec.cam_matrix, spec.yfov, np.float32(spec.aspect), None, 72)
    assert (hs.camera.w, hs.camera.h) == (want.w, want.h) == (128, 72)
    for f in ("origin", "lower_left_corner", "right", "up"):
        np.testing.assert_allclose(list(getattr(hs.camera, f)), list(getattr(want, f)), rtol=3e-6, atol=1e-5)
    # ---- materials: factors, stb-style linearisation, MASK/BLEND alpha, sampler clamp ----
    assert len(hs.materials) == len(spec.materials) and hs.num_images == 3
    m0 = hs.materials[1]
    assert m0.base_color.texels.shape == (1, 1, 3) and (m0.base_color.u_min, m0.base_color.u_max) == (0, 0)
    np.testing.assert_array_equal(m0.base_color.texels.reshape(3), np.float32([0.65, 0.05, 0.05]))
    np.testing.assert_array_equal(hs.materials[3].emissive.texels.reshape(3), np.float32([15, 15, 15]))
    img8 = gltf_io._to8(spec.materials[4].base_color.texels)                       # what the writer stored
    lin = np.power((img8.astype(np.float32) / np.float32(255.0)).astype(np.float64), float(np.float32(2.2))).astype(np.float32)   # stb: gamma is 2.2f
    assert np.array_equal(hs.materials[4].base_color.texels, lin)
    assert hs.materials[4].base_color.u_min == np.iinfo(np.int32).min                # REPEAT: no clamp range
    m5 = hs.materials[5]
    assert (m5.base_color.u_min, m5.base_color.u_max, m5.base_color.v_max) == (0, 63, 63)   # CLAMP_TO_EDGE
    assert set(np.unique(m5.transparency.texels)) == {0.0, 1.0}                      # MASK: alpha > cutoff
    m6 = hs.materials[6]
    assert m6.transparency.texels.shape == (1, 1, 1)
    np.testing.assert_array_equal(m6.transparency.texels.reshape(1), np.float32([128.0]) / np.float32(255.0))   # BLEND: alpha as is


def test_loader_size_rules_and_errors(tmp_path):
    spec = scenes.cornell(subdiv=1)
    path = str(tmp_path / "s.gltf")
    gltf_io.write_gltf(spec, path)
    for kw, msg in (({}, "OutputImgSizeIsNotSpecified"), ({"width": 10, "height": 10}, "CameraHasAspectRatio"),
                    ({"height": 10, "camera": "nope"}, "CameraNotFound")):
        with pytest.raises(rtc.RtcError, match=msg):
            rtc.HostSceneData(path, **kw)
    assert rtc.HostSceneData(path, width=160).camera.h == 90          # trunc(w / aspectRatio), src/stage1.zig:341
    with pytest.raises(rtc.RtcError) as e:
        rtc.HostSceneData(str(tmp_path / "missing.gltf"), height=4)
    assert e.value.code == abi.RTC_ERR_IO
    open(str(tmp_path / "bad.gltf"), "w").write("{ not json")
    with pytest.raises(rtc.RtcError) as e:
        rtc.HostSceneData(str(tmp_path / "bad.gltf"), height=4)
    assert e.value.code == abi.RTC_ERR_FORMAT
    doc = json.load(open(path))
    del doc["meshes"][0]["primitives"][0]["material"]                 # primitive.material.? (src/stage1.zig:239)
    json.dump(doc, open(str(tmp_path / "nomat.gltf"), "w"))
    with pytest.raises(rtc.RtcError, match="without material"):
        rtc.HostSceneData(str(tmp_path / "nomat.gltf"), height=4)


def test_loader_index_types_strides_alignment_match_oracle(orc, tmp_path):
    """u8 / u16 / u32 indices, interleaved byteStride, accessors at unaligned offsets, no TEXCOORD_0, rotated
    nested nodes: the host loader against the oracle's loadTriangles + zgltf transforms, bit for bit."""
    path, doc, want = handmade(tmp_path, np.random.default_rng(5))
    hs = rtc.HostSceneData(path, height=32)
    wp, wd = oracle_soup(orc, doc, want)
    assert len(hs.positions) == 50 + 900 + 211
    assert np.array_equal(hs.positions.view(np.uint32), wp.view(np.uint32))
    assert np.array_equal(hs.tri_data.view(np.uint8), wd.view(np.uint8))
    assert (hs.camera.w, hs.camera.h) == (48, 32)
    bad, _, _ = handmade(tmp_path, np.random.default_rng(6), bad_index=True)
    with pytest.raises(rtc.RtcError) as e:
        rtc.HostSceneData(bad, height=32)
    assert e.value.code == abi.RTC_ERR_FORMAT


def test_device_assembly_needs_a_device(tmp_path):
    """rtc_gltf_load_device / rtc_scene_create_from_gltf have no host fallback."""
    if rtc.device_count() > 0:
        pytest.skip("a GPU is present")
    path, _, _ = handmade(tmp_path, np.random.default_rng(5))
    with pytest.raises(rtc.RtcError) as e:
        rtc.HostSceneData(path, height=32, device=0)
    assert e.value.code == abi.RTC_ERR_NO_DEVICE
    with pytest.raises(rtc.RtcError) as e:
        rtc.Scene.from_gltf(path, (8, 8, 8), height=32)
    assert e.value.code == abi.RTC_ERR_NO_DEVICE
    with pytest.raises(rtc.RtcError) as e:                 # file errors come first, as in the reference's order of work
        rtc.Scene.from_gltf(str(tmp_path / "nope.gltf"), (8, 8, 8), height=32)
    assert e.value.code == abi.RTC_ERR_IO


def test_config_contract(tmp_path):
    p = str(tmp_path / "config.json")
    json.dump(CONFIG, open(p, "w"))
    assert rtc.load_config(p) == {"grid_resolution": (16, 16, 16), "num_threads": None, "num_samples": 3, "max_bounce": 4}
    ref_cfg = os.path.join(ROOT, "tests", "golden", "config.json")
    assert rtc.load_config(ref_cfg) == {"grid_resolution": (128, 128, 128), "num_threads": None, "num_samples": 3, "max_bounce": 4}
    for bad in ({k: v for k, v in CONFIG.items() if k != "max_bounce"}, dict(CONFIG, extra=1), dict(CONFIG, num_samples=70000),
                dict(CONFIG, grid_resolution=[1, 2])):
        json.dump(bad, open(p, "w"))
        with pytest.raises(rtc.RtcError) as e:
            rtc.load_config(p)
        assert e.value.code == abi.RTC_ERR_FORMAT


def test_png_roundtrip_and_decoder_colour_types(tmp_path):
    rng = np.random.RandomState(3)
    rgb = rng.randint(0, 256, (13, 7, 3)).astype(np.uint8)
    rtc.png_write(str(tmp_path / "a.png"), rgb)
    assert np.array_equal(rtc.png_read(str(tmp_path / "a.png")), rgb)
    grey = rng.randint(0, 256, (5, 9, 1)).astype(np.uint8)
    open(str(tmp_path / "g.png"), "wb").write(gltf_io.png_encode(grey))
    assert np.array_equal(rtc.png_read(str(tmp_path / "g.png")), np.repeat(grey, 3, axis=2))
    rgba = rng.randint(0, 256, (4, 4, 4)).astype(np.uint8)
    open(str(tmp_path / "q.png"), "wb").write(gltf_io.png_encode(rgba))
    assert np.array_equal(rtc.png_read(str(tmp_path / "q.png")), rgba[:, :, :3])


def test_png_encoder_slices_are_one_valid_stream(tmp_path, monkeypatch):
    """A frame larger than several 1 MiB deflate slices: the file decodes with an independent inflater (Python's
    zlib: one stream, Adler-32 trailer checked) to the same pixels, and its bytes do not depend on the thread count."""
    import struct
    import zlib
    rng = np.random.RandomState(4)
    h, w = 900, 1400                                       # 3.8 MB of rows -> 4 slices
    rgb = (np.linspace(0, 255, w)[None, :, None] * np.ones((h, 1, 3))).astype(np.uint8)
    rgb[::3] = rng.randint(0, 256, (len(rgb[::3]), w, 3)).astype(np.uint8)      # noisy rows next to flat ones
    files = []
    for threads in ("1", "3", "16"):
        monkeypatch.setenv("RTC_PNG_THREADS", threads)
        path = str(tmp_path / ("t%s.png" % threads))
        rtc.png_write(path, rgb)
        files.append(open(path, "rb").read())
        assert np.array_equal(rtc.png_read(path), rgb)
    assert files[0] == files[1] == files[2]
    data, p, idat = files[0], 8, b""
    while p < len(data):
        n, typ = struct.unpack(">I4s", data[p:p + 8])
        body = data[p + 8:p + 8 + n]
        assert struct.unpack(">I", data[p + 8 + n:p + 12 + n])[0] == zlib.crc32(typ + body)
        if typ == b"IDAT":
            idat += body
        p += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 1 + 3 * w)
    assert np.all(raw[:, 0] == 0) and np.array_equal(raw[:, 1:].reshape(h, w, 3), rgb)


def _run_cli(cwd, *args, env=None):
    return subprocess.run([CLI] + list(args), cwd=str(cwd), capture_output=True, text=True, timeout=300,
                          env=dict(os.environ, **env) if env else None)


def test_cli_command_line_contract(tmp_path):
    assert os.path.exists(CLI), "bin/rtc is built by csrc/Makefile (__graft_entry__.build())"
    r = _run_cli(tmp_path, "--bogus", "1")
    assert r.returncode != 0 and "bogus" in r.stderr and "InvalidArguments" in r.stderr
    r = _run_cli(tmp_path, "--height", "72")                           # config.json is read from the cwd
    assert r.returncode != 0 and "config.json" in r.stderr
    json.dump(CONFIG, open(str(tmp_path / "config.json"), "w"))
    r = _run_cli(tmp_path, "--height", "72")                           # default --in is input.gltf
    assert r.returncode != 0 and "input.gltf" in r.stderr
    r = _run_cli(tmp_path, "--height=70000")                           # ?u16
    assert r.returncode != 0 and "height" in r.stderr


def test_cli_stops_at_the_device_boundary_without_gpu(tmp_path):
    if rtc.device_count() > 0:
        pytest.skip("a GPU is present")
    gltf_io.write_gltf(scenes.cornell(subdiv=1), str(tmp_path / "input.gltf"))
    json.dump(CONFIG, open(str(tmp_path / "config.json"), "w"))
    r = _run_cli(tmp_path, "--height=36")
    assert r.returncode != 0
    assert "Grid resolution: { 16, 16, 16 }" in r.stderr and "no CPU fallback" in r.stderr and "Rendered in" not in r.stderr
    r = _run_cli(tmp_path, "--height=36", env={"RTC_HOST_LOAD": "1"})          # stage1 on the host: runs, then stops the same way
    assert r.returncode != 0
    assert "Loaded in" in r.stderr and "Preprocessed in" in r.stderr and "Materials count: 7" in r.stderr
    assert "no CPU fallback" in r.stderr and "Rendered in" not in r.stderr
    assert not os.path.exists(str(tmp_path / "output.png"))


@pytest.mark.gpu
def test_cli_renders_the_same_frame_as_the_binding(tmp_path):
    spec = scenes.cornell(subdiv=3)
    gltf_io.write_gltf(spec, str(tmp_path / "room.glb"), node_translation=(0.1, 0.2, -0.3), node_scale=0.5, glb=True)
    json.dump(CONFIG, open(str(tmp_path / "config.json"), "w"))
    r = _run_cli(tmp_path, "--in", "room.glb", "--out=frame.png", "--camera", "Camera 1", "--height", "72", "--seed", "5")
    assert r.returncode == 0, r.stderr
    # the reference's six phase lines, in its order (src/main.zig:106,113,119,127,141,142)
    pos = [r.stderr.find("info: %s in " % p) for p in ("Loaded", "Preprocessed", "Compiled", "Rendered", "Saved", "Done")]
    assert all(p >= 0 for p in pos) and pos == sorted(pos), r.stderr
    got = rtc.png_read(str(tmp_path / "frame.png"))
    for env in ({"RTC_HOST_LOAD": "1"}, {"RTC_HOST_BUILD": "1"}):      # stage1 / stage1+2 on the host: same frame
        r = _run_cli(tmp_path, "--in", "room.glb", "--out=host.png", "--camera", "Camera 1", "--height", "72", "--seed", "5", env=env)
        assert r.returncode == 0, r.stderr
        for phase in ("Loaded in", "Preprocessed in", "Compiled in", "Rendered in", "Saved in", "Done in"):
            assert phase in r.stderr
        assert np.array_equal(rtc.png_read(str(tmp_path / "host.png")), got)
    hs = rtc.HostSceneData(str(tmp_path / "room.glb"), height=72)
    geo = rtc.Geometry(hs.positions, hs.tri_data, CONFIG["grid_resolution"])
    scene = rtc.Scene(geo, hs.materials, device=0)
    want = scene.render(hs.camera, rtc.make_params(num_samples=3, max_bounce=4, seed=5))["rgb"]
    scene.close()
    assert got.shape == (72, 128, 3) and np.array_equal(got, want)
