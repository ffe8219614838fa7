"""stage1's triangle assembly on the device (rtc_gltf_load_device / rtc_scene_create_from_gltf)
against the host loader and the oracle's restatement of loadTriangles (src/stage1.zig:217-259): the soup bit for
bit (positions, normals, texcoords, material, source index) for every index type, strided and unaligned
accessors and primitives without TEXCOORD_0; then a render from the scene built straight from the file against a
render from the soup the host loader produced."""
import numpy as np
import pytest

from gltf_cases import handmade, oracle_soup
from zig_raytracing_contest_b200 import abi, gltf_io, rtc, scenes

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def _same_host_scene(a, b):
    assert a.positions.shape == b.positions.shape
    assert np.array_equal(_bits(a.positions), _bits(b.positions))
    assert np.array_equal(a.tri_data.view(np.uint8), b.tri_data.view(np.uint8))
    assert bytes(a.camera) == bytes(b.camera)
    assert len(a.materials) == len(b.materials) and a.num_images == b.num_images
    for ma, mb in zip(a.materials, b.materials):
        for f in ("base_color", "emissive", "transparency"):
            ta, tb = getattr(ma, f), getattr(mb, f)
            assert np.array_equal(_bits(ta.texels), _bits(tb.texels))
            assert (ta.u_min, ta.u_max, ta.v_min, ta.v_max) == (tb.u_min, tb.u_max, tb.v_min, tb.v_max)


@pytest.mark.parametrize("scene", ["atrium", "cornell"])
@pytest.mark.parametrize("flavour", ["gltf", "glb", "embed"])
def test_device_assembly_matches_host_loader(tmp_path, flavour, scene):
    """Triangles AND texels (device linearisation: colour * factor, MASK / BLEND alpha) against the host loader."""
    spec = scenes.atrium(detail=0.3) if scene == "atrium" else scenes.cornell(subdiv=6)
    path = str(tmp_path / ("scene.glb" if flavour == "glb" else "scene.gltf"))
    gltf_io.write_gltf(spec, path, node_translation=(0.25, -0.5, 1.0), node_scale=2.0, glb=flavour == "glb", embed=flavour == "embed")
    host = rtc.HostSceneData(path, camera="Camera 1", height=72, num_threads=3)
    dev = rtc.HostSceneData(path, camera="Camera 1", height=72, num_threads=3, device=0)
    assert len(host.positions) == len(spec.positions) > 400
    assert any(m.base_color.texels.shape[0] > 1 for m in host.materials) and any(m.transparency.texels.shape[0] > 1 for m in host.materials)
    _same_host_scene(host, dev)


def test_jpeg_textures_through_the_device_path(tmp_path):
    """JPEG-textured materials: decoded on the host (csrc/rtc_jpeg.h), linearised on the device."""
    pytest.importorskip("PIL.Image")
    from test_ref_stb import _jpeg_cases, _scene_with_images
    cases = [c for c in _jpeg_cases() if "131x67" in c[0] or "restart" in c[0]]
    for name, data in cases:
        open(str(tmp_path / (name + ".png")), "wb").write(data)
    path = _scene_with_images(tmp_path, [n for n, _ in cases])
    _same_host_scene(rtc.HostSceneData(path, height=8, num_threads=2), rtc.HostSceneData(path, height=8, num_threads=2, device=0))


def test_device_assembly_index_types_strides_alignment(orc, tmp_path):
    path, doc, want = handmade(tmp_path, np.random.default_rng(5))
    host = rtc.HostSceneData(path, height=32)
    dev = rtc.HostSceneData(path, height=32, device=0)
    assert len(host.positions) == 50 + 900 + 211
    _same_host_scene(host, dev)
    assert np.array_equal(dev.tri_data["src_tri"], np.arange(len(dev.tri_data), dtype=np.uint32))
    assert np.all(dev.tri_data["texcoord"][50:950] == 0.0)
    wp, wd = oracle_soup(orc, doc, want)
    assert np.array_equal(_bits(dev.positions), _bits(wp))
    for f in ("normal", "texcoord"):
        assert np.array_equal(_bits(dev.tri_data[f]), _bits(wd[f])), f
    assert np.array_equal(dev.tri_data["material"], wd["material"])


def test_scene_from_gltf_renders_like_the_host_path(orc, tmp_path):
    spec = scenes.atrium(detail=0.35)
    path = str(tmp_path / "atrium.glb")
    gltf_io.write_gltf(spec, path, node_translation=(0.1, 0.2, -0.3), node_scale=0.5, glb=True)
    res = (64, 64, 64)
    hs = rtc.HostSceneData(path, height=108, num_threads=2)
    a = rtc.Scene.from_soup(hs.positions, hs.tri_data, res, hs.materials, device=0)
    b = rtc.Scene.from_gltf(path, res, height=108, num_threads=2, device=0)
    assert bytes(b.camera) == bytes(hs.camera)
    ia, ib = a.build_info, b.build_info
    for k in ("num_cells", "num_tris", "num_src_tris", "empty_cells", "min_tris", "max_tris"):
        assert getattr(ia, k) == getattr(ib, k), k
    assert bytes(ia.grid) == bytes(ib.grid)
    assert ib.assemble_ms > 0.0 and ia.assemble_ms == 0.0 and ib.total_ms > ib.device_ms
    for prm in (rtc.make_params(mode=abi.RTC_MODE_PRIMARY), rtc.make_params(num_samples=5, max_bounce=4, seed=11)):
        ra = a.render(hs.camera, prm, want_radiance=True, want_hits=True)
        rb = b.render(b.camera, prm, want_radiance=True, want_hits=True)
        assert np.array_equal(_bits(ra["radiance"]), _bits(rb["radiance"]))
        assert np.array_equal(ra["hit_id"], rb["hit_id"]) and np.array_equal(ra["rgb"], rb["rgb"])
    # the oracle on the host-built geometry of the host-loaded soup
    geo = rtc.Geometry(hs.positions, hs.tri_data, res)
    c = rtc.Scene(geo, hs.materials, device=0)
    prm = rtc.make_params(num_samples=2, max_bounce=3, seed=3)
    ro = orc.render(c.desc, hs.camera, prm, want_radiance=True)
    rb = b.render(b.camera, prm, want_radiance=True)
    assert np.array_equal(_bits(rb["radiance"]), _bits(ro["radiance"]))
    a.close(); b.close(); c.close()


def test_from_gltf_errors(tmp_path):
    path, _, _ = handmade(tmp_path, np.random.default_rng(6), bad_index=True)
    for load in (lambda: rtc.HostSceneData(path, height=32), lambda: rtc.HostSceneData(path, height=32, device=0),
                 lambda: rtc.Scene.from_gltf(path, (8, 8, 8), height=32, device=0)):
        with pytest.raises(rtc.RtcError) as e:
            load()
        assert e.value.code == abi.RTC_ERR_FORMAT and "index out of range" in str(e.value)
    with pytest.raises(rtc.RtcError) as e:
        rtc.Scene.from_gltf(str(tmp_path / "missing.glb"), (8, 8, 8), height=32, device=0)
    assert e.value.code == abi.RTC_ERR_IO
    good, _, _ = handmade(tmp_path, np.random.default_rng(7))
    with pytest.raises(rtc.RtcError) as e:
        rtc.Scene.from_gltf(good, (0, 8, 8), height=32, device=0)
    assert e.value.code == abi.RTC_ERR_INVALID_ARGUMENT
    s = rtc.Scene.from_gltf(good, (8, 8, 8), height=32, device=0)      # and the good file renders
    out = s.render(s.camera, rtc.make_params(num_samples=1, max_bounce=1, seed=1))
    assert out["rgb"].shape == (32, 48, 3)
    s.close()
