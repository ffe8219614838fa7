"""CPU-side checks of the drop-in boundary: librtc.so loads, exports every symbol include/rtc.h
declares, refuses to compute without a GPU (no CPU fallback), and its host-side stage2
(rtc_geometry_build) is bit-identical to the oracle's restatement of src/stage2.zig."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from zig_raytracing_contest_b200 import abi, rtc, scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "rtc.h")).read()
    declared = set(re.findall(r"\b(rtc_[a-z_]+)\s*\(", header))
    assert declared == set(abi.EXPORTED_SYMBOLS)
    L = rtc.lib()
    for name in declared:
        assert hasattr(L, name), name
    assert L.rtc_abi_version() == abi.RTC_ABI_VERSION


def test_struct_sizes_match_header_layout():
    assert C.sizeof(abi.Grid) == 48 and C.sizeof(abi.Cell) == 8 and C.sizeof(abi.TriPos) == 48
    assert C.sizeof(abi.TriData) == 68 and C.sizeof(abi.Texture) == 40 and C.sizeof(abi.Material) == 120
    assert C.sizeof(abi.Camera) == 56 and C.sizeof(abi.Params) == 56 and C.sizeof(abi.Stats) == 80


def _has_gpu():
    return rtc.device_count() > 0


def test_no_cpu_fallback_without_device():
    if _has_gpu():
        pytest.skip("a GPU is present")
    spec = scenes.cornell(subdiv=1)
    geo = rtc.Geometry(spec.positions, spec.tri_data, (4, 4, 4))
    with pytest.raises(rtc.RtcError) as e:
        rtc.Scene(geo, spec.materials)
    assert e.value.code == abi.RTC_ERR_NO_DEVICE
    with pytest.raises(rtc.RtcError):
        rtc.resolve(1, 1, 1)


def test_invalid_arguments_are_reported():
    L = rtc.lib()
    g = abi.Geometry()
    res = np.array([0, 4, 4], np.uint32)
    assert L.rtc_geometry_build(None, None, 0, res.ctypes.data, C.byref(g)) == abi.RTC_ERR_INVALID_ARGUMENT
    assert b"resolution" in L.rtc_last_error()
    assert L.rtc_scene_create(None, 0, None) == abi.RTC_ERR_INVALID_ARGUMENT


@pytest.mark.parametrize("name,res", [("cornell", (16, 16, 16)), ("cornell", (7, 5, 3)), ("atrium", (64, 64, 64))])
def test_geometry_build_matches_oracle(orc, name, res):
    spec = scenes.SCENES[name]() if name != "atrium" else scenes.atrium(detail=0.3)
    g = rtc.Geometry(spec.positions, spec.tri_data, res)
    o = orc.Geometry(spec.positions, spec.tri_data, res)
    assert bytes(g.grid) == bytes(o.grid)
    assert np.array_equal(g.cells, o.cells)
    assert np.array_equal(g.tri_pos.view(np.uint32), o.tri_pos.view(np.uint32))
    assert g.tri_data.tobytes() == o.tri_data.tobytes()
    assert (g.empty_cells, g.min_tris, g.max_tris) == (o.empty_cells, o.min_tris, o.max_tris)
    assert np.array_equal(g.tri_data["src_tri"], o.indices.astype(np.uint32))


def test_geometry_build_degenerate_inputs(orc):
    # empty soup, a zero-area triangle, and a triangle exactly on cell boundaries
    g = rtc.Geometry(np.zeros((0, 3, 3), np.float32), np.zeros(0, rtc.TRI_DATA_DTYPE), (4, 4, 4))
    assert g.num_tris == 0 and g.empty_cells == 64
    pos = np.array([[[0, 0, 0], [1, 0, 0], [0, 1, 0]],
                    [[0.5, 0.5, 0.5], [0.5, 0.5, 0.5], [0.5, 0.5, 0.5]],
                    [[0, 0, 1], [1, 0, 1], [0, 1, 1]]], np.float32)
    td = np.zeros(3, rtc.TRI_DATA_DTYPE)
    g = rtc.Geometry(pos, td, (4, 4, 4))
    o = orc.Geometry(pos, td, (4, 4, 4))
    assert np.array_equal(g.cells, o.cells) and g.tri_data.tobytes() == o.tri_data.tobytes()


def test_load_camera_matches_oracle(orc):
    spec = scenes.cornell(subdiv=1)
    a = spec.camera(height=1080)
    b = orc.load_camera(spec.cam_matrix, spec.yfov, spec.aspect, None, 1080)
    assert (a.w, a.h) == (b.w, b.h) == (1920, 1080)
    for f in ("origin", "lower_left_corner", "right", "up"):
        np.testing.assert_allclose(list(getattr(a, f)), list(getattr(b, f)), rtol=2e-6, atol=1e-4)
    with pytest.raises(ValueError):
        spec.camera()
    with pytest.raises(ValueError):
        spec.camera(width=10, height=10)      # camera has an aspect ratio (src/stage1.zig:326-329)
