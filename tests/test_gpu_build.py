"""Parity of stage2 on the device (rtc_geometry_build_device / rtc_scene_create_from_soup)
against the CPU oracle's Geometry.build + bakeInto (src/stage2.zig:44-164): grid, every Cell{begin,end},
and the baked triangle arrays bit for bit — the order inside a cell is what Scene.traceRay's tie-break
depends on — then a render from the device-built scene against a render from the uploaded one."""
import numpy as np
import pytest

from zig_raytracing_contest_b200 import abi, rtc, scenes

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def _same_geometry(g, o):
    for k in ("bbox_min", "bbox_max", "cell_size"):
        assert np.array_equal(np.array(getattr(g.grid, k), np.float32), np.array(getattr(o.grid, k), np.float32)), k
    assert list(g.grid.resolution) == list(o.grid.resolution)
    assert np.array_equal(g.cells, o.cells)
    assert g.tri_pos.shape == o.tri_pos.shape
    assert np.array_equal(_bits(g.tri_pos), _bits(o.tri_pos))
    assert np.array_equal(g.tri_data["src_tri"], o.indices.astype(np.uint32))
    for k in ("normal", "texcoord"):
        assert np.array_equal(_bits(g.tri_data[k]), _bits(o.tri_data[k])), k
    assert np.array_equal(g.tri_data["material"], o.tri_data["material"])
    assert (g.num_src_tris, g.empty_cells, g.max_tris) == (o.num_src_tris, o.empty_cells, o.max_tris)
    if g.num_tris:
        assert g.min_tris == o.min_tris


@pytest.mark.parametrize("name,res", [("cornell", (32, 32, 32)), ("cornell", (7, 13, 5)), ("atrium", (64, 64, 64)),
                                      ("terrain", (128, 128, 128)), ("terrain", (40, 9, 33))])
def test_device_build_matches_oracle(orc, name, res):
    spec = {"cornell": lambda: scenes.cornell(), "atrium": lambda: scenes.atrium(detail=0.35),
            "terrain": lambda: scenes.terrain(120_000)}[name]()
    g = rtc.Geometry(spec.positions, spec.tri_data, res, device=0)
    o = orc.Geometry(spec.positions, spec.tri_data, res)
    _same_geometry(g, o)
    # and the host-thread build agrees as well (three producers, one result)
    h = rtc.Geometry(spec.positions, spec.tri_data, res)
    assert np.array_equal(h.cells, g.cells) and np.array_equal(_bits(h.tri_pos), _bits(g.tri_pos))


def test_device_build_edge_cases(orc):
    """Empty soup; one triangle in a 1x1x1 grid; triangles lying exactly on cell faces (axis-aligned, at
    multiples of the cell size); a zero-area triangle (normalize() of a zero edge = inf/NaN axes,
    src/linalg.zig:517-563 evaluated as written); one triangle spanning the whole grid (the deferred,
    one-CTA-per-triangle path)."""
    td = lambda n: np.zeros(n, rtc.TRI_DATA_DTYPE)
    cases = [
        (np.zeros((0, 3, 3), np.float32), (4, 4, 4)),
        (np.array([[[-1, -1, 0], [1, -1, 0], [0, 1, 0]]], np.float32), (1, 1, 1)),
        (np.array([[[0, 0, 0], [8, 0, 0], [0, 8, 0]], [[0, 0, 4], [8, 0, 4], [0, 8, 4]], [[0, 0, 8], [8, 8, 8], [0, 8, 8]],
                   [[2, 2, 2], [2, 6, 2], [2, 2, 6]]], np.float32), (8, 8, 8)),
        (np.array([[[0, 0, 0], [4, 4, 4], [1, 3, 2]], [[1, 1, 1], [1, 1, 1], [1, 1, 1]], [[2, 2, 2], [3, 3, 3], [2, 2, 2]]], np.float32), (6, 6, 6)),
        (np.array([[[0, 0, 0], [100, 3, 90], [5, 100, 100]], [[10, 10, 10], [11, 10, 10], [10, 11, 10]]], np.float32), (96, 80, 112)),
    ]
    for pos, res in cases:
        g = rtc.Geometry(pos, td(len(pos)), res, device=0)
        o = orc.Geometry(pos, td(len(pos)), res)
        _same_geometry(g, o)


def test_scene_from_soup_renders_identically(orc):
    spec = scenes.atrium(detail=0.35)
    res = (64, 64, 64)
    cam = spec.camera(height=108)
    geo = rtc.Geometry(spec.positions, spec.tri_data, res)
    a = rtc.Scene(geo, spec.materials, device=0)
    b = rtc.Scene.from_soup(spec.positions, spec.tri_data, res, spec.materials, device=0)
    bi = b.build_info
    assert (bi.num_tris, bi.num_src_tris, bi.empty_cells, bi.max_tris) == (geo.num_tris, geo.num_src_tris, geo.empty_cells, geo.max_tris)
    assert bi.device_ms > 0.0 and bi.total_ms >= bi.device_ms
    for prm in (rtc.make_params(mode=abi.RTC_MODE_PRIMARY), rtc.make_params(num_samples=6, max_bounce=4, seed=21)):
        ra = a.render(cam, prm, want_radiance=True, want_hits=True, want_stats=True)
        rb = b.render(cam, prm, want_radiance=True, want_hits=True, want_stats=True)
        assert np.array_equal(_bits(ra["radiance"]), _bits(rb["radiance"]))
        assert np.array_equal(ra["hit_id"], rb["hit_id"]) and np.array_equal(ra["rgb"], rb["rgb"])
        for k in ("rays", "cells_visited", "tris_tested", "hits_shaded", "paths"):
            assert ra["stats"][k] == rb["stats"][k], k
    # and against the oracle directly
    prm = rtc.make_params(num_samples=3, max_bounce=4, seed=4)
    rb = b.render(cam, prm, want_radiance=True)
    ro = orc.render(a.desc, cam, prm, want_radiance=True)
    assert np.array_equal(_bits(rb["radiance"]), _bits(ro["radiance"]))
    a.close(); b.close()


def test_from_soup_errors():
    spec = scenes.cornell()
    td = spec.tri_data.copy()
    td["material"][5] = 999
    with pytest.raises(rtc.RtcError) as e:
        rtc.Scene.from_soup(spec.positions, td, (8, 8, 8), spec.materials, device=0)
    assert e.value.code == abi.RTC_ERR_INVALID_ARGUMENT
    with pytest.raises(rtc.RtcError) as e:
        rtc.Scene.from_soup(spec.positions, spec.tri_data, (2048, 8, 8), spec.materials, device=0)
    assert e.value.code == abi.RTC_ERR_UNSUPPORTED
