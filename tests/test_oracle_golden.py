"""Pins the CPU oracle against every golden vector the reference's own tests hold for the
hot path: the 13 unit tests of /root/reference/src/linalg.zig. Expected values
are the literals of those tests; the tolerance is the one each test uses (1e-4) unless the
reference asserts equality."""
import math

import numpy as np
import pytest

INF = float("inf")


def norm(v):
    v = np.asarray(v, np.float32)
    l = np.float32(np.sqrt(np.float32(np.float32(v[0] * v[0] + v[1] * v[1]) + v[2] * v[2])))
    return v * (np.float32(1.0) / l)


def test_rgb_size_is_3():
    # linalg.zig:9-11 — the framebuffer element is 3 bytes; rtc_render's rgb_out contract.
    assert np.dtype(np.uint8).itemsize * 3 == 3


def test_cross_product(orc):
    # linalg.zig:231-236
    assert orc.cross([1, -8, 12], [4, 6, 3]).tolist() == [-96.0, 45.0, 38.0]


def test_vector_length(orc):
    # linalg.zig:238-241
    assert orc.length([1.5, 100.0, -21.1]) == pytest.approx(102.21281720019266, abs=1e-4)


def test_bbox_ray_intersection_1(orc):
    # linalg.zig:352-364
    hit, t = orc.bbox_ray_intersection([-1] * 3, [1] * 3, [0, 0, 5], [0, 0, -1])
    assert hit and t == pytest.approx(4.0, abs=1e-4)


def test_bbox_ray_intersection_2(orc):
    # linalg.zig:366-378
    hit, t = orc.bbox_ray_intersection([1] * 3, [2] * 3, [0, 0, 0], orc.normalize([1, 1, 1]))
    assert hit and t == pytest.approx(math.sqrt(3.0), abs=1e-4)


def test_bbox_ray_intersection_3_origin_inside_gives_negative_t(orc):
    # linalg.zig:380-392
    hit, t = orc.bbox_ray_intersection([-1] * 3, [3] * 3, [0, 0, 0], orc.normalize([1, 1, 0]))
    assert hit and t == pytest.approx(-math.sqrt(2.0), abs=1e-4)


def test_bbox_ray_intersection_4_miss(orc):
    # linalg.zig:394-405
    hit, _ = orc.bbox_ray_intersection([-1] * 3, [3] * 3, [5, 5, 5], orc.normalize([1, 1, 0]))
    assert not hit


def test_decrement_via_add():
    # linalg.zig:565-569 — the DDA's negative step is a wrapping add of 0xFFFFFFFF
    x = np.uint32(5)
    with np.errstate(over="ignore"):
        x = np.uint32((int(x) + 0xFFFFFFFF) & 0xFFFFFFFF)
    assert x == 4


def test_grid_get_cell_bbox(orc):
    # linalg.zig:571-581
    g = orc.grid_init([0, 0, 0], [5, 5, 5], [5, 5, 5])
    lo, hi = orc.grid_cell_bbox(g, 0, 1, 4)
    assert lo.tolist() == [0, 1, 4] and hi.tolist() == [1, 2, 5]


def _check_walk(orc, orig, direction, cells, ts):
    g = orc.grid_init([0, 0, 0], [5, 5, 5], [5, 5, 5])
    got = orc.grid_walk(g, orig, direction)
    assert got is not None
    got_cells, got_ts = got
    assert got_cells.tolist() == [list(c) for c in cells]
    assert len(got_ts) == len(ts)
    for a, b in zip(got_ts.tolist(), ts):
        if b == INF:
            assert a == INF
        else:
            assert a == pytest.approx(b, abs=1e-4)


def test_grid_trace_ray_1(orc):
    # linalg.zig:583-607
    _check_walk(orc, [0.5, 0.5, 0.5], orc.normalize([2, 1, 0]),
                [(0, 0, 0), (1, 0, 0), (1, 1, 0), (2, 1, 0), (3, 1, 0), (3, 2, 0), (4, 2, 0)],
                [0.559017002, 1.11803400, 1.67705106, 2.79508495, 3.35410213, 3.91311883, INF])


def test_grid_trace_ray_2(orc):
    # linalg.zig:609-629 (axis-aligned, -y: +-inf lanes)
    _check_walk(orc, [0.5, 10.0, 0.5], [0, -1, 0],
                [(0, 4, 0), (0, 3, 0), (0, 2, 0), (0, 1, 0), (0, 0, 0)], [6, 7, 8, 9, INF])


def test_grid_trace_ray_3(orc):
    # linalg.zig:631-651
    _check_walk(orc, [0.5, -5.0, 0.5], [0, 1, 0],
                [(0, 0, 0), (0, 1, 0), (0, 2, 0), (0, 3, 0), (0, 4, 0)], [6, 7, 8, 9, INF])


def test_grid_trace_ray_4_exact_diagonal_tie_order(orc):
    # linalg.zig:653-681 — ties step y before x
    _check_walk(orc, [0.5, 0.5, 0.5], orc.normalize([1, 1, 0]),
                [(0, 0, 0), (0, 1, 0), (1, 1, 0), (1, 2, 0), (2, 2, 0), (2, 3, 0), (3, 3, 0), (3, 4, 0), (4, 4, 0)],
                [0.707106769, 0.707106769, 2.12132024, 2.12132024, 3.53553390, 3.53553390,
                 4.94974756, 4.94974756, INF])


# ---- published known-answer vectors for the third-party pieces the oracle restates ----------
def test_philox4x32_10_known_answers(orc):
    # Random123 kat_vectors (philox4x32 10 rounds): zero, all-ones, and the pi-digits vector.
    assert [hex(x) for x in orc.philox4x32_10([0, 0, 0, 0], [0, 0])] == \
        ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    assert [hex(x) for x in orc.philox4x32_10([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2)] == \
        ["0x408f276d", "0x41c83b0e", "0xa20bc7c6", "0x6d5451fd"]
    assert [hex(x) for x in orc.philox4x32_10([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344],
                                              [0xa4093822, 0x299f31d0])] == \
        ["0xd16cfe09", "0x94fdcceb", "0x5001e420", "0x24126ea1"]


def test_unit_vector_has_no_pole_atom(orc):
    """The sphere map takes u1 at its bin centre: z = +-1 (where normal + unit vector cancels exactly on an
    axis-aligned wall and normalize() yields NaN) is never produced; vectors are unit length and isotropic."""
    top = 1.0 - 2.0 ** -24
    for u1, z in ((0.0, top), (top, -top), (0.5, -2.0 ** -24)):
        v = orc.unit_vector(u1, 0.3)
        assert v[2] == np.float32(z) and abs(v[2]) < 1.0
        assert abs(float(np.linalg.norm(v.astype(np.float64))) - 1.0) < 1e-6
        assert v[0] != 0.0 or v[1] != 0.0
    rng = np.random.RandomState(3)
    k = rng.randint(0, 1 << 24, size=(20000, 2))
    vs = np.array([orc.unit_vector(a / 16777216.0, b / 16777216.0) for a, b in k], np.float64)
    assert np.abs(np.linalg.norm(vs, axis=1) - 1.0).max() < 1e-6
    assert np.abs(vs.mean(axis=0)).max() < 0.02 and np.abs((vs ** 2).mean(axis=0) - 1.0 / 3.0).max() < 0.01
