"""The two-level walk of csrc/rtc_trav.cuh (an empty 4x4x4 block taken in one step, the other axes' in-block
crossings consumed lazily) against the cell-by-cell walk, ON THE CPU: the product's own __host__ __device__
traversal functions compiled for the host (tests/native/trav_check.cu). Same triangle-holding cells in the same
order, same iterator state in each of them, same crossing value leaving each of them, same end of the walk —
bit for bit — over camera rays, bounce rays, axis-parallel rays (+-inf lanes), exact diagonals (three-way ties),
rays that start inside the grid, random occupancies and random termination thresholds."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from zig_raytracing_contest_b200 import rtc, scenes

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def checker():
    so = os.path.join(HERE, "native", "libtrav_check.so")
    src = os.path.join(HERE, "native", "trav_check.cu")
    hdr = os.path.join(HERE, "..", "zig_raytracing_contest_b200", "csrc", "rtc_trav.cuh")
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.run(["nvcc", "-O2", "-std=c++17", "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-fmad=false",
                        "-Xcompiler", "-fPIC,-ffp-contract=off", "-o", so, src], check=True, capture_output=True)
    L = C.CDLL(so)
    L.trav_check.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                             C.POINTER(C.c_int), C.c_void_p, C.POINTER(C.c_uint64)]
    return L


def _iters(orc, grid, rays):
    """Grid.traceRay's iterator for every ray that meets the grid."""
    L = orc.lib()
    us, fs = [], []
    u, f = np.zeros(9, np.uint32), np.zeros(6, np.float32)
    for o, d in rays:
        o, d = np.ascontiguousarray(o, np.float32), np.ascontiguousarray(d, np.float32)
        if L.orc_grid_iter_init(C.byref(grid), o.ctypes.data_as(C.c_void_p), d.ctypes.data_as(C.c_void_p),
                                u.ctypes.data_as(C.c_void_p), f.ctypes.data_as(C.c_void_p)):
            us.append(u.copy()); fs.append(f.copy())
    return np.array(us, np.uint32), np.array(fs, np.float32)


def _check(checker, res, cells, us, fs, best_t):
    res = np.ascontiguousarray(res, np.uint32)
    cells = np.ascontiguousarray(cells, np.uint32)
    best_t = np.ascontiguousarray(best_t, np.float32)
    first, steps, events = C.c_int(-1), np.zeros(2, np.uint64), C.c_uint64(0)
    bad = checker.trav_check(res.ctypes.data, cells.ctypes.data, len(us), us.ctypes.data, fs.ctypes.data, best_t.ctypes.data,
                             C.byref(first), steps.ctypes.data, C.byref(events))
    return bad, first.value, steps, events.value


def _rays(orc, spec, geo, cam, n, seed):
    rng = np.random.RandomState(seed)
    desc, keep = rtc.make_scene_desc(geo, spec.materials)
    rays = []
    for _ in range(n):
        o, d = orc.camera_get_ray(cam, rng.uniform(0, cam.w), rng.uniform(0, cam.h))
        for _ in range(3):
            rays.append((o, d))
            idx, t, u, v = orc.trace_ray(desc, o, d)
            if idx < 0:
                break
            tp = geo.tri_pos[idx]
            nrm = np.cross(tp[1, :3], tp[2, :3]); nrm /= np.linalg.norm(nrm)
            r = rng.normal(size=3); r /= np.linalg.norm(r)
            nd = nrm + r
            o, d = (o + d * np.float32(t)).astype(np.float32), (nd / np.linalg.norm(nd)).astype(np.float32)
    return rays


@pytest.mark.parametrize("name,res", [("cornell", (32, 32, 32)), ("atrium", (64, 64, 64)), ("atrium", (128, 128, 128)),
                                      ("terrain", (128, 64, 128)), ("cornell", (30, 32, 32))])
def test_two_level_walk_equals_cell_walk(orc, checker, name, res):
    spec = {"cornell": lambda: scenes.cornell(), "atrium": lambda: scenes.atrium(detail=0.35),
            "terrain": lambda: scenes.terrain(60_000)}[name]()
    geo = rtc.Geometry(spec.positions, spec.tri_data, res)
    rays = _rays(orc, spec, geo, spec.camera(height=90), 700, 5)
    us, fs = _iters(orc, geo.grid, rays)
    assert len(us) > 500
    rng = np.random.RandomState(1)
    diag = np.linalg.norm(np.array(geo.grid.bbox_max) - np.array(geo.grid.bbox_min))
    for best in (np.full(len(us), np.inf), rng.uniform(0, diag, len(us)), fs[:, 3:].min(axis=1) + rng.uniform(0, 0.2 * diag, len(us))):
        bad, first, steps, events = _check(checker, res, geo.cells, us, fs, best)
        assert bad == 0, "ray %d differs" % first
        assert events > 0
    if all(r % 4 == 0 for r in res):
        bad, first, steps, events = _check(checker, res, geo.cells, us, fs, np.full(len(us), np.inf))
        assert steps[1] < steps[0]                # the point of it: fewer loop iterations ...
        if name == "atrium":
            assert steps[1] < 0.8 * steps[0]      # ... far fewer where there is open space
        print(name, res, "loop iterations: cell", steps[0], "two-level", steps[1])
    else:
        assert steps[1] == steps[0]               # grids that are not 4-aligned walk cell by cell


def test_two_level_walk_special_rays(orc, checker):
    """Axis-parallel rays (t_delta = +inf, t_next = +-inf), exact diagonals (every crossing a three-way tie),
    origins on cell and block faces, origins inside the grid, -0.0 direction components (t_next = -inf:
    those rays must stay in cell mode), and random sparse occupancies."""
    rng = np.random.RandomState(7)
    for res in [(8, 8, 8), (16, 12, 20), (4, 4, 4), (64, 64, 64)]:
        grid = orc.grid_init((0.0, 0.0, 0.0), tuple(float(r) for r in res), res)
        for fill in (0.0, 0.002, 0.02, 0.3):
            occ = rng.uniform(size=res[0] * res[1] * res[2]) < fill
            cells = np.zeros((occ.size, 2), np.uint32)
            cells[:, 1] = np.cumsum(occ); cells[:, 0] = cells[:, 1] - occ
            rays = []
            for _ in range(400):
                o = rng.uniform(-2, max(res) + 2, 3).astype(np.float32)
                d = rng.normal(size=3)
                kind = rng.randint(6)
                if kind == 0: d[rng.randint(3)] = 0.0
                if kind == 1: d[[rng.randint(3), rng.randint(3)]] = 0.0
                if kind == 2: d = np.sign(d) + (d == 0); o = np.round(o)          # exact diagonal from a lattice point
                if kind == 3: o = np.round(o * 0.25) * 4.0                          # origin on block faces
                if kind == 4: d[rng.randint(3)] = -0.0
                if kind == 5: o = rng.uniform(0.01, min(res) - 0.01, 3).astype(np.float32)   # inside the grid
                if not np.any(d): continue
                d = (d / np.linalg.norm(d)).astype(np.float32) if kind != 2 else (d / np.float32(np.sqrt(3.0))).astype(np.float32)
                rays.append((o, d))
            us, fs = _iters(orc, grid, rays)
            if len(us) == 0: continue
            for best in (np.full(len(us), np.inf), rng.uniform(0, 2.0 * max(res), len(us))):
                bad, first, steps, events = _check(checker, res, cells, us, fs, best)
                assert bad == 0, (res, fill, first, rays[first] if first < len(rays) else None)
