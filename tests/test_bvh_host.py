"""RTC_ACCEL_BVH on the CPU: the product's BVH builder (csrc/rtc_bvh_build.cpp) and its __host__ __device__
traversal (csrc/rtc_bvh.cuh) compiled for the host (tests/native/bvh_check.cu), against a brute-force loop over all
triangles with the same triangle test (Triangle.rayIntersection, src/linalg.zig:696-722) and the same acceptance
rule. The BVH may only change WHICH triangles are tested, never the answer: ids and t must agree bit for bit on
camera rays, bounce rays from surfaces, axis-parallel rays, rays along shared edges and vertices, degenerate
triangles and tiny scenes."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from zig_raytracing_contest_b200 import rtc, scenes

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "..", "zig_raytracing_contest_b200", "csrc")


@pytest.fixture(scope="module")
def checker():
    so = os.path.join(HERE, "native", "libbvh_check.so")
    srcs = [os.path.join(HERE, "native", "bvh_check.cu"), os.path.join(CSRC, "rtc_bvh_build.cpp")]
    deps = srcs + [os.path.join(CSRC, f) for f in ("rtc_bvh.cuh", "rtc_bvh_host.h", "rtc_trav.cuh", "rtc_device.cuh")]
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(p) for p in deps):
        subprocess.run(["nvcc", "-O2", "-std=c++17", "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-fmad=false",
                        "-Xcompiler", "-fPIC,-ffp-contract=off", "-o", so] + srcs + ["-lpthread"], check=True, capture_output=True)
    L = C.CDLL(so)
    L.bvh_check.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_int] + [C.c_void_p] * 5
    return L


def tri_records(positions):
    """[N,3,3] world-space vertices -> [N,12] {v0 _ e1 _ e2 _} (Triangle.init, src/linalg.zig:688-694) in fp32."""
    p = np.ascontiguousarray(positions, np.float32).reshape(-1, 3, 3)
    out = np.zeros((len(p), 12), np.float32)
    out[:, 0:3] = p[:, 0]
    out[:, 4:7] = p[:, 1] - p[:, 0]
    out[:, 8:11] = p[:, 2] - p[:, 0]
    return out


def run(checker, tris, rays, threads=4):
    tris = np.ascontiguousarray(tris, np.float32)
    rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
    m = len(rays)
    tri, t = np.zeros(m, np.int32), np.zeros(m, np.float32)
    btri, bt = np.zeros(m, np.int32), np.zeros(m, np.float32)
    stats = np.zeros(8, np.uint64)
    bad = checker.bvh_check(tris.ctypes.data, len(tris), rays.ctypes.data, m, threads, tri.ctypes.data, t.ctypes.data,
                            btri.ctypes.data, bt.ctypes.data, stats.ctypes.data)
    return bad, tri, t, btri, bt, stats


def camera_rays(spec, height, n, seed):
    cam = spec.camera(height=height)
    rng = np.random.RandomState(seed)
    o = np.array(cam.origin[:], np.float32)
    llc, right, up = (np.array(v[:], np.float32) for v in (cam.lower_left_corner, cam.right, cam.up))
    xs, ys = rng.uniform(0, cam.w, n).astype(np.float32), rng.uniform(0, cam.h, n).astype(np.float32)
    d = llc[None] + right[None] * xs[:, None] + up[None] * ys[:, None]
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    return np.concatenate([np.broadcast_to(o, d.shape), d], axis=1)


def bounce_rays(tris, rays, tri, t, seed):
    """Diffuse-like rays leaving the hit points of `rays` (the incoherent case, origins ON surfaces)."""
    rng = np.random.RandomState(seed)
    hit = tri >= 0
    o = rays[hit, :3] + rays[hit, 3:] * t[hit, None]
    e1, e2 = tris[tri[hit], 4:7], tris[tri[hit], 8:11]
    nrm = np.cross(e1, e2)
    nrm /= np.maximum(np.linalg.norm(nrm, axis=1, keepdims=True), 1e-30)
    d = nrm + rng.normal(size=nrm.shape) * 0.8
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return np.concatenate([o, d], axis=1).astype(np.float32)


@pytest.mark.parametrize("name,kw,height", [("cornell", {"subdiv": 6}, 60), ("atrium", {"detail": 0.25}, 48),
                                            ("terrain", {"num_tris": 40000}, 48)])
def test_bvh_equals_brute_force(checker, name, kw, height):
    spec = scenes.SCENES[name](**kw)
    tris = tri_records(spec.positions)
    rays = camera_rays(spec, height, 1500, 5)
    bad, tri, t, btri, bt, stats = run(checker, tris, rays)
    assert bad == 0, (bad, np.nonzero((tri != btri) | (t.view(np.uint32) != bt.view(np.uint32)))[0][:5])
    assert (tri >= 0).sum() > 200
    rays2 = bounce_rays(tris, rays, tri, t, 9)
    bad2, tri2, t2, *_ = run(checker, tris, rays2)
    assert bad2 == 0
    assert stats[1] < 40 and stats[4] < 40                                  # depth and stack use far below kBvhStack
    # the point of the structure: far fewer triangle tests than the list is long
    assert stats[3] / len(rays) < 0.02 * len(tris) + 8


def test_bvh_edge_rays(checker):
    """Axis-parallel rays (zero direction components), rays through shared edges and vertices of a regular lattice
    (exact-t ties between neighbours -> smaller source id), rays starting on a surface, rays that miss everything."""
    m = scenes.Mesh()
    scenes.quad(m, (0, 0, 0), (0, 0, 1), (1, 0, 0), 8, 8, 0)       # floor lattice facing +y: vertices at k/8
    scenes.quad(m, (0, 1, 0), (1, 0, 0), (0, 0, 1), 8, 8, 0)       # ceiling facing -y
    pos, _ = m.finish()
    tris = tri_records(pos)
    rays = []
    for i in range(9):
        for j in range(9):
            rays.append([i / 8, 0.5, j / 8, 0, -1, 0])             # straight down onto every lattice vertex
            rays.append([i / 8, 0.5, j / 8, 0, 1, 0])              # straight up
            rays.append([i / 8 + 1 / 16, 0.5, j / 8, 0, -1, 0])    # onto edge midpoints
    rays += [[0.5, 0.5, 0.5, 1, 0, 0], [0.5, 0.5, 0.5, 0, 0, -1], [5, 5, 5, 0, 1, 0], [0.5, 0.0, 0.5, 0, 1, 0],
             [0.3, 0.5, 0.3, -0.0, -1, 0.0]]
    bad, tri, t, btri, bt, stats = run(checker, tris, np.array(rays, np.float32))
    assert bad == 0
    assert (tri >= 0).sum() >= 150


def test_bvh_degenerate_and_tiny(checker):
    rng = np.random.RandomState(3)
    rays = np.concatenate([rng.uniform(-1, 2, (300, 3)), rng.normal(size=(300, 3))], axis=1).astype(np.float32)
    rays[:, 3:] /= np.linalg.norm(rays[:, 3:], axis=1, keepdims=True)
    # one triangle; three triangles (single leaf); zero-area and repeated triangles; NaN / inf edges
    one = tri_records(np.array([[[0, 0, 0], [1, 0, 0], [0, 1, 0]]], np.float32))
    assert run(checker, one, rays)[0] == 0
    soup = rng.uniform(0, 1, (40, 3, 3)).astype(np.float32)
    soup[5] = soup[5, 0]                                            # a point
    soup[6, 2] = soup[6, 1]                                         # a segment
    soup[7] = soup[8]                                               # a duplicate (exact tie on every hit)
    t = tri_records(soup)
    t[9, 4] = np.inf; t[10, 8] = np.nan
    bad, tri, *_ = run(checker, t, rays)
    assert bad == 0 and (tri >= 0).sum() > 5
    assert run(checker, tri_records(soup[:3]), rays)[0] == 0
    # many identical centroids (median-split fallback)
    same = np.repeat(soup[:1], 50, axis=0)
    assert run(checker, tri_records(same), rays)[0] == 0
    # empty scene: every ray misses
    bad, tri, *_ = run(checker, np.zeros((0, 12), np.float32), rays[:10])
    assert bad == 0 and (tri == -1).all()


def test_bvh_random_soup_many_rays(checker):
    """Random intersecting triangles of mixed sizes (long thin ones span many boxes), incoherent rays from inside."""
    rng = np.random.RandomState(11)
    c = rng.uniform(0, 10, (3000, 1, 3))
    size = rng.choice([0.05, 0.3, 3.0], (3000, 1, 1), p=[0.6, 0.3, 0.1])
    soup = (c + rng.normal(size=(3000, 3, 3)) * size).astype(np.float32)
    tris = tri_records(soup)
    rays = np.concatenate([rng.uniform(0, 10, (3000, 3)), rng.normal(size=(3000, 3))], axis=1).astype(np.float32)
    rays[:, 3:] /= np.linalg.norm(rays[:, 3:], axis=1, keepdims=True)
    bad, tri, t, btri, bt, stats = run(checker, tris, rays, threads=1)
    assert bad == 0
    bad1, tri1, t1, *_ = run(checker, tris, rays, threads=8)                 # the tree may differ, the answers may not
    assert bad1 == 0 and np.array_equal(tri, tri1) and np.array_equal(t.view(np.uint32), t1.view(np.uint32))


def test_bvh_long_ranges_split_by_several_threads(checker):
    """Ranges of 2^17 triangles or more are binned and partitioned by several threads (stable, out of place): the answers
    equal brute force and the TREE — node count, depth, node visits and triangle tests of the same rays — is the same
    whatever the number of threads."""
    spec = scenes.terrain(300_000)
    tris = tri_records(spec.positions)
    assert len(tris) >= (1 << 17) * 2
    rays = camera_rays(spec, 40, 400, 7)
    out = {}
    for threads in (1, 3, 8):
        bad, tri, t, btri, bt, stats = run(checker, tris, rays, threads=threads)
        assert bad == 0
        out[threads] = (tri, t, tuple(int(x) for x in stats[:4]))
    assert (out[1][0] >= 0).sum() > 100
    for threads in (3, 8):
        assert np.array_equal(out[threads][0], out[1][0]) and np.array_equal(out[threads][1].view(np.uint32), out[1][1].view(np.uint32))
        assert out[threads][2] == out[1][2]
