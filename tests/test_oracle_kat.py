"""Authored known-answer tests for the half of the oracle the reference's own 13 unit tests do not pin
(SURVEY §8(c)): Triangle.rayIntersection (src/linalg.zig:696-722), Triangle.Data.interpolate (src/stage3.zig:53-71),
Texture.sample (src/stage3.zig:82-123), Vec3.toRGB (src/linalg.zig:150-159), getEnvColor (src/stage3.zig:144-150),
Camera.getRay (src/stage3.zig:27-35), and the decision structure of traceRayRecursive / traceRay
(src/stage3.zig:152-220). Every expected value is derived here, independently of the oracle, from the cited lines:
 * inputs made of small dyadic rationals, for which every fp32 operation of the reference is exact — the expected
   t, u, v / texel / colour is then an exact rational (fractions.Fraction) and the oracle must return exactly it;
 * general inputs evaluated in float64 — the DECISION (hit / miss / which branch) must be equal, values within a
   stated number of fp32 ulps.
CPU only (the CUDA path is compared with the oracle in tests/test_gpu_*.py)."""
from fractions import Fraction as Fr

import numpy as np
import pytest

from zig_raytracing_contest_b200 import abi, rtc

F = np.float32


def f32(x):
    return float(np.float32(x))


def ulps(a, b):
    a, b = np.float32(a), np.float32(b)
    return abs(int(a.view(np.int32)) - int(b.view(np.int32)))


# ---- Triangle.rayIntersection ---------------------------------------------------------------------------------
def mt_exact(v0, v1, v2, o, d):
    """Moeller-Trumbore as the reference writes it (linalg.zig:696-722) over exact rationals.
    Returns None (det < 1e-8 or outside) or (t, u, v) as Fractions."""
    R = lambda a: [Fr(x) for x in a]
    v0, v1, v2, o, d = R(v0), R(v1), R(v2), R(o), R(d)
    sub = lambda a, b: [a[i] - b[i] for i in range(3)]
    dot = lambda a, b: a[0] * b[0] + a[1] * b[1] + a[2] * b[2]
    cross = lambda a, b: [a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]]
    e1, e2 = sub(v1, v0), sub(v2, v0)
    p = cross(d, e2)
    det = dot(e1, p)
    if det < Fr(f32(0.00000001)):                 # epsilon is an f32 literal in the reference's comparison
        return None
    tv = sub(o, v0)
    u = dot(tv, p) / det
    if u < 0 or u > 1:
        return None
    q = cross(tv, e1)
    v = dot(d, q) / det
    if v < 0 or u + v > 1:
        return None
    return dot(e2, q) / det, u, v


UNIT = ([0, 0, 0], [1, 0, 0], [0, 1, 0])           # e1 x e2 = +z: visible from z > 0

EXACT_TRI_CASES = [
    # (name, triangle, origin, direction, expected (t,u,v) or None) — all dyadic, all fp32 operations exact
    ("interior, straight down", UNIT, [0.25, 0.25, 1], [0, 0, -1], (1, 0.25, 0.25)),
    ("back face is culled (det = -1)", UNIT, [0.25, 0.25, -1], [0, 0, 1], None),
    ("ray parallel to the plane (det = 0)", UNIT, [0.25, 0.25, 1], [1, 0, 0], None),
    ("edge u = 0 is inside", UNIT, [0, 0.5, 1], [0, 0, -1], (1, 0, 0.5)),
    ("edge v = 0 is inside", UNIT, [0.5, 0, 1], [0, 0, -1], (1, 0.5, 0)),
    ("vertex v1: u = 1, v = 0 is inside", UNIT, [1, 0, 1], [0, 0, -1], (1, 1, 0)),
    ("vertex v2: u = 0, v = 1 is inside", UNIT, [0, 1, 1], [0, 0, -1], (1, 0, 1)),
    ("vertex v0", UNIT, [0, 0, 1], [0, 0, -1], (1, 0, 0)),
    ("hypotenuse u + v = 1 is inside", UNIT, [0.5, 0.5, 1], [0, 0, -1], (1, 0.5, 0.5)),
    ("just beyond the hypotenuse", UNIT, [0.5, 0.5 + 2.0 ** -10, 1], [0, 0, -1], None),
    ("u < 0 by 2^-10", UNIT, [-2.0 ** -10, 0.5, 1], [0, 0, -1], None),
    ("v < 0 by 2^-10", UNIT, [0.5, -2.0 ** -10, 1], [0, 0, -1], None),
    ("u > 1", UNIT, [1.5, 0, 1], [0, 0, -1], None),
    ("plane behind the origin: t = -1 is RETURNED (no t test here, traceRay does it)", UNIT, [0.25, 0.25, -1], [0, 0, -1], (-1, 0.25, 0.25)),
    ("origin on the plane: t = 0", UNIT, [0.25, 0.5, 0], [0, 0, -1], (0, 0.25, 0.5)),
    ("unnormalised direction halves t", UNIT, [0.25, 0.25, 1], [0, 0, -2], (0.5, 0.25, 0.25)),
    ("oblique dyadic ray", UNIT, [0.5, 0.25, 2], [-0.25, 0, -2], (1, 0.25, 0.25)),
    ("translated, scaled triangle", ([4, 2, -8], [8, 2, -8], [4, 4, -8]), [5, 2.5, 0], [0, 0, -4], (2, 0.25, 0.25)),
    ("triangle in the x = 2 plane facing +x", ([2, 0, 0], [2, 1, 0], [2, 0, 1]), [4, 0.5, 0.25], [-1, 0, 0], (2, 0.5, 0.25)),
    ("absolute epsilon: a front-facing 2^-14 triangle has det = 2^-28 < 1e-8 and is MISSED",
     ([0, 0, 0], [2.0 ** -14, 0, 0], [0, 2.0 ** -14, 0]), [2.0 ** -16, 2.0 ** -16, 1], [0, 0, -1], None),
    ("... while 2^-13 (det = 2^-26 = 1.49e-8) is hit",
     ([0, 0, 0], [2.0 ** -13, 0, 0], [0, 2.0 ** -13, 0]), [2.0 ** -15, 2.0 ** -15, 1], [0, 0, -1], (1, 0.25, 0.25)),
]


@pytest.mark.parametrize("name,tri,o,d,want", EXACT_TRI_CASES, ids=[c[0] for c in EXACT_TRI_CASES])
def test_triangle_exact_cases(orc, name, tri, o, d, want):
    v0, v1, v2 = (np.array(v, F) for v in tri)
    # the independent derivation agrees with the table (the table is hand-derived, the function re-derives it)
    ex = mt_exact(tri[0], tri[1], tri[2], o, d)
    assert (ex is None) == (want is None)
    if want is not None:
        assert tuple(Fr(x) for x in want) == ex
    got = orc.triangle_ray_intersection(v0, v1 - v0, v2 - v0, np.array(o, F), np.array(d, F))
    if want is None:
        assert got is None
    else:
        assert got is not None
        assert tuple(np.float32(g) for g in got) == tuple(np.float32(w) for w in want)     # exactly, not approximately


def test_triangle_general_cases_against_float64(orc):
    """300 random triangles and rays: decision equal to the exact-rational evaluation whenever that evaluation is
    not within 1e-5 of a threshold; t, u, v within 16 ulp of it (fp32 evaluation of ~20 operations)."""
    rng = np.random.RandomState(42)
    decided = hits = 0
    for _ in range(300):
        tri = rng.uniform(-2, 2, (3, 3)).astype(F)
        o = rng.uniform(-3, 3, 3).astype(F)
        c = tri.mean(axis=0) + rng.normal(size=3) * 0.7
        d = (c - o).astype(F)
        d = (d / np.linalg.norm(d)).astype(F)
        e1, e2 = (tri[1] - tri[0]).astype(F), (tri[2] - tri[0]).astype(F)
        # exact evaluation ON THE SAME fp32 INPUTS (v0, e1, e2 as the reference stores them)
        ex = mt_exact(tri[0].tolist(), (tri[0].astype(np.float64) + e1).tolist(), (tri[0].astype(np.float64) + e2).tolist(), o.tolist(), d.tolist())
        got = orc.triangle_ray_intersection(tri[0], e1, e2, o, d)
        # margin of the exact evaluation to the nearest threshold
        R = lambda a: [Fr(float(x)) for x in a]
        dotf = lambda a, b: a[0] * b[0] + a[1] * b[1] + a[2] * b[2]
        crossf = lambda a, b: [a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]]
        det = dotf(R(e1), crossf(R(d), R(e2)))
        margin = abs(float(det) - 1e-8)
        if ex is not None or float(det) >= 1e-8:
            tv = [Fr(float(o[i])) - Fr(float(tri[0][i])) for i in range(3)]
            u = dotf(tv, crossf(R(d), R(e2))) / det
            v = dotf(R(d), crossf(tv, R(e1))) / det
            margin = min(margin, abs(float(u)), abs(float(u) - 1), abs(float(v)), abs(float(u + v) - 1))
        if margin < 1e-5:
            continue
        decided += 1
        assert (got is None) == (ex is None)
        if ex is not None:
            hits += 1
            for g, w in zip(got, ex):
                assert abs(g - float(w)) <= 16 * 1.2e-7 * max(1.0, abs(float(w)))
    assert decided > 250 and hits > 40


def test_trace_ray_rejects_t_not_positive_and_keeps_first_on_ties(orc):
    """Scene.traceRay (stage3.zig:168-177): `nearest.t > hit.t and hit.t > 0` — t = 0 and t < 0 are no hits; of two
    coincident triangles the first tested (lower baked index) wins; the nearer of two stacked triangles wins."""
    M = rtc.MaterialData
    pos = np.array([[[0, 0, 0], [1, 0, 0], [0, 1, 0]],            # z = 0
                    [[0, 0, 0], [1, 0, 0], [0, 1, 0]],            # coincident copy
                    [[0, 0, 0.5], [1, 0, 0.5], [0, 1, 0.5]]], F)  # nearer to a viewer at z = 1
    td = np.zeros(3, rtc.TRI_DATA_DTYPE)
    geo = orc.Geometry(pos, td, (2, 2, 2))
    desc, keep = rtc.make_scene_desc(geo, [M()])
    src = lambda i: -1 if i < 0 else int(geo.tri_data["src_tri"][i])
    i, t, u, v = orc.trace_ray(desc, [0.25, 0.25, 1], [0, 0, -1])
    assert (src(i), t, u, v) == (2, 0.5, 0.25, 0.25)
    i, t, u, v = orc.trace_ray(desc, [0.25, 0.25, 0.25], [0, 0, -1])       # below the z = 0.5 triangle
    assert (src(i), t) == (0, 0.25)                                         # exact tie 0 / 1: first tested
    i, t, u, v = orc.trace_ray(desc, [0.25, 0.25, 0.0], [0, 0, -1])        # ON the z = 0 plane: t = 0 is not a hit
    assert i < 0 and t == np.inf
    i, t, u, v = orc.trace_ray(desc, [0.25, 0.25, -1], [0, 0, -1])         # everything behind: t < 0
    assert i < 0
    ib, tb, *_ = orc.trace_ray(desc, [0.25, 0.25, 1], [0, 0, -1], brute=True)
    assert (src(ib), tb) == (2, 0.5)


# ---- Triangle.Data.interpolate ----------------------------------------------------------------------------------
def test_interpolate_exact_and_association(orc):
    rec = np.zeros(1, rtc.TRI_DATA_DTYPE)
    rec["normal"][0] = [[1, 0, 0], [0, 1, 0], [0, 0, 1]]
    rec["texcoord"][0] = [[0, 0], [1, 0], [0, 1]]
    n, t = orc.interpolate(rec, 0.25, 0.5)
    assert n.tolist() == [0.25, 0.25, 0.5] and t.tolist() == [0.25, 0.5]       # w = 1 - u - v = 0.25
    n, t = orc.interpolate(rec, 0.0, 0.0)
    assert n.tolist() == [1, 0, 0] and t.tolist() == [0, 0]
    n, t = orc.interpolate(rec, 1.0, 0.0)
    assert n.tolist() == [0, 1, 0] and t.tolist() == [1, 0]
    # the result is NOT renormalised (stage3.zig:203): |n| = sqrt(0.375) here
    n, _ = orc.interpolate(rec, 0.25, 0.5)
    assert abs(float(np.dot(n, n)) - 0.375) < 1e-7
    # association (v0*w + v1*u) + v2*v: with w = 0.5, u = v = 0.25 the products are 1e8, -1e8, 1 — left to right
    # gives exactly 1, any other order gives 0 (1 is below half an ulp of 1e8)
    rec["normal"][0] = [[2e8, 2e8, 0], [-4e8, -4e8, 0], [4, 4, 0]]
    rec["texcoord"][0] = [[2e8, 2e8], [-4e8, -4e8], [4, 4]]
    n, t = orc.interpolate(rec, 0.25, 0.25)
    assert n.tolist() == [1.0, 1.0, 0.0] and t.tolist() == [1.0, 1.0]
    # 1 - u - v is (1 - u) - v: u = 2^-24 vanishes against 1 only if subtracted first from 1
    rec["normal"][0] = [[1, 0, 0], [0, 0, 0], [0, 0, 0]]
    n, _ = orc.interpolate(rec, 2.0 ** -25, 1.0)
    assert n[0] == np.float32(np.float32(np.float32(1.0) - np.float32(2.0 ** -25)) - np.float32(1.0))


# ---- Texture.sample ----------------------------------------------------------------------------------------------
def tex(vals, clamp=False):
    return rtc.TextureData(np.array(vals, F), clamp, clamp)


TEX22 = [[0, 1], [2, 3]]        # row 0 (v = 0): 0 1 ; row 1: 2 3

TEX_CASES = [
    # (name, clamp, u, v, expected) — frac() is taken of the UNSCALED coordinate (stage3.zig:118-121, the
    # reference's quirk: a bilinear filter would use frac(w*u)); lerp(a,b,t) = fma(b-a, t, a)
    ("texel centre by the quirk: weights 0.25", False, 0.25, 0.25, 0.75),      # r1 = .25, r2 = 2.25 -> .75 (a filter: 1.5)
    ("u = v = 0: first texel", False, 0.0, 0.0, 0.0),
    ("REPEAT: u + 1 samples the same texels", False, 1.25, 0.25, 0.75),
    ("REPEAT: integer u wraps to column 0, weight 0", False, 1.0, 0.0, 0.0),
    ("REPEAT: x2 wraps around the right edge", False, 0.75, 0.0, 0.25),        # ui = 1, x2 = 2 mod 2 = 0: lerp(1, 0, .75)
    ("negative u: floored mod, |frac|", False, -0.25, 0.25, 1.25),              # x1 = 1, x2 = 0: r1 = .75, r2 = 2.75
    ("negative v", False, 0.25, -0.25, 1.75),                                   # y1 = 1, y2 = 0: r1 = 2.25, r2 = .25 -> fma(-2,.25,2.25)
    ("CLAMP_TO_EDGE right: both taps are column 1", True, 1.25, 0.25, 1.5),    # r1 = 1, r2 = 3
    ("CLAMP_TO_EDGE left", True, -0.75, 0.25, 0.5),                             # both taps column 0: r1 = 0, r2 = 2 (frac(v) = .25)
    ("CLAMP_TO_EDGE top", True, 0.25, 2.5, 2.25),                               # both rows 1: r1 = r2 = lerp(2, 3, .25)
]


@pytest.mark.parametrize("name,clamp,u,v,want", TEX_CASES, ids=[c[0] for c in TEX_CASES])
def test_texture_sample_f32(orc, name, clamp, u, v, want):
    t = tex(TEX22, clamp)
    got = orc.texture_sample(t.as_struct(), u, v)
    assert got.shape == (1,) and got[0] == np.float32(want)


def test_texture_sample_vec3_and_dummy(orc):
    t = rtc.TextureData(np.array([[[0, 10, 100], [1, 11, 101]], [[2, 12, 102], [3, 13, 103]]], F))
    assert orc.texture_sample(t.as_struct(), 0.25, 0.25).tolist() == [0.75, 10.75, 100.75]
    # dummy 1x1 textures (stage1.zig:411-425): every uv, any wrap, returns the texel
    d = rtc.constant_texture([0.5, 0.25, 0.125])
    for u, v in [(0, 0), (0.3, 0.9), (-7.25, 123.5), (1e6, -1e6)]:
        assert orc.texture_sample(d.as_struct(), u, v).tolist() == [0.5, 0.25, 0.125]
    # non-square: w = 4, h = 1; u = 0.5 -> ui = 2, x2 = 3; frac(0.5) = 0.5
    t = tex([[0, 1, 4, 8]])
    assert orc.texture_sample(t.as_struct(), 0.5, 0.0)[0] == 6.0
    # w*u rounds in fp32 before the floor: u = 0.3 -> f32(4 * f32(0.3)) = 1.2 -> ui = 1
    assert orc.texture_sample(t.as_struct(), 0.3, 0.0)[0] == np.float32(np.float32(3.0) * np.float32(0.3) + np.float32(1.0))


# ---- Vec3.toRGB ---------------------------------------------------------------------------------------------------
def test_to_rgb_bins(orc):
    assert orc.to_rgb([0.0, 1.0, 4.0]).tolist() == [0, 255, 255]              # pow(0) = 0; upper clamp 0.999999 * 256 = 255.99974
    assert orc.to_rgb([1e-10, 0.999999, 1e30]).tolist() == [0, 255, 255]      # tiny: 256 * 1e-10^(1/2.2) = 0.007
    # mid-bin inputs: c = ((k + 0.5) / 256)^2.2 lands in bin k with half a bin of margin on either side
    for k in (0, 1, 2, 17, 64, 127, 128, 200, 254, 255):
        c = ((k + 0.5) / 256.0) ** 2.2
        assert orc.to_rgb([c, c, c]).tolist() == [k, k, k], k
    # bin edges: within 1 LSB whatever pow's last ulp is
    for k in (1, 64, 128, 255):
        c = (k / 256.0) ** 2.2
        assert abs(int(orc.to_rgb([c, c, c])[0]) - k) <= 1
    # `clamp` is an upper clamp only (linalg.zig:58-60) and @min returns the non-NaN operand (LLVM minnum, the
    # oracle's stated definition): pow(negative, 1/2.2) = NaN -> min(NaN, 0.999999) = 0.999999 -> 255
    assert orc.to_rgb([-1.0, np.nan, 0.25])[:2].tolist() == [255, 255]
    # monotone over the whole range
    xs = np.linspace(0, 1.2, 2000)
    q = np.array([orc.to_rgb([x, x, x])[0] for x in xs], int)
    assert (np.diff(q) >= 0).all() and q[0] == 0 and q[-1] == 255


# ---- getEnvColor, Camera.getRay --------------------------------------------------------------------------------
def test_env_color(orc):
    assert orc.env_color([0, 1, 0]).tolist() == [0.5, f32(0.7), 1.0]                       # t = 1: pure (0.5, 0.7, 1)
    assert orc.env_color([0, -1, 0]).tolist() == [1.0, 1.0, 1.0]                           # t = 0: white
    h = orc.env_color([1, 0, 0])                                                           # t = 0.5
    assert h.tolist() == [0.75, f32(f32(0.5) + f32(0.7) * f32(0.5)), 1.0]
    # unnormalised directions are used as they are: dir.y = 3 -> t = 2 -> ones*(-1) + c*2
    assert orc.env_color([0, 3, 0]).tolist() == [0.0, f32(f32(-1.0) + f32(f32(0.7) * f32(2.0))), 1.0]


def test_camera_get_ray(orc):
    cam = abi.Camera()
    cam.w, cam.h = 4, 2
    cam.origin[:] = [1, 2, 3]
    cam.lower_left_corner[:] = [-2, 1, -4]          # llc + right*x + up*y, then normalize = scale(1/length)
    cam.right[:] = [1, 0, 0]
    cam.up[:] = [0, -1, 0]                          # up points DOWN the image (stage1.zig:354-355)
    o, d = orc.camera_get_ray(cam, 2.0, 1.0)        # (0, 0, -4) -> (0, 0, -1)
    assert o.tolist() == [1, 2, 3] and d.tolist() == [0, 0, -1]
    o, d = orc.camera_get_ray(cam, 5.0, 1.0)        # (3, 0, -4): length 5 -> scale by f32(1/5)
    inv = np.float32(1.0) / np.float32(5.0)
    assert d.tolist() == [float(np.float32(3.0) * inv), 0.0, float(np.float32(-4.0) * inv)]
    o, d = orc.camera_get_ray(cam, 2.0, -2.0)       # y above the image: (0, 3, -4)
    assert d.tolist() == [0.0, float(np.float32(3.0) * inv), float(np.float32(-4.0) * inv)]


# ---- traceRayRecursive's decision structure through Scene.render -------------------------------------------
def _render(orc, pos, mats, mat_of, cam, spp, bounce, seed=5, mode=abi.RTC_MODE_PATH):
    td = np.zeros(len(pos), rtc.TRI_DATA_DTYPE)
    td["material"] = mat_of
    td["normal"][:] = (0, 0, 1)
    geo = orc.Geometry(np.array(pos, F), td, (2, 2, 2))
    desc, keep = rtc.make_scene_desc(geo, mats)
    prm = rtc.make_params(num_samples=spp, max_bounce=bounce, seed=seed, mode=mode)
    return orc.render(desc, cam, prm, num_threads=1, want_radiance=True)["radiance"]


def _cam():
    cam = abi.Camera()
    cam.w, cam.h = 2, 2
    cam.origin[:] = [0, 0, 4]
    cam.lower_left_corner[:] = [-0.5, 0.5, -8]      # a narrow pencil of rays around -z, all inside the big triangles
    cam.right[:] = [0.5, 0, 0]
    cam.up[:] = [0, -0.5, 0]
    return cam


BIG = [[-40, -40, 0], [40, -40, 0], [0, 60, 0]]                    # faces +z (towards the camera)


def test_recursion_depth_zero_miss_and_emissive(orc):
    M = rtc.MaterialData
    cam = _cam()
    glow = M(base_color=(0, 0, 0), emissive=(1, 2, 3))
    # depth 0 -> zeroes (stage3.zig:189-191)
    assert (_render(orc, [BIG], [glow], [0], cam, 4, 0) == 0).all()
    # albedo 0: emissive + 0 * recurse = emissive exactly, for every sample (4 spp: the 1/S scale is exact)
    r = _render(orc, [BIG], [glow], [0], cam, 4, 3)
    assert (r == np.array([1, 2, 3], F)).all()
    # a miss returns getEnvColor(ray.dir) (stage3.zig:194-196): with the triangle facing away every ray misses; the
    # pixel is the mean of 4 sky colours, each between white (down) and (0.5, 0.7, 1) (up), blue channel exactly 1
    away = [BIG[0], BIG[2], BIG[1]]
    r = _render(orc, [away], [glow], [0], cam, 4, 3)
    assert (r[..., 2] == 1.0).all() and (r[..., 0] > 0.7).all() and (r[..., 0] < 0.8).all()


def test_alpha_pass_through_consumes_depth_without_attenuation(orc):
    """stage3.zig:207-213: `if random.float(f32) > transparency` -> the ray goes on in the SAME direction from
    at(t + eps) with depth - 1 and NO albedo factor. transparency = 0 always passes (U > 0 but for U = 0)."""
    M = rtc.MaterialData
    cam = _cam()
    ghost = M(base_color=(0.5, 0.5, 0.5), emissive=(9, 9, 9), transparency=0.0)      # its emissive is never added
    glow = M(base_color=(0, 0, 0), emissive=(1, 2, 3))
    front = [[-40, -40, 1], [40, -40, 1], [0, 60, 1]]
    # ghost passes, glow is shaded -> exactly its emissive, unattenuated. The continued ray starts at
    # at(t + floatEps): in fp32 t + 1.19e-7 == t for t ~ 3, so the new origin lies ON the ghost (to rounding) and
    # may hit it again at a tiny t > 0 — each such re-hit passes again and consumes a level (the reference's
    # behaviour, restated as it is). With 2 levels some samples therefore end at depth 0; 6 levels suffice here.
    r2 = _render(orc, [front, BIG], [ghost, glow], [0, 1], cam, 4, 2)
    assert np.isin(r2 * 4, np.array([0, 1, 2, 3, 4], F)[:, None] * np.array([1, 2, 3], F)).all()    # k of 4 samples reached glow
    r = _render(orc, [front, BIG], [ghost, glow], [0, 1], cam, 4, 6)
    assert (r == np.array([1, 2, 3], F)).all()
    # depth 1: the pass-through used the only level -> zeroes
    r = _render(orc, [front, BIG], [ghost, glow], [0, 1], cam, 4, 1)
    assert (r == 0).all()
    # transparency = 1 never passes (U in [0,1) is never > 1): the front triangle is shaded
    solid = M(base_color=(0, 0, 0), emissive=(4, 5, 6), transparency=1.0)
    r = _render(orc, [front, BIG], [solid, glow], [0, 1], cam, 4, 2)
    assert (r == np.array([4, 5, 6], F)).all()


def test_scatter_returns_emissive_plus_albedo_times_recursion(orc):
    """stage3.zig:214-219 inside a closed emissive room: whatever direction the scattered ray takes it ends on a wall
    with albedo 0 and emissive E, so the pixel is exactly e + a * E."""
    from zig_raytracing_contest_b200 import scenes
    M = rtc.MaterialData
    m = scenes.Mesh()
    scenes.box(m, (-64, -64, -64), (64, 64, 64), 1, 1, inward=True)           # room: material 1
    pos, td_room = m.finish()
    pos = np.concatenate([np.array([BIG], F), pos])
    mat_of = np.concatenate([[0], td_room["material"]])
    card = M(base_color=(0.5, 0.25, 1.0), emissive=(0.125, 0.0, 2.0))
    wall = M(base_color=(0, 0, 0), emissive=(2, 4, 8))
    r = _render(orc, pos, [card, wall], mat_of, _cam(), 4, 2)
    assert (r == np.array([0.125 + 0.5 * 2, 0.0 + 0.25 * 4, 2.0 + 1.0 * 8], F)).all()
    # one level only: the scattered ray is not traced (depth - 1 == 0 -> zeroes): e + a * 0
    r = _render(orc, pos, [card, wall], mat_of, _cam(), 4, 1)
    assert (r == np.array([0.125, 0.0, 2.0], F)).all()
    # the deterministic slice (RTC_MODE_PRIMARY = depth 1 through the pixel centre) agrees
    r = _render(orc, pos, [card, wall], mat_of, _cam(), 1, 1, mode=abi.RTC_MODE_PRIMARY)
    assert (r == np.array([0.125, 0.0, 2.0], F)).all()
