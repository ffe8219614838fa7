"""RTC_ACCEL_BVH (csrc/rtc_bvh.cuh, rtc_bvh_build.cpp, k_trace_bvh) against the CPU oracle, through the C ABI, by the
relaxed metric SURVEY §8(c) gives a structure that is not the reference's grid: hit ids equal to the oracle's
EXCEPT ENUMERATED near-ties, t,u,v bit-exact on equal ids (the triangle test is the same code with the same inputs).
Every differing pixel is classified, none may stay unexplained:
  near-tie   : the two winners' t differ by <= 4 ulp (exact-t ties go to the smaller source id here, to the first
               visited cell in the reference);
  grid-edge  : the BVH's answer is the brute-force answer over ALL triangles (orc_trace_ray_brute) — the grid's
               binning / crossing arithmetic lost the hit, the BVH did not.
The stochastic mode shares the counter RNG with the oracle, so a path is bit-identical until one of its rays meets
such a pixel-level difference: the fraction of bit-identical pixels and the RMSE are asserted."""
import numpy as np
import pytest

from zig_raytracing_contest_b200 import abi, rtc, scenes

pytestmark = pytest.mark.gpu

BVH = abi.RTC_ACCEL_BVH


@pytest.fixture(scope="module")
def setups(orc):
    out = {}
    for name, spec, res, hgt in [("cornell", scenes.cornell(), (32, 32, 32), 144),
                                 ("atrium", scenes.atrium(detail=0.35), (64, 64, 64), 108),
                                 ("terrain", scenes.terrain(120_000), (128, 128, 128), 90)]:
        geo = rtc.Geometry(spec.positions, spec.tri_data, res)
        scene = rtc.Scene(geo, spec.materials, device=0)
        info = scene.build_bvh()
        assert info["num_tris"] == len(np.unique(geo.tri_data["src_tri"])) and info["num_nodes"] >= 1
        out[name] = (spec, geo, scene, spec.camera(height=hgt))
    yield out
    for v in out.values():
        v[2].close()


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def _ulps(a, b):
    a, b = np.float32(a), np.float32(b)
    if not (np.isfinite(a) and np.isfinite(b)):
        return 0 if a == b else 1 << 30
    return abs(int(a.view(np.int32)) - int(b.view(np.int32)))


def classify(orc, scene, geo, cam, g, o, limit=400):
    """Explains every pixel whose hit id differs. Returns dict(near_tie, grid_edge, unexplained, total)."""
    ys, xs = np.nonzero(g["hit_id"] != o["hit_id"])
    out = {"near_tie": 0, "grid_edge": 0, "unexplained": 0, "total": len(ys)}
    for y, x in list(zip(ys, xs))[:limit]:
        if _ulps(g["hit_tuv"][y, x, 0], o["hit_tuv"][y, x, 0]) <= 4:
            out["near_tie"] += 1
            continue
        org, d = orc.camera_get_ray(cam, x + 0.5, y + 0.5)
        idx, t, u, v = orc.trace_ray(scene.desc, org, d, brute=True)
        want = abi.MISS_ID if idx < 0 else int(geo.tri_data["src_tri"][idx])
        tg = g["hit_tuv"][y, x, 0]
        if (want == int(g["hit_id"][y, x]) and (idx < 0 or np.float32(t) == tg)) or (idx >= 0 and _ulps(t, tg) <= 4):
            out["grid_edge"] += 1
        else:
            out["unexplained"] += 1
    return out


@pytest.mark.parametrize("name", ["cornell", "atrium", "terrain"])
def test_bvh_primary_hit_ids(orc, setups, name):
    spec, geo, scene, cam = setups[name]
    g = scene.render(cam, rtc.make_params(mode=abi.RTC_MODE_PRIMARY, accel=BVH), want_hits=True, want_radiance=True, want_stats=True)
    o = orc.render(scene.desc, cam, rtc.make_params(mode=abi.RTC_MODE_PRIMARY), want_hits=True, want_radiance=True, want_stats=True)
    same = g["hit_id"] == o["hit_id"]
    assert (o["hit_id"] != abi.MISS_ID).mean() > 0.2
    # same triangle -> same t, u, v, same emissive/sky colour, bit for bit
    assert np.array_equal(_bits(g["hit_tuv"])[same], _bits(o["hit_tuv"])[same])
    assert np.array_equal(_bits(g["radiance"])[same], _bits(o["radiance"])[same])
    c = classify(orc, scene, geo, cam, g, o)
    print(name, "bvh vs oracle:", c, "of", same.size, "| per ray: nodes %.2f tests %.2f (grid: cells %.2f tests %.2f)" % (
        g["stats"]["cells_visited"] / g["stats"]["rays"], g["stats"]["tris_tested"] / g["stats"]["rays"],
        o["stats"]["cells_visited"] / o["stats"]["rays"], o["stats"]["tris_tested"] / o["stats"]["rays"]))
    assert c["unexplained"] == 0, c
    assert c["total"] <= 2e-3 * same.size, c                  # stated near-tie fraction (measured: see DESIGN.md)
    assert g["stats"]["rays"] == o["stats"]["rays"] == cam.w * cam.h
    assert g["stats"]["tris_tested"] < o["stats"]["tris_tested"]
    # production kernel (no counters) = instrumented kernel
    p = scene.render(cam, rtc.make_params(mode=abi.RTC_MODE_PRIMARY, accel=BVH), want_hits=True, want_radiance=True)
    assert np.array_equal(p["hit_id"], g["hit_id"]) and np.array_equal(_bits(p["hit_tuv"]), _bits(g["hit_tuv"]))


@pytest.mark.parametrize("name,spp,bounce", [("cornell", 8, 4), ("atrium", 8, 4), ("terrain", 6, 4), ("cornell", 5, 9)])
def test_bvh_path_radiance(orc, setups, name, spp, bounce):
    spec, geo, scene, cam = setups[name]
    g = scene.render(cam, rtc.make_params(num_samples=spp, max_bounce=bounce, seed=77, accel=BVH), want_radiance=True, want_stats=True)
    o = orc.render(scene.desc, cam, rtc.make_params(num_samples=spp, max_bounce=bounce, seed=77), want_radiance=True, want_stats=True)
    exact = (_bits(g["radiance"]) == _bits(o["radiance"])).all(axis=-1)
    rmse = float(np.sqrt(np.mean((g["radiance"] - o["radiance"]) ** 2)))
    o2 = orc.render(scene.desc, cam, rtc.make_params(num_samples=spp, max_bounce=bounce, seed=78), want_radiance=True)
    noise = float(np.sqrt(np.mean((o2["radiance"] - o["radiance"]) ** 2)))
    print(name, spp, bounce, "bit-identical pixels %.5f  rmse %.3e  MC noise floor %.3e  rays %d vs %d" % (
        exact.mean(), rmse, noise, g["stats"]["rays"], o["stats"]["rays"]))
    assert exact.mean() >= 0.98
    assert rmse <= 0.1 * noise
    assert g["stats"]["paths"] == o["stats"]["paths"] == cam.w * cam.h * spp
    assert abs(g["stats"]["rays"] - o["stats"]["rays"]) <= 1e-3 * o["stats"]["rays"]
    assert np.abs(g["rgb"].astype(int) - o["rgb"].astype(int))[exact].max() <= 1
    # deterministic: the same call again, and the frame dealt as strips over 3 "ranks", give the same bits
    g2 = scene.render(cam, rtc.make_params(num_samples=spp, max_bounce=bounce, seed=77, accel=BVH), want_radiance=True)
    assert np.array_equal(_bits(g2["radiance"]), _bits(g["radiance"]))
    img = np.zeros((cam.h, cam.w, 3), np.uint8)
    for r in range(3):
        scene.render(cam, rtc.make_params(num_samples=spp, max_bounce=bounce, seed=77, accel=BVH, tile_rank=r, tile_world=3, tile_rows=5), img=img)
    assert np.array_equal(img, g["rgb"])


def test_bvh_fast_math_and_device_built_scene(orc, setups):
    """The FMA-contracting build walks the same tree; a scene built on the device from the soup gives the same BVH
    answers as the uploaded one (the BVH is built from the scene's own device-resident records either way)."""
    spec, geo, scene, cam = setups["atrium"]
    prm = rtc.make_params(mode=abi.RTC_MODE_PRIMARY, accel=BVH)
    g = scene.render(cam, prm, want_hits=True)
    f = scene.render(cam, rtc.make_params(mode=abi.RTC_MODE_PRIMARY, accel=BVH, math=abi.RTC_MATH_FAST), want_hits=True)
    assert (f["hit_id"] != g["hit_id"]).mean() <= 1e-3
    built = rtc.Scene.from_soup(spec.positions, spec.tri_data, (64, 64, 64), spec.materials, device=0)
    b = built.render(cam, prm, want_hits=True)
    assert np.array_equal(b["hit_id"], g["hit_id"]) and np.array_equal(_bits(b["hit_tuv"]), _bits(g["hit_tuv"]))
    built.close()


def test_bvh_edge_scenes(orc):
    """One triangle; an empty scene; a camera outside looking away; exact-tie soup (duplicated triangles)."""
    M = rtc.MaterialData
    pos = np.array([[[0, 0, -2], [1, 0, -2], [0, 1, -2]]], np.float32)
    td = np.zeros(1, rtc.TRI_DATA_DTYPE); td["normal"][:] = (0, 0, 1)
    cam = scenes.load_camera(scenes.look_at((0.3, 0.3, 1), (0.3, 0.3, -2)), 0.8, 1.0, height=32)
    for p, t in ((pos, td), (np.zeros((0, 3, 3), np.float32), np.zeros(0, rtc.TRI_DATA_DTYPE)),
                 (np.repeat(pos, 7, axis=0), np.repeat(td, 7))):
        geo = rtc.Geometry(p, t, (4, 4, 4))
        sc = rtc.Scene(geo, [M(emissive=(1, 2, 3))], device=0)
        prm = rtc.make_params(mode=abi.RTC_MODE_PRIMARY, accel=BVH)
        g = sc.render(cam, prm, want_hits=True, want_radiance=True)
        o = orc.render(sc.desc, cam, rtc.make_params(mode=abi.RTC_MODE_PRIMARY), want_hits=True, want_radiance=True)
        # duplicates: every hit is an exact tie; the smaller source id wins in both structures (source order)
        assert np.array_equal(g["hit_id"], o["hit_id"])
        assert np.array_equal(_bits(g["hit_tuv"]), _bits(o["hit_tuv"]))
        assert np.array_equal(_bits(g["radiance"]), _bits(o["radiance"]))
        away = scenes.load_camera(scenes.look_at((0.3, 0.3, 1), (0.3, 0.3, 5)), 0.8, 1.0, height=16)
        ga = sc.render(away, prm, want_hits=True)
        assert (ga["hit_id"] == abi.MISS_ID).all()
        sc.close()
