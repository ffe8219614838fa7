"""Parity of the CUDA render path (through the C ABI) against the CPU oracle on the same seeded
inputs. Bars: primary-ray hit ids and t,u,v bit-exact in the exact build; f32
radiance bit-exact in the exact build (same counter RNG, same operation order); RGB8 within
1 LSB (powf differs between libm and CUDA); fast build: hit ids equal except near-ties (fraction
asserted), radiance within 1e-3 relative RMSE."""
import numpy as np
import pytest

from zig_raytracing_contest_b200 import abi, rtc, scenes, shard

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def setups(orc):
    out = {}
    for name, spec, res, hgt in [("cornell", scenes.cornell(), (32, 32, 32), 144),
                                 ("atrium", scenes.atrium(detail=0.35), (64, 64, 64), 108),
                                 ("terrain", scenes.terrain(120_000), (128, 128, 128), 90)]:
        geo = rtc.Geometry(spec.positions, spec.tri_data, res)
        out[name] = (spec, geo, rtc.Scene(geo, spec.materials, device=0), spec.camera(height=hgt))
    yield out
    for v in out.values():
        v[2].close()


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


@pytest.mark.parametrize("name", ["cornell", "atrium", "terrain"])
def test_primary_hit_ids_bit_exact(orc, setups, name):
    spec, geo, scene, cam = setups[name]
    prm = rtc.make_params(mode=abi.RTC_MODE_PRIMARY)
    g = scene.render(cam, prm, want_hits=True, want_radiance=True, want_stats=True)
    o = orc.render(scene.desc, cam, prm, want_hits=True, want_radiance=True, want_stats=True)
    assert np.array_equal(g["hit_id"], o["hit_id"])
    assert (o["hit_id"] != abi.MISS_ID).mean() > 0.2          # the scene is actually in view
    assert np.array_equal(_bits(g["hit_tuv"]), _bits(o["hit_tuv"]))
    assert np.array_equal(_bits(g["radiance"]), _bits(o["radiance"]))
    assert np.abs(g["rgb"].astype(int) - o["rgb"].astype(int)).max() <= 1
    for k in ("rays", "cells_visited", "tris_tested", "hits_shaded", "paths"):
        assert g["stats"][k] == o["stats"][k], k
    # the production kernels (no counters: empty blocks taken in one step, csrc/rtc_trav.cuh) give the same bits
    p = scene.render(cam, prm, want_hits=True, want_radiance=True)
    assert np.array_equal(p["hit_id"], o["hit_id"]) and np.array_equal(_bits(p["hit_tuv"]), _bits(o["hit_tuv"]))
    assert np.array_equal(_bits(p["radiance"]), _bits(o["radiance"]))


@pytest.mark.parametrize("name", ["cornell", "atrium"])
def test_primary_matches_brute_force(orc, setups, name):
    """cfg2: accelerated traversal vs a loop over ALL baked triangles (no grid), on a pixel subset."""
    spec, geo, scene, cam = setups[name]
    g = scene.render(cam, rtc.make_params(mode=abi.RTC_MODE_PRIMARY), want_hits=True)
    rng = np.random.RandomState(0)
    ties = 0
    for _ in range(150 if name == "cornell" else 25):
        x, y = rng.randint(cam.w), rng.randint(cam.h)
        o, d = orc.camera_get_ray(cam, x + 0.5, y + 0.5)
        idx, t, u, v = orc.trace_ray(scene.desc, o, d, brute=True)
        want = abi.MISS_ID if idx < 0 else int(geo.tri_data["src_tri"][idx])
        if want != int(g["hit_id"][y, x]):
            ties += 1                                       # only an exact-t tie may differ (first tested wins)
            assert g["hit_tuv"][y, x, 0] == np.float32(t)
    assert ties <= 2


@pytest.mark.parametrize("name,spp,bounce", [("cornell", 3, 4), ("cornell", 40, 6), ("atrium", 8, 4),
                                             ("terrain", 5, 4), ("cornell", 33, 1), ("atrium", 2, 12)])
def test_path_radiance_bit_exact(orc, setups, name, spp, bounce):
    spec, geo, scene, cam = setups[name]
    prm = rtc.make_params(num_samples=spp, max_bounce=bounce, seed=0x1234567890 + spp)
    g = scene.render(cam, prm, want_radiance=True, want_stats=True)
    o = orc.render(scene.desc, cam, prm, want_radiance=True, want_stats=True)
    assert np.array_equal(_bits(g["radiance"]), _bits(o["radiance"]))
    assert np.abs(g["rgb"].astype(int) - o["rgb"].astype(int)).max() <= 1
    assert (g["rgb"] != o["rgb"]).mean() < 0.02
    for k in ("rays", "cells_visited", "tris_tested", "hits_shaded", "paths"):
        assert g["stats"][k] == o["stats"][k], k
    assert g["stats"]["paths"] == cam.w * cam.h * spp
    # production kernels (no counters, two-level walk): same bits
    p = scene.render(cam, prm, want_radiance=True)
    assert np.array_equal(_bits(p["radiance"]), _bits(o["radiance"]))
    assert np.array_equal(p["rgb"], g["rgb"])


def test_fast_math_within_tolerance(orc, setups):
    spec, geo, scene, cam = setups["atrium"]
    o1 = orc.render(scene.desc, cam, rtc.make_params(mode=abi.RTC_MODE_PRIMARY), want_hits=True)
    g1 = scene.render(cam, rtc.make_params(mode=abi.RTC_MODE_PRIMARY, math=abi.RTC_MATH_FAST), want_hits=True)
    print("fast math: primary hit ids differing from the oracle: %d of %d" % ((g1["hit_id"] != o1["hit_id"]).sum(), g1["hit_id"].size))
    assert (g1["hit_id"] != o1["hit_id"]).mean() <= 1e-3        # stated near-tie fraction for the fast build
    prm = rtc.make_params(num_samples=16, max_bounce=4, seed=9, math=abi.RTC_MATH_FAST)
    g = scene.render(cam, prm, want_radiance=True)
    o = orc.render(scene.desc, cam, rtc.make_params(num_samples=16, max_bounce=4, seed=9), want_radiance=True)
    # The fast build is NOT path-identical: alpha pass-through restarts the ray ON the surface
    # (stage3.zig:208-211, t + floatEps), so whether the same triangle is re-hit at t ~ +-1e-7 depends on
    # rounding, and contraction changes rounding. The check is therefore statistical: the difference to
    # the oracle stays below the oracle's own Monte-Carlo noise floor (two seeds) and the image mean agrees.
    rmse = np.sqrt(np.mean((g["radiance"] - o["radiance"]) ** 2))
    o2 = orc.render(scene.desc, cam, rtc.make_params(num_samples=16, max_bounce=4, seed=10), want_radiance=True)
    noise = np.sqrt(np.mean((o2["radiance"] - o["radiance"]) ** 2))
    print("fast math: radiance rmse vs oracle %.4e, Monte-Carlo noise floor (two seeds) %.4e, mean %.4f vs %.4f" % (
        rmse, noise, g["radiance"].mean(), o["radiance"].mean()))
    assert rmse <= 0.5 * noise, (rmse, noise)
    assert abs(g["radiance"].mean() - o["radiance"].mean()) <= 5e-3 * o["radiance"].mean()


def test_sample_sharding_and_row_sharding(orc, setups):
    """§8(e): samples [0,S) split over two calls into one accumulator == one call, within fp32 sum
    reordering; row bands from two 'ranks' tile the frame exactly."""
    spec, geo, scene, cam = setups["cornell"]
    full = scene.render(cam, rtc.make_params(num_samples=8, seed=3), want_radiance=True)
    img = np.zeros((cam.h, cam.w, 3), np.uint8)
    rad = []
    for r in range(2):
        out = scene.render(cam, rtc.make_params(num_samples=8, seed=3, tile_rank=r, tile_world=2), img=img, want_radiance=True)
        rad.append(out["radiance"])
    assert np.array_equal(img, full["rgb"])
    h2 = cam.h // 2
    assert np.array_equal(_bits(rad[0][:h2]), _bits(full["radiance"][:h2]))
    assert np.array_equal(_bits(rad[1][h2:]), _bits(full["radiance"][h2:]))
    # strips of k rows dealt round-robin over 3 "ranks" (the last strip is short when k does not divide h)
    for rows in (1, 5, 16, 1000):
        img2 = np.zeros((cam.h, cam.w, 3), np.uint8)
        rad2 = np.zeros((cam.h, cam.w, 3), np.float32)
        paths = 0
        for r in range(3):
            out = scene.render(cam, rtc.make_params(num_samples=8, seed=3, tile_rank=r, tile_world=3, tile_rows=rows), img=img2,
                               want_radiance=True, want_stats=True)
            paths += out["stats"]["paths"]
            for a, b in shard.strips_of(r, 3, cam.h, rows):
                rad2[a:b] = out["radiance"][a:b]
        assert paths == cam.w * cam.h * 8                          # every pixel rendered by exactly one rank
        assert np.array_equal(img2, full["rgb"])
        assert np.array_equal(_bits(rad2), _bits(full["radiance"]))
    import torch
    acc = torch.zeros(cam.h * cam.w, 4, device="cuda:0")
    for r in range(2):
        scene.render_accumulate(cam, rtc.make_params(num_samples=4, seed=3, sample_offset=4 * r, total_samples=8), acc.data_ptr())
    torch.cuda.synchronize()
    d_rad = torch.zeros(cam.h * cam.w, 3, device="cuda:0")
    d_rgb = torch.zeros(cam.h * cam.w, 3, dtype=torch.uint8, device="cuda:0")
    rtc.resolve(acc.data_ptr(), cam.h * cam.w, 8, d_rgb.data_ptr(), d_rad.data_ptr(), device=0)
    torch.cuda.synchronize()
    assert float(acc[:, 3].min()) == 8.0 and float(acc[:, 3].max()) == 8.0
    np.testing.assert_allclose(d_rad.cpu().numpy().reshape(cam.h, cam.w, 3), full["radiance"], rtol=2e-6, atol=1e-7)
    assert np.abs(d_rgb.cpu().numpy().reshape(cam.h, cam.w, 3).astype(int) - full["rgb"].astype(int)).max() <= 1


def test_edge_cases(orc):
    """Empty scene, camera inside / outside / behind the grid, axis-parallel rays (±inf DDA lanes,
    src/linalg.zig:609-651), single-cell grid, max_bounce = 0, and error codes."""
    M = rtc.MaterialData
    td0 = np.zeros(0, rtc.TRI_DATA_DTYPE)
    # one emissive triangle facing +z, grid 1x1x1 and 5x5x5
    pos = np.array([[[-1, -1, 0], [1, -1, 0], [0, 1, 0]]], np.float32)
    td = np.zeros(1, rtc.TRI_DATA_DTYPE)
    td["normal"][:] = (0, 0, 1)
    mats = [M(base_color=(0.5, 0.5, 0.5), emissive=(2.0, 1.0, 0.5))]
    cams = []
    for eye, tgt in [((0, 0, 3), (0, 0, 0)), ((0, 0, 0.5), (0, 0, 0)), ((0, 0, -3), (0, 0, -6)), ((0, 0, -3), (0, 0, 0))]:
        cams.append(scenes.load_camera(scenes.look_at(eye, tgt), 0.8, 1.0, None, 33))
    axis = abi.Camera()                       # every ray exactly along -z (up/right of zero length in x,y jitter)
    axis.w, axis.h = 8, 8
    axis.origin[:] = (0.1, -0.2, 5.0); axis.lower_left_corner[:] = (0.0, 0.0, -1.0)
    axis.right[:] = (0.0, 0.0, 0.0); axis.up[:] = (0.0, 0.0, 0.0)
    cams.append(axis)
    for res in [(1, 1, 1), (5, 5, 5), (3, 1, 7)]:
        geo = rtc.Geometry(pos, td, res)
        scene = rtc.Scene(geo, mats, device=0)
        for cam in cams:
            for prm in (rtc.make_params(mode=abi.RTC_MODE_PRIMARY), rtc.make_params(num_samples=5, max_bounce=3, seed=2),
                        rtc.make_params(num_samples=2, max_bounce=0, seed=2)):
                g = scene.render(cam, prm, want_radiance=True, want_hits=True)
                o = orc.render(scene.desc, cam, prm, want_radiance=True, want_hits=True)
                assert np.array_equal(_bits(g["radiance"]), _bits(o["radiance"]))
                assert np.array_equal(g["hit_id"], o["hit_id"])
        with pytest.raises(rtc.RtcError) as e:
            scene.render(cams[0], rtc.make_params(num_samples=1, max_bounce=65))
        assert e.value.code == abi.RTC_ERR_UNSUPPORTED
        scene.close()
    # empty scene: every pixel is sky
    geo = rtc.Geometry(np.zeros((0, 3, 3), np.float32), td0, (4, 4, 4))
    scene = rtc.Scene(geo, mats, device=0)
    g = scene.render(cams[0], rtc.make_params(num_samples=4, seed=1), want_radiance=True, want_stats=True)
    o = orc.render(scene.desc, cams[0], rtc.make_params(num_samples=4, seed=1), want_radiance=True)
    assert np.array_equal(_bits(g["radiance"]), _bits(o["radiance"])) and g["stats"]["hits_shaded"] == 0
    scene.close()


@pytest.mark.parametrize("seed,res", [(0, (8, 8, 8)), (1, (12, 4, 20)), (2, (7, 5, 3)), (3, (16, 16, 4)), (4, (1, 9, 2)), (5, (32, 32, 32))])
def test_random_soups_with_exact_ties(orc, seed, res):
    """Random triangle soups built to provoke what procedural scenes avoid: exact duplicates and coplanar overlapping
    triangles (equal t: the first tested must win — cell order, then source order), triangles spanning the whole grid,
    slivers, zero-area triangles, axis-aligned quads on integer coordinates seen by rays through their shared edges,
    grids that are and are not multiples of the 4x4x4 block (two-level walk on / off), cameras inside and outside.
    Hit ids, t/u/v, radiance and the traversal counters against the oracle, for the uploaded and the device-built scene."""
    rng = np.random.RandomState(100 + seed)
    tris = []
    c = rng.uniform(-1, 1, (60, 1, 3))
    tris.append(c + rng.normal(0, 0.25, (60, 3, 3)))                                   # ordinary
    tris.append(rng.uniform(-1.2, 1.2, (6, 3, 3)))                                     # spanning many cells
    base = c[:8] + rng.normal(0, 0.3, (8, 3, 3))
    tris += [base, base.copy(), base[:, [1, 2, 0]]]                                    # duplicates (and a rotated vertex order)
    q = np.array([[[0, 0, 0], [1, 0, 0], [1, 1, 0]], [[0, 0, 0], [1, 1, 0], [0, 1, 0]]], np.float64)
    for ox in (-1, 0):
        for oy in (-1, 0):
            tris.append(q + (ox, oy, -0.5))                                            # a 2x2 wall of unit quads at z = -0.5
    tris.append(q * 1.5 + (-0.75, -0.75, -0.5))                                        # coplanar with the wall, overlapping it
    p = rng.uniform(-1, 1, (5, 1, 3))
    tris.append(np.concatenate([p, p, p + rng.normal(0, 0.2, (5, 1, 3))], axis=1))     # zero area
    tris.append(np.concatenate([p, p + 1e-4, p + rng.normal(0, 0.5, (5, 1, 3))], axis=1))   # slivers
    pos = np.concatenate(tris).astype(np.float32)
    n = len(pos)
    td = np.zeros(n, rtc.TRI_DATA_DTYPE)
    e1, e2 = pos[:, 1] - pos[:, 0], pos[:, 2] - pos[:, 0]
    nrm = np.cross(e1, e2)
    ln = np.linalg.norm(nrm, axis=1, keepdims=True)
    nrm = np.where(ln > 0, nrm / np.maximum(ln, 1e-30), (0, 0, 1)).astype(np.float32)
    td["normal"][:] = nrm[:, None, :]
    td["texcoord"][:] = rng.uniform(0, 1, (n, 3, 2))
    td["material"][:] = rng.randint(0, 3, n)
    M = rtc.MaterialData
    mats = [M(base_color=(0.7, 0.6, 0.5)), M(base_color=(0.2, 0.5, 0.8), emissive=(1.5, 1.2, 0.9)),
            M(base_color=rtc.TextureData(rng.uniform(0, 1, (8, 8, 3)).astype(np.float32)),
              transparency=rtc.TextureData((rng.uniform(0, 1, (4, 4, 1)) > 0.4).astype(np.float32)))]
    geo = rtc.Geometry(pos, td, res)
    a = rtc.Scene(geo, mats, device=0)
    b = rtc.Scene.from_soup(pos, td, res, mats, device=0)
    ogeo = orc.Geometry(pos, td, res)
    assert np.array_equal(geo.cells, ogeo.cells) and np.array_equal(geo.tri_data["src_tri"], ogeo.indices.astype(np.uint32))
    cams = [scenes.load_camera(scenes.look_at((0.0, 0.0, 3.0), (0, 0, -0.5)), 0.9, 1.0, None, 64),     # pixel centres meet the wall's edges
            scenes.load_camera(scenes.look_at((0.3, -0.2, 0.1), (1, 0.5, -1)), 1.2, 1.5, None, 48),    # inside the soup
            scenes.load_camera(scenes.look_at((2.5, 2.0, -2.0), (0, 0, 0)), 0.7, 1.0, None, 48)]
    for cam in cams:
        for prm in (rtc.make_params(mode=abi.RTC_MODE_PRIMARY), rtc.make_params(num_samples=3, max_bounce=4, seed=seed)):
            o = orc.render(a.desc, cam, prm, want_radiance=True, want_hits=True, want_stats=True)
            primary = prm.mode == abi.RTC_MODE_PRIMARY
            for scene in (a, b):
                g = scene.render(cam, prm, want_radiance=True, want_hits=True, want_stats=True)
                if primary:
                    assert np.array_equal(g["hit_id"], o["hit_id"])
                    assert np.array_equal(_bits(g["hit_tuv"]), _bits(o["hit_tuv"]))
                assert np.array_equal(_bits(g["radiance"]), _bits(o["radiance"]))
                for k in ("rays", "cells_visited", "tris_tested", "hits_shaded", "paths"):
                    assert g["stats"][k] == o["stats"][k], k
            # the production kernels (no counters: two-level walk where the grid allows it) give the same frame
            g = a.render(cam, prm, want_radiance=True, want_hits=True)
            assert np.array_equal(_bits(g["radiance"]), _bits(o["radiance"]))
            if primary:
                assert np.array_equal(g["hit_id"], o["hit_id"])
                assert (o["hit_id"] != 0xFFFFFFFF).sum() > 100
    a.close(); b.close()
