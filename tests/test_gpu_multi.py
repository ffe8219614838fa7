"""Multi-GPU behind the C ABI (SURVEY §8(b) num_gpus / shard, §8(e)), on a box with >= 2 B200s (skipped on one).
The reference's caller is one blocking call (src/main.zig:126); the tests drive the same shape:
  * rtc_render(num_gpus = N) from ONE process: strips -> the frame is BYTE-IDENTICAL to the one-GPU frame (and its
    radiance bit-identical, hence to the oracle); sample ranges -> radiance equal to fp32 summation order, RGB8 +-1;
  * bin/rtc --gpus N writes the same PNG as --gpus 1 (strips);
  * one process per GPU: rtc_comm + rtc_render_sharded under torchrun-style ranks (two processes spawned here, the
    communicator id handed over through a file) gives rank 0 the same frame.
On a one-GPU box the error path is checked instead: num_gpus = 2 fails with RTC_ERR_NO_DEVICE, nothing rendered."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from zig_raytracing_contest_b200 import abi, gltf_io, rtc, scenes

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def _ngpu():
    return rtc.device_count()


@pytest.fixture(scope="module")
def setup():
    spec = scenes.atrium(detail=0.35)
    geo = rtc.Geometry(spec.positions, spec.tri_data, (64, 64, 64))
    scene = rtc.Scene(geo, spec.materials, device=0)
    yield spec, geo, scene, spec.camera(height=135)        # 240 x 135: strips of 8 rows leave a short last strip
    scene.close()


def test_num_gpus_beyond_the_box_is_an_error(setup):
    spec, geo, scene, cam = setup
    n = _ngpu() + 1
    if n > abi.RTC_MAX_GPUS:
        pytest.skip("box has RTC_MAX_GPUS devices")
    with pytest.raises(rtc.RtcError) as e:
        scene.render(cam, rtc.make_params(num_samples=2, num_gpus=n))
    assert e.value.code == abi.RTC_ERR_NO_DEVICE
    with pytest.raises(rtc.RtcError) as e:
        scene.prepare_gpus(n)
    assert e.value.code == abi.RTC_ERR_NO_DEVICE
    scene.prepare_gpus(1, abi.RTC_ACCEL_BVH)                # one device: builds the BVH, nothing else
    assert scene.build_bvh()["num_tris"] == len(np.unique(geo.tri_data["src_tri"]))
    with pytest.raises(rtc.RtcError) as e:                  # sharding fields and num_gpus exclude each other
        scene.render(cam, rtc.make_params(num_samples=2, num_gpus=2, tile_rank=0, tile_world=2))
    assert e.value.code == abi.RTC_ERR_INVALID_ARGUMENT


@pytest.mark.parametrize("accel", [abi.RTC_ACCEL_GRID, abi.RTC_ACCEL_BVH])
def test_inprocess_strips_equal_one_gpu(orc, setup, accel):
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    spec, geo, scene, cam = setup
    n = min(_ngpu(), 4)
    scene.prepare_gpus(n, accel)                            # replicas ahead of the first render (idempotent)
    one = scene.render(cam, rtc.make_params(num_samples=6, seed=3, accel=accel), want_radiance=True, want_stats=True, instrumented=False)
    many = scene.render(cam, rtc.make_params(num_samples=6, seed=3, accel=accel, num_gpus=n, shard=abi.RTC_SHARD_STRIPS),
                        want_radiance=True, want_stats=True, instrumented=False)
    assert np.array_equal(many["rgb"], one["rgb"])
    assert np.array_equal(_bits(many["radiance"]), _bits(one["radiance"]))
    assert many["stats"]["rays"] == one["stats"]["rays"] and many["stats"]["paths"] == one["stats"]["paths"] == cam.w * cam.h * 6
    if accel == abi.RTC_ACCEL_GRID:
        o = orc.render(scene.desc, cam, rtc.make_params(num_samples=6, seed=3), want_radiance=True)
        assert np.array_equal(_bits(many["radiance"]), _bits(o["radiance"]))
    # odd strip height, and the deterministic slice (hit buffers come back from every device)
    m3 = scene.render(cam, rtc.make_params(num_samples=6, seed=3, accel=accel, num_gpus=n, tile_rows=5))
    assert np.array_equal(m3["rgb"], one["rgb"])
    p1 = scene.render(cam, rtc.make_params(mode=abi.RTC_MODE_PRIMARY, accel=accel), want_hits=True)
    pn = scene.render(cam, rtc.make_params(mode=abi.RTC_MODE_PRIMARY, accel=accel, num_gpus=n), want_hits=True)
    assert np.array_equal(pn["hit_id"], p1["hit_id"]) and np.array_equal(_bits(pn["hit_tuv"]), _bits(p1["hit_tuv"]))
    assert np.array_equal(pn["rgb"], p1["rgb"])


def test_inprocess_sample_ranges(setup):
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    spec, geo, scene, cam = setup
    n = min(_ngpu(), 4)
    one = scene.render(cam, rtc.make_params(num_samples=9, seed=4), want_radiance=True, want_stats=True, instrumented=False)
    many = scene.render(cam, rtc.make_params(num_samples=9, seed=4, num_gpus=n, shard=abi.RTC_SHARD_SAMPLES),
                        want_radiance=True, want_stats=True, instrumented=False)
    # the same 9 samples per pixel, summed in a different association: fp32 reordering only
    np.testing.assert_allclose(many["radiance"], one["radiance"], rtol=2e-6, atol=1e-7)
    assert np.abs(many["rgb"].astype(int) - one["rgb"].astype(int)).max() <= 1
    assert many["stats"]["rays"] == one["stats"]["rays"] and many["stats"]["paths"] == one["stats"]["paths"]
    again = scene.render(cam, rtc.make_params(num_samples=9, seed=4, num_gpus=n, shard=abi.RTC_SHARD_SAMPLES), want_radiance=True)
    assert np.array_equal(_bits(again["radiance"]), _bits(many["radiance"]))      # fixed summation order: deterministic


def test_cli_gpus_writes_the_same_png(tmp_path):
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    spec = scenes.cornell(subdiv=3)
    gltf_io.write_gltf(spec, str(tmp_path / "room.glb"), glb=True)
    json.dump({"grid_resolution": [16, 16, 16], "num_threads": 2, "num_samples": 5, "max_bounce": 4}, open(str(tmp_path / "config.json"), "w"))
    exe = os.path.join(ROOT, "zig_raytracing_contest_b200", "bin", "rtc")
    outs = []
    for extra in ([], ["--gpus", "2"], ["--gpus=2", "--accel", "bvh"]):
        r = subprocess.run([exe, "--in", "room.glb", "--out", "o%d.png" % len(outs), "--height", "90", "--seed", "7"] + extra,
                           cwd=str(tmp_path), capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr
        outs.append(rtc.png_read(str(tmp_path / ("o%d.png" % len(outs)))))
    assert np.array_equal(outs[0], outs[1])
    assert (outs[2] != outs[0]).mean() < 1e-3                # BVH: equal but for enumerated near-ties


RANK_SCRIPT = r'''
import os, sys, time
import numpy as np
sys.path.insert(0, %(root)r)
from zig_raytracing_contest_b200 import abi, rtc, scenes
rank, world, tmp = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
idf = os.path.join(tmp, "id.bin")
if rank == 0:
    open(idf + ".tmp", "wb").write(rtc.comm_unique_id()); os.rename(idf + ".tmp", idf)
while not os.path.exists(idf):
    time.sleep(0.05)
comm = rtc.Comm(open(idf, "rb").read(), rank, world, device=rank)
spec = scenes.atrium(detail=0.35)
geo = rtc.Geometry(spec.positions, spec.tri_data, (64, 64, 64))
scene = rtc.Scene(geo, spec.materials, device=rank)
cam = spec.camera(height=135)
for k, (shard, accel) in enumerate([(abi.RTC_SHARD_STRIPS, abi.RTC_ACCEL_GRID), (abi.RTC_SHARD_SAMPLES, abi.RTC_ACCEL_GRID),
                                    (abi.RTC_SHARD_STRIPS, abi.RTC_ACCEL_BVH)]):
    prm = rtc.make_params(num_samples=6, seed=3, shard=shard, accel=accel, tile_rows=8)
    out = scene.render_sharded(comm, cam, prm, want_stats=True)
    scene.render_sharded(comm, cam, prm, to_host=False)                 # the device-resident variant runs too
    if rank == 0:
        np.save(os.path.join(tmp, "frame%%d.npy" %% k), out["rgb"])
    np.save(os.path.join(tmp, "rays%%d_%%d.npy" %% (k, rank)), np.array([out["stats"]["rays"], out["stats"]["paths"]]))
scene.close(); comm.close()
'''


def test_one_process_per_gpu_comm(setup, tmp_path):
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    spec, geo, scene, cam = setup
    world = 2
    script = tmp_path / "rank.py"
    script.write_text(RANK_SCRIPT % {"root": ROOT})
    procs = [subprocess.Popen([sys.executable, str(script), str(r), str(world), str(tmp_path)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(world)]
    outs = [p.communicate(timeout=600)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), "\n".join(outs)
    one = scene.render(cam, rtc.make_params(num_samples=6, seed=3), want_stats=True, instrumented=False)
    assert np.array_equal(np.load(str(tmp_path / "frame0.npy")), one["rgb"])                     # strips: byte-identical
    assert np.abs(np.load(str(tmp_path / "frame1.npy")).astype(int) - one["rgb"].astype(int)).max() <= 1    # samples: +-1 LSB
    bvh1 = scene.render(cam, rtc.make_params(num_samples=6, seed=3, accel=abi.RTC_ACCEL_BVH))
    assert np.array_equal(np.load(str(tmp_path / "frame2.npy")), bvh1["rgb"])
    for k in range(3):
        tot = sum(np.load(str(tmp_path / ("rays%d_%d.npy" % (k, r)))) for r in range(world))
        assert tot[1] == cam.w * cam.h * 6
        if k < 2:
            assert tot[0] == one["stats"]["rays"]
