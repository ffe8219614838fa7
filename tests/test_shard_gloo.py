"""The N>1 host path on CPU: two processes over gloo shard one frame by sample range (one
reduce(SUM) of the accumulators) and by row band (one gather), with the CPU oracle standing in for the
device render. Checks the partition arithmetic of zig_raytracing_contest_b200/shard.py, that the
counter RNG makes the sharded frame equal the unsharded one, and the collective plumbing bench.py uses."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from zig_raytracing_contest_b200 import abi, rtc, scenes, shard


def test_sample_shard_covers_every_sample_once():
    for total in (1, 3, 64, 65, 4096):
        for world in (1, 2, 3, 4, 8):
            seen = []
            for r in range(world):
                off, cnt = shard.sample_shard(r, world, total)
                seen += list(range(off, off + cnt))
            assert seen == list(range(total))
            counts = [shard.sample_shard(r, world, total)[1] for r in range(world)]
            assert max(counts) - min(counts) <= 1
    with pytest.raises(ValueError):
        shard.sample_shard(2, 2, 8)


def test_row_band_tiles_the_frame():
    for h in (1, 7, 1080, 2160):
        for world in (1, 2, 3, 8):
            edges = [shard.row_band(r, world, h) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == h
            assert all(edges[i][1] == edges[i + 1][0] for i in range(world - 1))


def test_strips_tile_the_frame():
    for h in (1, 7, 27, 1080, 2160):
        for world in (1, 2, 3, 8):
            for rows in (1, 4, 8, 4000):
                got = sorted(sum((shard.strips_of(r, world, h, rows) for r in range(world)), []))
                assert got[0][0] == 0 and got[-1][1] == h
                assert all(got[i][1] == got[i + 1][0] for i in range(len(got) - 1))
                assert all(b - a == rows for a, b in got[:-1])


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle
    spec = scenes.cornell(subdiv=2)
    geo = rtc.Geometry(spec.positions, spec.tri_data, (8, 8, 8))
    cam = spec.camera(height=27)
    desc, keep = rtc.make_scene_desc(geo, spec.materials)
    S, P = 6, cam.w * cam.h
    # ---- sample ranges + reduce(SUM) ----
    off, cnt = shard.sample_shard(rank, world, S)
    prm = rtc.make_params(num_samples=cnt, max_bounce=3, seed=11, sample_offset=off, total_samples=S)
    part = oracle.render(desc, cam, prm, num_threads=1)["radiance"]          # = (sum over my samples) / S
    acc = torch.zeros(P, 4)
    acc[:, :3] = torch.from_numpy(part.reshape(P, 3)) * S
    acc[:, 3] = cnt
    shard.reduce_accumulators(acc, dst=0)
    # ---- row bands + gather ----
    y0, y1 = shard.row_band(rank, world, cam.h)
    full_prm = rtc.make_params(num_samples=S, max_bounce=3, seed=11)
    band = oracle.render(desc, cam, full_prm, num_threads=1, rows=(y0, y1))["rgb"][y0:y1]
    frame = shard.gather_row_bands(torch.from_numpy(band.copy()), cam.w, cam.h, dst=0)
    # ---- row strips dealt round-robin + gather (cam.h = 27 rows, strips of 4: the last strip is short) ----
    mine = torch.zeros(cam.h, cam.w, 3, dtype=torch.uint8)
    for a, b in shard.strips_of(rank, world, cam.h, 4):
        mine[a:b] = torch.from_numpy(oracle.render(desc, cam, full_prm, num_threads=1, rows=(a, b))["rgb"][a:b].copy())
    frame2 = shard.gather_strips(mine, 4, dst=0)
    # a NON-CONTIGUOUS frame whose strips tile it exactly (27 rows = 3 strips of 9): reshape of a strided tensor copies,
    # so the gathered strips must still land in the frame that is returned
    strided = torch.zeros(cam.h, cam.w * 2, 3, dtype=torch.uint8)[:, ::2]
    assert not strided.is_contiguous()
    for a, b in shard.strips_of(rank, world, cam.h, 9):
        strided[a:b] = torch.from_numpy(oracle.render(desc, cam, full_prm, num_threads=1, rows=(a, b))["rgb"][a:b].copy())
    frame3 = shard.gather_strips(strided, 9, dst=0)
    # what bench.py / tools/cfg5_sweep.py ask of torch.distributed at N > 1: the library's communicator id from rank 0 to
    # every rank (made by librtc.so's run-time-loaded NCCL when that loads here, else a stand-in), and a max over ranks
    def make_id():
        try:
            return rtc.comm_unique_id()
        except (rtc.RtcError, ImportError, OSError):
            return bytes(range(128))
    cid = shard.broadcast_comm_id(make_id, src=0)
    slowest = shard.max_over_ranks(10.0 + rank)
    np.save(os.path.join(out_dir, "cid%d.npy" % rank), np.frombuffer(cid, np.uint8))
    np.save(os.path.join(out_dir, "max%d.npy" % rank), np.array([slowest]))
    if rank == 0:
        ref = oracle.render(desc, cam, full_prm, num_threads=1)
        np.save(os.path.join(out_dir, "acc.npy"), acc.numpy())
        np.save(os.path.join(out_dir, "frame.npy"), frame.numpy())
        np.save(os.path.join(out_dir, "frame2.npy"), frame2.numpy())
        np.save(os.path.join(out_dir, "frame3.npy"), frame3.numpy())
        np.save(os.path.join(out_dir, "ref_rad.npy"), ref["radiance"])
        np.save(os.path.join(out_dir, "ref_rgb.npy"), ref["rgb"])
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_reproduce_the_single_frame(tmp_path, orc):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    acc = np.load(tmp_path / "acc.npy")
    ref_rad = np.load(tmp_path / "ref_rad.npy").reshape(-1, 3)
    assert (acc[:, 3] == 6).all()                                    # every sample rendered exactly once
    np.testing.assert_allclose(acc[:, :3] / 6.0, ref_rad, rtol=3e-6, atol=1e-7)   # fp32 sum reordering only
    assert np.array_equal(np.load(tmp_path / "frame.npy"), np.load(tmp_path / "ref_rgb.npy"))
    assert np.array_equal(np.load(tmp_path / "frame2.npy"), np.load(tmp_path / "ref_rgb.npy"))
    assert np.array_equal(np.load(tmp_path / "frame3.npy"), np.load(tmp_path / "ref_rgb.npy"))
    ids = [np.load(tmp_path / ("cid%d.npy" % r)) for r in range(world)]
    assert ids[0].shape == (128,) and np.array_equal(ids[0], ids[1])         # every rank holds rank 0's communicator id
    assert [float(np.load(tmp_path / ("max%d.npy" % r))[0]) for r in range(world)] == [11.0, 11.0]
