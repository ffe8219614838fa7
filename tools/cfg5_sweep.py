#!/usr/bin/env python
"""BASELINE.json configs[4]: spp sweep at 3840x2160 on the 5 M-triangle scene across 1/2/4/8 GPUs, the frame's samples
sharded over the ranks and the framebuffers combined with an NCCL reduce — through the library's own collective entry
point (rtc_render_sharded, RTC_SHARD_SAMPLES; add --strips for the strip partition). One process per GPU:

    python tools/cfg5_sweep.py --out gpurun_out/cfg5_n1.json                                     (N = 1)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 ... tools/cfg5_sweep.py --out ...

The scene (and the BVH) is built once per rank, then every (accel, spp) is one warm-up frame (spp <= 256) and one frame
timed on the device (CUDA events on the stream the library launches on, frame left on rank 0's device, max over ranks)
plus one end-to-end frame (host RGB8 out on rank 0, host clock between barriers). Rows go to --out as JSON.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", required=True)
    ap.add_argument("--spp", type=int, nargs="+", default=[16, 64, 256, 1024, 4096])
    ap.add_argument("--grid-spp", type=int, nargs="*", default=[16, 64, 256], help="spp values also run on the exact grid path")
    ap.add_argument("--tris", type=int, default=5_000_000)
    ap.add_argument("--height", type=int, default=2160)
    ap.add_argument("--strips", action="store_true", help="strip partition + send/recv instead of sample ranges + reduce")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from zig_raytracing_contest_b200 import abi, rtc, scenes, shard
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    spec = scenes.terrain(args.tris)
    t0 = time.perf_counter()
    scene = rtc.Scene.from_soup(spec.positions, spec.tri_data, (128, 128, 128), spec.materials, device=local)   # built on the device
    build_s = time.perf_counter() - t0
    bvh = scene.build_bvh()
    cam = spec.camera(height=args.height)
    comm = None
    if world > 1:
        comm = rtc.Comm(shard.broadcast_comm_id(rtc.comm_unique_id), rank, world, device=local)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    scene.set_stream(stream.cuda_stream)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    img = np.zeros((cam.h, cam.w, 3), np.uint8)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(x):
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    def allsum(x):
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t[0])

    # one-off allocations (the pinned host frame, the wavefront scratch, NCCL's channels) happen here, not in a timed row
    scene.render_sharded(comm, cam, rtc.make_params(num_samples=1, max_bounce=4, seed=1, accel=abi.RTC_ACCEL_BVH,
                                                    shard=abi.RTC_SHARD_STRIPS if args.strips else abi.RTC_SHARD_SAMPLES),
                         img=img if rank == 0 else None, to_host=True)
    rows = []
    shard_mode = abi.RTC_SHARD_STRIPS if args.strips else abi.RTC_SHARD_SAMPLES
    plan = [("bvh", s) for s in args.spp] + [("grid", s) for s in args.grid_spp]
    for accel_name, spp in plan:
        accel = abi.RTC_ACCEL_BVH if accel_name == "bvh" else abi.RTC_ACCEL_GRID
        prm = rtc.make_params(num_samples=spp, max_bounce=4, seed=1, accel=accel, shard=shard_mode, tile_rows=8)
        if spp <= 256:
            scene.render_sharded(comm, cam, prm, to_host=False)
        barrier()
        scene.stats()
        flush.zero_()
        torch.cuda.synchronize()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        st = scene.render_sharded(comm, cam, prm, to_host=False, want_stats=True)["stats"]
        e1.record()
        torch.cuda.synchronize()
        ms = allmax(e0.elapsed_time(e1))
        kern_ms = allmax(st["kernel_ms"])
        rays = allsum(float(st["rays"]))
        barrier()
        t0 = time.perf_counter()
        scene.render_sharded(comm, cam, prm, img=img if rank == 0 else None, to_host=True)
        barrier()
        e2e_s = allmax(time.perf_counter() - t0)
        if rank == 0:
            row = {"accel": accel_name, "n_gpus": world, "spp_total": spp, "Mrays_per_s": round(rays / (ms * 1e-3) / 1e6, 1),
                   "ms_per_frame": round(ms, 3), "render_kernels_ms_max_rank": round(kern_ms, 3),
                   "combine_and_gaps_ms": round(ms - kern_ms, 3),
                   "e2e_Mrays_per_s": round(rays / e2e_s / 1e6, 1), "e2e_ms": round(e2e_s * 1e3, 3), "rays": int(rays)}
            rows.append(row)
            print(json.dumps(row), file=sys.stderr, flush=True)
    if rank == 0:
        out = {"workload": "cfg5: terrain %d triangles, %dx%d, max_bounce 4, seed 1" % (spec.num_tris, cam.w, cam.h),
               "sharding": "strips of 8 rows + ncclSend/ncclRecv of RGB8" if args.strips else "sample ranges + ncclReduce(sum) of rgb accumulators (99.5 MB at 4K) to rank 0",
               "entry_point": "rtc_render_sharded (librtc.so drives NCCL itself)", "n_gpus": world, "scene_build_s": round(build_s, 2), "bvh": bvh,
               "timing": "one frame per row, CUDA events on the launch stream, max over ranks; e2e = host clock between barriers, host RGB8 frame out on rank 0",
               "rows": rows}
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        json.dump(out, open(args.out, "w"), indent=1)
    scene.close()
    if comm is not None:
        comm.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
