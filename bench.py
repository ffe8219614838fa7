#!/usr/bin/env python
"""bench.py — Mrays/s and ms/frame of the render path on N B200s, beside the CPU restatement.

A step = one frame of the workload (default: BASELINE.json configs[2], the path-traced 1920x1080
64 spp render; configs[0]/[1] are ~2 M-ray primary-visibility parity cases, not bench lines).
N > 1 (torchrun, one rank per GPU): the scene is replicated, every rank renders its own sample
range of the same frame (weak scaling: 64 spp per GPU), the float4 accumulators are summed to
rank 0 with one NCCL reduce and rank 0 resolves to RGB8 (SURVEY §8(e) partitioning B).

  value : rays traced by all ranks / device time (CUDA events on the launch stream, max over ranks),
          scene resident in HBM
  e2e   : the same through the blocking C-ABI call rtc_render() with host camera/params in and the
          host RGB8 frame out (D2H inside the timed region)
  --impl reference : the CPU restatement of the reference (oracle/, all host threads) on a bounded
          sample of the same frame
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (scene, scene kwargs, grid, height, spp, max_bounce)
    "cfg3_atrium_1080p_64spp": ("atrium", {}, (128, 128, 128), 1080, 64, 4),
    "cfg4_terrain5M_4k_256spp": ("terrain", {"num_tris": 5_000_000}, (128, 128, 128), 2160, 256, 4),
    "smoke_cornell_360p_8spp": ("cornell", {}, (32, 32, 32), 360, 8, 4),
}
DEFAULT_WORKLOAD = "cfg3_atrium_1080p_64spp"
SEED = 1
OVERRIDE = {}      # --spp / --height (profiling runs only; the default line uses the workload as named)
# dram__bytes_read.sum + dram__bytes_write.sum per k_trace launch from the committed `ncu --set full` capture
# (profiles/); None until a capture of the current kernel exists.
TRAFFIC = {}
try:
    TRAFFIC = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["per_launch_bytes"]
except (OSError, KeyError, ValueError):
    pass


def build_workload(name):
    from zig_raytracing_contest_b200 import rtc, scenes
    scene_name, kw, grid, height, spp, bounce = WORKLOADS[name]
    spp = OVERRIDE.get("spp") or spp
    height = OVERRIDE.get("height") or height
    grid = OVERRIDE.get("grid") or grid
    spec = scenes.SCENES[scene_name](**kw)
    t0 = time.time()
    geo = rtc.Geometry(spec.positions, spec.tri_data, grid)
    build_s = time.time() - t0
    cam = spec.camera(height=height)
    return spec, geo, cam, spp, bounce, build_s


def sample_bands(h, n_rows, bands=8):
    """n_rows rows of the frame as `bands` evenly spaced contiguous bands -> [(y0,y1)]."""
    bands = max(1, min(bands, n_rows))
    per = max(1, n_rows // bands)
    out = []
    for b in range(bands):
        y0 = int((b + 0.5) * h / bands - per / 2)
        y0 = max(0, min(h - per, y0))
        out.append((y0, y0 + per))
    return out


def cpu_render_sample(desc, cam, spp, bounce, n_rows, threads):
    """Times the oracle's Scene.render on a bounded sample of rows. Returns (rays, seconds, rows)."""
    from oracle import oracle
    from zig_raytracing_contest_b200 import rtc
    prm = rtc.make_params(num_samples=spp, max_bounce=bounce, seed=SEED)
    rays, secs, rows = 0, 0.0, 0
    for y0, y1 in sample_bands(cam.h, n_rows):
        t0 = time.perf_counter()
        out = oracle.render(desc, cam, prm, num_threads=threads, rows=(y0, y1), want_radiance=False, want_stats=True)
        secs += time.perf_counter() - t0
        rays += out["stats"]["rays"]
        rows += y1 - y0
    return rays, secs, rows


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            self.proc.terminate()
            self.t.join(timeout=2)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def run_reference(args, rank, world):
    """--impl reference: the CPU restatement on the host cores (rank 0 only)."""
    if rank != 0:
        return
    from zig_raytracing_contest_b200 import rtc
    spec, geo, cam, spp, bounce, _ = build_workload(args.workload)
    desc, keep = rtc.make_scene_desc(geo, spec.materials)
    threads = os.cpu_count() or 1
    n_rows = args.cpu_rows
    if n_rows <= 0:
        rays, secs, rows = cpu_render_sample(desc, cam, spp, bounce, 8, threads)       # calibrate ~ to 8 s per step
        n_rows = int(max(8, min(cam.h, rows * 8.0 / max(secs, 1e-3))))
    for _ in range(args.warmup):
        cpu_render_sample(desc, cam, spp, bounce, max(8, n_rows // 8), threads)
    tot_rays, tot_secs, rows = 0, 0.0, 0
    for _ in range(args.steps):
        rays, secs, rows = cpu_render_sample(desc, cam, spp, bounce, n_rows, threads)
        tot_rays += rays; tot_secs += secs
    mrays = tot_rays / tot_secs / 1e6
    frame_ms = tot_secs / args.steps * (cam.h / rows) * 1e3
    sample = "%d of %d rows (%d evenly spaced bands), all %d spp, per step" % (rows, cam.h, len(sample_bands(cam.h, n_rows)), spp)
    emit(json.dumps({
        "impl": "reference", "metric": "Mrays/s", "value": round(mrays, 4), "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(frame_ms, 2), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.workload, spec, geo, cam, spp, bounce, max(1, args.gpus)),
        "cpu_baseline": {"value": round(mrays, 4), "unit": "Mrays/s", "cores": threads, "kind": "port", "sample": sample,
                         "note": "C restatement of the Zig reference (no Zig toolchain on this box); ms_per_step is the sample extrapolated to the full frame"},
        "e2e": {"value": round(mrays, 4), "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def workload_config(name, spec, geo, cam, spp, bounce, world, tiles=0):
    if OVERRIDE.get("spp") or OVERRIDE.get("height") or OVERRIDE.get("grid"):
        name = "%s [override spp=%s height=%s grid=%s]" % (name, OVERRIDE.get("spp") or "-", OVERRIDE.get("height") or "-", OVERRIDE.get("grid") or "-")
    sharding = "none"
    if world > 1:
        sharding = ("strips of %d rows dealt round-robin, scene replicated, one NCCL gather of the RGB8 strips" % tiles) if tiles \
            else "sample ranges, scene replicated, NCCL reduce(sum) of float4 accumulators"
    return {"workload": name, "scene": spec.name, "source_triangles": spec.num_tris, "baked_triangles": geo.num_tris,
            "grid": [int(x) for x in geo.grid.resolution], "width": int(cam.w), "height": int(cam.h),
            "spp_per_gpu": spp, "spp_total": spp if tiles else spp * world, "max_bounce": bounce, "seed": SEED,
            "sharding": sharding,
            "l2": "flushed between timed steps by writing a 512 MiB buffer", "math": "exact (no FMA contraction)"}


def run_gpu(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from zig_raytracing_contest_b200 import abi, rtc, shard
    if rtc.device_count() < 1:
        raise SystemExit("bench.py: no sm_100 GPU visible and the render path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    spec, geo, cam, spp, bounce, build_s = build_workload(args.workload)
    t0 = time.perf_counter()
    scene = rtc.Scene(geo, spec.materials, device=local_rank)
    upload_s = time.perf_counter() - t0
    device_build = None
    if rank == 0:
        # the same scene built and baked ON the device from the triangle soup (not timed as part
        # of a step; reported beside the host-thread build it replaces)
        for _ in range(2):                          # the first call pays the lazy loading of the build kernels
            sb = rtc.Scene.from_soup(spec.positions, spec.tri_data, geo.grid.resolution, spec.materials, device=local_rank)
            bi = sb.build_info
            sb.close()
        assert (int(bi.num_tris), int(bi.empty_cells)) == (geo.num_tris, geo.empty_cells), "device build disagrees with the host build"
        device_build = {"kernels_ms": round(float(bi.device_ms), 3), "call_ms": round(float(bi.total_ms), 1),
                        "note": "rtc_scene_create_from_soup, second call: grid-build kernels (CUDA events) / whole call incl. H2D of the soup and textures"}
    P = cam.w * cam.h
    math = abi.RTC_MATH_FAST if args.math == "fast" else abi.RTC_MATH_EXACT
    accel = abi.RTC_ACCEL_BVH if args.accel == "bvh" else abi.RTC_ACCEL_GRID
    bvh_info = scene.build_bvh() if accel == abi.RTC_ACCEL_BVH else None
    tiles = args.tile_rows if (args.sharding == "tiles" and world > 1) else 0
    if tiles:
        total = spp                                 # strong scaling: the frame is fixed, its pixels are dealt over the ranks
        prm = rtc.make_params(num_samples=spp, max_bounce=bounce, seed=SEED, total_samples=total, math=math, accel=accel,
                              tile_rank=rank, tile_world=world, tile_rows=tiles)
    else:
        total = spp * world                         # weak scaling: the frame's sample count grows with N
        s_off, s_cnt = shard.sample_shard(rank, world, total)
        prm = rtc.make_params(num_samples=s_cnt, max_bounce=bounce, seed=SEED, sample_offset=s_off, total_samples=total, math=math, accel=accel)
    acc = torch.zeros(P, 4, device=dev)
    d_rgb = torch.zeros(P * 3, dtype=torch.uint8, device=dev)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream().cuda_stream

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(ev_k0=None, ev_k1=None):
        acc.zero_()
        if ev_k0: ev_k0.record()
        scene.render_accumulate(cam, prm, acc.data_ptr(), stream)
        if ev_k1: ev_k1.record()
        if tiles:
            rtc.resolve(acc.data_ptr(), P, total, d_rgb.data_ptr(), None, device=local_rank, stream_ptr=stream)
            frame = shard.gather_strips(d_rgb.view(cam.h, cam.w, 3), tiles, dst=0)      # the path's only collective
            if rank == 0 and frame.data_ptr() != d_rgb.data_ptr():
                d_rgb.copy_(frame.reshape(-1))
            return
        shard.reduce_accumulators(acc, dst=0)       # the path's only collective (no-op at N = 1)
        if rank == 0:
            rtc.resolve(acc.data_ptr(), P, total, d_rgb.data_ptr(), None, device=local_rank, stream_ptr=stream)

    for _ in range(args.warmup):
        step()
    barrier()
    scene.stats()                                   # clears the ray counters
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(args.steps)]
    barrier()
    with ClockSampler(local_rank) as clocks:
        for k in range(args.steps):
            flush.zero_()                           # evict the scene from the 126 MB L2 between steps
            ev[k][0].record()
            step(ev[k][1], ev[k][2])
            ev[k][3].record()
        barrier()
    step_ms = sum(e[0].elapsed_time(e[3]) for e in ev)
    kern_ms = sum(e[1].elapsed_time(e[2]) for e in ev) / args.steps
    st = scene.stats()
    t = torch.tensor([step_ms], device=dev, dtype=torch.float64)
    r = torch.tensor([float(st["rays"]), float(st["paths"]), float(st["launches"])], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(r, op=dist.ReduceOp.SUM)
    step_ms = float(t[0]) / args.steps
    trace_ms = st["trace_ms"] / args.steps                     # rank 0's k_trace launches, CUDA events per launch
    trace_launches = st["trace_launches"] / args.steps
    rays_per_step = float(r[0]) / args.steps
    launches = int(r[2]) + args.steps * (world if tiles else 1)               # + the resolve(s) per step
    value = rays_per_step / (step_ms * 1e-3) / 1e6

    # ---- e2e: the blocking public call, host buffers, D2H of the frame inside the timed region ----
    img = np.zeros((cam.h, cam.w, 3), np.uint8)
    host_acc = torch.zeros(P * 3, dtype=torch.uint8).pin_memory() if world > 1 else None

    def e2e_step():
        if world == 1:
            scene.render(cam, prm, img=img)
        else:
            step()
            if rank == 0:
                host_acc.copy_(d_rgb, non_blocking=True)
                torch.cuda.synchronize()
                img.reshape(-1)[:] = host_acc.numpy()
    e2e_step()
    barrier()
    flush.zero_(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    e2e_s = (time.perf_counter() - t0) / args.steps
    te = torch.tensor([e2e_s], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_s = float(te[0])
    e2e_val = rays_per_step / e2e_s / 1e6

    out = None
    if rank == 0:
        # instrumented launch of the same frame: counts behind the algorithmic bytes
        prm_s = rtc.make_params(num_samples=spp, max_bounce=bounce, seed=SEED, sample_offset=0, total_samples=total,
                                math=math, accel=accel, collect_stats=True, tile_rank=rank if tiles else 0, tile_world=world if tiles else 0,
                                tile_rows=tiles)
        acc.zero_(); torch.cuda.synchronize()
        scene.stats()
        scene.render_accumulate(cam, prm_s, acc.data_ptr(), stream)
        cs = scene.stats()
        algo_bytes = rtc.bytes_per_frame(cs, P)
        rec = 80 if bvh_info else 8                                         # BVH: 80-byte nodes; grid: 8-byte Cell records
        trav_bytes = rec * cs["cells_visited"] + 48 * cs["tris_tested"]    # k_trace's share: Cell / node + {v0,e1,e2} records
        if bvh_info:
            algo_bytes += (80 - 8) * cs["cells_visited"]
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        achieved = trav_bytes / (max(trace_ms, 1e-9) * 1e-3) / 1e9
        cfg = workload_config(args.workload, spec, geo, cam, spp, bounce, world, tiles)
        cfg["math"] = "exact (no FMA contraction)" if math == abi.RTC_MATH_EXACT else "fast (FMA contraction)"
        cfg["host_build_s"] = round(build_s, 3); cfg["scene_upload_s"] = round(upload_s, 3)
        cfg["device_build"] = device_build
        cfg["accel"] = "bvh8 (RTC_ACCEL_BVH, opt-in)" if bvh_info else "grid (the reference's structure)"
        if bvh_info:
            cfg["bvh"] = bvh_info
        cfg["per_ray"] = {"nodes" if bvh_info else "cells": round(cs["cells_visited"] / cs["rays"], 2), "tris": round(cs["tris_tested"] / cs["rays"], 2),
                          "bytes": round(algo_bytes / cs["rays"], 1), "rays_per_path": round(cs["rays"] / cs["paths"], 3)}
        out = {
            "metric": "Mrays/s", "value": round(value, 2), "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(step_ms, 3), "higher_is_better": True, "scaling": "strong" if tiles else "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg,
            "e2e": {"value": round(e2e_val, 2), "unit": "Mrays/s", "ms_per_frame": round(e2e_s * 1e3, 3),
                    "h2d_bytes_per_step": 56 + 48, "d2h_bytes_per_step": P * 3,
                    "note": "rtc_render(): host camera+params in, host RGB8 frame out; scene resident (uploaded once, like the reference's bakeInto before render)"},
            "gpu_launches": launches,
            "kernel_ms": {"render_kernels_per_step": round(kern_ms, 3), "k_trace_per_step": round(trace_ms, 3),
                          "k_trace_share": round(trace_ms / max(kern_ms, 1e-9), 4)},
            "roofline": {"bound": "hbm", "kernel": "k_trace_bvh" if bvh_info else "k_trace", "achieved": round(achieved, 2), "peak": peak, "unit": "GB/s",
                         "frac": round(achieved / peak, 5), "traffic": TRAFFIC.get(args.workload),
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 (of fallback)",
                         "launches_per_step": round(trace_launches, 1),
                         "avg_launch_ms": round(trace_ms / max(trace_launches, 1), 4),
                         "algorithmic_bytes_per_launch": int(trav_bytes / max(trace_launches, 1)),
                         "whole_frame": {"algorithmic_bytes": int(algo_bytes), "achieved": round(algo_bytes / (kern_ms * 1e-3) / 1e9, 2)},
                         "note": "k_trace (traversal + intersection), one launch per bounce per batch; bytes = 8 per visited cell + 48 per "
                                 "triangle test, counted by an instrumented run of the same frame; duration = CUDA events around every "
                                 "k_trace launch of the timed steps. whole_frame adds 224 B per shaded hit + 3 B per pixel over all render kernels. "
                                 "The scene is L2-resident, so this is algorithmic bytes against the HBM peak, not DRAM traffic"},
            "clocks": clocks.summary(),
        }
        if world == 1 and not args.no_cpu:
            desc = scene.desc
            threads = os.cpu_count() or 1
            rays, secs, rows = cpu_render_sample(desc, cam, spp, bounce, 8, threads)
            n_rows = int(max(8, min(cam.h, rows * args.cpu_seconds / max(secs, 1e-3))))
            rays, secs, rows = cpu_render_sample(desc, cam, spp, bounce, n_rows, threads)
            out["cpu_baseline"] = {"value": round(rays / secs / 1e6, 4), "unit": "Mrays/s", "cores": threads, "kind": "port",
                                   "sample": "%d of %d rows (%d evenly spaced bands), all %d spp; %.1f s" % (rows, cam.h, len(sample_bands(cam.h, n_rows)), spp, secs),
                                   "ms_per_frame_extrapolated": round(secs * cam.h / rows * 1e3, 1)}
        emit(json.dumps(out))
    scene.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


_RESULT_FD = None


def emit(line):
    """The ONE JSON line goes to the process' real stdout; everything else that writes to fd 1 while the bench runs
    (NCCL prints its version banner there) has been pointed at stderr."""
    with os.fdopen(os.dup(_RESULT_FD if _RESULT_FD is not None else 1), "w") as f:
        f.write(line + "\n")


def main():
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--math", default="exact", choices=["exact", "fast"])
    ap.add_argument("--accel", default="grid", choices=["grid", "bvh"],
                    help="grid: the reference's structure, bit-exact (default); bvh: RTC_ACCEL_BVH (same triangle test, enumerated near-ties)")
    ap.add_argument("--sharding", default="samples", choices=["samples", "tiles"],
                    help="N > 1: sample ranges + reduce (weak scaling, default) or row strips + gather (strong scaling, cfg4)")
    ap.add_argument("--tile-rows", type=int, default=8, help="--sharding tiles: rows per strip")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--cpu-rows", type=int, default=0, help="--impl reference: rows per step (0 = calibrate to ~8 s)")
    ap.add_argument("--spp", type=int, default=0, help="override the workload's spp (profiling only)")
    ap.add_argument("--height", type=int, default=0, help="override the workload's height (profiling only)")
    ap.add_argument("--grid", type=int, nargs=3, default=None,
                    help="override grid_resolution (a config.json key of the reference, src/main.zig:56-61)")
    args = ap.parse_args()
    OVERRIDE.update(spp=args.spp, height=args.height, grid=tuple(args.grid) if args.grid else None)
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_gpu(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
