#!/usr/bin/env python
"""bench.py — Mrays/s and ms/frame of the render path on N B200s, beside the CPU restatement.

A step = one frame of the workload (default: BASELINE.json configs[2], the path-traced 1920x1080 64 spp render;
configs[0]/[1] are ~2 M-ray primary-visibility parity cases, not bench lines).

N > 1 (torchrun, one rank per GPU): ONE FIXED FRAME is sharded over the ranks (strong scaling): the scene is
replicated, every rank calls the library's collective entry point rtc_render_sharded() — row strips dealt round-robin
+ ncclSend/ncclRecv of the finished RGB8 strips (default), or sample ranges + ncclReduce(sum) of the rgb accumulators
(--sharding samples). NCCL is driven by librtc.so itself; torch.distributed only carries the communicator id, the
barriers and the max-over-ranks of the timings. `--scaling weak` keeps the per-GPU work fixed instead (spp per GPU).

  value : rays traced by all ranks / device time of the steps (CUDA events on the stream the library launches on,
          max over ranks), scene resident in HBM, frame left in device memory of rank 0
  e2e   : the same frame through the blocking C-ABI call with host camera/params in and the host RGB8 frame out
          (rtc_render at N = 1, rtc_render_sharded at N > 1; D2H inside the timed region)
  --impl reference : the CPU restatement of the reference (oracle/, all host threads) on a bounded sample of the
          same frame; maps oracle/liboracle.so only
  --inprocess N : (no torchrun) the single-process multi-GPU path of the C ABI, rtc_render with num_gpus = N
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
OVERRIDE = {}      # --spp / --height / --grid (profiling and sweeps; the default line uses the workload as named)


def load_json(*path):
    try:
        return json.load(open(os.path.join(ROOT, *path)))
    except (OSError, ValueError):
        return {}


# Counters of the dominant kernel read from the committed `ncu --set full` capture of this same command
# (profiles/counters.json, written by tools/ncu_counters.py): DRAM bytes per launch, issue-active, lanes per
# instruction, L1 / L2 throughput. Keyed "<workload>/<accel>".
COUNTERS = load_json("profiles", "counters.json")


def scene_spec(name):
    from zig_raytracing_contest_b200 import scenes
    scene_name, kw, grid, height, spp, bounce = WORKLOADS[name]
    spp = OVERRIDE.get("spp") or spp
    height = OVERRIDE.get("height") or height
    grid = OVERRIDE.get("grid") or grid
    spec = scenes.SCENES[scene_name](**kw)
    return spec, grid, spec.camera(height=height), spp, bounce


def build_workload(name):
    """The product's own stage2 (rtc_geometry_build on the host threads) — GPU arm only."""
    from zig_raytracing_contest_b200 import rtc
    spec, grid, cam, spp, bounce = scene_spec(name)
    t0 = time.time()
    geo = rtc.Geometry(spec.positions, spec.tri_data, grid)
    return spec, geo, cam, spp, bounce, time.time() - t0


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
    from zig_raytracing_contest_b200 import rtc       # make_params: a ctypes struct, no library is loaded
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


def workload_config(name, spec, grid, baked, cam, spp_total, bounce, world, scaling, sharding, tile_rows, accel, math):
    if OVERRIDE.get("spp") or OVERRIDE.get("height") or OVERRIDE.get("grid"):
        name = "%s [override spp=%s height=%s grid=%s]" % (name, OVERRIDE.get("spp") or "-", OVERRIDE.get("height") or "-", OVERRIDE.get("grid") or "-")
    how = "none"
    if world > 1:
        how = ("strips of %d rows dealt round-robin, scene replicated; every rank resolves its strips, rank 0 receives them (ncclSend/ncclRecv inside librtc.so)" % tile_rows) \
            if sharding == "strips" else "sample ranges, scene replicated; one ncclReduce(sum) of the rgb accumulators to rank 0 (inside librtc.so)"
    return {"workload": name, "scene": spec.name, "source_triangles": spec.num_tris, "baked_triangles": baked,
            "grid": [int(x) for x in grid], "width": int(cam.w), "height": int(cam.h),
            "spp_total": spp_total, "max_bounce": bounce, "seed": SEED, "scaling": scaling, "sharding": how,
            "accel": accel, "math": math,
            "l2": "flushed between timed steps by writing a 512 MiB buffer"}


def run_reference(args, rank, world):
    """--impl reference: the CPU restatement on the host cores (rank 0 only). Builds the scene with the ORACLE's own
    stage2 (orc_geometry_build): nothing of the product is loaded in this process."""
    if rank != 0:
        return
    from oracle import oracle
    from zig_raytracing_contest_b200 import rtc       # Python-side struct helpers only; rtc.lib() is never called
    spec, grid, cam, spp, bounce = scene_spec(args.workload)
    geo = oracle.Geometry(spec.positions, spec.tri_data, grid)
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
    step_ms = tot_secs / args.steps * 1e3
    bands = len(sample_bands(cam.h, n_rows))
    sample = "%d of %d rows (%d evenly spaced bands), all %d spp, per step" % (rows, cam.h, bands, spp)
    assert "librtc" not in open("/proc/self/maps").read(), "the reference arm must not map the product"
    emit(json.dumps({
        "impl": "reference", "metric": "Mrays/s", "value": round(mrays, 4), "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(step_ms, 2), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.workload, spec, grid, int(geo.tri_pos.shape[0]), cam, spp, bounce, 1, "strong", "none", 0,
                                  "grid (the reference's structure)", "exact (no FMA contraction)"),
        "cpu_baseline": {"value": round(mrays, 4), "unit": "Mrays/s", "cores": threads, "kind": "port", "sample": sample,
                         "rows_timed": rows, "rows_total": int(cam.h),
                         "frame_ms_extrapolated": round(step_ms * cam.h / rows, 1),
                         "note": "C restatement of the Zig reference (no Zig toolchain on this box); a step is the row sample, ms_per_step its measured time"},
        "e2e": {"value": round(mrays, 4), "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def roofline(name, accel_key, cs, trace_ms, trace_launches, kern_ms, algo_bytes, bvh):
    """The contract's object (algorithmic bytes of the dominant kernel / its CUDA-event duration against the measured
    HBM peak) plus what ncu says the kernel is actually limited by."""
    peaks = load_json("MEASURED_PEAKS.json")
    peak = float(peaks.get("hbm_gbs", 6650.0))
    rec = 80 if bvh else 8                                             # BVH: 80-byte nodes; grid: 8-byte Cell records
    trav_bytes = rec * cs["cells_visited"] + 48 * cs["tris_tested"]    # the traversal kernel's share
    achieved = trav_bytes / (max(trace_ms, 1e-9) * 1e-3) / 1e9
    avg_ms = trace_ms / max(trace_launches, 1)
    out = {"bound": "hbm", "kernel": "k_trace_bvh" if bvh else "k_trace", "achieved": round(achieved, 2), "peak": peak, "unit": "GB/s",
           "frac": round(achieved / peak, 5), "traffic": None,
           "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 (of fallback)",
           "launches_per_step": round(trace_launches, 1), "avg_launch_ms": round(avg_ms, 4),
           "algorithmic_bytes_per_launch": int(trav_bytes / max(trace_launches, 1)),
           "whole_frame": {"algorithmic_bytes": int(algo_bytes), "achieved": round(algo_bytes / (max(kern_ms, 1e-9) * 1e-3) / 1e9, 2)},
           "limiter": "issue",
           "note": "achieved = ALGORITHMIC bytes (%d B per %s + 48 B per triangle test, counted by an instrumented run of the same frame) / "
                   "CUDA-event time of the traversal launches. The scene is L2-resident: this is the rate at which HBM would have to feed the "
                   "algorithm, NOT DRAM traffic (see dram) and not the limiter. The kernel is instruction-issue bound (see issue): "
                   "useful_lane_issue = issue-active x active lanes / 32 is the fraction of the SMs' lane-issue slots doing work."
                   % (rec, "visited node" if bvh else "visited cell")}
    c = COUNTERS.get("%s/%s" % (name, accel_key))
    if c:
        ncu_ms = c.get("per_launch", {}).get("ncu_ms") or []
        if c.get("dram_bytes_per_launch") and ncu_ms:
            cap_ms = sum(ncu_ms) / len(ncu_ms)
            # the capture's bytes are scaled by the ratio of the launch durations to give bytes per TIMED launch: ~1 for a capture
            # of one-batch launches (BVH, profiles/r2z_*), ~6.4 where the captured launches were the 6 GiB-stage batches (grid)
            out["traffic"] = int(c["dram_bytes_per_launch"] * avg_ms / cap_ms)
            gbs = c["dram_bytes_per_launch"] / (cap_ms * 1e-3) / 1e9
            out["dram"] = {"achieved": round(gbs, 1), "frac": round(gbs / peak, 4), "unit": "GB/s",
                           "note": "measured dram__bytes_read+write of the captured launches / their own duration under ncu (%.2f ms per launch); "
                                   "traffic = those bytes scaled to the duration of a timed launch (x%.2f)" % (cap_ms, avg_ms / cap_ms)}
        out["issue"] = {k: c[k] for k in ("issue_active", "lanes_per_inst", "useful_lane_issue", "l1_throughput", "l2_throughput",
                                          "warps_active", "registers") if k in c}
        out["issue"]["source"] = c.get("source")
    return out


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
    if rank == 0 and not args.lean:
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
    bvh = args.accel == "bvh"
    accel = abi.RTC_ACCEL_BVH if bvh else abi.RTC_ACCEL_GRID
    bvh_info = scene.build_bvh() if bvh else None
    inproc = args.inprocess if world == 1 else 0
    nshards = inproc if inproc > 1 else world
    weak = args.scaling == "weak" and nshards > 1
    sharding = "samples" if weak else args.sharding
    total = spp * nshards if weak else spp
    shard_mode = abi.RTC_SHARD_SAMPLES if sharding == "samples" else abi.RTC_SHARD_STRIPS
    prm = rtc.make_params(num_samples=total, max_bounce=bounce, seed=SEED, math=math, accel=accel, shard=shard_mode,
                          tile_rows=args.tile_rows, num_gpus=inproc if inproc > 1 else 0)
    comm = None
    if world > 1:
        # the library's own NCCL communicator; torch.distributed only hands the id around
        comm = rtc.Comm(shard.broadcast_comm_id(rtc.comm_unique_id), rank, world, device=local_rank)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    # the library launches on the stream the events are recorded on: a stream of this process, handed to the scene
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    scene.set_stream(stream.cuda_stream)
    img = np.zeros((cam.h, cam.w, 3), np.uint8)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        if inproc > 1:
            scene.render(cam, prm, img=img)         # one process, N devices: the blocking call is the only entry point
        else:
            scene.render_sharded(comm, cam, prm, to_host=False)

    for _ in range(args.warmup):
        step()
    barrier()
    scene.stats()                                   # clears the ray counters
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(args.steps)]
    wall = 0.0
    barrier()
    with ClockSampler(local_rank) as clocks:
        for k in range(args.steps):
            flush.zero_()                           # evict the scene from the 126 MB L2 between steps
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            ev[k][0].record()
            step()                                  # blocking: every kernel and the collective of this step are done
            ev[k][1].record()
            torch.cuda.synchronize()
            wall += time.perf_counter() - t0
        barrier()
    step_ms = sum(e[0].elapsed_time(e[1]) for e in ev)
    if inproc > 1:
        step_ms = wall * 1e3                        # devices other than this rank's are invisible to these events: host clock
    st = scene.stats()
    t = torch.tensor([step_ms], device=dev, dtype=torch.float64)
    r = torch.tensor([float(st["rays"]), float(st["paths"]), float(st["launches"])], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(r, op=dist.ReduceOp.SUM)
    step_ms = float(t[0]) / args.steps
    trace_ms = st["trace_ms"] / args.steps                     # rank 0's traversal launches, CUDA events per launch
    trace_launches = st["trace_launches"] / args.steps
    # one more frame with the per-call brackets: render kernels alone, and the frame up to the end of the collective
    one = scene.render(cam, prm, img=img, want_stats=True, instrumented=False)["stats"] if inproc > 1 else \
        scene.render_sharded(comm, cam, prm, to_host=False, want_stats=True)["stats"]
    if inproc > 1:
        trace_ms, trace_launches = trace_ms / inproc, trace_launches / inproc      # summed over the devices: per device
    rays_per_step = float(r[0]) / args.steps
    launches = int(round(float(r[2]) / args.steps))
    value = rays_per_step / (step_ms * 1e-3) / 1e6
    kern_ms = one["kernel_ms"]

    # ---- e2e: the blocking public call, host buffers, D2H of the frame inside the timed region ----
    def e2e_step():
        if world == 1:
            scene.render(cam, prm, img=img)         # rtc_render: the reference-facing call (src/main.zig:126)
        else:
            scene.render_sharded(comm, cam, prm, img=img if rank == 0 else None, to_host=True)
    e2e_step()
    barrier()
    flush.zero_(); torch.cuda.synchronize()
    barrier()
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

    if rank == 0:
        # instrumented run of the same frame on this device alone: counts behind the algorithmic bytes
        prm_s = rtc.make_params(num_samples=spp, max_bounce=bounce, seed=SEED, math=math, accel=accel, collect_stats=True)
        scene.stats()
        cs = scene.render_sharded(None, cam, prm_s, to_host=False, want_stats=True)["stats"]
        algo_bytes = rtc.bytes_per_frame(cs, P) + ((80 - 8) * cs["cells_visited"] if bvh else 0)
        math_s = "exact (no FMA contraction)" if math == abi.RTC_MATH_EXACT else "fast (FMA contraction)"
        accel_s = "bvh8 (RTC_ACCEL_BVH: the reference's triangle test, 8-wide quantised BVH)" if bvh else "grid (the reference's structure, bit-exact)"
        cfg = workload_config(args.workload, spec, geo.grid.resolution, geo.num_tris, cam, total, bounce, nshards,
                              "weak" if weak else "strong", sharding, args.tile_rows, accel_s, math_s)
        cfg["host_build_s"] = round(build_s, 3); cfg["scene_upload_s"] = round(upload_s, 3)
        cfg["device_build"] = device_build
        if bvh_info:
            cfg["bvh"] = bvh_info
        if inproc > 1:
            cfg["process_model"] = "one process, %d devices: rtc_render(num_gpus=%d); timed on the host clock around the blocking call" % (inproc, inproc)
        cfg["per_ray"] = {"nodes" if bvh else "cells": round(cs["cells_visited"] / cs["rays"], 2), "tris": round(cs["tris_tested"] / cs["rays"], 2),
                          "bytes": round(algo_bytes / cs["rays"], 1), "rays_per_path": round(cs["rays"] / cs["paths"], 3)}
        # rank 0 traces 1/N of a strong-scaled frame: scale its kernel times to the whole frame's counts
        share = nshards if (nshards > 1 and not weak) else 1
        out = {
            "metric": "Mrays/s", "value": round(value, 2), "unit": "Mrays/s", "n_gpus": nshards, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(step_ms, 3), "higher_is_better": True, "scaling": "weak" if weak else "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg,
            "e2e": {"value": round(e2e_val, 2), "unit": "Mrays/s", "ms_per_frame": round(e2e_s * 1e3, 3),
                    "h2d_bytes_per_step": 56 + 56, "d2h_bytes_per_step": P * 3,
                    "note": "blocking C-ABI call: host camera+params in, host RGB8 frame out; scene resident (uploaded once, like the reference's bakeInto before render)"},
            "gpu_launches": launches,
            "kernel_ms": {"render_kernels_per_step": round(kern_ms, 3), "frame_to_end_of_combine": round(one["frame_ms"], 3),
                          "k_trace_per_step": round(trace_ms, 3), "k_trace_share": round(trace_ms / max(kern_ms, 1e-9), 4)},
            "roofline": roofline(args.workload, "bvh" if bvh else "grid", cs, trace_ms * share, trace_launches, kern_ms * share, algo_bytes, bvh),
            "clocks": clocks.summary(),
        }
        if world == 1 and inproc <= 1 and not args.no_cpu:
            desc = scene.desc
            threads = os.cpu_count() or 1
            rays, secs, rows = cpu_render_sample(desc, cam, spp, bounce, 8, threads)
            n_rows = int(max(8, min(cam.h, rows * args.cpu_seconds / max(secs, 1e-3))))
            rays, secs, rows = cpu_render_sample(desc, cam, spp, bounce, n_rows, threads)
            out["cpu_baseline"] = {"value": round(rays / secs / 1e6, 4), "unit": "Mrays/s", "cores": threads, "kind": "port",
                                   "sample": "%d of %d rows (%d evenly spaced bands), all %d spp; %.1f s" % (rows, cam.h, len(sample_bands(cam.h, n_rows)), spp, secs),
                                   "rows_timed": rows, "rows_total": int(cam.h),
                                   "ms_per_frame_extrapolated": round(secs * cam.h / rows * 1e3, 1)}
        if world == 1 and inproc <= 1 and bvh and not args.lean:
            # the reference's own structure (the bit-exact default of the library) on the same frame, same run
            pg = rtc.make_params(num_samples=spp, max_bounce=bounce, seed=SEED, math=math, accel=abi.RTC_ACCEL_GRID)
            for _ in range(2):
                scene.render_sharded(None, cam, pg, to_host=False)
            flush.zero_(); torch.cuda.synchronize()
            g = scene.render_sharded(None, cam, pg, to_host=False, want_stats=True)["stats"]
            out["grid_exact"] = {"value": round(g["rays"] / (g["frame_ms"] * 1e-3) / 1e6, 2), "unit": "Mrays/s", "ms_per_step": round(g["frame_ms"], 3),
                                 "note": "RTC_ACCEL_GRID (the reference's uniform grid, results bit-identical to the oracle), one frame after 2 warm-ups"}
        emit(json.dumps(out))
    scene.close()
    if comm is not None:
        comm.close()
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
    ap.add_argument("--accel", default="bvh", choices=["grid", "bvh"],
                    help="bvh (default): RTC_ACCEL_BVH, the reference's triangle test under an 8-wide BVH (hit ids equal to the oracle's except "
                         "enumerated near-ties; tests/test_gpu_bvh.py); grid: the reference's structure, bit-exact")
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"],
                    help="N > 1: strong = one fixed frame sharded over the GPUs (default); weak = the workload's spp PER GPU (sample ranges)")
    ap.add_argument("--sharding", default="strips", choices=["strips", "samples", "tiles"],
                    help="N > 1: row strips + send/recv of RGB8 strips (default), or sample ranges + ncclReduce of the accumulators")
    ap.add_argument("--tile-rows", type=int, default=8, help="--sharding strips: rows per strip")
    ap.add_argument("--inprocess", type=int, default=0, help="without torchrun: rtc_render(num_gpus=N) from this one process")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--lean", action="store_true", help="skip the side measurements (device build, grid_exact): sweeps")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--cpu-rows", type=int, default=0, help="--impl reference: rows per step (0 = calibrate to ~8 s)")
    ap.add_argument("--spp", type=int, default=0, help="override the workload's spp (profiling / sweeps)")
    ap.add_argument("--height", type=int, default=0, help="override the workload's height (profiling only)")
    ap.add_argument("--grid", type=int, nargs=3, default=None,
                    help="override grid_resolution (a config.json key of the reference, src/main.zig:56-61)")
    args = ap.parse_args()
    if args.sharding == "tiles":
        args.sharding = "strips"
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
