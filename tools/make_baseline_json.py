#!/usr/bin/env python
"""Rebuilds BASELINE.json's `measured_b200` from the committed bench lines and sweep tables under profiles/ (round 2
block first, the round-1 runs kept under `round1_runs`), so that every number in it can be traced to a file.
usage: tools/make_baseline_json.py"""
import glob
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load(rel):
    try:
        return json.load(open(os.path.join(ROOT, rel)))
    except (OSError, ValueError):
        return None


def bench_row(rel, label):
    d = load(rel)
    if not d:
        return None
    cfg = d["config"]
    row = {"run": label, "file": rel, "n_gpus": d["n_gpus"], "scaling": d.get("scaling"), "accel": cfg.get("accel", "grid"),
           "Mrays_per_s": d["value"], "ms_per_frame": d["ms_per_step"], "e2e_Mrays_per_s": d["e2e"]["value"],
           "spp_total": cfg.get("spp_total"), "grid": cfg.get("grid")}
    rf = d.get("roofline", {})
    row["roofline_frac_algorithmic_bytes_vs_hbm"] = rf.get("frac")
    if "issue" in rf:
        row["issue"] = {k: v for k, v in rf["issue"].items() if k != "source"}
    if "dram" in rf:
        row["dram_frac_of_hbm_peak"] = rf["dram"]["frac"]
    if "cpu_baseline" in d:
        row["cpu_port_Mrays_per_s"] = d["cpu_baseline"]["value"]
        row["cpu_cores"] = d["cpu_baseline"]["cores"]
    if "grid_exact" in d:
        row["grid_exact_Mrays_per_s_same_run"] = d["grid_exact"]["value"]
    return row


def main():
    path = os.path.join(ROOT, "BASELINE.json")
    base = json.load(open(path))
    old = base.get("measured_b200", {})
    r1 = old.get("round1_runs") or old.get("runs", [])
    runs = []
    for rel, label in [
        ("profiles/r2z_bench.json", "cfg3 1 GPU, default bench line (RTC_ACCEL_BVH)"),
        ("profiles/r2z_bench_grid.json", "cfg3 1 GPU, --accel grid (the reference's structure, bit-exact)"),
        ("profiles/r2a_bench_fast.json", "cfg3 1 GPU, grid, --math fast (FMA contraction; round-2 start kernel)"),
        ("profiles/r2z_cfg4_bvh.json", "cfg4 scene 1 GPU at 8 spp, BVH"),
        ("profiles/r2a_cfg4_128x128x128.json", "cfg4 scene 1 GPU at 8 spp, grid 128^3 (the reference's config.json)"),
        ("profiles/r2a_cfg4_512x128x512.json", "cfg4 scene 1 GPU at 8 spp, grid with tuned grid_resolution 512x128x512 (a config.json key)"),
        ("profiles/r2a_cfg3_256x96x128.json", "cfg3 1 GPU at 16 spp, grid with tuned grid_resolution 256x96x128"),
        ("profiles/r2z_bench_n2.json", "cfg3 2 GPUs strong, strips over rtc_render_sharded (NCCL send/recv), final kernels"),
        ("profiles/r2z_bench_n4.json", "cfg3 4 GPUs strong, strips over rtc_render_sharded, final kernels"),
        ("profiles/r2w_bench_n2_samples.json", "cfg3 2 GPUs strong, sample ranges over rtc_render_sharded (ncclReduce of rgb)"),
        ("profiles/r2w_inproc2.json", "cfg3 2 GPUs strong, ONE process: rtc_render(num_gpus=2), host clock incl. D2H"),
        ("profiles/r2_n8_cfg3_strips.json", "cfg3 8 GPUs strong, strips"),
        ("profiles/r2_n8_cfg3_samples.json", "cfg3 8 GPUs strong, sample ranges"),
        ("profiles/r2_n8_cfg3_inproc.json", "cfg3 8 GPUs strong, ONE process: rtc_render(num_gpus=8), strips, host clock incl. D2H + host copy"),
        ("profiles/r2_n8_cfg3_inproc_samples.json", "cfg3 8 GPUs strong, ONE process: rtc_render(num_gpus=8), sample ranges combined over NVLink peer loads"),
        ("profiles/r2n_inproc4.json", "cfg3 4 GPUs strong, ONE process: rtc_render(num_gpus=4), strips"),
        ("profiles/r2_n8_cfg4_strips.json", "cfg4 (5 M triangles, 4K, 256 spp) 8 GPUs strong, strips, BVH"),
    ]:
        row = bench_row(rel, label)
        if row:
            runs.append(row)
    sweep = []
    # the latest sweep (r2u: final kernels and scratch policy); the earlier one (r2h / r2i) stays under profiles/ as history
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r2u_cfg5_n*.json"))) or sorted(glob.glob(os.path.join(ROOT, "profiles", "r2[hi]_cfg5_n*.json")))
    for rel in files:
        d = load(os.path.relpath(rel, ROOT))
        if not d:
            continue
        for r in d["rows"]:
            sweep.append({"file": os.path.relpath(rel, ROOT), "accel": r["accel"], "n_gpus": r["n_gpus"], "spp_total": r["spp_total"],
                          "Mrays_per_s": r["Mrays_per_s"], "ms_per_frame": r["ms_per_frame"],
                          "combine_and_gaps_ms": r["combine_and_gaps_ms"], "e2e_Mrays_per_s": r["e2e_Mrays_per_s"], "e2e_ms": r["e2e_ms"]})
    sweep.sort(key=lambda r: (r["accel"], r["spp_total"], r["n_gpus"]))
    # strong-scaling efficiency of the sweep against its own N = 1 row
    one = {(r["accel"], r["spp_total"]): r["ms_per_frame"] for r in sweep if r["n_gpus"] == 1}
    for r in sweep:
        t1 = one.get((r["accel"], r["spp_total"]))
        if t1:
            r["speedup_vs_1gpu"] = round(t1 / r["ms_per_frame"], 3)
            r["efficiency"] = round(t1 / r["ms_per_frame"] / r["n_gpus"], 4)
    base["measured_b200"] = {
        "note": "measured by bench.py / tools/cfg5_sweep.py on B200 boxes of this pool; full JSON lines under profiles/. Round 2 first; "
                "roofline_frac is ALGORITHMIC bytes against the measured HBM peak (6448.7 GB/s) as SURVEY 8(d) defines it — the traversal "
                "kernels are instruction-issue bound (issue: issue-active, lanes per instruction, useful lane-issue, L1/L2 throughput from "
                "the committed ncu captures, profiles/counters.json), DRAM runs at 0.03-0.05 of its peak",
        "runs": runs,
        "cfg5_spp_sweep": {"workload": "5 M-triangle terrain, 3840x2160, max_bounce 4; one fixed frame sharded over the GPUs by sample ranges, "
                                       "rgb accumulators combined by ncclReduce inside librtc.so (rtc_render_sharded)",
                           "what_bends_the_curve": "combine_and_gaps_ms (ncclReduce of 99.5 MB of rgb sums + resolve + launch gaps) is 0.5-0.8 ms at every N and "
                                                   "every spp: 6.7 % of the 9.3 ms frame at 16 spp on 8 GPUs, 0.4 % at 256 spp; the rest of the loss at 16 spp on 8 GPUs is the "
                                                   "render itself at 2 spp per rank (kernel tails and launch gaps of 11 launches over 8.7 ms); the e2e column adds the 24.9 MB D2H "
                                                   "+ host copy on rank 0 (3 ms)",
                           "rows": sweep},
        "counters": load("profiles/counters.json"),
        "round1_runs": r1,
        "callers": old.get("callers"),
    }
    json.dump(base, open(path, "w"), indent=2)
    print("runs:", len(runs), "sweep rows:", len(sweep))


if __name__ == "__main__":
    main()
