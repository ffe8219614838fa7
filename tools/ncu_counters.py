#!/usr/bin/env python
"""profiles/counters.json: per (workload, accel) the counters of the traversal kernel that bench.py attaches to its
roofline object — read from the committed tools/ncu_summary.py tables of `ncu --set full` captures of bench.py's own
command. usage: tools/ncu_counters.py KEY=profiles/summary.txt[:column-substring] ...   (KEY = workload/accel)"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WANT = {
    "dram__bytes_read.sum": "dram_read", "dram__bytes_write.sum": "dram_write",
    "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active",
    "smsp__thread_inst_executed_per_inst_executed.ratio": "lanes_per_inst",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed": "l1_throughput",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed": "l2_throughput",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active",
    "launch__registers_per_thread": "registers",
    "gpu__time_duration.sum": "ncu_ms",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active": "alu_pipe",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active": "fma_pipe",
}
UNIT = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}


def parse(path, pick):
    lines = open(os.path.join(ROOT, path)).read().splitlines()
    hdr = next(l for l in lines if l.startswith("metric"))
    # column names are "<8 chars>#k"; split on the '#k' markers
    names = hdr[len("metric"):].split("#")
    cols = []
    for i, nm in enumerate(names[:-1]):
        cols.append(nm.strip().split()[-1] if nm.strip() else "")
    keep = [i for i, nm in enumerate(names[:-1]) if pick in nm]
    out = {}
    for l in lines:
        for metric, key in WANT.items():
            if l.startswith(metric + " ["):
                unit = l[l.index("[") + 1:l.index("]")]
                vals = [float(x) for x in l[l.index("]") + 1:].split()]
                vals = [vals[i] for i in keep]
                out[key] = [v * UNIT.get(unit, 1.0) if key.startswith("dram") else v for v in vals]
    return out


def main():
    dst = os.path.join(ROOT, "profiles", "counters.json")
    try:
        table = json.load(open(dst))
    except (OSError, ValueError):
        table = {}
    for arg in sys.argv[1:]:
        key, src = arg.split("=", 1)
        pick = "k_t"
        if ":" in src:
            src, pick = src.split(":", 1)
        c = parse(src, pick)
        n = len(c["issue_active"])
        avg = lambda k: sum(c[k]) / len(c[k])
        issue, lanes = avg("issue_active") / 100.0, avg("lanes_per_inst")
        table[key] = {
            "source": "%s (ncu --set full of bench.py's command: %d launches of the traversal kernel, one per bounce of one batch)" % (src, n),
            "dram_bytes_per_launch": int(avg("dram_read") + avg("dram_write")),
            "issue_active": round(issue, 4), "lanes_per_inst": round(lanes, 2), "useful_lane_issue": round(issue * lanes / 32.0, 4),
            "l1_throughput": round(avg("l1_throughput") / 100.0, 4), "l2_throughput": round(avg("l2_throughput") / 100.0, 4),
            "warps_active": round(avg("warps_active") / 100.0, 4), "registers": int(c["registers"][0]),
            "alu_pipe": round(avg("alu_pipe") / 100.0, 4), "fma_pipe": round(avg("fma_pipe") / 100.0, 4),
            "per_launch": {"issue_active_pct": c["issue_active"], "lanes_per_inst": c["lanes_per_inst"], "ncu_ms": c["ncu_ms"]},
        }
    json.dump(table, open(dst, "w"), indent=1)
    print(json.dumps(table, indent=1)[:3000])


if __name__ == "__main__":
    main()
