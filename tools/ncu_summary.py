#!/usr/bin/env python
"""One table per ncu report: the counters DESIGN.md and bench.py's roofline quote, one column per launch.
usage: tools/ncu_summary.py REPORT.ncu-rep [title...]   (reads with `ncu -i ... --page raw --csv`)"""
import csv
import io
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
]


def main():
    rep = sys.argv[1]
    if len(sys.argv) > 2:
        print(" ".join(sys.argv[2:]))
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    name_i = hdr.index("Kernel Name")
    print("%-92s %s" % ("metric", " ".join("%14s" % ("%s#%d" % (d[name_i][:8], k)) for k, d in enumerate(data))))
    for w in WANT:
        if w not in hdr:
            continue
        i = hdr.index(w)
        vals = []
        for d in data:
            try:
                vals.append("%14.6g" % float(d[i].replace(",", "")))
            except ValueError:
                vals.append("%14s" % d[i])
        print("%-92s %s" % ("%s [%s]" % (w, units[i]), " ".join(vals)))


if __name__ == "__main__":
    main()
