#!/usr/bin/env python
"""Fills the measured tables of DESIGN.md (§6, marker @@TABLE@@ or the block between the table markers) and BASELINE.md
(§5) from BASELINE.json's measured_b200 block (itself generated from profiles/ by tools/make_baseline_json.py)."""
import json
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BEGIN, END = "<!-- measured:begin -->", "<!-- measured:end -->"


def main():
    m = json.load(open(os.path.join(ROOT, "BASELINE.json")))["measured_b200"]
    lines = ["| run | GPUs | Mrays/s (e2e) | ms/frame | evidence |", "|---|---|---|---|---|"]
    for r in m["runs"]:
        extra = []
        if r.get("grid_exact_Mrays_per_s_same_run"):
            extra.append("grid in the same run: %.0f" % r["grid_exact_Mrays_per_s_same_run"])
        if r.get("cpu_port_Mrays_per_s"):
            extra.append("CPU restatement on %d host cores: %.1f" % (r["cpu_cores"], r["cpu_port_Mrays_per_s"]))
        if r.get("issue"):
            extra.append("useful lane-issue %.2f" % r["issue"]["useful_lane_issue"])
        lines.append("| %s | %d | %.0f (%.0f) | %.1f | `%s`%s |" % (r["run"], r["n_gpus"], r["Mrays_per_s"], r["e2e_Mrays_per_s"], r["ms_per_frame"],
                                                                  r["file"], ("; " + "; ".join(extra)) if extra else ""))
    sw = m["cfg5_spp_sweep"]["rows"]
    lines += ["", "cfg5 — spp sweep at 3840×2160 on the 5 M-triangle scene, ONE fixed frame sharded by sample ranges, rgb accumulators combined by `ncclReduce` inside "
              "`librtc.so` (`tools/cfg5_sweep.py`; ms per frame on the device, max over ranks; in brackets the strong-scaling efficiency against the same row at 1 GPU; "
              "`profiles/r2u_cfg5_n{1,2,4,8}.json`; the same sweep with the kernels of three commits earlier: `r2h_cfg5_n1.json`, `r2i_cfg5_n{2,4,8}.json`):", "",
              "| path | spp | 1 GPU | 2 GPUs | 4 GPUs | 8 GPUs | 8 GPUs: Mrays/s | combine + gaps at 8 | e2e at 8 (host frame out) |", "|---|---|---|---|---|---|---|---|---|"]
    for accel in ("bvh", "grid"):
        for spp in sorted({r["spp_total"] for r in sw if r["accel"] == accel}):
            cells = []
            r8 = None
            for n in (1, 2, 4, 8):
                r = next((x for x in sw if x["accel"] == accel and x["spp_total"] == spp and x["n_gpus"] == n), None)
                if not r:
                    cells.append("—")
                    continue
                cells.append("%.1f ms" % r["ms_per_frame"] + (" (%.3f)" % r["efficiency"] if n > 1 and "efficiency" in r else ""))
                if n == 8:
                    r8 = r
            lines.append("| %s | %d | %s | %s | %s | %s | %s | %s | %s |" % (
                accel, spp, cells[0], cells[1], cells[2], cells[3], "%.0f" % r8["Mrays_per_s"] if r8 else "—",
                "%.2f ms" % r8["combine_and_gaps_ms"] if r8 else "—", "%.1f ms" % r8["e2e_ms"] if r8 else "—"))
    lines += ["", m["cfg5_spp_sweep"]["what_bends_the_curve"] + "."]
    block = BEGIN + "\n" + "\n".join(lines) + "\n" + END
    for name, marker in (("DESIGN.md", "@@TABLE@@"), ("BASELINE.md", "| config | GPUs | path | Mrays/s | ms/frame | note |\n|---|---|---|---|---|---|\n@@ROWS@@")):
        p = os.path.join(ROOT, name)
        s = open(p).read()
        if BEGIN in s:
            s = re.sub(re.escape(BEGIN) + ".*?" + re.escape(END), lambda _: block, s, flags=re.S)
        elif marker in s:
            s = s.replace(marker, block)
        else:
            raise SystemExit("no table marker in " + name)
        open(p, "w").write(s)
    print("\n".join(lines))


if __name__ == "__main__":
    main()
