#!/usr/bin/env python
"""Per-source-line view of an ncu report: joins `ncu --page source --csv` (per SASS instruction:
stall samples, executed instructions, active threads) with nvdisasm's line info of the same cubin.

usage: tools/ncu_lines.py REPORT.ncu-rep [--so zig_raytracing_contest_b200/csrc/build] [--top 40]
"""
import argparse
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile


def sass_lines(so, kernel_mangled_substr):
    """[(sass text, file, line, inlined_chain)] for the first function whose name contains the substring."""
    # Both arithmetic builds come from rtc_kernels.cu, so their cubins collide by name when extracted
    # from librtc.so: read the per-build object files instead.
    objs = [so] if not os.path.isdir(so) else [os.path.join(so, f) for f in sorted(os.listdir(so)) if f.endswith(".o")]
    out = []
    cubins = []
    for o in objs:
        tmp = tempfile.mkdtemp()
        subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(o)], cwd=tmp, capture_output=True)
        cubins += [os.path.join(tmp, f) for f in sorted(os.listdir(tmp)) if f.endswith(".cubin")]
    for f in cubins:
        txt = subprocess.run(["nvdisasm", "--print-line-info-inline", f], capture_output=True, text=True).stdout
        in_fn, cur, fresh = False, ("?", 0), True
        for ln in txt.splitlines():
            if ln.startswith(".text."):
                in_fn = kernel_mangled_substr in ln
                continue
            if not in_fn:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', ln)
            if m:
                # a group of //## lines precedes the instructions it covers: innermost frame first, then its
                # callers. Key = innermost frame, plus the innermost frame inside the kernel file if different.
                here = "%s:%s" % (os.path.basename(m.group(1)), m.group(2))
                if fresh:
                    cur, fresh = (here, ""), False
                elif not cur[1] and here.endswith(tuple("0123456789")) and here.startswith("rtc_kernels.cu") and not cur[0].startswith("rtc_kernels.cu"):
                    cur = (cur[0], here)
                continue
            m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
            if m:
                fresh = True
                out.append((m.group(2).strip(), cur[0], cur[1]))
        if out:
            return out
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("--so", default="zig_raytracing_contest_b200/csrc/build")
    ap.add_argument("--top", type=int, default=40)
    ap.add_argument("--launch", type=int, default=0, help="which captured launch of the report (0-based)")
    ap.add_argument("--kernel", default=None, help="substring of the mangled kernel name (default: from the report)")
    a = ap.parse_args()
    raw = subprocess.run(["ncu", "-i", a.report, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]       # one block per captured launch
    if a.launch >= len(starts):
        sys.exit("report holds %d launches" % len(starts))
    lo = starts[a.launch]
    hi = starts[a.launch + 1] if a.launch + 1 < len(starts) else len(rows)
    rows = rows[lo:hi]
    kname = rows[0][1]
    hdr = rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    body = [r for r in rows[2:] if len(r) == len(hdr)]
    if a.kernel is None:
        # "void rtc::exact::k_path<(bool)0, (int)8>(...)" -> "5exact6k_pathILb0ELi8E"
        m = re.search(r"rtc::(\w+)::(\w+)(<[^>]*>)?", kname)
        ns, fn, targs = m.group(1), m.group(2), m.group(3) or ""
        mang = "%d%s%d%s" % (len(ns), ns, len(fn), fn)
        if targs:
            mang += "I"
            for t in targs[1:-1].split(","):
                t = t.strip()
                mm = re.match(r"\((\w+)\)(\d+)", t)
                mang += "L%s%sE" % ({"bool": "b", "int": "i"}[mm.group(1)], mm.group(2))
            mang += "E"
        a.kernel = mang
    sl = sass_lines(a.so, a.kernel)
    if len(sl) != len(body):
        print("warning: %d SASS instructions in the .so vs %d in the report (rebuilt since?)" % (len(sl), len(body)), file=sys.stderr)
    agg = collections.defaultdict(lambda: [0, 0, 0, 0, 0])   # samples, inst, thread inst, long_sb, n sass
    tot = [0, 0, 0]
    for i, r in enumerate(body):
        f, l = (sl[i][1], sl[i][2]) if i < len(sl) else ("?", 0)
        s = int(r[col["# Samples"]]); ie = int(r[col["Instructions Executed"]]); te = int(r[col["Thread Instructions Executed"]])
        lsb = int(r[col["stall_long_sb"]]) if "stall_long_sb" in col else 0
        g = agg[(f, l)]
        g[0] += s; g[1] += ie; g[2] += te; g[3] += lsb; g[4] += 1
        tot[0] += s; tot[1] += ie; tot[2] += te
    print("kernel:", kname)
    print("total samples %d, warp instructions %d, avg active threads %.2f" % (tot[0], tot[1], tot[2] / max(1, tot[1])))
    print("%-40s %8s %7s %12s %7s %7s %6s %5s" % ("file:line", "samples", "%", "warp inst", "%", "thr/inst", "longsb", "sass"))
    for (f, l), g in sorted(agg.items(), key=lambda kv: -kv[1][0])[:a.top]:
        print("%-40s %8d %6.2f%% %12d %6.2f%% %7.1f %6d %5d" % ("%s%s" % (f, " < " + l.split(":")[1] if l else ""), g[0], 100.0 * g[0] / max(1, tot[0]), g[1],
                                                         100.0 * g[1] / max(1, tot[1]), g[2] / max(1, g[1]), g[3], g[4]))


if __name__ == "__main__":
    main()
