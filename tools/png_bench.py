"""Times rtc_png_write on a 4K frame (noisy gradient, like a path-traced image) with 1 thread and all threads.
    python tools/png_bench.py [--out gpurun_out/png.json]"""
import argparse, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from zig_raytracing_contest_b200 import rtc  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--out", default=None)
a = ap.parse_args()
rng = np.random.RandomState(1)
img = (rng.randint(0, 64, (2160, 3840, 3)) + np.linspace(0, 190, 3840)[None, :, None]).astype(np.uint8)
res = {"what": "rtc_png_write of a 3840x2160 RGB8 frame (noisy gradient), best of 3", "cores": os.cpu_count(), "runs": {}}
for th in ("1", "2", "4", "8", str(os.cpu_count())):
    os.environ["RTC_PNG_THREADS"] = th
    best = 1e9
    for _ in range(3):
        t = time.perf_counter(); rtc.png_write("/tmp/rtc_png_bench.png", img); best = min(best, time.perf_counter() - t)
    assert np.array_equal(rtc.png_read("/tmp/rtc_png_bench.png"), img)
    res["runs"][th] = {"seconds": round(best, 4), "bytes": os.path.getsize("/tmp/rtc_png_bench.png")}
os.remove("/tmp/rtc_png_bench.png")
txt = json.dumps(res)
print(txt)
if a.out:
    open(a.out, "w").write(txt + "\n")
