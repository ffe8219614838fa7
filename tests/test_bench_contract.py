"""bench.py's reference arm on the CPU (no GPU needed): the JSON line carries the contract's keys, a step is the measured
row sample (not an extrapolation), and the process never maps the product — it builds its scene with the oracle's own
stage2 and times the oracle only."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_line_and_isolation():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "smoke_cornell_360p_8spp",
                        "--steps", "2", "--warmup", "0", "--cpu-rows", "16"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1                                         # ONE JSON line on stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "Mrays/s" and d["unit"] == "Mrays/s" and d["higher_is_better"] is True
    assert d["steps"] == 2 and d["value"] > 0 and d["ms_per_step"] > 0
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"]
    assert cb["rows_timed"] == 16 and cb["rows_total"] == 360
    # ms_per_step is the MEASURED time of the sample; the whole-frame figure is a separate, labelled key
    assert abs(cb["frame_ms_extrapolated"] - d["ms_per_step"] * 360 / 16) <= 0.1 * cb["frame_ms_extrapolated"]
    assert d["e2e"] == {"value": d["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["workload"] == "smoke_cornell_360p_8spp" and d["config"]["accel"].startswith("grid")
    # bench.py asserts itself that librtc.so is not mapped into the reference arm's process (run_reference); a run that
    # got here passed that assertion


def test_reference_arm_other_ranks_do_nothing():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=120, cwd=ROOT, env=env)
    assert r.returncode == 0 and r.stdout.strip() == ""
