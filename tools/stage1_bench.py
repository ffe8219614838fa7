"""Times stage1 + stage2 from a .glb on disk: the host loader (rtc_gltf_load) + device grid build from the uploaded
soup (rtc_scene_create_from_soup), against the one-call device path (rtc_scene_create_from_gltf: triangle assembly
+ grid build on the GPU). Checks first that both scenes render the same primary-visibility frame bit for bit.
    python tools/stage1_bench.py [--tris 2000000] [--threads 8] [--out gpurun_out/stage1.json]"""
import argparse
import ctypes as C
import json
import os
import sys
import tempfile
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from zig_raytracing_contest_b200 import abi, gltf_io, rtc, scenes  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tris", type=int, default=2_000_000)
    ap.add_argument("--threads", type=int, default=8)
    ap.add_argument("--res", type=int, default=128)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    spec = scenes.terrain(num_tris=a.tris)
    tmp = tempfile.mkdtemp(prefix="rtc_stage1_")
    path = os.path.join(tmp, "terrain.glb")
    gltf_io.write_gltf(spec, path, node_translation=(0.5, 0.25, -1.0), node_scale=1.5, glb=True)
    size = os.path.getsize(path)
    res = (a.res,) * 3
    L = rtc.lib()
    resv = np.asarray(res, np.uint32)

    def host_path():
        hs = abi.HostScene()
        t0 = time.perf_counter()
        rtc._check(L.rtc_gltf_load(os.fsencode(path), None, 0, 270, a.threads, C.byref(hs)))
        t1 = time.perf_counter()
        h, bi = C.c_void_p(), abi.BuildInfo()
        rtc._check(L.rtc_scene_create_from_soup(hs.positions, hs.tri_data, hs.num_tris, C.c_void_p(resv.ctypes.data),
                                                hs.materials, hs.num_materials, 0, C.byref(h), C.byref(bi)))
        t2 = time.perf_counter()
        cam = abi.Camera.from_buffer_copy(hs.camera)
        L.rtc_host_scene_free(C.byref(hs))
        return h, cam, bi, {"load_s": t1 - t0, "from_soup_s": t2 - t1, "total_s": t2 - t0}

    def device_path():
        t0 = time.perf_counter()
        s = rtc.Scene.from_gltf(path, res, height=270, num_threads=a.threads, device=0)
        t1 = time.perf_counter()
        return s, {"total_s": t1 - t0, "assemble_kernel_ms": s.build_info.assemble_ms, "grid_kernels_ms": s.build_info.device_ms,
                   "call_ms": s.build_info.total_ms}

    runs = {"host": [], "device": []}
    frames = []
    for rep in range(a.reps):
        h, cam, bi, th = host_path()
        s, td = device_path()
        runs["host"].append(th); runs["device"].append(td)
        if rep == 0:
            sa = rtc.Scene.__new__(rtc.Scene)
            sa.desc, sa._keep, sa._h = None, None, h
            prm = rtc.make_params(mode=abi.RTC_MODE_PRIMARY)
            fa = sa.render(cam, prm, want_hits=True)
            fb = s.render(s.camera, prm, want_hits=True)
            assert bytes(cam) == bytes(s.camera)
            assert np.array_equal(fa["rgb"], fb["rgb"]) and np.array_equal(fa["hit_id"], fb["hit_id"])
            assert (bi.num_tris, bi.empty_cells, bi.max_tris) == (s.build_info.num_tris, s.build_info.empty_cells, s.build_info.max_tris)
            frames.append(int(fa["hit_id"].max()))
            sa.close()
        else:
            L.rtc_scene_destroy(h)
        s.close()
    best = lambda k, f: min(r[f] for r in runs[k])
    line = {"what": "stage1+stage2 from a .glb: host loader + device grid build vs device assembly + device grid build",
            "file_bytes": size, "source_triangles": int(len(spec.positions)), "grid": list(res), "host_threads": a.threads,
            "parity": "primary-visibility frame and hit ids identical",
            "host_path_s": {"load": best("host", "load_s"), "from_soup": best("host", "from_soup_s"), "total": best("host", "total_s")},
            "device_path_s": {"total": best("device", "total_s")},
            "device_kernels_ms": {"k_assemble": best("device", "assemble_kernel_ms"), "grid_build": best("device", "grid_kernels_ms")},
            "runs": runs}
    txt = json.dumps(line)
    print(txt)
    if a.out:
        open(a.out, "w").write(txt + "\n")
    for f in os.listdir(tmp):
        os.remove(os.path.join(tmp, f))
    os.rmdir(tmp)


if __name__ == "__main__":
    main()
