import sys, ctypes as C, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
from oracle import oracle as orc
from zig_raytracing_contest_b200 import rtc, scenes
import test_trav_twolevel as T
import subprocess
subprocess.run(["nvcc", "-O2", "-std=c++17", "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-fmad=false", "-Xcompiler", "-fPIC,-ffp-contract=off", "-o", "/root/repo/tests/native/libtrav_check.so", "/root/repo/tests/native/trav_check.cu"], check=True)
L = C.CDLL("/root/repo/tests/native/libtrav_check.so")
L.trav_check.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int), C.c_void_p, C.POINTER(C.c_uint64)]
L.trav_dump.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float]
spec = scenes.cornell(); res = (32, 32, 32)
geo = rtc.Geometry(spec.positions, spec.tri_data, res)
rays = T._rays(orc, spec, geo, spec.camera(height=90), 700, 5)
us, fs = T._iters(orc, geo.grid, rays)
best = np.full(len(us), np.inf, np.float32)
bad, first, steps, ev = T._check(L, res, geo.cells, us, fs, best)
print("bad", bad, "first", first, steps)
r = np.ascontiguousarray(res, np.uint32); cells = np.ascontiguousarray(geo.cells, np.uint32)
u = np.ascontiguousarray(us[first]); f = np.ascontiguousarray(fs[first])
L.trav_dump(r.ctypes.data, cells.ctypes.data, u.ctypes.data, f.ctypes.data, C.c_float(np.inf))
