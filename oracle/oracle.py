"""TEST INFRASTRUCTURE, NOT PRODUCT — ctypes front end of oracle/liboracle.so.

Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline / --impl reference) import
this module. It never touches the GPU and never reads /root/reference at run time.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from zig_raytracing_contest_b200 import abi   # struct layouts of include/rtc.h only

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

TRI_DATA_DTYPE = np.dtype([("normal", "<f4", (3, 3)), ("texcoord", "<f4", (3, 2)),
                           ("material", "<u4"), ("src_tri", "<u4")])
assert TRI_DATA_DTYPE.itemsize == C.sizeof(abi.TriData) == 68


class OrcGeometry(C.Structure):
    _fields_ = [("grid", abi.Grid),
                ("cells", C.POINTER(abi.Cell)), ("num_cells", C.c_uint64),
                ("indices", C.POINTER(C.c_uint64)),
                ("tri_pos", C.POINTER(abi.TriPos)), ("tri_data", C.POINTER(abi.TriData)),
                ("num_tris", C.c_uint64), ("num_src_tris", C.c_uint32),
                ("empty_cells", C.c_uint32), ("min_tris", C.c_uint32), ("max_tris", C.c_uint32)]


def _cpu_has_v3():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    flags = set(line.split(":", 1)[1].split())
                    return {"avx2", "fma", "bmi2"} <= flags
    except OSError:
        pass
    return False


def build(force=False):
    """Compile liboracle.so with gcc (oracle/Makefile)."""
    so = os.path.join(_HERE, "liboracle.so")
    if force and os.path.exists(so):
        os.remove(so)
    args = ["make", "-C", _HERE, "liboracle.so"]
    if not _cpu_has_v3():
        args.append("PORTABLE=1")
    subprocess.run(args, check=True, capture_output=True)
    return so


def lib():
    global _LIB
    if _LIB is not None:
        return _LIB
    so = os.path.join(_HERE, "liboracle.so")
    if not os.path.exists(so) or not _cpu_has_v3():
        # the prebuilt file targets x86-64-v3; rebuild portable when the host cannot run it
        so = build(force=os.path.exists(so) and not _cpu_has_v3())
    L = C.CDLL(so)
    f3 = C.POINTER(C.c_float)
    L.orc_length.restype = C.c_float
    L.orc_trace_ray.restype = C.c_int64
    L.orc_trace_ray_brute.restype = C.c_int64
    L.orc_geometry_build.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.POINTER(OrcGeometry)]
    L.orc_render.argtypes = [C.POINTER(abi.SceneDesc), C.POINTER(abi.Camera), C.POINTER(abi.Params),
                             C.c_int, C.c_uint32, C.c_uint32, C.c_void_p, C.POINTER(abi.Aux)]
    L.orc_rng_block.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, f3]
    L.orc_unit_vector.argtypes = [C.c_float, C.c_float, f3]
    L.orc_texture_sample.argtypes = [C.POINTER(abi.Texture), C.c_float, C.c_float, f3]
    L.orc_camera_get_ray.argtypes = [C.POINTER(abi.Camera), C.c_float, C.c_float, f3, f3]
    L.orc_load_camera.argtypes = [f3, C.c_float, C.c_float, C.c_uint32, C.c_uint32, C.POINTER(abi.Camera)]
    _LIB = L
    return L


def _f(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


# ---- linalg unit level ---------------------------------------------------------------------
def cross(a, b):
    a, b, o = _f(a), _f(b), np.zeros(3, np.float32)
    lib().orc_cross(_fp(a), _fp(b), _fp(o))
    return o


def length(v):
    v = _f(v)
    return float(lib().orc_length(_fp(v)))


def normalize(v):
    v, o = _f(v), np.zeros(3, np.float32)
    lib().orc_normalize(_fp(v), _fp(o))
    return o


def bbox_ray_intersection(bmin, bmax, orig, direction):
    t = C.c_float(np.nan)
    a, b, o, d = _f(bmin), _f(bmax), _f(orig), _f(direction)
    hit = lib().orc_bbox_ray_intersection(_fp(a), _fp(b), _fp(o), _fp(d), C.byref(t))
    return bool(hit), t.value


def grid_init(bmin, bmax, res):
    g = abi.Grid()
    a, b = _f(bmin), _f(bmax)
    r = np.ascontiguousarray(res, dtype=np.uint32)
    lib().orc_grid_init(_fp(a), _fp(b), r.ctypes.data_as(C.POINTER(C.c_uint32)), C.byref(g))
    return g


def grid_cell_bbox(g, x, y, z):
    lo, hi = np.zeros(3, np.float32), np.zeros(3, np.float32)
    lib().orc_grid_cell_bbox(C.byref(g), C.c_uint32(x), C.c_uint32(y), C.c_uint32(z), _fp(lo), _fp(hi))
    return lo, hi


def grid_walk(g, orig, direction, max_steps=4096):
    cells = np.zeros((max_steps, 3), np.uint32)
    ts = np.zeros(max_steps, np.float32)
    o, d = _f(orig), _f(direction)
    n = lib().orc_grid_walk(C.byref(g), _fp(o), _fp(d), C.c_int(max_steps),
                            cells.ctypes.data_as(C.POINTER(C.c_uint32)), _fp(ts))
    if n < 0:
        return None
    return cells[:n].copy(), ts[:n].copy()


def triangle_ray_intersection(v0, e1, e2, orig, direction):
    t, u, v = C.c_float(), C.c_float(), C.c_float()
    a, b, c, o, d = _f(v0), _f(e1), _f(e2), _f(orig), _f(direction)
    hit = lib().orc_triangle_ray_intersection(_fp(a), _fp(b), _fp(c), _fp(o), _fp(d),
                                              C.byref(t), C.byref(u), C.byref(v))
    return (t.value, u.value, v.value) if hit else None


def triangle_aabb(tri, bmin, bmax):
    t, a, b = _f(tri).reshape(9), _f(bmin), _f(bmax)
    return bool(lib().orc_triangle_aabb(_fp(t), _fp(a), _fp(b)))


def texture_sample(tex, u, v):
    out = np.zeros(tex.channels, np.float32)
    lib().orc_texture_sample(C.byref(tex), C.c_float(u), C.c_float(v), _fp(out))
    return out


def interpolate(tri_data_record, u, v):
    """Triangle.Data.interpolate for the normal and the texcoord of one TRI_DATA_DTYPE record."""
    rec = np.ascontiguousarray(tri_data_record, dtype=TRI_DATA_DTYPE).reshape(1)
    n, t = np.zeros(3, np.float32), np.zeros(2, np.float32)
    lib().orc_interpolate(rec.ctypes.data_as(C.c_void_p), C.c_float(u), C.c_float(v), _fp(n), _fp(t))
    return n, t


def to_rgb(c):
    c = _f(c)
    out = np.zeros(3, np.uint8)
    lib().orc_to_rgb(_fp(c), out.ctypes.data_as(C.POINTER(C.c_uint8)))
    return out


def env_color(d):
    d, o = _f(d), np.zeros(3, np.float32)
    lib().orc_env_color(_fp(d), _fp(o))
    return o


def camera_get_ray(cam, x, y):
    o, d = np.zeros(3, np.float32), np.zeros(3, np.float32)
    lib().orc_camera_get_ray(C.byref(cam), C.c_float(x), C.c_float(y), _fp(o), _fp(d))
    return o, d


# ---- RNG -------------------------------------------------------------------------------------
def philox4x32_10(ctr, key):
    c = np.ascontiguousarray(ctr, dtype=np.uint32)
    k = np.ascontiguousarray(key, dtype=np.uint32)
    o = np.zeros(4, np.uint32)
    p = C.POINTER(C.c_uint32)
    lib().orc_philox4x32_10(c.ctypes.data_as(p), k.ctypes.data_as(p), o.ctypes.data_as(p))
    return o


def rng_block(seed, pixel, sample, block):
    o = np.zeros(4, np.float32)
    lib().orc_rng_block(seed, pixel, sample, block, _fp(o))
    return o


def unit_vector(u1, u2):
    o = np.zeros(3, np.float32)
    lib().orc_unit_vector(C.c_float(u1), C.c_float(u2), _fp(o))
    return o


# ---- stage1 math -----------------------------------------------------------------------------
def mat4_mul(left, right):
    a, b, o = _f(left).reshape(16), _f(right).reshape(16), np.zeros(16, np.float32)
    lib().orc_mat4_mul(_fp(a), _fp(b), _fp(o))
    return o


def mat4_recompose(t, r, s):
    t, r, s, o = _f(t), _f(r), _f(s), np.zeros(16, np.float32)
    lib().orc_mat4_recompose(_fp(t), _fp(r), _fp(s), _fp(o))
    return o


def assemble_triangles(matrix, positions, normals, texcoords, indices, material):
    """stage1.loadTriangles for one primitive -> (positions[n,3,3], tri_data[n])."""
    m = _f(matrix).reshape(16)
    p, n = _f(positions), _f(normals)
    t = _f(texcoords) if texcoords is not None else None
    idx = np.ascontiguousarray(indices, dtype=np.uint16)
    nt = idx.size // 3
    out_p = np.zeros((nt, 3, 3), np.float32)
    out_d = np.zeros(nt, TRI_DATA_DTYPE)
    lib().orc_assemble_triangles(_fp(m), _fp(p), _fp(n), _fp(t) if t is not None else None,
                                 idx.ctypes.data_as(C.POINTER(C.c_uint16)), C.c_uint32(nt),
                                 C.c_uint32(material), _fp(out_p), out_d.ctypes.data_as(C.c_void_p))
    return out_p, out_d


def load_camera(matrix, yfov, aspect_ratio, width, height):
    cam = abi.Camera()
    m = _f(matrix).reshape(16)
    rc = lib().orc_load_camera(_fp(m), yfov, aspect_ratio or 0.0, width or 0, height or 0, C.byref(cam))
    if rc != 0:
        raise ValueError({-1: "OutputImgSizeIsNotSpecified", -2: "CameraHasAspectRatio",
                          -3: "CameraHasntAspectRatio"}[rc])
    return cam


# ---- stage2 ----------------------------------------------------------------------------------
class Geometry:
    """Geometry.build + bakeInto on the CPU oracle; numpy views are copies (owned here)."""

    def __init__(self, positions, tri_data, resolution):
        positions = _f(positions).reshape(-1, 9)
        tri_data = np.ascontiguousarray(tri_data, dtype=TRI_DATA_DTYPE)
        res = np.ascontiguousarray(resolution, dtype=np.uint32)
        g = OrcGeometry()
        rc = lib().orc_geometry_build(positions.ctypes.data, tri_data.ctypes.data,
                                      positions.shape[0], res.ctypes.data, C.byref(g))
        if rc != 0:
            raise MemoryError("orc_geometry_build failed")
        try:
            nc, nt = int(g.num_cells), int(g.num_tris)
            self.grid = abi.Grid.from_buffer_copy(g.grid)
            self.cells = np.ctypeslib.as_array(C.cast(g.cells, C.POINTER(C.c_uint32)), (nc, 2)).copy()
            self.indices = (np.ctypeslib.as_array(g.indices, (nt,)).copy() if nt else np.zeros(0, np.uint64))
            if nt:
                self.tri_pos = np.ctypeslib.as_array(C.cast(g.tri_pos, C.POINTER(C.c_float)), (nt, 3, 4)).copy()
                raw = np.ctypeslib.as_array(C.cast(g.tri_data, C.POINTER(C.c_uint8)), (nt * 68,)).copy()
                self.tri_data = raw.view(TRI_DATA_DTYPE)
            else:
                self.tri_pos = np.zeros((0, 3, 4), np.float32)
                self.tri_data = np.zeros(0, TRI_DATA_DTYPE)
            self.num_src_tris = int(g.num_src_tris)
            self.empty_cells, self.min_tris, self.max_tris = int(g.empty_cells), int(g.min_tris), int(g.max_tris)
        finally:
            lib().orc_geometry_free(C.byref(g))


# ---- stage3 ----------------------------------------------------------------------------------
def trace_ray(desc, orig, direction, brute=False, stats=None):
    t, u, v = C.c_float(), C.c_float(), C.c_float()
    o, d = _f(orig), _f(direction)
    if brute:
        idx = lib().orc_trace_ray_brute(C.byref(desc), _fp(o), _fp(d), C.byref(t), C.byref(u), C.byref(v))
    else:
        idx = lib().orc_trace_ray(C.byref(desc), _fp(o), _fp(d), C.byref(t), C.byref(u), C.byref(v),
                                  C.byref(stats) if stats is not None else None)
    return int(idx), t.value, u.value, v.value


def render(desc, camera, params, num_threads=None, rows=None, want_hits=False, want_radiance=True,
           want_stats=False):
    """Scene.render on the CPU oracle. Returns dict(rgb, radiance, hit_id, hit_tuv, stats)."""
    if num_threads is None:
        num_threads = os.cpu_count() or 1
    w, h = int(camera.w), int(camera.h)
    y0, y1 = rows if rows is not None else (0, h)
    rgb = np.zeros((h, w, 3), np.uint8)
    aux = abi.Aux()
    out = {"rgb": rgb}
    if want_radiance:
        out["radiance"] = np.zeros((h, w, 3), np.float32)
        aux.radiance = _fp(out["radiance"])
    if want_hits:
        out["hit_id"] = np.full((h, w), abi.MISS_ID, np.uint32)
        out["hit_tuv"] = np.zeros((h, w, 3), np.float32)
        aux.hit_id = out["hit_id"].ctypes.data_as(C.POINTER(C.c_uint32))
        aux.hit_tuv = _fp(out["hit_tuv"])
    st = abi.Stats()
    if want_stats:
        aux.stats = C.pointer(st)
    rc = lib().orc_render(C.byref(desc), C.byref(camera), C.byref(params), int(num_threads),
                          int(y0), int(y1), rgb.ctypes.data, C.byref(aux))
    if rc != 0:
        raise RuntimeError("orc_render failed rc=%d" % rc)
    if want_stats:
        out["stats"] = st.as_dict()
    return out
