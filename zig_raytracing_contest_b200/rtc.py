"""Host side of the render path: ctypes binding of librtc.so (include/rtc.h) and a Python mirror
of the reference's call sequence around it:

    geometry.build(); geometry.bakeInto(&scene)      src/main.zig:117-118  -> Geometry(...)
    scene.render(threads, camera, img)               src/main.zig:126      -> Scene.render(camera, ...)

There is no CPU fallback: a missing library raises at import of `lib()`, and every compute call
fails with RTC_ERR_NO_DEVICE when no sm_100 GPU is present.
"""
import ctypes as C
import os

import numpy as np

from . import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RTC_LIB_PATH") or os.path.join(_HERE, "librtc.so")   # override: kernel tuning builds only
_LIB = None

TRI_DATA_DTYPE = np.dtype([("normal", "<f4", (3, 3)), ("texcoord", "<f4", (3, 2)),
                           ("material", "<u4"), ("src_tri", "<u4")])


class RtcError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("rtc error %d: %s" % (code, message))
        self.code = code


def lib():
    """Load librtc.so (built in-tree by csrc/Makefile / __graft_entry__.build())."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise ImportError("%s is missing: run `make -C zig_raytracing_contest_b200/csrc` "
                          "(there is no CPU fallback)" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    L.rtc_last_error.restype = C.c_char_p
    L.rtc_abi_version.restype = C.c_uint32
    L.rtc_device_count.restype = C.c_int
    L.rtc_scene_create.argtypes = [C.POINTER(abi.SceneDesc), C.c_int, C.POINTER(C.c_void_p)]
    L.rtc_scene_destroy.argtypes = [C.c_void_p]
    L.rtc_scene_destroy.restype = None
    L.rtc_render.argtypes = [C.c_void_p, C.POINTER(abi.Camera), C.POINTER(abi.Params), C.c_void_p,
                             C.POINTER(abi.Aux)]
    L.rtc_render_accumulate.argtypes = [C.c_void_p, C.POINTER(abi.Camera), C.POINTER(abi.Params),
                                        C.c_void_p, C.c_void_p]
    L.rtc_render_accumulate_static.argtypes = [C.c_void_p, C.POINTER(abi.Camera), C.POINTER(abi.Params),
                                               C.c_void_p, C.c_void_p]
    L.rtc_resolve.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    L.rtc_scene_stats.argtypes = [C.c_void_p, C.POINTER(abi.Stats)]
    L.rtc_scene_build_bvh.argtypes = [C.c_void_p, C.c_int, C.POINTER(abi.BvhInfo)]
    L.rtc_scene_set_stream.argtypes = [C.c_void_p, C.c_void_p]
    L.rtc_scene_prepare_gpus.argtypes = [C.c_void_p, C.c_int, C.c_uint32]
    L.rtc_comm_unique_id.argtypes = [C.c_void_p]
    L.rtc_comm_create.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
    L.rtc_comm_destroy.argtypes = [C.c_void_p]
    L.rtc_comm_destroy.restype = None
    L.rtc_render_sharded.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(abi.Camera), C.POINTER(abi.Params), C.c_void_p,
                                     C.POINTER(abi.Aux)]
    L.rtc_geometry_build.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.POINTER(abi.Geometry)]
    L.rtc_geometry_build_device.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_int,
                                            C.POINTER(abi.Geometry)]
    L.rtc_scene_create_from_soup.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p,
                                             C.POINTER(abi.Material), C.c_uint32, C.c_int,
                                             C.POINTER(C.c_void_p), C.POINTER(abi.BuildInfo)]
    L.rtc_geometry_free.argtypes = [C.POINTER(abi.Geometry)]
    L.rtc_geometry_free.restype = None
    L.rtc_config_load.argtypes = [C.c_char_p, C.POINTER(abi.Config)]
    L.rtc_gltf_load.argtypes = [C.c_char_p, C.c_char_p, C.c_uint32, C.c_uint32, C.c_int, C.POINTER(abi.HostScene)]
    L.rtc_gltf_load_device.argtypes = [C.c_char_p, C.c_char_p, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.POINTER(abi.HostScene)]
    L.rtc_scene_create_from_gltf.argtypes = [C.c_char_p, C.c_char_p, C.c_uint32, C.c_uint32, C.c_int, C.c_void_p, C.c_int,
                                             C.POINTER(C.c_void_p), C.POINTER(abi.Camera), C.POINTER(abi.BuildInfo)]
    L.rtc_host_scene_free.argtypes = [C.POINTER(abi.HostScene)]
    L.rtc_host_scene_free.restype = None
    L.rtc_png_write.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, C.c_void_p]
    L.rtc_png_read.argtypes = [C.c_char_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_void_p)]
    if L.rtc_abi_version() != abi.RTC_ABI_VERSION and not os.environ.get("RTC_SKIP_ABI_CHECK"):
        raise ImportError("librtc.so ABI %d != binding ABI %d" % (L.rtc_abi_version(), abi.RTC_ABI_VERSION))
    _LIB = L
    return L


def _check(rc):
    if rc != abi.RTC_OK:
        raise RtcError(rc, lib().rtc_last_error().decode("utf-8", "replace"))


def device_count():
    return int(lib().rtc_device_count())


# ---- main.zig's file contracts + stage1 ---------------------------------------------------------------
def load_config(path="config.json"):
    """main.Config (src/main.zig:56-69). Returns dict(grid_resolution, num_threads|None, num_samples, max_bounce)."""
    c = abi.Config()
    _check(lib().rtc_config_load(os.fsencode(path), C.byref(c)))
    return {"grid_resolution": tuple(int(x) for x in c.grid_resolution), "num_threads": int(c.num_threads) or None,
            "num_samples": int(c.num_samples), "max_bounce": int(c.max_bounce)}


class HostSceneData:
    """What stage1 leaves in memory (src/main.zig:104-112), copied out of rtc_gltf_load into numpy:
    camera (abi.Camera), positions [N,3,3], tri_data [N] (TRI_DATA_DTYPE), materials [MaterialData].
    device=None: the host loader; device=k (-1 = current): the triangles are assembled on that GPU
    (rtc_gltf_load_device), bit-identical results."""

    def __init__(self, path, camera=None, width=None, height=None, num_threads=1, device=None):
        hs = abi.HostScene()
        cam = camera.encode() if camera is not None else None
        if device is None:
            _check(lib().rtc_gltf_load(os.fsencode(path), cam, int(width or 0), int(height or 0), int(num_threads), C.byref(hs)))
        else:
            _check(lib().rtc_gltf_load_device(os.fsencode(path), cam, int(width or 0), int(height or 0), int(num_threads),
                                              int(device), C.byref(hs)))
        try:
            n = int(hs.num_tris)
            self.camera = abi.Camera.from_buffer_copy(hs.camera)
            self.num_images = int(hs.num_images)
            if n:
                self.positions = np.ctypeslib.as_array(hs.positions, (n, 3, 3)).copy()
                self.tri_data = np.ctypeslib.as_array(C.cast(hs.tri_data, C.POINTER(C.c_uint8)), (n * 68,)).copy().view(TRI_DATA_DTYPE)
            else:
                self.positions, self.tri_data = np.zeros((0, 3, 3), np.float32), np.zeros(0, TRI_DATA_DTYPE)
            self.materials = []
            for i in range(int(hs.num_materials)):
                m = hs.materials[i]
                tex = []
                for t in (m.base_color, m.emissive, m.transparency):
                    arr = np.ctypeslib.as_array(t.texels, (t.h, t.w, t.channels)).copy()
                    td = TextureData(arr)
                    td.u_min, td.u_max, td.v_min, td.v_max = t.u_min, t.u_max, t.v_min, t.v_max
                    tex.append(td)
                self.materials.append(MaterialData(*tex))
        finally:
            lib().rtc_host_scene_free(C.byref(hs))


def png_write(path, rgb):
    """stbi_write_png(path, w, h, 3, img, 3*w) (src/main.zig:130-135)."""
    rgb = np.ascontiguousarray(rgb, np.uint8)
    h, w, c = rgb.shape
    if c != 3:
        raise ValueError("rgb must be [h,w,3]")
    _check(lib().rtc_png_write(os.fsencode(path), w, h, rgb.ctypes.data))


def png_read(path):
    w, h, p = C.c_uint32(), C.c_uint32(), C.c_void_p()
    _check(lib().rtc_png_read(os.fsencode(path), C.byref(w), C.byref(h), C.byref(p)))
    try:
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), (h.value, w.value, 3)).copy()
    finally:
        C.CDLL(None).free(p)


# ---- stage2: Geometry.build + bakeInto ---------------------------------------------------------
class Geometry:
    """Uniform-grid build + bake of a triangle soup (src/stage2.zig:131-164) on the host threads.

    positions: [N,3,3] world-space vertices; tri_data: N records of TRI_DATA_DTYPE (src_tri
    ignored on input); resolution: config.json's grid_resolution.
    """

    def __init__(self, positions, tri_data, resolution, device=None):
        """device=None: the host-thread build (rtc_geometry_build); an int: the same build on that GPU
        (rtc_geometry_build_device, -1 = current device), bit-identical outputs."""
        positions = np.ascontiguousarray(positions, dtype=np.float32).reshape(-1, 9)
        tri_data = np.ascontiguousarray(tri_data, dtype=TRI_DATA_DTYPE)
        if tri_data.shape[0] != positions.shape[0]:
            raise ValueError("positions and tri_data disagree on the triangle count")
        res = np.ascontiguousarray(resolution, dtype=np.uint32)
        g = abi.Geometry()
        if device is None:
            _check(lib().rtc_geometry_build(C.c_void_p(positions.ctypes.data), C.c_void_p(tri_data.ctypes.data),
                                            C.c_uint32(positions.shape[0]), C.c_void_p(res.ctypes.data), C.byref(g)))
        else:
            _check(lib().rtc_geometry_build_device(C.c_void_p(positions.ctypes.data), C.c_void_p(tri_data.ctypes.data),
                                                   C.c_uint32(positions.shape[0]), C.c_void_p(res.ctypes.data),
                                                   C.c_int(device), C.byref(g)))
        try:
            nc, nt = int(g.num_cells), int(g.num_tris)
            self.grid = abi.Grid.from_buffer_copy(g.grid)
            self.cells = np.ctypeslib.as_array(C.cast(g.cells, C.POINTER(C.c_uint32)), (nc, 2)).copy()
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
            lib().rtc_geometry_free(C.byref(g))

    @property
    def num_tris(self):
        return int(self.tri_pos.shape[0])


# ---- materials -----------------------------------------------------------------------------------
class TextureData:
    """stage3.Texture(T) (src/stage3.zig:75-80): linear float texels + the clamp range that
    stage1 derives from the sampler's wrap mode (src/stage1.zig:383-397)."""

    def __init__(self, texels, clamp_u=False, clamp_v=False):
        t = np.ascontiguousarray(texels, dtype=np.float32)
        if t.ndim == 2:
            t = t[:, :, None]
        self.texels = t
        self.h, self.w, self.channels = t.shape
        i32 = np.iinfo(np.int32)
        self.u_min, self.u_max = (0, self.w - 1) if clamp_u else (i32.min, i32.max)
        self.v_min, self.v_max = (0, self.h - 1) if clamp_v else (i32.min, i32.max)

    def as_struct(self):
        return abi.Texture(self.texels.ctypes.data_as(C.POINTER(C.c_float)), self.w, self.h,
                           self.u_min, self.u_max, self.v_min, self.v_max, self.channels)


def constant_texture(value):
    v = np.atleast_1d(np.asarray(value, dtype=np.float32))
    return TextureData(v.reshape(1, 1, -1))


class MaterialData:
    """stage3.Material (src/stage3.zig:125-129)."""

    def __init__(self, base_color=(0.8, 0.8, 0.8), emissive=(0.0, 0.0, 0.0), transparency=1.0):
        def tex(x, ch):
            if isinstance(x, TextureData):
                return x
            return constant_texture(x if ch == 3 else [x])
        self.base_color, self.emissive, self.transparency = tex(base_color, 3), tex(emissive, 3), tex(transparency, 1)

    def as_struct(self):
        return abi.Material(self.base_color.as_struct(), self.emissive.as_struct(), self.transparency.as_struct())


def make_scene_desc(geometry, materials):
    """rtc_scene_desc over numpy-owned arrays. Returns (desc, keepalive)."""
    mats = (abi.Material * len(materials))(*[m.as_struct() for m in materials])
    desc = abi.SceneDesc()
    desc.grid = geometry.grid
    desc.cells = geometry.cells.ctypes.data_as(C.POINTER(abi.Cell))
    desc.num_cells = geometry.cells.shape[0]
    desc.tri_pos = geometry.tri_pos.ctypes.data_as(C.POINTER(abi.TriPos))
    desc.tri_data = geometry.tri_data.ctypes.data_as(C.POINTER(abi.TriData))
    desc.num_tris = geometry.tri_pos.shape[0]
    desc.materials = mats
    desc.num_materials = len(materials)
    desc.num_src_tris = geometry.num_src_tris
    return desc, (mats, geometry, materials)


def make_params(num_samples=3, max_bounce=4, mode=abi.RTC_MODE_PATH, seed=0, sample_offset=0,
                total_samples=0, math=abi.RTC_MATH_EXACT, tile_rank=0, tile_world=0, collect_stats=False,
                tile_rows=0, accel=abi.RTC_ACCEL_GRID, num_gpus=0, shard=abi.RTC_SHARD_STRIPS):
    """main.Config's num_samples / max_bounce (src/main.zig:56-61; config.json defaults 3 and 4)."""
    p = abi.Params()
    p.num_samples, p.max_bounce, p.mode, p.seed = num_samples, max_bounce, mode, seed
    p.sample_offset, p.total_samples = sample_offset, total_samples
    p.accel, p.math = accel, math
    p.tile_rank, p.tile_world, p.collect_stats = tile_rank, tile_world, 1 if collect_stats else 0
    p.tile_rows = tile_rows
    p.num_gpus, p.shard = num_gpus, shard
    return p


# ---- one process per GPU: the library's own NCCL communicator ---------------------------------------------
def comm_unique_id():
    """rank 0: the 128-byte id every rank passes to Comm (hand it over with MPI / torch.distributed / a file)."""
    buf = (C.c_uint8 * abi.RTC_COMM_ID_BYTES)()
    _check(lib().rtc_comm_unique_id(buf))
    return bytes(buf)


class Comm:
    """rtc_comm: one rank of a frame sharded over processes (collective create: ncclCommInitRank)."""

    def __init__(self, unique_id, rank, world, device=-1):
        if len(unique_id) != abi.RTC_COMM_ID_BYTES:
            raise ValueError("unique_id must be %d bytes" % abi.RTC_COMM_ID_BYTES)
        self._h = C.c_void_p()
        self.rank, self.world = int(rank), int(world)
        buf = (C.c_uint8 * abi.RTC_COMM_ID_BYTES).from_buffer_copy(unique_id)
        _check(lib().rtc_comm_create(buf, self.rank, self.world, int(device), C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            lib().rtc_comm_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close


# ---- stage3: Scene ---------------------------------------------------------------------------------
class Scene:
    """Device-resident stage3.Scene. `render` is the reference's Scene.render (src/stage3.zig:247)."""

    def __init__(self, geometry, materials, device=-1):
        self.desc, self._keep = make_scene_desc(geometry, materials)
        self._h = C.c_void_p()
        _check(lib().rtc_scene_create(C.byref(self.desc), device, C.byref(self._h)))

    @classmethod
    def from_soup(cls, positions, tri_data, resolution, materials, device=-1):
        """geometry.build() + bakeInto + upload in one step with the grid built ON the device
        (rtc_scene_create_from_soup). `build_info` holds what stage2 logs (src/stage2.zig:97-100,126-128)."""
        positions = np.ascontiguousarray(positions, dtype=np.float32).reshape(-1, 9)
        tri_data = np.ascontiguousarray(tri_data, dtype=TRI_DATA_DTYPE)
        if tri_data.shape[0] != positions.shape[0]:
            raise ValueError("positions and tri_data disagree on the triangle count")
        res = np.ascontiguousarray(resolution, dtype=np.uint32)
        mats = (abi.Material * len(materials))(*[m.as_struct() for m in materials])
        self = cls.__new__(cls)
        self.desc, self._keep = None, (mats, materials)
        self._h = C.c_void_p()
        self.build_info = abi.BuildInfo()
        _check(lib().rtc_scene_create_from_soup(C.c_void_p(positions.ctypes.data), C.c_void_p(tri_data.ctypes.data),
                                                C.c_uint32(positions.shape[0]), C.c_void_p(res.ctypes.data),
                                                mats, C.c_uint32(len(materials)), C.c_int(device),
                                                C.byref(self._h), C.byref(self.build_info)))
        return self

    @classmethod
    def from_gltf(cls, path, resolution, camera=None, width=None, height=None, num_threads=1, device=-1):
        """loadGltfFile .. bakeInto (src/main.zig:104-118) in one call, triangle assembly and grid build on the
        device (rtc_scene_create_from_gltf). Sets `camera` (abi.Camera) and `build_info`."""
        res = np.ascontiguousarray(resolution, dtype=np.uint32)
        self = cls.__new__(cls)
        self.desc, self._keep = None, None
        self._h = C.c_void_p()
        self.build_info = abi.BuildInfo()
        self.camera = abi.Camera()
        _check(lib().rtc_scene_create_from_gltf(os.fsencode(path), camera.encode() if camera is not None else None,
                                                int(width or 0), int(height or 0), int(num_threads),
                                                C.c_void_p(res.ctypes.data), C.c_int(device),
                                                C.byref(self._h), C.byref(self.camera), C.byref(self.build_info)))
        return self

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            lib().rtc_scene_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def render(self, camera, params, img=None, want_radiance=False, want_hits=False, want_stats=False, instrumented=True):
        """Blocking render into img ([h,w,3] uint8, row 0 = top). Returns a dict with rgb and the
        optional aux outputs. want_stats runs the instrumented kernels (cell / triangle-test counters) unless
        instrumented=False (rays, paths, hits and the timings only)."""
        w, h = int(camera.w), int(camera.h)
        if img is None:
            img = np.zeros((h, w, 3), np.uint8)
        if img.shape != (h, w, 3) or img.dtype != np.uint8 or not img.flags.c_contiguous:
            raise ValueError("img must be a contiguous [h,w,3] uint8 array")
        out = {"rgb": img}
        aux = abi.Aux()
        if want_radiance:
            out["radiance"] = np.zeros((h, w, 3), np.float32)
            aux.radiance = out["radiance"].ctypes.data_as(C.POINTER(C.c_float))
        if want_hits:
            out["hit_id"] = np.full((h, w), abi.MISS_ID, np.uint32)
            out["hit_tuv"] = np.zeros((h, w, 3), np.float32)
            aux.hit_id = out["hit_id"].ctypes.data_as(C.POINTER(C.c_uint32))
            aux.hit_tuv = out["hit_tuv"].ctypes.data_as(C.POINTER(C.c_float))
        st = abi.Stats()
        if want_stats:
            aux.stats = C.pointer(st)
            if instrumented:
                params = abi.Params.from_buffer_copy(params)
                params.collect_stats = 1             # instrumented kernels: cells / triangle-test counters
        _check(lib().rtc_render(self._h, C.byref(camera), C.byref(params), img.ctypes.data, C.byref(aux)))
        if want_stats:
            out["stats"] = st.as_dict()
        return out

    def render_sharded(self, comm, camera, params, img=None, want_stats=False, to_host=True):
        """rtc_render_sharded: this rank's share of one frame; the frame arrives on rank 0 (in `img` when to_host,
        else it stays in rank 0's device memory). comm=None: this process alone. Returns dict(rgb?, stats?)."""
        w, h = int(camera.w), int(camera.h)
        rank = comm.rank if comm is not None else 0
        out = {}
        ptr = None
        if to_host and rank == 0:
            if img is None:
                img = np.zeros((h, w, 3), np.uint8)
            if img.shape != (h, w, 3) or img.dtype != np.uint8 or not img.flags.c_contiguous:
                raise ValueError("img must be a contiguous [h,w,3] uint8 array")
            out["rgb"] = img
            ptr = img.ctypes.data
        aux = abi.Aux()
        st = abi.Stats()
        if want_stats:
            aux.stats = C.pointer(st)
        _check(lib().rtc_render_sharded(self._h, comm._h if comm is not None else None, C.byref(camera), C.byref(params),
                                        C.c_void_p(ptr), C.byref(aux)))
        if want_stats:
            out["stats"] = st.as_dict()
        return out

    def prepare_gpus(self, num_gpus, accel=abi.RTC_ACCEL_GRID):
        """Replicas on the other devices (and the BVH) ahead of the first render that needs them."""
        _check(lib().rtc_scene_prepare_gpus(self._h, int(num_gpus), int(accel)))

    def set_stream(self, stream_ptr=None):
        """Every later call on this scene launches on this cudaStream_t (None: back to the scene's own stream)."""
        _check(lib().rtc_scene_set_stream(self._h, C.c_void_p(stream_ptr or 0)))

    def render_accumulate(self, camera, params, d_accum_ptr, stream_ptr=None):
        """Adds params.num_samples samples per pixel into a device float4 accumulator (async)."""
        _check(lib().rtc_render_accumulate(self._h, C.byref(camera), C.byref(params),
                                           C.c_void_p(d_accum_ptr), C.c_void_p(stream_ptr or 0)))

    def render_accumulate_static(self, camera, params, d_accum_ptr, stream_ptr=None):
        """render_accumulate with the wavefront scratch in librtc.so's static device array: allocates nothing; at most
        abi.RTC_STATIC_SCRATCH_BYTES / (152 + 24 * max_bounce) paths per call (rtc_render_accumulate_static)."""
        _check(lib().rtc_render_accumulate_static(self._h, C.byref(camera), C.byref(params),
                                                  C.c_void_p(d_accum_ptr), C.c_void_p(stream_ptr or 0)))

    def build_bvh(self, num_threads=0):
        """RTC_ACCEL_BVH's structure, ahead of the first render that asks for it (rtc_scene_build_bvh)."""
        info = abi.BvhInfo()
        _check(lib().rtc_scene_build_bvh(self._h, int(num_threads), C.byref(info)))
        return {k: getattr(info, k) for k, _ in info._fields_ if k != "reserved"}

    def stats(self):
        st = abi.Stats()
        _check(lib().rtc_scene_stats(self._h, C.byref(st)))
        return st.as_dict()


def resolve(d_accum_ptr, num_pixels, total_samples, d_rgb_ptr=None, d_radiance_ptr=None, device=-1, stream_ptr=None):
    """pixel.mul(inv_num_samples).toRGB() on device buffers (src/stage3.zig:242)."""
    _check(lib().rtc_resolve(C.c_void_p(d_accum_ptr), num_pixels, total_samples, C.c_void_p(d_rgb_ptr or 0),
                             C.c_void_p(d_radiance_ptr or 0), device, C.c_void_p(stream_ptr or 0)))


def bytes_per_frame(stats, num_pixels):
    """algorithmic bytes at the reference's record sizes — 8 B per visited cell,
    48 B per tested triangle, 224 B per shaded hit, 3 B per stored pixel."""
    return 8 * stats["cells_visited"] + 48 * stats["tris_tested"] + 224 * stats["hits_shaded"] + 3 * num_pixels
