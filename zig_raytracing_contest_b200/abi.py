"""ctypes mirror of include/rtc.h (struct layouts and constants only; no code).

Each class names the rtc.h struct it mirrors; rtc.h in turn cites the reference type
(src/stage3.zig, src/linalg.zig) it restates with a fixed layout.
"""
import ctypes as C

RTC_ABI_VERSION = 9

RTC_OK = 0
RTC_ERR_INVALID_ARGUMENT = 1
RTC_ERR_NO_DEVICE = 2
RTC_ERR_CUDA = 3
RTC_ERR_OUT_OF_MEMORY = 4
RTC_ERR_IO = 5
RTC_ERR_FORMAT = 6
RTC_ERR_UNSUPPORTED = 7

RTC_MODE_PATH = 0
RTC_MODE_PRIMARY = 1
RTC_ACCEL_GRID = 0
RTC_ACCEL_BVH = 1
RTC_SHARD_STRIPS = 0
RTC_SHARD_SAMPLES = 1
RTC_MAX_GPUS = 16
RTC_STATIC_SCRATCH_BYTES = 64 << 20       # rtc_render_accumulate_static's static device array
RTC_COMM_ID_BYTES = 128
RTC_MATH_EXACT = 0
RTC_MATH_FAST = 1

MISS_ID = 0xFFFFFFFF

c_float3 = C.c_float * 3
c_float4 = C.c_float * 4
c_uint3 = C.c_uint32 * 3


class Grid(C.Structure):          # rtc_grid
    _fields_ = [("bbox_min", c_float3), ("bbox_max", c_float3),
                ("resolution", c_uint3), ("cell_size", c_float3)]


class Cell(C.Structure):          # rtc_cell
    _fields_ = [("begin", C.c_uint32), ("end", C.c_uint32)]


class TriPos(C.Structure):        # rtc_tri_pos
    _fields_ = [("v0", c_float4), ("e1", c_float4), ("e2", c_float4)]


class TriData(C.Structure):       # rtc_tri_data
    _fields_ = [("normal", (C.c_float * 3) * 3), ("texcoord", (C.c_float * 2) * 3),
                ("material", C.c_uint32), ("src_tri", C.c_uint32)]


class Texture(C.Structure):       # rtc_texture
    _fields_ = [("texels", C.POINTER(C.c_float)),
                ("w", C.c_int32), ("h", C.c_int32),
                ("u_min", C.c_int32), ("u_max", C.c_int32),
                ("v_min", C.c_int32), ("v_max", C.c_int32),
                ("channels", C.c_int32)]


class Material(C.Structure):      # rtc_material
    _fields_ = [("base_color", Texture), ("emissive", Texture), ("transparency", Texture)]


class SceneDesc(C.Structure):     # rtc_scene_desc
    _fields_ = [("grid", Grid),
                ("cells", C.POINTER(Cell)), ("num_cells", C.c_uint64),
                ("tri_pos", C.POINTER(TriPos)), ("tri_data", C.POINTER(TriData)),
                ("num_tris", C.c_uint64),
                ("materials", C.POINTER(Material)),
                ("num_materials", C.c_uint32), ("num_src_tris", C.c_uint32)]


class Camera(C.Structure):        # rtc_camera
    _fields_ = [("w", C.c_uint32), ("h", C.c_uint32),
                ("origin", c_float3), ("lower_left_corner", c_float3),
                ("right", c_float3), ("up", c_float3)]


class Params(C.Structure):        # rtc_params
    _fields_ = [("num_samples", C.c_uint16), ("max_bounce", C.c_uint16),
                ("mode", C.c_uint32), ("seed", C.c_uint64),
                ("sample_offset", C.c_uint32), ("total_samples", C.c_uint32),
                ("accel", C.c_uint32), ("math", C.c_uint32),
                ("tile_rank", C.c_uint32), ("tile_world", C.c_uint32),
                ("collect_stats", C.c_uint32), ("tile_rows", C.c_uint32),
                ("num_gpus", C.c_int32), ("shard", C.c_uint32)]


class Stats(C.Structure):         # rtc_stats
    _fields_ = [("rays", C.c_uint64), ("cells_visited", C.c_uint64),
                ("tris_tested", C.c_uint64), ("hits_shaded", C.c_uint64),
                ("paths", C.c_uint64), ("kernel_ms", C.c_double), ("launches", C.c_uint64),
                ("trace_ms", C.c_double), ("trace_launches", C.c_uint32), ("trace_launches_untimed", C.c_uint32),
                ("frame_ms", C.c_double)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class Aux(C.Structure):           # rtc_aux
    _fields_ = [("hit_id", C.POINTER(C.c_uint32)), ("hit_tuv", C.POINTER(C.c_float)),
                ("radiance", C.POINTER(C.c_float)), ("stats", C.POINTER(Stats))]


class BvhInfo(C.Structure):       # rtc_bvh_info
    _fields_ = [("num_nodes", C.c_uint32), ("num_tris", C.c_uint32), ("depth", C.c_uint32), ("reserved", C.c_uint32),
                ("bytes", C.c_uint64), ("build_ms", C.c_double)]


class Geometry(C.Structure):      # rtc_geometry
    _fields_ = [("grid", Grid),
                ("cells", C.POINTER(Cell)), ("num_cells", C.c_uint64),
                ("tri_pos", C.POINTER(TriPos)), ("tri_data", C.POINTER(TriData)),
                ("num_tris", C.c_uint64), ("num_src_tris", C.c_uint32),
                ("empty_cells", C.c_uint32), ("min_tris", C.c_uint32), ("max_tris", C.c_uint32),
                ("owner", C.c_void_p)]


class BuildInfo(C.Structure):     # rtc_build_info
    _fields_ = [("grid", Grid), ("num_cells", C.c_uint64), ("num_tris", C.c_uint64),
                ("num_src_tris", C.c_uint32), ("empty_cells", C.c_uint32),
                ("min_tris", C.c_uint32), ("max_tris", C.c_uint32),
                ("device_ms", C.c_float), ("assemble_ms", C.c_float), ("total_ms", C.c_double)]


class Config(C.Structure):        # rtc_config
    _fields_ = [("grid_resolution", c_uint3), ("num_threads", C.c_uint32),
                ("num_samples", C.c_uint16), ("max_bounce", C.c_uint16)]


class HostScene(C.Structure):     # rtc_host_scene
    _fields_ = [("camera", Camera),
                ("positions", C.POINTER(C.c_float)), ("tri_data", C.POINTER(TriData)),
                ("num_tris", C.c_uint32),
                ("materials", C.POINTER(Material)), ("num_materials", C.c_uint32),
                ("num_images", C.c_uint32), ("owner", C.c_void_p)]


# Every symbol include/rtc.h declares (checked by tests/test_abi.py against the built library).
EXPORTED_SYMBOLS = [
    "rtc_scene_create", "rtc_scene_destroy", "rtc_render", "rtc_render_accumulate", "rtc_render_accumulate_static",
    "rtc_resolve", "rtc_scene_stats", "rtc_scene_build_bvh", "rtc_scene_set_stream", "rtc_scene_prepare_gpus",
    "rtc_comm_unique_id", "rtc_comm_create", "rtc_comm_destroy", "rtc_render_sharded", "rtc_geometry_build", "rtc_geometry_free",
    "rtc_geometry_build_device", "rtc_scene_create_from_soup",
    "rtc_scene_create_from_gltf", "rtc_gltf_load_device",
    "rtc_last_error", "rtc_abi_version", "rtc_device_count",
    "rtc_config_load", "rtc_gltf_load", "rtc_host_scene_free", "rtc_png_write", "rtc_png_read",
]
