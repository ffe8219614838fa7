/*
 * rtc.h — C ABI of the B200-native render path for zig_raytracing_contest.
 *
 * This is the drop-in boundary for the reference's one data-parallel hot path:
 *     try scene.render(threads, camera, img);          (src/main.zig:126)
 *     pub fn render(self: Scene, threads: []std.Thread,
 *                   camera: Camera, img: []RGB) !void  (src/stage3.zig:247)
 * The reference's only existing FFI seam is Zig -> C for stb
 * (src/c.zig:2-5, src/c_impl.c:1-5, linked by build.zig:29-48); this header is
 * what a maintainer adds to that @cImport block.
 *
 * Conventions
 *   - plain C, explicit layouts, no Zig @Vector / auto-layout structs cross the ABI;
 *   - every function returns 0 on success, a nonzero rtc_status otherwise, and
 *     leaves a message retrievable with rtc_last_error() (thread-local);
 *   - the caller owns every host pointer it passes; the library copies at
 *     rtc_scene_create and never retains host pointers;
 *   - no CPU fallback: every entry point that computes fails with
 *     RTC_ERR_NO_DEVICE when no sm_100 device is usable.
 */
#ifndef RTC_H_
#define RTC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTC_ABI_VERSION 9u

typedef enum rtc_status {
    RTC_OK = 0,
    RTC_ERR_INVALID_ARGUMENT = 1,
    RTC_ERR_NO_DEVICE = 2,
    RTC_ERR_CUDA = 3,
    RTC_ERR_OUT_OF_MEMORY = 4,
    RTC_ERR_IO = 5,
    RTC_ERR_FORMAT = 6,      /* glTF / config.json contract violation */
    RTC_ERR_UNSUPPORTED = 7
} rtc_status;

/* ---- data the hot path consumes ------------------------------- */

/* linalg.Grid (src/linalg.zig:407-410): bbox + resolution + cell_size = size/res. */
typedef struct rtc_grid {
    float    bbox_min[3];
    float    bbox_max[3];
    uint32_t resolution[3];
    float    cell_size[3];
} rtc_grid;

/* stage3.Cell (src/stage3.zig:131-134): [begin,end) into the baked triangle arrays. */
typedef struct rtc_cell {
    uint32_t begin;
    uint32_t end;
} rtc_cell;

/* linalg.Triangle {v0,e1,e2} (src/linalg.zig:683-694). Vec3 = @Vector(3,f32) is
 * 16 bytes in memory, so the reference's record is 48 bytes; [3] of each row is
 * ignored on input. */
typedef struct rtc_tri_pos {
    float v0[4];
    float e1[4];
    float e2[4];
} rtc_tri_pos;

/* stage3.Triangle.Data (src/stage3.zig:42-50) restated with a fixed layout, plus the
 * source-triangle id that stage2's indices[] holds (src/stage2.zig:151-152) and the
 * reference drops at bake time. src_tri is what parity tests compare. */
typedef struct rtc_tri_data {
    float    normal[3][3];   /* per vertex, world space, as loaded (not renormalised) */
    float    texcoord[3][2];
    uint32_t material;
    uint32_t src_tri;
} rtc_tri_data;

/* stage3.Texture(T) (src/stage3.zig:82-123). channels = 3 (Vec3 texel, tightly packed
 * rgb floats, linear light, factor pre-multiplied) or 1 (f32 texel). */
typedef struct rtc_texture {
    const float* texels;     /* w*h*channels floats, row-major, row 0 first */
    int32_t w, h;
    int32_t u_min, u_max;    /* clamp range applied before the floored mod */
    int32_t v_min, v_max;
    int32_t channels;
} rtc_texture;

/* stage3.Material (src/stage3.zig:125-129). */
typedef struct rtc_material {
    rtc_texture base_color;    /* channels = 3 */
    rtc_texture emissive;      /* channels = 3 */
    rtc_texture transparency;  /* channels = 1; value is the scatter probability */
} rtc_material;

/* stage3.Scene (src/stage3.zig:136-142) after stage2.bakeInto (src/stage2.zig:137-164). */
typedef struct rtc_scene_desc {
    rtc_grid            grid;
    const rtc_cell*     cells;        /* prod(resolution) entries, x fastest (linalg.zig:429) */
    uint64_t            num_cells;
    const rtc_tri_pos*  tri_pos;      /* num_tris baked (duplicated, cell-ordered) records */
    const rtc_tri_data* tri_data;     /* num_tris records, same order */
    uint64_t            num_tris;
    const rtc_material* materials;
    uint32_t            num_materials;
    uint32_t            num_src_tris; /* 1 + max(src_tri); 0 = derive */
} rtc_scene_desc;

/* stage3.Camera (src/stage3.zig:19-35). right/up are per-pixel steps; up points DOWN
 * the image (src/stage1.zig:354-355) so row 0 is the top row. */
typedef struct rtc_camera {
    uint32_t w, h;
    float origin[3];
    float lower_left_corner[3];
    float right[3];
    float up[3];
} rtc_camera;

/* Render modes. The reference has exactly one (PATH); PRIMARY is the deterministic
 * slice used for the hit-id parity gate. */
enum {
    RTC_MODE_PATH = 0,     /* renderWorker + traceRayRecursive (src/stage3.zig:188-245) */
    RTC_MODE_PRIMARY = 1   /* one ray through each pixel centre, depth 1, no RNG:
                              hit -> emissive, miss -> sky; fills rtc_aux hit buffers */
};

/* Traversal selector. Both run the reference's Triangle.rayIntersection (src/linalg.zig:696-722)
 * with identical arithmetic, so t, u, v of a hit are the same bits whichever structure found the triangle.
 *   GRID: the reference's structure and visiting order — results bit-identical to the oracle. The default.
 *   BVH : opt-in 8-wide BVH with quantised 80-byte nodes over the source triangles (no duplication). Same nearest hit
 *         except (a) exact-t ties between two triangles go to the smaller source id instead of the grid's
 *         first-visited, (b) hits the grid's binning / crossing arithmetic loses or finds at the edge of a cell.
 *         tests/test_gpu_bvh.py enumerates both against the oracle ("bit-exact except enumerated
 *         near-ties"). Built on first use, or ahead of time with rtc_scene_build_bvh. */
enum {
    RTC_ACCEL_GRID = 0,    /* uniform grid + 3D-DDA in the reference's cell order; resolution <= 1024 per axis */
    RTC_ACCEL_BVH = 1
};

/* Arithmetic contract. EXACT: no FMA contraction anywhere the reference has separate
 * mul/add, IEEE div/sqrt — bit-identical to the oracle. FAST: same kernels compiled
 * with contraction allowed; hit ids may differ on near-tie pixels only. */
enum {
    RTC_MATH_EXACT = 0,
    RTC_MATH_FAST = 1
};

/* Partition of ONE frame over several GPUs; the scene is replicated on every device. The reference's own
 * split is contiguous pixel ranges per worker thread with no communication (src/stage3.zig:228-233,247-255). */
enum {
    RTC_SHARD_STRIPS = 0,   /* strips of tile_rows (0 = 8) full rows dealt round-robin over the devices (sky rows and
                               geometry rows spread over all of them). Every device resolves its own strips to RGB8.
                               Frame bit-identical to the one-GPU frame. */
    RTC_SHARD_SAMPLES = 1   /* device g renders samples [g*S/N, (g+1)*S/N) of every pixel; the float accumulators are
                               summed in device order. Equal to the one-GPU frame to fp32 summation order (RGB8 +-1). */
};
#define RTC_MAX_GPUS 16

/* main.Config (src/main.zig:56-61) + what the CLI adds. num_samples/max_bounce are the
 * two hot-loop trip counts (src/stage3.zig:237,239). */
typedef struct rtc_params {
    uint16_t num_samples;    /* samples rendered by THIS call */
    uint16_t max_bounce;
    uint32_t mode;           /* RTC_MODE_* */
    uint64_t seed;           /* counter-RNG key; stream = (seed, pixel, sample, draw) */
    uint32_t sample_offset;  /* index of this call's first sample in the frame (sample sharding) */
    uint32_t total_samples;  /* frame total used for the 1/S scale; 0 = num_samples */
    uint32_t accel;          /* RTC_ACCEL_* */
    uint32_t math;           /* RTC_MATH_* */
    uint32_t tile_rank;      /* pixel sharding: this call renders the pixels of rank tile_rank of   */
    uint32_t tile_world;     /* tile_world (0/1 = all pixels); see tile_rows                        */
    uint32_t collect_stats;  /* nonzero: run the instrumented kernels, fill rtc_aux.stats           */
    uint32_t tile_rows;      /* 0: one contiguous band of rows per rank, rows [h*r/n, h*(r+1)/n);   */
                             /* k>0: strips of k full rows dealt round-robin, strip i -> rank i % n */
    int32_t  num_gpus;       /* rtc_render: 0 / 1 = the scene's device. N > 1: this ONE blocking call (the shape of   */
                             /* src/main.zig:126) renders the frame on N devices of this process — the scene's device */
                             /* first, then the other sm_100 devices in index order; replicas are made on first use   */
                             /* by peer copies. With N > 1 the tile_* / sample_offset / total_samples fields must be 0 */
    uint32_t shard;          /* RTC_SHARD_*: the partition num_gpus > 1 and rtc_render_sharded use                    */
} rtc_params;

/* Limits narrower than the reference's types, all reported as RTC_ERR_UNSUPPORTED: max_bounce <= 64 (the reference's
 * field is u16; the kernels keep one queue counter per bounce), grid_resolution <= 1024 per axis (packed cell word),
 * fewer than 2^31 pixels, fewer than 2^32 baked triangles. */

/* Counters behind SURVEY §8(d)'s bytes-per-ray figure. */
typedef struct rtc_stats {
    uint64_t rays;           /* traceRay invocations (camera + bounce + alpha continuation) */
    uint64_t cells_visited;  /* 8-byte Cell records fetched (RTC_ACCEL_BVH: 80-byte nodes)   */
    uint64_t tris_tested;    /* 48-byte {v0,e1,e2} records fetched                           */
    uint64_t hits_shaded;    /* hits that fetched triangle data + 3 textures                 */
    uint64_t paths;          /* camera rays                                                  */
    double   kernel_ms;      /* device time of the render kernels of the last call           */
    uint64_t launches;       /* kernels launched on the device since the counters were last read */
    double   trace_ms;       /* summed device time of the traversal kernel launches in that span  */
    uint32_t trace_launches; /* ... how many launches that sum covers                            */
    uint32_t trace_launches_untimed;
    double   frame_ms;       /* device time of the whole call on this device's stream: render kernels + combining the  */
                             /* shards + resolve, up to (not including) the copy of the frame to the host              */
} rtc_stats;

/* Optional extra outputs (any pointer may be NULL). Host pointers. */
typedef struct rtc_aux {
    uint32_t*  hit_id;    /* w*h: source-triangle id of the primary hit, 0xFFFFFFFF = miss (PRIMARY) */
    float*     hit_tuv;   /* w*h*3: t,u,v of the primary hit (PRIMARY)                               */
    float*     radiance;  /* w*h*3: linear mean radiance before toRGB                                */
    rtc_stats* stats;
} rtc_aux;

typedef struct rtc_scene rtc_scene;   /* opaque; owns device memory */

/* ---- the render path ------------------------------------------------------------- */

/* Uploads a baked scene (what the reference holds in `scene` at src/main.zig:118).
 * device < 0 selects the current CUDA device. */
int rtc_scene_create(const rtc_scene_desc* desc, int device, rtc_scene** out);

void rtc_scene_destroy(rtc_scene* scene);

/* Replaces scene.render(threads, camera, img) (src/main.zig:126, src/stage3.zig:247).
 * rgb_out: w*h*3 bytes, row 0 = top, the exact buffer stbi_write_png is handed at
 * src/main.zig:130-135. Blocking, like the reference (threads joined at stage3.zig:253). */
int rtc_render(rtc_scene* scene, const rtc_camera* camera, const rtc_params* params,
               uint8_t* rgb_out, rtc_aux* aux);

/* ---- one process per GPU (MPI- / torchrun-style hosts): the same sharded frame, combined with NCCL ------------- */

/* A communicator over `world` processes, each driving one device. rank 0 makes an id, the host hands the bytes to every
 * rank by its own means (MPI_Bcast, torch.distributed, a file), every rank calls rtc_comm_create (collective:
 * ncclCommInitRank). NCCL is loaded at run time (libnccl.so.2); RTC_ERR_UNSUPPORTED when it cannot be. */
typedef struct rtc_comm rtc_comm;
#define RTC_COMM_ID_BYTES 128
int  rtc_comm_unique_id(uint8_t id_out[RTC_COMM_ID_BYTES]);
int  rtc_comm_create(const uint8_t id[RTC_COMM_ID_BYTES], int rank, int world, int device, rtc_comm** out);
void rtc_comm_destroy(rtc_comm* comm);

/* scene.render for a frame sharded over the ranks of `comm` (NULL = this process alone). Collective and blocking: every
 * rank passes its replica of the scene, the same camera and the same params (params->shard picks the partition, num_gpus
 * and the tile_* / sample_* fields are ignored). STRIPS: every rank resolves its strips, rank 0 receives them
 * (ncclSend / ncclRecv) and assembles the frame. SAMPLES: one ncclReduce(sum) of the rgb accumulators to rank 0, which
 * resolves. rank 0 gets the frame in rgb_out (host, w*h*3); with rgb_out NULL the finished RGB8 frame stays in device
 * memory of rank 0 and nothing is copied to the host. aux (may be NULL): stats are per rank; radiance / hits on rank 0 are
 * not filled by this entry point. */
int  rtc_render_sharded(rtc_scene* scene, rtc_comm* comm, const rtc_camera* camera, const rtc_params* params,
                        uint8_t* rgb_out, rtc_aux* aux);

/* What the first rtc_render with num_gpus = N (and / or accel) would do before it renders, ahead of time: CUDA contexts
 * on the other devices, the scene's replicas there (peer copies), peer access, the BVH. A host that brackets its render
 * phase (the reference's "Rendered in", src/main.zig:125-127) calls this in its "Compiled in" phase; optional. */
int  rtc_scene_prepare_gpus(rtc_scene* scene, int num_gpus, uint32_t accel);

/* Makes `stream` (a cudaStream_t; NULL = back to the scene's own stream) the stream every later call on this scene
 * launches on — a host that brackets frames with its own CUDA events, or orders them against its own work, passes its
 * stream here. A scene allows ONE stream in flight at a time: its ray queues, counters and timer events are per scene,
 * so calls on one scene must not overlap (this holds for rtc_render_accumulate's stream argument too). */
int  rtc_scene_set_stream(rtc_scene* scene, void* stream);

/* ---- device-resident pieces of the same path (multi-GPU plumbing, benchmarking) --- */

/* Adds this call's samples into d_accum (device, w*h float4: rgb sum + sample count),
 * asynchronously on `stream` (a cudaStream_t, NULL = default stream). Calls on one scene must not overlap (see
 * rtc_scene_set_stream): the wavefront scratch is per scene. */
int rtc_render_accumulate(rtc_scene* scene, const rtc_camera* camera, const rtc_params* params,
                          void* d_accum, void* stream);

/* rtc_render_accumulate for callers that cannot allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc). NO COUNTERPART IN THE REFERENCE.
 * Same arguments, same kernels and bit-identical sums, but the wavefront scratch (ray queues, hit records, throughput
 * stacks, queue counters) is a static device array of RTC_STATIC_SCRATCH_BYTES inside librtc.so — every GPU that loads
 * the library holds those bytes until the process exits — and the call makes no cudaMalloc / cudaFree / allocating
 * library call. The caller passes the inputs (scene, camera, params) and the output (d_accum, device, w*h float4).
 * Limit: this call's paths = (pixels of the call's band or strips) x num_samples must fit the array in ONE batch at
 * 152 + 24 * max_bounce bytes per path: 270 k paths at max_bounce 4 (e.g. 256 x 256 at 4 samples), 39 k at max_bounce 64;
 * more -> RTC_ERR_UNSUPPORTED with the numbers in rtc_last_error(), nothing rendered (split the frame with
 * tile_rank / tile_world or sample_offset, or use rtc_render_accumulate). RTC_ACCEL_BVH needs rtc_scene_build_bvh to
 * have run (the build allocates); num_gpus > 1 is rejected. The array is shared by EVERY scene and every call on a
 * device and keeps the last call's contents (the call clears the part that must start at zero, on `stream`): calls
 * that overlap on one GPU — two streams, two scenes, two threads — race; order them on one stream or with events. */
#define RTC_STATIC_SCRATCH_BYTES (64u << 20)
int rtc_render_accumulate_static(rtc_scene* scene, const rtc_camera* camera, const rtc_params* params,
                                 void* d_accum, void* stream);

/* pixel.mul(inv_num_samples).toRGB() (src/stage3.zig:242, src/linalg.zig:150-159):
 * d_accum (w*h float4) -> d_rgb (w*h*3 bytes) and/or d_radiance (w*h*3 floats). */
int rtc_resolve(const void* d_accum, uint32_t num_pixels, uint32_t total_samples,
                void* d_rgb, void* d_radiance, int device, void* stream);

/* Fetches and clears the ray/traversal counters accumulated on the scene's device — and, after num_gpus > 1 renders,
 * on its replicas' devices (summed) — since the last call (synchronises those devices). */
int rtc_scene_stats(rtc_scene* scene, rtc_stats* out);

/* RTC_ACCEL_BVH's structure: built on the host threads (num_threads < 1 = all) from the scene's own device-resident
 * {v0,e1,e2} records — whichever way the scene came to exist — and uploaded next to the grid. Idempotent. */
typedef struct rtc_bvh_info {
    uint32_t num_nodes;      /* 80-byte nodes                                   */
    uint32_t num_tris;       /* triangles referenced (each source triangle once) */
    uint32_t depth;          /* levels of inner nodes                            */
    uint32_t reserved;
    uint64_t bytes;          /* device bytes of nodes + triangle records         */
    double   build_ms;       /* wall time of download + build + upload           */
} rtc_bvh_info;
int rtc_scene_build_bvh(rtc_scene* scene, int num_threads, rtc_bvh_info* info /* may be NULL */);

/* ---- stage2 as the path's producer ------------------------------ */

/* Geometry.build + bakeInto (src/stage2.zig:131-164) from a triangle soup:
 * positions = num_tris*9 floats (3 world-space vertices), data = num_tris records whose
 * src_tri is ignored on input. Outputs are library-owned and live until rtc_geometry_free. */
typedef struct rtc_geometry {
    rtc_grid      grid;
    rtc_cell*     cells;
    uint64_t      num_cells;
    rtc_tri_pos*  tri_pos;
    rtc_tri_data* tri_data;
    uint64_t      num_tris;      /* baked count (sum over cells) */
    uint32_t      num_src_tris;
    uint32_t      empty_cells;
    uint32_t      min_tris, max_tris;   /* over non-empty cells (src/stage2.zig:76-100) */
    void*         owner;
} rtc_geometry;

int  rtc_geometry_build(const float* positions, const rtc_tri_data* data, uint32_t num_tris,
                        const uint32_t grid_resolution[3], rtc_geometry* out);
void rtc_geometry_free(rtc_geometry* g);

/* The same producer on the GPU (sm_100a kernels: bounding box, per-triangle cell binning with the
 * reference's triangle/box overlap test, scan, sort, bake). Outputs are bit-identical to
 * rtc_geometry_build's and are freed with rtc_geometry_free. device < 0 = current device. */
int  rtc_geometry_build_device(const float* positions, const rtc_tri_data* data, uint32_t num_tris,
                               const uint32_t grid_resolution[3], int device, rtc_geometry* out);

/* What the device build reports (the reference logs the same figures, src/stage2.zig:97-100,126-128). */
typedef struct rtc_build_info {
    rtc_grid grid;
    uint64_t num_cells;
    uint64_t num_tris;       /* baked count */
    uint32_t num_src_tris;
    uint32_t empty_cells;
    uint32_t min_tris, max_tris;
    float    device_ms;      /* CUDA-event time of the grid-build kernels (bbox, binning, scan, sort)         */
    float    assemble_ms;    /* CUDA-event time of triangle assembly + texel linearisation (rtc_scene_create_from_gltf), else 0 */
    double   total_ms;       /* wall time of the whole call: upload + build + bake + textures    */
} rtc_build_info;

/* geometry.build() + geometry.bakeInto(&scene) (src/main.zig:117-118) + rtc_scene_create in one step, with
 * the grid built and baked ON the device: the soup is uploaded once and the baked arrays never exist on the
 * host. The resulting scene renders bit-identically to rtc_scene_create(rtc_geometry_build(...)).
 * info may be NULL. */
int  rtc_scene_create_from_soup(const float* positions, const rtc_tri_data* data, uint32_t num_tris,
                                const uint32_t grid_resolution[3],
                                const rtc_material* materials, uint32_t num_materials,
                                int device, rtc_scene** out, rtc_build_info* info);

/* ---- stage1 + main.zig's file contracts as the path's callers ---------
 * Host-only (no GPU needed). These are what a host that is NOT the Zig program uses to honour the
 * reference's input/output contract; the Zig host keeps zgltf + stb and needs none of them. */

/* main.Config (src/main.zig:56-69): config.json, read from the cwd by the reference. */
typedef struct rtc_config {
    uint32_t grid_resolution[3];
    uint32_t num_threads;      /* 0 = JSON null = all cores (src/main.zig:90) */
    uint16_t num_samples;
    uint16_t max_bounce;
} rtc_config;
int rtc_config_load(const char* path, rtc_config* out);

/* What stage1 leaves in memory (src/main.zig:104-112): the camera, the material textures and the
 * world-space triangle soup. Library-owned until rtc_host_scene_free. */
typedef struct rtc_host_scene {
    rtc_camera    camera;
    float*        positions;     /* num_tris * 9: three world-space vertices per triangle          */
    rtc_tri_data* tri_data;      /* num_tris records (src_tri = index in this soup)                */
    uint32_t      num_tris;
    rtc_material* materials;     /* num_materials, texels owned by the library                      */
    uint32_t      num_materials;
    uint32_t      num_images;
    void*         owner;
} rtc_host_scene;

/* loadGltfFile + loadCamera + loadMaterials + loadGeometry (src/stage1.zig:75,261,309,485) for a
 * .gltf (+ external buffers / PNG images / data: URIs) or .glb file. camera_name NULL = the first
 * camera; width/height 0 = not given (size rules of src/stage1.zig:322-342). PNG textures are
 * linearised like stb's stbi_loadf (pow(c/255.0f, 2.2f) in double, alpha linear). PNG (all colour types and
 * depths, Adam7) and JPEG (baseline / progressive Huffman; grey, YCbCr, RGB, CMYK / YCCK) decode to the bytes stb produces;
 * other formats are RTC_ERR_UNSUPPORTED. */
int  rtc_gltf_load(const char* path, const char* camera_name, uint32_t width, uint32_t height,
                   int num_threads, rtc_host_scene* out);
void rtc_host_scene_free(rtc_host_scene* scene);

/* The same stage1 with loadTriangles' per-triangle loop (src/stage1.zig:235-255: index fetch, node transform
 * of positions and normals, normalize, texcoords) ON the device: the file's buffers and one table row per
 * primitive are uploaded, one sm_100a kernel writes the soup; texture images go up as decoded 8-bit RGBA and are
 * linearised there (stb's pow(c/255.0f, 2.2f) table times the material factor, MASK / BLEND alpha; src/stage1.zig:427-472).
 * Needs a GPU; results are bit-identical to rtc_gltf_load's (an out-of-range vertex index is the same RTC_ERR_FORMAT).
 *
 * rtc_scene_create_from_gltf = loadGltfFile .. bakeInto (src/main.zig:104-118) in one call: the soup never
 * exists on the host, the grid is built and baked on the device from it (as rtc_scene_create_from_soup), and
 * the scene renders bit-identically to rtc_gltf_load + rtc_geometry_build + rtc_scene_create. camera receives
 * loadCamera's result (may be NULL); info may be NULL.
 * rtc_gltf_load_device returns the device-assembled soup to the host in rtc_gltf_load's layout (parity checks,
 * hosts that want the soup); free with rtc_host_scene_free. */
int  rtc_scene_create_from_gltf(const char* path, const char* camera_name, uint32_t width, uint32_t height,
                                int num_threads, const uint32_t grid_resolution[3], int device,
                                rtc_scene** out, rtc_camera* camera, rtc_build_info* info);
int  rtc_gltf_load_device(const char* path, const char* camera_name, uint32_t width, uint32_t height,
                          int num_threads, int device, rtc_host_scene* out);

/* stbi_write_png(path, w, h, 3, rgb, 3*w) of src/main.zig:130-135 (8-bit RGB, row 0 = top).
 * Returns RTC_OK / RTC_ERR_IO (stb returns 1 / 0). Default: zlib slices on all host threads (same pixels, smaller
 * file, ~10x faster at 4K); environment RTC_PNG_STB=1: the very bytes stb writes (its filter choice and deflate). */
int  rtc_png_write(const char* path, uint32_t w, uint32_t h, const uint8_t* rgb);
/* Decodes a PNG file to 8-bit RGB (the inverse, for tests and tools). rgb_out: malloc'ed, free with free(). */
int  rtc_png_read(const char* path, uint32_t* w, uint32_t* h, uint8_t** rgb_out);

/* ---- misc ------------------------------------------------------------------------ */

const char* rtc_last_error(void);
uint32_t    rtc_abi_version(void);
/* Number of usable sm_100 devices (0 when none; never an error). */
int         rtc_device_count(void);

#ifdef __cplusplus
}
#endif
#endif /* RTC_H_ */
