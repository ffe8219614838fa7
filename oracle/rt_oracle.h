/*
 * rt_oracle.h — TEST INFRASTRUCTURE, NOT PRODUCT.
 *
 * CPU restatement (plain C) of the reference's per-pixel path-tracing hot path and of
 * the host code that produces its inputs. Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library; the product
 * (zig_raytracing_contest_b200/csrc, include/rtc.h) never links or calls it.
 *
 * Follows, function by function:
 *   /root/reference/src/linalg.zig  :13-222 (Vec), :280-350 (Ray, Bbox), :407-498 (Grid,
 *                                    Iterator), :500-563 (SAT), :683-723 (Triangle)
 *   /root/reference/src/stage3.zig  whole file (Camera, Texture, Scene.traceRay,
 *                                    traceRayRecursive, renderWorker, render)
 *   /root/reference/src/stage2.zig  :44-164 (grid build + bake)
 *   /root/reference/src/stage1.zig  :217-259 (triangle assembly), :309-371 (camera)
 *   /root/reference/libs/zgltf/src/helpers.zig :17-99, main.zig :248-280 (node transforms)
 *
 * PARITY PINNING. The reference is Zig and no Zig toolchain exists on the build or GPU
 * machines, so the reference binary cannot be run. What the reference's own tests pin
 * (13 unit tests in src/linalg.zig: slab test, DDA walks incl. tie order and +-inf
 * lanes, cross, length, getCellBbox, RGB size) is checked in tests/test_oracle_golden.py.
 * Everything else (Moller-Trumbore, textures, shading, tonemap, any image) has no
 * golden vector in the reference: for those pieces PARITY IS UNPINNED and this file is
 * the specification by construction from the cited lines. The reference's one C unit
 * (src/c_impl.c, stb) does build here (oracle/Makefile `ref` -> oracle/_ref/libc_libs.so);
 * it pins texture decode + linearisation and PNG output (tests/test_ref_stb.py), which
 * are inputs and outputs of the path, not the path.
 *
 * Arithmetic living in the un-vendored Zig standard library (version unpinned, 0.11 era):
 *   std.rand.DefaultPrng (Xoshiro256++), Random.float, Random.floatNorm (ziggurat):
 *       replaced by a counter-based Philox4x32-10 keyed by (seed; pixel, sample, block)
 *       and an exact-arithmetic uniform-sphere map; the reference's stream depends on the
 *       worker-thread count (stage3.zig:225-229) and is not reproducible by design.
 *   std.math.pow : C powf (+-1 LSB in RGB8 against the real reference at bin edges).
 *   std.math.lerp: @mulAdd(b-a, t, a) == fmaf(b-a, t, a).
 *
 * Undefined corners of the reference that this restatement DEFINES:
 *   - @intFromFloat of NaN / out-of-range: saturating, NaN -> 0 (CUDA cvt.rzi semantics);
 *   - grid wholly behind the ray (slab tmax < 0): miss (A6);
 *   - @max/@min with a NaN operand return the other operand (fmaxf/fminf; LLVM maxnum).
 *
 * Build: see oracle/Makefile (-O3 -ffp-contract=off -fno-fast-math; explicit fmaf only
 * where the reference calls @mulAdd).
 */
#ifndef RT_ORACLE_H_
#define RT_ORACLE_H_

#include <stdint.h>
#include "../include/rtc.h"   /* struct layouts only; no product code is linked */

#ifdef __cplusplus
extern "C" {
#endif

/* ---- linalg.zig unit-level entry points (golden-vector tests) ---- */
void  orc_cross(const float a[3], const float b[3], float out[3]);
float orc_length(const float v[3]);
void  orc_normalize(const float v[3], float out[3]);
int   orc_bbox_ray_intersection(const float bmin[3], const float bmax[3],
                                const float orig[3], const float dir[3], float* t);
void  orc_grid_init(const float bmin[3], const float bmax[3], const uint32_t res[3], rtc_grid* g);
void  orc_grid_cell_bbox(const rtc_grid* g, uint32_t x, uint32_t y, uint32_t z,
                         float out_min[3], float out_max[3]);
void  orc_grid_cell_idx(const rtc_grid* g, const float p[3], uint32_t out[3]);
/* Grid.traceRay + repeated Iterator.next: writes the cell BEFORE each next() and the
 * value next() returned; stops after the next() that returned +inf or at max_steps.
 * Returns the number of (cell, t) pairs written, or -1 when traceRay returns null. */
int   orc_grid_walk(const rtc_grid* g, const float orig[3], const float dir[3],
                    int max_steps, uint32_t* cells_xyz, float* t_next);
/* The Iterator Grid.traceRay constructs: out_u = cell[3], exit[3], step[3]; out_f = t_delta[3], t_next[3]. */
int   orc_grid_iter_init(const rtc_grid* g, const float orig[3], const float dir[3], uint32_t out_u[9], float out_f[6]);
int   orc_triangle_ray_intersection(const float v0[3], const float e1[3], const float e2[3],
                                    const float orig[3], const float dir[3],
                                    float* t, float* u, float* v);
int   orc_triangle_aabb(const float tri[9], const float bmin[3], const float bmax[3]);
void  orc_texture_sample(const rtc_texture* tex, float u, float v, float* out /* channels */);
/* Triangle.Data.interpolate (stage3.zig:53-71) for both field types */
void  orc_interpolate(const rtc_tri_data* d, float u, float v, float normal_out[3], float uv_out[2]);
void  orc_to_rgb(const float c[3], uint8_t out[3]);
void  orc_env_color(const float dir[3], float out[3]);
void  orc_camera_get_ray(const rtc_camera* cam, float x, float y, float orig[3], float dir[3]);

/* ---- RNG (shared definition with the CUDA path; see header comment) ---- */
void  orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
void  orc_rng_block(uint64_t seed, uint32_t pixel, uint32_t sample, uint32_t block, float out[4]);
void  orc_unit_vector(float u1, float u2, float out[3]);

/* ---- stage1 math ---- */
void  orc_mat4_mul(const float left[16], const float right[16], float out[16]);
void  orc_mat4_recompose(const float t[3], const float r[4], const float s[3], float out[16]);
/* loadTriangles inner loop for one primitive (stage1.zig:235-255): indices are u16. */
void  orc_assemble_triangles(const float matrix[16], const float* positions, const float* normals,
                             const float* texcoords /* may be NULL */, const uint16_t* indices,
                             uint32_t num_tris, uint32_t material,
                             float* out_positions /* 9/tri */, rtc_tri_data* out_data);
/* loadCamera (stage1.zig:309-371). width/height: 0 = not given. aspect <= 0 = absent.
 * Returns 0 or a negative error mirroring the reference's error set. */
int   orc_load_camera(const float matrix[16], float yfov, float aspect_ratio,
                      uint32_t width, uint32_t height, rtc_camera* out);

/* ---- stage2 ---- */
typedef struct orc_geometry {
    rtc_grid      grid;
    rtc_cell*     cells;
    uint64_t      num_cells;
    uint64_t*     indices;     /* stage2 indices[] (source triangle per baked slot) */
    rtc_tri_pos*  tri_pos;
    rtc_tri_data* tri_data;
    uint64_t      num_tris;
    uint32_t      num_src_tris;
    uint32_t      empty_cells, min_tris, max_tris;
} orc_geometry;
int   orc_geometry_build(const float* positions, const rtc_tri_data* data, uint32_t num_tris,
                         const uint32_t res[3], orc_geometry* out);
void  orc_geometry_free(orc_geometry* g);

/* ---- stage3 ---- */
/* Scene.traceRay (stage3.zig:152-186): returns baked triangle index or -1. */
int64_t orc_trace_ray(const rtc_scene_desc* scene, const float orig[3], const float dir[3],
                      float* t, float* u, float* v, rtc_stats* stats /* may be NULL */);
/* Brute-force nearest hit over ALL baked triangles in source order (cfg2's check). */
int64_t orc_trace_ray_brute(const rtc_scene_desc* scene, const float orig[3], const float dir[3],
                            float* t, float* u, float* v);
/* Scene.render with the reference's static contiguous partition over num_threads
 * (stage3.zig:228-233,247-256). Restricting to rows [y0,y1) (y1=0: all) renders a bounded
 * sample of the frame with identical per-pixel results. */
int   orc_render(const rtc_scene_desc* scene, const rtc_camera* camera, const rtc_params* params,
                 int num_threads, uint32_t y0, uint32_t y1, uint8_t* rgb_out, rtc_aux* aux);

#ifdef __cplusplus
}
#endif
#endif
