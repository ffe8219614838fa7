/*
 * rt_oracle.c — TEST INFRASTRUCTURE, NOT PRODUCT. See rt_oracle.h for scope, the
 * reference lines followed, the parity-pinning statement and the defined UB corners.
 *
 * Every function below names the reference lines it restates. Operation order is the
 * reference's; the file must be compiled with -ffp-contract=off so that no mul/add pair
 * is fused except where the reference itself calls @mulAdd (explicit fmaf here).
 */
#include "rt_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------
 * Vec3 (linalg.zig:13-222). @Vector(3,f32): lane-wise ops; @reduce(.Add) keeps lane order.
 * ---------------------------------------------------------------------------------------- */
typedef struct { float x, y, z; } v3;

static inline v3 v3_make(float x, float y, float z) { v3 r = { x, y, z }; return r; }
static inline v3 v3_splat(float s) { return v3_make(s, s, s); }
static inline v3 v3_from(const float a[3]) { return v3_make(a[0], a[1], a[2]); }
static inline void v3_to(v3 a, float out[3]) { out[0] = a.x; out[1] = a.y; out[2] = a.z; }
/* linalg.zig:181-187 */
static inline v3 v3_add(v3 a, v3 b) { return v3_make(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 v3_sub(v3 a, v3 b) { return v3_make(a.x - b.x, a.y - b.y, a.z - b.z); }
/* linalg.zig:197-203 */
static inline v3 v3_mul(v3 a, v3 b) { return v3_make(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline v3 v3_div(v3 a, v3 b) { return v3_make(a.x / b.x, a.y / b.y, a.z / b.z); }
/* linalg.zig:62-64 */
static inline v3 v3_scale(v3 a, float s) { return v3_make(a.x * s, a.y * s, a.z * s); }
/* linalg.zig:50-56: @max/@min -> maxnum/minnum */
static inline v3 v3_max(v3 a, v3 b) { return v3_make(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }
static inline v3 v3_min(v3 a, v3 b) { return v3_make(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)); }
static inline v3 v3_abs(v3 a) { return v3_make(fabsf(a.x), fabsf(a.y), fabsf(a.z)); }
/* linalg.zig:189-191: ordered 3-term sum */
static inline float v3_dot(v3 a, v3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
/* linalg.zig:172-179: a.yzx*b.zxy - a.zxy*b.yzx */
static inline v3 v3_cross(v3 a, v3 b)
{
    return v3_make(a.y * b.z - a.z * b.y,
                   a.z * b.x - a.x * b.z,
                   a.x * b.y - a.y * b.x);
}
/* linalg.zig:119-125 */
static inline float v3_length(v3 a) { return sqrtf(v3_dot(a, a)); }
static inline v3 v3_normalize(v3 a) { return v3_scale(a, 1.0f / v3_length(a)); }

/* Defined semantics for @intFromFloat (rt_oracle.h): saturating, NaN -> 0. */
static inline uint32_t f2u_sat(float f)
{
    if (!(f > 0.0f)) return 0u;                 /* NaN, negatives, zero */
    if (f >= 4294967296.0f) return 0xFFFFFFFFu;
    return (uint32_t)f;
}
static inline int32_t f2i_sat(float f)
{
    if (f != f) return 0;
    if (f >= 2147483648.0f) return INT32_MAX;
    if (f <= -2147483648.0f) return INT32_MIN;
    return (int32_t)f;
}

void orc_cross(const float a[3], const float b[3], float out[3]) { v3_to(v3_cross(v3_from(a), v3_from(b)), out); }
float orc_length(const float v[3]) { return v3_length(v3_from(v)); }
void orc_normalize(const float v[3], float out[3]) { v3_to(v3_normalize(v3_from(v)), out); }

/* ------------------------------------------------------------------------------------------
 * Ray, Bbox (linalg.zig:280-350)
 * ---------------------------------------------------------------------------------------- */
typedef struct { v3 orig, dir; } ray_t;

/* linalg.zig:284-286 */
static inline v3 ray_at(ray_t r, float t) { return v3_add(r.orig, v3_scale(r.dir, t)); }

/* linalg.zig:324-349. Also reports tmax so Grid.traceRay can apply the defined
 * "grid behind the ray" rule; the reference computes tmax and discards it. */
static int bbox_ray_intersection(v3 bmin, v3 bmax, ray_t ray, float* t, float* t_far)
{
    const int sx = ray.dir.x < 0.0f, sy = ray.dir.y < 0.0f, sz = ray.dir.z < 0.0f;
    const v3 near_ = v3_make(sx ? bmax.x : bmin.x, sy ? bmax.y : bmin.y, sz ? bmax.z : bmin.z);
    const v3 far_  = v3_make(sx ? bmin.x : bmax.x, sy ? bmin.y : bmax.y, sz ? bmin.z : bmax.z);
    const v3 mn = v3_div(v3_sub(near_, ray.orig), ray.dir);
    const v3 mx = v3_div(v3_sub(far_, ray.orig), ray.dir);

    float tmin = mn.x;
    float tmax = mx.x;
    if ((tmin > mx.y) || (tmax < mn.y)) return 0;
    tmin = fmaxf(tmin, mn.y);
    tmax = fminf(tmax, mx.y);
    if ((tmin > mx.z) || (tmax < mn.z)) return 0;
    tmin = fmaxf(tmin, mn.z);
    tmax = fminf(tmax, mx.z);
    *t = tmin;
    if (t_far) *t_far = tmax;
    return 1;
}

int orc_bbox_ray_intersection(const float bmin[3], const float bmax[3],
                              const float orig[3], const float dir[3], float* t)
{
    ray_t r = { v3_from(orig), v3_from(dir) };
    return bbox_ray_intersection(v3_from(bmin), v3_from(bmax), r, t, NULL);
}

/* ------------------------------------------------------------------------------------------
 * Grid + Iterator (linalg.zig:407-498)
 * ---------------------------------------------------------------------------------------- */
/* linalg.zig:412-418 */
void orc_grid_init(const float bmin[3], const float bmax[3], const uint32_t res[3], rtc_grid* g)
{
    for (int i = 0; i < 3; ++i) {
        g->bbox_min[i] = bmin[i];
        g->bbox_max[i] = bmax[i];
        g->resolution[i] = res[i];
        g->cell_size[i] = (bmax[i] - bmin[i]) / (float)res[i];
    }
}

/* linalg.zig:424-427 */
void orc_grid_cell_idx(const rtc_grid* g, const float p[3], uint32_t out[3])
{
    for (int i = 0; i < 3; ++i) {
        uint32_t c = f2u_sat((p[i] - g->bbox_min[i]) / g->cell_size[i]);
        uint32_t lim = g->resolution[i] - 1u;
        out[i] = c < lim ? c : lim;
    }
}

/* linalg.zig:429-431 */
static inline uint64_t grid_linear(const rtc_grid* g, uint64_t x, uint64_t y, uint64_t z)
{
    return z * g->resolution[0] * g->resolution[1] + y * g->resolution[0] + x;
}

/* linalg.zig:433-441 */
void orc_grid_cell_bbox(const rtc_grid* g, uint32_t x, uint32_t y, uint32_t z,
                        float out_min[3], float out_max[3])
{
    const float f[3] = { (float)x, (float)y, (float)z };
    for (int i = 0; i < 3; ++i) {
        out_min[i] = g->bbox_min[i] + g->cell_size[i] * f[i];
        out_max[i] = out_min[i] + g->cell_size[i];
    }
}

typedef struct {
    uint32_t cell[3];
    uint32_t exit_[3];
    uint32_t step[3];
    float t_delta[3];
    float t_next_crossing[3];
} grid_iter;

/* linalg.zig:443-469. Returns 0 for the reference's `null`. */
static int grid_trace_ray(const rtc_grid* g, ray_t ray, grid_iter* it)
{
    float t_hit, t_far;
    const v3 bmin = v3_from(g->bbox_min), bmax = v3_from(g->bbox_max);
    if (!bbox_ray_intersection(bmin, bmax, ray, &t_hit, &t_far)) return 0;
    if (t_far < 0.0f) return 0;   /* DEFINED corner: grid wholly behind the ray (see header) */
    t_hit = fmaxf(0.0f, t_hit);

    const float dir[3] = { ray.dir.x, ray.dir.y, ray.dir.z };
    const v3 cs = v3_from(g->cell_size);
    const v3 t_delta = v3_abs(v3_div(cs, ray.dir));
    const v3 local = v3_sub(ray_at(ray, t_hit), bmin);
    const v3 cellf = v3_div(local, cs);
    const float cellf_a[3] = { cellf.x, cellf.y, cellf.z };
    uint32_t cell[3];
    float next_cell[3];
    for (int i = 0; i < 3; ++i) {
        const int sign = dir[i] < 0.0f;
        const uint32_t lim = g->resolution[i] - 1u;
        uint32_t c = f2u_sat(cellf_a[i]);
        cell[i] = c < lim ? c : lim;
        it->cell[i] = cell[i];
        it->step[i] = sign ? 0xFFFFFFFFu : 1u;
        it->exit_[i] = sign ? 0u : lim;
        next_cell[i] = (float)(cell[i] + (sign ? 0u : 1u));
    }
    const v3 tn = v3_add(v3_splat(t_hit),
                         v3_div(v3_sub(v3_mul(v3_from(next_cell), cs), local), ray.dir));
    it->t_delta[0] = t_delta.x; it->t_delta[1] = t_delta.y; it->t_delta[2] = t_delta.z;
    it->t_next_crossing[0] = tn.x; it->t_next_crossing[1] = tn.y; it->t_next_crossing[2] = tn.z;
    return 1;
}

/* linalg.zig:478-497 */
static float grid_iter_next(grid_iter* it)
{
    static const uint8_t map[8] = { 2, 1, 2, 1, 2, 2, 0, 0 };
    const unsigned k =
        ((unsigned)(it->t_next_crossing[0] < it->t_next_crossing[1]) << 2) +
        ((unsigned)(it->t_next_crossing[0] < it->t_next_crossing[2]) << 1) +
        ((unsigned)(it->t_next_crossing[1] < it->t_next_crossing[2]));
    const unsigned axis = map[k];
    if (it->cell[axis] == it->exit_[axis]) return INFINITY;
    const float t = it->t_next_crossing[axis];
    it->cell[axis] += it->step[axis];            /* wrapping add (linalg.zig:493, test :565-569) */
    it->t_next_crossing[axis] += it->t_delta[axis];
    return t;
}

/* Grid.traceRay's Iterator as constructed (linalg.zig:443-469): 5 x 3 words per ray =
 * cell, exit, step (1 / 0xFFFFFFFF), t_delta, t_next_crossing. Returns 0 when traceRay returns null. */
int orc_grid_iter_init(const rtc_grid* g, const float orig[3], const float dir[3], uint32_t out_u[9], float out_f[6])
{
    ray_t r = { v3_from(orig), v3_from(dir) };
    grid_iter it;
    if (!grid_trace_ray(g, r, &it)) return 0;
    for (int i = 0; i < 3; ++i) {
        out_u[i] = it.cell[i]; out_u[3 + i] = it.exit_[i]; out_u[6 + i] = it.step[i];
        out_f[i] = it.t_delta[i]; out_f[3 + i] = it.t_next_crossing[i];
    }
    return 1;
}

int orc_grid_walk(const rtc_grid* g, const float orig[3], const float dir[3],
                  int max_steps, uint32_t* cells_xyz, float* t_next)
{
    ray_t r = { v3_from(orig), v3_from(dir) };
    grid_iter it;
    if (!grid_trace_ray(g, r, &it)) return -1;
    int n = 0;
    while (n < max_steps) {
        cells_xyz[3 * n + 0] = it.cell[0];
        cells_xyz[3 * n + 1] = it.cell[1];
        cells_xyz[3 * n + 2] = it.cell[2];
        const float t = grid_iter_next(&it);
        t_next[n] = t;
        ++n;
        if (t == INFINITY) break;
    }
    return n;
}

/* ------------------------------------------------------------------------------------------
 * Triangle (linalg.zig:683-723) and triangle/AABB SAT (linalg.zig:500-563)
 * ---------------------------------------------------------------------------------------- */
/* linalg.zig:696-722 */
static inline int triangle_ray_intersection(v3 v0, v3 e1, v3 e2, ray_t ray, float* t, float* u_, float* v_)
{
    const v3 pvec = v3_cross(ray.dir, e2);
    const float det = v3_dot(e1, pvec);
    const float epsilon = 0.00000001f;
    if (det < epsilon) return 0;
    const float inv_det = 1.0f / det;
    const v3 tvec = v3_sub(ray.orig, v0);
    const float u = v3_dot(tvec, pvec) * inv_det;
    if (u < 0.0f || u > 1.0f) return 0;
    const v3 qvec = v3_cross(tvec, e1);
    const float v = v3_dot(ray.dir, qvec) * inv_det;
    if (v < 0.0f || u + v > 1.0f) return 0;
    *t = v3_dot(e2, qvec) * inv_det;
    *u_ = u;
    *v_ = v;
    return 1;
}

int orc_triangle_ray_intersection(const float v0[3], const float e1[3], const float e2[3],
                                  const float orig[3], const float dir[3],
                                  float* t, float* u, float* v)
{
    ray_t r = { v3_from(orig), v3_from(dir) };
    return triangle_ray_intersection(v3_from(v0), v3_from(e1), v3_from(e2), r, t, u, v);
}

/* linalg.zig:500-515 */
static int sat_axis(v3 a, v3 b, v3 c, v3 extents, v3 axis)
{
    const float p0 = v3_dot(a, axis);
    const float p1 = v3_dot(b, axis);
    const float p2 = v3_dot(c, axis);
    const float r = extents.x * fabsf(v3_dot(v3_make(1, 0, 0), axis)) +
                    extents.y * fabsf(v3_dot(v3_make(0, 1, 0), axis)) +
                    extents.z * fabsf(v3_dot(v3_make(0, 0, 1), axis));
    const float maxP = fmaxf(p0, fmaxf(p1, p2));
    const float minP = fminf(p0, fminf(p1, p2));
    return !(fmaxf(-maxP, minP) > r);
}

/* linalg.zig:517-563 */
static int triangle_aabb(v3 t0, v3 t1, v3 t2, v3 bmin, v3 bmax)
{
    const v3 center = v3_scale(v3_add(bmax, bmin), 0.5f);      /* linalg.zig:315-317 */
    const v3 extents = v3_scale(v3_sub(bmax, bmin), 0.5f);     /* linalg.zig:319-321 */
    const v3 a = v3_sub(t0, center);
    const v3 b = v3_sub(t1, center);
    const v3 c = v3_sub(t2, center);
    const v3 ab = v3_normalize(v3_sub(b, a));
    const v3 bc = v3_normalize(v3_sub(c, b));
    const v3 ca = v3_normalize(v3_sub(a, c));
    const v3 a00 = v3_make(0.0f, -ab.z, ab.y);
    const v3 a01 = v3_make(0.0f, -bc.z, bc.y);
    const v3 a02 = v3_make(0.0f, -ca.z, ca.y);
    const v3 a10 = v3_make(ab.z, 0.0f, -ab.x);
    const v3 a11 = v3_make(bc.z, 0.0f, -bc.x);
    const v3 a12 = v3_make(ca.z, 0.0f, -ca.x);
    const v3 a20 = v3_make(-ab.y, ab.x, 0.0f);
    const v3 a21 = v3_make(-bc.y, bc.x, 0.0f);
    const v3 a22 = v3_make(-ca.y, ca.x, 0.0f);
    if (!sat_axis(a, b, c, extents, a00) ||
        !sat_axis(a, b, c, extents, a01) ||
        !sat_axis(a, b, c, extents, a02) ||
        !sat_axis(a, b, c, extents, a10) ||
        !sat_axis(a, b, c, extents, a11) ||
        !sat_axis(a, b, c, extents, a12) ||
        !sat_axis(a, b, c, extents, a20) ||
        !sat_axis(a, b, c, extents, a21) ||
        !sat_axis(a, b, c, extents, a22) ||
        !sat_axis(a, b, c, extents, v3_make(1, 0, 0)) ||
        !sat_axis(a, b, c, extents, v3_make(0, 1, 0)) ||
        !sat_axis(a, b, c, extents, v3_make(0, 0, 1)) ||
        !sat_axis(a, b, c, extents, v3_cross(ab, bc)))
        return 0;
    return 1;
}

int orc_triangle_aabb(const float tri[9], const float bmin[3], const float bmax[3])
{
    return triangle_aabb(v3_from(tri), v3_from(tri + 3), v3_from(tri + 6), v3_from(bmin), v3_from(bmax));
}

/* ------------------------------------------------------------------------------------------
 * Texture(T).sample (stage3.zig:82-123)
 * ---------------------------------------------------------------------------------------- */
static inline float tex_frac(float v) { return fabsf(v - truncf(v)); }               /* :93-95 */
static inline int32_t clamp_i32(int32_t v, int32_t lo, int32_t hi)                   /* std.math.clamp */
{
    const int32_t m = v < hi ? v : hi;
    return lo > m ? lo : m;
}
static inline int32_t floored_mod(int32_t a, int32_t n)                              /* @mod */
{
    int32_t r = a % n;
    return r < 0 ? r + n : r;
}
static inline int32_t wrap_inc(int32_t a) { return (int32_t)((uint32_t)a + 1u); }    /* ui + 1, wrapping */
static inline float lerp_f(float a, float b, float t) { return fmaf(b - a, t, a); }  /* std.math.lerp = @mulAdd */

void orc_texture_sample(const rtc_texture* tex, float u, float v, float* out)
{
    const float wf = (float)tex->w, hf = (float)tex->h;
    const int32_t ui = f2i_sat(floorf(wf * u));
    const int32_t vi = f2i_sat(floorf(hf * v));
    const int32_t x1 = floored_mod(clamp_i32(ui, tex->u_min, tex->u_max), tex->w);
    const int32_t y1 = floored_mod(clamp_i32(vi, tex->v_min, tex->v_max), tex->h);
    const int32_t x2 = floored_mod(clamp_i32(wrap_inc(ui), tex->u_min, tex->u_max), tex->w);
    const int32_t y2 = floored_mod(clamp_i32(wrap_inc(vi), tex->v_min, tex->v_max), tex->h);
    const float fu = tex_frac(u), fv = tex_frac(v);
    const int ch = tex->channels;
    const float* p11 = tex->texels + ((size_t)y1 * (size_t)tex->w + (size_t)x1) * (size_t)ch;
    const float* p21 = tex->texels + ((size_t)y1 * (size_t)tex->w + (size_t)x2) * (size_t)ch;
    const float* p12 = tex->texels + ((size_t)y2 * (size_t)tex->w + (size_t)x1) * (size_t)ch;
    const float* p22 = tex->texels + ((size_t)y2 * (size_t)tex->w + (size_t)x2) * (size_t)ch;
    for (int c = 0; c < ch; ++c) {
        const float r1 = lerp_f(p11[c], p21[c], fu);
        const float r2 = lerp_f(p12[c], p22[c], fu);
        out[c] = lerp_f(r1, r2, fv);
    }
}

/* ------------------------------------------------------------------------------------------
 * toRGB (linalg.zig:150-159), sky (stage3.zig:144-150), camera ray (stage3.zig:27-35)
 * ---------------------------------------------------------------------------------------- */
void orc_to_rgb(const float c[3], uint8_t out[3])
{
    const float p = (float)(1.0 / 2.2);                 /* comptime 1.0/gamma coerced to f32 */
    const float hi = fmaxf(0.0f, 0.999999f);            /* clamp() = min(self, max(lo, hi)) (:58-60) */
    for (int i = 0; i < 3; ++i) {
        const float g = fminf(powf(c[i], p), hi) * 256.0f;
        const uint32_t q = f2u_sat(g);
        out[i] = (uint8_t)(q > 255u ? 255u : q);
    }
}

void orc_env_color(const float dir[3], float out[3])
{
    const float t = 0.5f * (dir[1] + 1.0f);
    const v3 c = v3_add(v3_scale(v3_splat(1.0f), 1.0f - t), v3_scale(v3_make(0.5f, 0.7f, 1.0f), t));
    v3_to(c, out);
}

static inline ray_t camera_get_ray(const rtc_camera* cam, float x, float y)
{
    ray_t r;
    r.orig = v3_from(cam->origin);
    r.dir = v3_normalize(v3_add(v3_add(v3_from(cam->lower_left_corner),
                                       v3_scale(v3_from(cam->right), x)),
                                v3_scale(v3_from(cam->up), y)));
    return r;
}

void orc_camera_get_ray(const rtc_camera* cam, float x, float y, float orig[3], float dir[3])
{
    ray_t r = camera_get_ray(cam, x, y);
    v3_to(r.orig, orig);
    v3_to(r.dir, dir);
}

/* ------------------------------------------------------------------------------------------
 * RNG. Philox4x32-10 (Salmon et al., SC'11; Random123 v1.x constants), counter =
 * (pixel, sample, block, 0), key = (seed lo, seed hi). Uniforms are the top 24 bits.
 * Draw layout per sample (order of use in stage3.zig:238,207,214):
 *   block 0      : [0] jitter x, [1] jitter y
 *   block 1 + L  : recursion level L of traceRayRecursive: [0] alpha test,
 *                  [1],[2] the two uniforms of randomUnitVector.
 * ---------------------------------------------------------------------------------------- */
void orc_philox4x32_10(const uint32_t ctr_in[4], const uint32_t key_in[2], uint32_t out[4])
{
    uint32_t c0 = ctr_in[0], c1 = ctr_in[1], c2 = ctr_in[2], c3 = ctr_in[3];
    uint32_t k0 = key_in[0], k1 = key_in[1];
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        const uint32_t n1 = (uint32_t)p1;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        const uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

void orc_rng_block(uint64_t seed, uint32_t pixel, uint32_t sample, uint32_t block, float out[4])
{
    const uint32_t ctr[4] = { pixel, sample, block, 0u };
    const uint32_t key[2] = { (uint32_t)seed, (uint32_t)(seed >> 32) };
    uint32_t r[4];
    orc_philox4x32_10(ctr, key, r);
    for (int i = 0; i < 4; ++i) out[i] = (float)(r[i] >> 8) * (1.0f / 16777216.0f);
}

/* Uniform direction on the unit sphere from two uniforms, using only operations that are
 * correctly rounded on both CPU and GPU (+,-,*,sqrt,fmaf), so both sides agree bit for bit.
 * Same distribution as linalg.zig:140-148 (normalised isotropic Gaussian).
 * z = 1-2*u1; (cos,sin)(2*pi*u2) by quadrant reduction + minimax polynomials on [-pi/4,pi/4]. */
void orc_unit_vector(float u1, float u2, float out[3])
{
    /* u1 = k / 2^24 is taken at the CENTRE of its bin, z = 1 - 2 (k + 1/2) / 2^24 = (1 - 2^-24) - 2 u1 (both
     * operations exact in f32), so the poles z = +-1 are never produced. With z = +-1 exactly, normal + unit
     * vector cancels to the zero vector on every axis-aligned wall facing the pole (probability 2^-24 per
     * bounce: a handful of NaN pixels per 1080p/64 spp frame); the reference's normalize(N(0,1)^3)
     * (src/linalg.zig:140-148) has no such atom. */
    const float z = 0.99999994f - 2.0f * u1;
    const float r = sqrtf(fmaxf(0.0f, 1.0f - z * z));
    const float t4 = u2 * 4.0f;
    const float kf = floorf(t4 + 0.5f);
    const float th = (t4 - kf) * 1.57079632679489661923f;
    const float s2 = th * th;
    float sp = fmaf(s2, -1.9515295891e-4f, 8.3321608736e-3f);
    sp = fmaf(sp, s2, -1.6666654611e-1f);
    const float sn = fmaf(sp * s2, th, th);
    float cp = fmaf(s2, 2.443315711809948e-5f, -1.388731625493765e-3f);
    cp = fmaf(cp, s2, 4.166664568298827e-2f);
    const float cs = fmaf(cp, s2 * s2, fmaf(s2, -0.5f, 1.0f));
    const int q = (int)kf & 3;
    float c, s;
    switch (q) {
        case 0: c = cs; s = sn; break;
        case 1: c = -sn; s = cs; break;
        case 2: c = -cs; s = -sn; break;
        default: c = sn; s = -cs; break;
    }
    out[0] = r * c;
    out[1] = r * s;
    out[2] = z;
}

/* ------------------------------------------------------------------------------------------
 * stage1 math: zgltf helpers.zig:17-99 (column-major Mat4 = [column][row]), getGlobalTransform
 * is a fold of orc_mat4_mul(parent, child) done by the caller (main.zig:267-280).
 * ---------------------------------------------------------------------------------------- */
void orc_mat4_mul(const float left[16], const float right[16], float out[16])
{
    float tmp[16];
    for (int column = 0; column < 4; ++column)
        for (int row = 0; row < 4; ++row) {
            float sum = 0.0f;
            for (int lc = 0; lc < 4; ++lc) sum += left[lc * 4 + row] * right[column * 4 + lc];
            tmp[column * 4 + row] = sum;
        }
    memcpy(out, tmp, sizeof tmp);
}

void orc_mat4_recompose(const float tr[3], const float q[4], const float sc[3], float out[16])
{
    float t[16] = { 1,0,0,0, 0,1,0,0, 0,0,1,0, 0,0,0,1 };
    float r[16] = { 1,0,0,0, 0,1,0,0, 0,0,1,0, 0,0,0,1 };
    float s[16] = { 1,0,0,0, 0,1,0,0, 0,0,1,0, 0,0,0,1 };
    t[12] = tr[0]; t[13] = tr[1]; t[14] = tr[2];
    const float x = q[0], y = q[1], z = q[2], w = q[3];
    const float xx = x * x, yy = y * y, zz = z * z;
    const float xy = x * y, xz = x * z, yz = y * z;
    const float wx = w * x, wy = w * y, wz = w * z;
    r[0] = 1.0f - 2.0f * (yy + zz); r[1] = 2.0f * (xy + wz);        r[2] = 2.0f * (xz - wy);
    r[4] = 2.0f * (xy - wz);        r[5] = 1.0f - 2.0f * (xx + zz); r[6] = 2.0f * (yz + wx);
    r[8] = 2.0f * (xz + wy);        r[9] = 2.0f * (yz - wx);        r[10] = 1.0f - 2.0f * (xx + yy);
    s[0] = sc[0]; s[5] = sc[1]; s[10] = sc[2];
    float rs[16];
    orc_mat4_mul(r, s, rs);
    orc_mat4_mul(t, rs, out);
}

/* linalg.zig:250-256 */
static inline v3 mat_col3(const float m[16], int column) { return v3_make(m[column * 4], m[column * 4 + 1], m[column * 4 + 2]); }
/* linalg.zig:258-266 */
static inline v3 mat_transform_position(const float m[16], v3 v)
{
    return v3_add(v3_add(v3_add(v3_scale(mat_col3(m, 0), v.x), v3_scale(mat_col3(m, 1), v.y)),
                         v3_scale(mat_col3(m, 2), v.z)), mat_col3(m, 3));
}
/* linalg.zig:268-275 */
static inline v3 mat_transform_direction(const float m[16], v3 v)
{
    return v3_add(v3_add(v3_scale(mat_col3(m, 0), v.x), v3_scale(mat_col3(m, 1), v.y)),
                  v3_scale(mat_col3(m, 2), v.z));
}

/* stage1.zig:235-255 */
void orc_assemble_triangles(const float matrix[16], const float* positions, const float* normals,
                            const float* texcoords, const uint16_t* indices,
                            uint32_t num_tris, uint32_t material,
                            float* out_positions, rtc_tri_data* out_data)
{
    for (uint32_t t = 0; t < num_tris; ++t) {
        rtc_tri_data* d = &out_data[t];
        d->material = material;
        d->src_tri = 0;
        for (int i = 0; i < 3; ++i) {
            const uint32_t vi = indices[t * 3 + i];
            const v3 p = mat_transform_position(matrix, v3_from(positions + 3 * vi));
            v3_to(p, out_positions + 9 * t + 3 * i);
            const v3 n = v3_normalize(mat_transform_direction(matrix, v3_from(normals + 3 * vi)));
            v3_to(n, d->normal[i]);
            d->texcoord[i][0] = texcoords ? texcoords[2 * vi] : 0.0f;
            d->texcoord[i][1] = texcoords ? texcoords[2 * vi + 1] : 0.0f;
        }
    }
}

/* stage1.zig:309-371 */
int orc_load_camera(const float matrix[16], float yfov, float aspect_ratio,
                    uint32_t width, uint32_t height, rtc_camera* out)
{
    uint32_t w = 0, h = 0;
    const int has_ar = aspect_ratio > 0.0f;
    if (width == 0 && height == 0) return -1;              /* OutputImgSizeIsNotSpecified */
    if (width != 0 && height != 0) {
        if (has_ar) return -2;                             /* CameraHasAspectRatio */
        w = width; h = height;
    } else {
        if (!has_ar) return -3;                            /* CameraHasntAspectRatio */
        w = width ? width : f2u_sat((float)height * aspect_ratio);
        h = height ? height : f2u_sat((float)width / aspect_ratio);
    }
    const float f_w = (float)w, f_h = (float)h;
    const v3 origin = mat_col3(matrix, 3);
    const v3 fwd = v3_normalize(v3_scale(mat_col3(matrix, 2), -1.0f));
    const v3 world_up = v3_make(0, 1, 0);
    const v3 right = v3_normalize(v3_cross(fwd, world_up));
    const v3 up = v3_cross(fwd, right);
    const float focal_length = (f_h / 2.0f) / tanf(yfov / 2.0f);
    const v3 llc = v3_sub(v3_sub(v3_scale(fwd, focal_length), v3_scale(right, f_w / 2.0f)),
                          v3_scale(up, f_h / 2.0f));
    out->w = w; out->h = h;
    v3_to(origin, out->origin);
    v3_to(llc, out->lower_left_corner);
    v3_to(right, out->right);
    v3_to(up, out->up);
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * stage2: Geometry.build + bakeInto (stage2.zig:44-164)
 * ---------------------------------------------------------------------------------------- */
typedef void (*cell_visitor)(void* ctx, uint64_t cell_index, uint32_t tri);

/* The triple loop shared by initCells (stage2.zig:63-80) and initIndices (:107-124). */
static void for_each_overlapped_cell(const rtc_grid* g, const float* positions, uint32_t num_tris,
                                     cell_visitor visit, void* ctx)
{
    for (uint32_t t = 0; t < num_tris; ++t) {
        const v3 p0 = v3_from(positions + 9 * t), p1 = v3_from(positions + 9 * t + 3), p2 = v3_from(positions + 9 * t + 6);
        float lo[3], hi[3];
        v3_to(v3_min(p0, v3_min(p1, p2)), lo);
        v3_to(v3_max(p0, v3_max(p1, p2)), hi);
        uint32_t cmin[3], cmax[3];
        orc_grid_cell_idx(g, lo, cmin);
        orc_grid_cell_idx(g, hi, cmax);
        for (uint32_t z = cmin[2]; z < cmax[2] + 1; ++z)
            for (uint32_t y = cmin[1]; y < cmax[1] + 1; ++y)
                for (uint32_t x = cmin[0]; x < cmax[0] + 1; ++x) {
                    float bmin[3], bmax[3];
                    orc_grid_cell_bbox(g, x, y, z, bmin, bmax);
                    if (triangle_aabb(p0, p1, p2, v3_from(bmin), v3_from(bmax)))
                        visit(ctx, grid_linear(g, x, y, z), t);
                }
    }
}

typedef struct { uint32_t* first; uint32_t* count; uint64_t* indices; } build_ctx;
static void visit_count(void* c, uint64_t cell, uint32_t tri) { (void)tri; ((build_ctx*)c)->count[cell] += 1; }
static void visit_fill(void* c, uint64_t cell, uint32_t tri)
{
    build_ctx* b = (build_ctx*)c;
    b->indices[b->first[cell] + b->count[cell]] = tri;
    b->count[cell] += 1;
}

int orc_geometry_build(const float* positions, const rtc_tri_data* data, uint32_t num_tris,
                       const uint32_t res[3], orc_geometry* out)
{
    memset(out, 0, sizeof *out);
    /* initGrid (stage2.zig:44-57) */
    v3 bmin = v3_splat(INFINITY), bmax = v3_splat(-INFINITY);
    for (uint32_t t = 0; t < num_tris; ++t)
        for (int i = 0; i < 3; ++i) {
            const v3 p = v3_from(positions + 9 * t + 3 * i);
            bmin = v3_min(bmin, p);
            bmax = v3_max(bmax, p);
        }
    float bmin_a[3], bmax_a[3];
    v3_to(bmin, bmin_a); v3_to(bmax, bmax_a);
    orc_grid_init(bmin_a, bmax_a, res, &out->grid);

    /* initCells (stage2.zig:59-102) */
    const uint64_t num_cells = (uint64_t)res[0] * res[1] * res[2];
    build_ctx ctx;
    ctx.first = (uint32_t*)calloc(num_cells, sizeof(uint32_t));
    ctx.count = (uint32_t*)calloc(num_cells, sizeof(uint32_t));
    ctx.indices = NULL;
    if (!ctx.first || !ctx.count) return -1;
    for_each_overlapped_cell(&out->grid, positions, num_tris, visit_count, &ctx);
    uint32_t min_t = 0xFFFFFFFFu, max_t = 0, empty = 0, total = 0;
    for (uint64_t c = 0; c < num_cells; ++c) {
        if (ctx.count[c] != 0) {
            if (ctx.count[c] < min_t) min_t = ctx.count[c];
            if (ctx.count[c] > max_t) max_t = ctx.count[c];
        } else {
            empty += 1;
        }
        ctx.first[c] = total;
        total += ctx.count[c];
        ctx.count[c] = 0;
    }
    /* initIndices (stage2.zig:104-129) */
    ctx.indices = (uint64_t*)malloc((total ? total : 1) * sizeof(uint64_t));
    for_each_overlapped_cell(&out->grid, positions, num_tris, visit_fill, &ctx);

    /* bakeInto (stage2.zig:137-164) */
    out->cells = (rtc_cell*)malloc(num_cells * sizeof(rtc_cell));
    for (uint64_t c = 0; c < num_cells; ++c) {
        out->cells[c].begin = ctx.first[c];
        out->cells[c].end = ctx.first[c] + ctx.count[c];
    }
    out->tri_pos = (rtc_tri_pos*)calloc(total ? total : 1, sizeof(rtc_tri_pos));
    out->tri_data = (rtc_tri_data*)calloc(total ? total : 1, sizeof(rtc_tri_data));
    for (uint64_t i = 0; i < total; ++i) {
        const uint64_t s = ctx.indices[i];
        const v3 p0 = v3_from(positions + 9 * s), p1 = v3_from(positions + 9 * s + 3), p2 = v3_from(positions + 9 * s + 6);
        v3_to(p0, out->tri_pos[i].v0);                      /* Triangle.init (linalg.zig:688-694) */
        v3_to(v3_sub(p1, p0), out->tri_pos[i].e1);
        v3_to(v3_sub(p2, p0), out->tri_pos[i].e2);
        out->tri_data[i] = data[s];
        out->tri_data[i].src_tri = (uint32_t)s;
    }
    out->num_cells = num_cells;
    out->indices = ctx.indices;
    out->num_tris = total;
    out->num_src_tris = num_tris;
    out->empty_cells = empty;
    out->min_tris = min_t;
    out->max_tris = max_t;
    free(ctx.first);
    free(ctx.count);
    return 0;
}

void orc_geometry_free(orc_geometry* g)
{
    free(g->cells); free(g->indices); free(g->tri_pos); free(g->tri_data);
    memset(g, 0, sizeof *g);
}

/* ------------------------------------------------------------------------------------------
 * stage3: Scene.traceRay, traceRayRecursive, renderWorker, render
 * ---------------------------------------------------------------------------------------- */
typedef struct { float t, u, v; int64_t triangle_idx; } hit_t;

/* stage3.zig:152-186 */
static hit_t scene_trace_ray(const rtc_scene_desc* scene, ray_t ray, rtc_stats* stats)
{
    hit_t nearest = { INFINITY, 0.0f, 0.0f, -1 };
    grid_iter it;
    if (stats) stats->rays += 1;
    if (grid_trace_ray(&scene->grid, ray, &it)) {
        for (;;) {
            const uint64_t cell_idx = grid_linear(&scene->grid, it.cell[0], it.cell[1], it.cell[2]);
            const rtc_cell cell = scene->cells[cell_idx];
            if (stats) { stats->cells_visited += 1; stats->tris_tested += cell.end - cell.begin; }
            for (uint32_t ti = cell.begin; ti < cell.end; ++ti) {
                const rtc_tri_pos* tp = &scene->tri_pos[ti];
                hit_t hit;
                hit.triangle_idx = ti;
                if (triangle_ray_intersection(v3_from(tp->v0), v3_from(tp->e1), v3_from(tp->e2), ray,
                                              &hit.t, &hit.u, &hit.v)) {
                    if (nearest.t > hit.t && hit.t > 0.0f) nearest = hit;
                }
            }
            const float t_next_crossing = grid_iter_next(&it);
            if (nearest.t <= t_next_crossing) break;
        }
    }
    return nearest;
}

int64_t orc_trace_ray(const rtc_scene_desc* scene, const float orig[3], const float dir[3],
                      float* t, float* u, float* v, rtc_stats* stats)
{
    ray_t r = { v3_from(orig), v3_from(dir) };
    hit_t h = scene_trace_ray(scene, r, stats);
    *t = h.t; *u = h.u; *v = h.v;
    return h.triangle_idx;
}

int64_t orc_trace_ray_brute(const rtc_scene_desc* scene, const float orig[3], const float dir[3],
                            float* t, float* u, float* v)
{
    ray_t r = { v3_from(orig), v3_from(dir) };
    hit_t nearest = { INFINITY, 0.0f, 0.0f, -1 };
    for (uint64_t ti = 0; ti < scene->num_tris; ++ti) {
        const rtc_tri_pos* tp = &scene->tri_pos[ti];
        hit_t hit;
        hit.triangle_idx = (int64_t)ti;
        if (triangle_ray_intersection(v3_from(tp->v0), v3_from(tp->e1), v3_from(tp->e2), r, &hit.t, &hit.u, &hit.v))
            if (nearest.t > hit.t && hit.t > 0.0f) nearest = hit;
    }
    *t = nearest.t; *u = nearest.u; *v = nearest.v;
    return nearest.triangle_idx;
}

/* Triangle.Data.interpolate (stage3.zig:53-71) */
static inline v3 interp_normal(const rtc_tri_data* d, float u, float v)
{
    const float w = 1.0f - u - v;
    return v3_add(v3_add(v3_scale(v3_from(d->normal[0]), w), v3_scale(v3_from(d->normal[1]), u)),
                  v3_scale(v3_from(d->normal[2]), v));
}
static inline void interp_texcoord(const rtc_tri_data* d, float u, float v, float out[2])
{
    const float w = 1.0f - u - v;
    out[0] = d->texcoord[0][0] * w + d->texcoord[1][0] * u + d->texcoord[2][0] * v;
    out[1] = d->texcoord[0][1] * w + d->texcoord[1][1] * u + d->texcoord[2][1] * v;
}

/* unit-level entry for the known-answer tests (tests/test_oracle_kat.py) */
void orc_interpolate(const rtc_tri_data* d, float u, float v, float normal_out[3], float uv_out[2])
{
    v3_to(interp_normal(d, u, v), normal_out);
    interp_texcoord(d, u, v, uv_out);
}

typedef struct {
    const rtc_scene_desc* scene;
    uint64_t seed;
    uint32_t pixel, sample;
    uint16_t max_bounce;
    rtc_stats* stats;
} path_ctx;

/* stage3.zig:188-220. `depth` counts down exactly as in the reference; the recursion level
 * (max_bounce - depth) selects the RNG block. */
static v3 trace_ray_recursive(const path_ctx* pc, ray_t ray, uint16_t depth)
{
    if (depth == 0) return v3_splat(0.0f);
    const hit_t hit = scene_trace_ray(pc->scene, ray, pc->stats);
    if (hit.t == INFINITY) {
        float c[3], d[3];
        v3_to(ray.dir, d);
        orc_env_color(d, c);
        return v3_from(c);
    }
    const rtc_tri_data* tri = &pc->scene->tri_data[hit.triangle_idx];
    const rtc_material* mat = &pc->scene->materials[tri->material];
    if (pc->stats) pc->stats->hits_shaded += 1;

    float uv[2], albedo[3], emissive[3], transparency;
    interp_texcoord(tri, hit.u, hit.v, uv);
    orc_texture_sample(&mat->base_color, uv[0], uv[1], albedo);
    orc_texture_sample(&mat->emissive, uv[0], uv[1], emissive);
    orc_texture_sample(&mat->transparency, uv[0], uv[1], &transparency);
    const v3 normal = interp_normal(tri, hit.u, hit.v);

    float rnd[4];
    orc_rng_block(pc->seed, pc->pixel, pc->sample, 1u + (uint32_t)(pc->max_bounce - depth), rnd);
    const float eps = 1.1920928955078125e-07f;             /* std.math.floatEps(f32) */
    if (rnd[0] > transparency) {
        ray_t nr = { ray_at(ray, hit.t + eps), ray.dir };
        return trace_ray_recursive(pc, nr, (uint16_t)(depth - 1));
    }
    float uvec[3];
    orc_unit_vector(rnd[1], rnd[2], uvec);
    const v3 scattered = v3_normalize(v3_add(normal, v3_from(uvec)));
    ray_t nr = { ray_at(ray, hit.t + eps), scattered };
    return v3_add(v3_from(emissive), v3_mul(v3_from(albedo), trace_ray_recursive(pc, nr, (uint16_t)(depth - 1))));
}

typedef struct {
    const rtc_scene_desc* scene;
    const rtc_camera* camera;
    const rtc_params* params;
    size_t thread_idx, thread_num;
    uint64_t first_pixel, num_pixels;   /* the pixel range this render call covers */
    uint8_t* rgb;
    rtc_aux* aux;
    rtc_stats stats;
} worker_args;

/* stage3.zig:222-245 */
static void* render_worker(void* p)
{
    worker_args* a = (worker_args*)p;
    const rtc_params* prm = a->params;
    const uint32_t total_samples = prm->total_samples ? prm->total_samples : prm->num_samples;
    const v3 inv_num_samples = v3_div(v3_splat(1.0f), v3_splat((float)total_samples));
    const uint64_t len = a->num_pixels;
    const uint64_t pixels_per_thread = (len + a->thread_num - 1) / a->thread_num;
    uint64_t i = pixels_per_thread * a->thread_idx;
    const int want_stats = a->aux && a->aux->stats;
    path_ctx pc;
    pc.scene = a->scene;
    pc.seed = prm->seed;
    pc.max_bounce = prm->max_bounce;
    pc.stats = want_stats ? &a->stats : NULL;
    for (uint64_t n = 0; n < pixels_per_thread; ++n, ++i) {
        if (i >= len) break;
        const uint64_t pix = a->first_pixel + i;
        const float x = (float)(pix % a->camera->w);
        const float y = (float)(pix / a->camera->w);
        v3 pixel = v3_splat(0.0f);
        pc.pixel = (uint32_t)pix;
        if (prm->mode == RTC_MODE_PRIMARY) {
            /* deterministic slice: centre ray, depth 1, alpha uniform := 0 (always scatter) */
            const ray_t ray = camera_get_ray(a->camera, x + 0.5f, y + 0.5f);
            const hit_t hit = scene_trace_ray(a->scene, ray, pc.stats);
            if (want_stats) a->stats.paths += 1;
            if (hit.t == INFINITY) {
                float c[3], d[3];
                v3_to(ray.dir, d);
                orc_env_color(d, c);
                pixel = v3_from(c);
                if (a->aux && a->aux->hit_id) a->aux->hit_id[pix] = 0xFFFFFFFFu;
                if (a->aux && a->aux->hit_tuv) { a->aux->hit_tuv[3 * pix] = INFINITY; a->aux->hit_tuv[3 * pix + 1] = 0; a->aux->hit_tuv[3 * pix + 2] = 0; }
            } else {
                const rtc_tri_data* tri = &a->scene->tri_data[hit.triangle_idx];
                const rtc_material* mat = &a->scene->materials[tri->material];
                float uv[2], emissive[3];
                interp_texcoord(tri, hit.u, hit.v, uv);
                orc_texture_sample(&mat->emissive, uv[0], uv[1], emissive);
                pixel = v3_from(emissive);
                if (want_stats) a->stats.hits_shaded += 1;
                if (a->aux && a->aux->hit_id) a->aux->hit_id[pix] = tri->src_tri;
                if (a->aux && a->aux->hit_tuv) { a->aux->hit_tuv[3 * pix] = hit.t; a->aux->hit_tuv[3 * pix + 1] = hit.u; a->aux->hit_tuv[3 * pix + 2] = hit.v; }
            }
            if (a->aux && a->aux->radiance) v3_to(pixel, a->aux->radiance + 3 * pix);
            float c[3];
            v3_to(pixel, c);
            orc_to_rgb(c, a->rgb + 3 * pix);
            continue;
        }
        for (uint32_t s = 0; s < prm->num_samples; ++s) {
            pc.sample = prm->sample_offset + s;
            float jit[4];
            orc_rng_block(pc.seed, pc.pixel, pc.sample, 0u, jit);
            const ray_t ray = camera_get_ray(a->camera, x + jit[0], y + jit[1]);
            if (want_stats) a->stats.paths += 1;
            const v3 ray_color = trace_ray_recursive(&pc, ray, prm->max_bounce);
            pixel = v3_add(pixel, ray_color);
        }
        const v3 mean = v3_mul(pixel, inv_num_samples);
        float c[3];
        v3_to(mean, c);
        if (a->aux && a->aux->radiance) v3_to(mean, a->aux->radiance + 3 * pix);
        orc_to_rgb(c, a->rgb + 3 * pix);
    }
    return NULL;
}

/* stage3.zig:247-256 */
int orc_render(const rtc_scene_desc* scene, const rtc_camera* camera, const rtc_params* params,
               int num_threads, uint32_t y0, uint32_t y1, uint8_t* rgb_out, rtc_aux* aux)
{
    if (!scene || !camera || !params || !rgb_out || num_threads < 1) return 1;
    if (y1 == 0 || y1 > camera->h) y1 = camera->h;
    if (y0 > y1) y0 = y1;
    worker_args* args = (worker_args*)calloc((size_t)num_threads, sizeof(worker_args));
    pthread_t* threads = (pthread_t*)calloc((size_t)num_threads, sizeof(pthread_t));
    pthread_attr_t attr;
    pthread_attr_init(&attr);
    pthread_attr_setstacksize(&attr, (size_t)256 << 20);   /* recursion depth = max_bounce (u16) */
    for (int i = 0; i < num_threads; ++i) {
        args[i].scene = scene; args[i].camera = camera; args[i].params = params;
        args[i].thread_idx = (size_t)i; args[i].thread_num = (size_t)num_threads;
        args[i].first_pixel = (uint64_t)y0 * camera->w;
        args[i].num_pixels = (uint64_t)(y1 - y0) * camera->w;
        args[i].rgb = rgb_out; args[i].aux = aux;
        if (pthread_create(&threads[i], &attr, render_worker, &args[i]) != 0) {
            for (int j = 0; j < i; ++j) pthread_join(threads[j], NULL);
            free(args); free(threads);
            return 2;
        }
    }
    for (int i = 0; i < num_threads; ++i) pthread_join(threads[i], NULL);
    pthread_attr_destroy(&attr);
    if (aux && aux->stats) {
        rtc_stats total;
        memset(&total, 0, sizeof total);
        for (int i = 0; i < num_threads; ++i) {
            total.rays += args[i].stats.rays;
            total.cells_visited += args[i].stats.cells_visited;
            total.tris_tested += args[i].stats.tris_tested;
            total.hits_shaded += args[i].stats.hits_shaded;
            total.paths += args[i].stats.paths;
        }
        *aux->stats = total;
    }
    free(args); free(threads);
    return 0;
}
