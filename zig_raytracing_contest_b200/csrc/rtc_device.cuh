// rtc_device.cuh — fp32 device mirror of the reference's linalg.zig math and stage3.zig shading.
//
// Every routine names the reference lines it mirrors (paths relative to the reference root).
// Operation order is the reference's. The translation unit that includes this header is
// compiled twice (csrc/Makefile):
//   RTC_MATH_EXACT : -fmad=false, IEEE div/sqrt  -> bit-identical to the CPU oracle
//   RTC_MATH_FAST  : -fmad=true                  -> contraction allowed
// Where the reference itself fuses (std.math.lerp == @mulAdd) __fmaf_rn is explicit.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rtc {

// ---- device-side scene layout (HBM) ---------------------------------------------------
// PACKED CELL. A cell (x,y,z) travels as one 32-bit word
//     x&3 | (y&3)<<2 | (z&3)<<4 | (x>>2)<<6 | (y>>2)<<(6+lx) | (z>>2)<<(6+lx+ly)
// i.e. the low 6 bits are the cell inside its 4x4x4 block and word>>6 is the block index, with the block
// grid padded to powers of two (lx, ly, lz = ceil(log2(blocks per axis))). mask[a] holds the bits of axis a.
// The word IS the index of the cell's record, so nothing is recomputed from coordinates on the device.
// cells     : uint2[64*NBP] {begin,end}, 8 B / cell, indexed by the packed cell (NBP = 2^(lx+ly+lz) blocks)
// fine      : uint64[NBP]   bit b = cell b of the block is non-empty. One 8-byte load covers every DDA step a
//             ray takes inside the block (the per-step occupancy probe was the top stall of the first kernel)
// coarse    : uint32[NBP/32] bit = block has any triangle; 4 KiB at 128^3 -> staged in shared memory per CTA
// tri_data  : float4[4*N]   per SOURCE triangle, 64 B = two sectors: n0.xyz n1.x | n1.yz n2.xy | n2.z uv0.xy uv1.x | uv1.y uv2.xy mat
// texels    : float4[w*h] (rgb, pad) for base colour / emissive, float[w*h] for transparency
struct DevTexture {
    const void* texels;
    int w, h, u_min, u_max, v_min, v_max;
    float wf, hf;
};
struct DevMaterial {
    DevTexture base, emissive, alpha;
};
struct DevScene {
    float bmin[3], bmax[3];
    uint32_t res[3];
    float cell[3];
    const uint2* cells;
    const float4* tri_v0;
    const float4* tri_e1;
    const float4* tri_e2;
    const float4* tri_data;
    const DevMaterial* materials;
    const unsigned long long* fine;
    const uint32_t* coarse;
    uint32_t mask[3];           // bits of the packed cell that belong to x, y, z
    uint32_t shift[3];          // bit position of the block coordinate of x, y, z (6, 6+lx, 6+lx+ly)
    uint32_t coarse_words;
    uint32_t num_materials;
};
static constexpr uint32_t kMaxGridRes = 1024;     // packed cell: 6 + 3*8 bits
static constexpr uint32_t kCoarseSmemMaxWords = 8192;   // 32 KiB: the coarse bitmap of a 256^3 grid
__host__ __device__ __forceinline__ uint32_t pack_cell(const uint32_t shift[3], uint32_t x, uint32_t y, uint32_t z)
{
    return (x & 3u) | ((y & 3u) << 2) | ((z & 3u) << 4) | ((x >> 2) << shift[0]) | ((y >> 2) << shift[1]) | ((z >> 2) << shift[2]);
}
struct DevCamera {
    uint32_t w, h;
    float origin[3], llc[3], right[3], up[3];
};
struct DevParams {
    uint32_t num_samples;    // samples of this call
    uint32_t max_bounce;
    uint32_t sample_offset;
    uint32_t seed_lo, seed_hi;
    // pixels of this call, as a list of num_pixels LOCAL indices pl -> frame pixel (slot_pixel):
    //   stride_px == 0 : the contiguous range [first_pixel, first_pixel + num_pixels)
    //   stride_px != 0 : strips of strip_px pixels dealt round-robin over the ranks; this rank's k-th strip
    //                    starts at first_pixel + k * stride_px (the frame's last strip may be short)
    uint32_t first_pixel;
    uint32_t num_pixels;
    uint32_t strip_px, stride_px;
};
static constexpr uint32_t kMaxBounce = 64;

// ---- wavefront scratch (HBM), sized for `cap` path slots; slot = sample_local * num_pixels + pixel_local ----
// q[2]      : two ray queues (ping-pong per bounce), 64 B / ray as four float4 arrays (rtc_wavefront.cu QueuedRay)
// hits      : uint2[cap]  per queue entry {t bits, baked triangle index | 0xFFFFFFFF}
// stack_e/a : per scattering level and slot {emissive.xyz, albedo.x} {albedo.yz}        24 B / level
// leaf      : float4[cap] value at the end of the path (sky / zeroes) + number of pushed levels
// counts    : uint32[2*(kMaxBounce+1)]  queue length per bounce, then the chunk dispenser head per bounce
struct RayQueue { float4 *a0, *a1, *a2, *a3; };
struct WaveScratch {
    RayQueue q[2];
    uint2* hits;
    float4* stack_e;
    float2* stack_a;
    float4* leaf;
    uint32_t* counts;
    uint32_t cap;
    uint32_t max_levels;
};
// one batch: pixels [first_pixel, +num_pixels) x samples [s0, s0+ns) (absolute sample indices)
struct WaveParams {
    uint32_t first_pixel, num_pixels;     // num_pixels local pixels starting at local index pl0
    uint32_t s0, ns;
    uint32_t max_bounce;
    uint32_t seed_lo, seed_hi;
    uint32_t pl0, strip_px, stride_px;    // DevParams' pixel list
};
// local pixel index of the batch -> frame pixel
__device__ __forceinline__ uint32_t slot_pixel(const WaveParams& wp, uint32_t pl)
{
    pl += wp.pl0;
    if (wp.stride_px == 0u) return wp.first_pixel + pl;
    const uint32_t k = pl / wp.strip_px;
    return wp.first_pixel + k * wp.stride_px + (pl - k * wp.strip_px);
}
// counters[]: 0 rays, 1 cells_visited, 2 tris_tested, 3 hits_shaded, 4 paths
enum { CNT_RAYS = 0, CNT_CELLS = 1, CNT_TRIS = 2, CNT_HITS = 3, CNT_PATHS = 4, CNT_N = 8 };

// ---- Vec3 (src/linalg.zig:13-222) -----------------------------------------------------
struct v3 { float x, y, z; };
__host__ __device__ __forceinline__ v3 mk(float x, float y, float z) { v3 r; r.x = x; r.y = y; r.z = z; return r; }
__host__ __device__ __forceinline__ v3 splat(float s) { return mk(s, s, s); }
__host__ __device__ __forceinline__ v3 from(const float* a) { return mk(a[0], a[1], a[2]); }
__host__ __device__ __forceinline__ v3 operator+(v3 a, v3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__host__ __device__ __forceinline__ v3 operator-(v3 a, v3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__host__ __device__ __forceinline__ v3 operator*(v3 a, v3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
__host__ __device__ __forceinline__ v3 operator/(v3 a, v3 b) { return mk(a.x / b.x, a.y / b.y, a.z / b.z); }
__host__ __device__ __forceinline__ v3 scale(v3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
// linalg.zig:189-191 — @reduce(.Add) keeps lane order: (x + y) + z
__host__ __device__ __forceinline__ float dot(v3 a, v3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
// linalg.zig:172-179 — a.yzx*b.zxy - a.zxy*b.yzx
__host__ __device__ __forceinline__ v3 cross(v3 a, v3 b)
{
    return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
// linalg.zig:119-125 — normalize = scale(1.0 / length)
__host__ __device__ __forceinline__ float length(v3 a) { return sqrtf(dot(a, a)); }
__host__ __device__ __forceinline__ v3 normalize(v3 a) { return scale(a, 1.0f / length(a)); }

struct Ray { v3 o, d; };
// linalg.zig:284-286
__host__ __device__ __forceinline__ v3 ray_at(const Ray& r, float t) { return r.o + scale(r.d, t); }

// @intFromFloat with the oracle's DEFINED semantics: truncation, saturating, NaN -> 0
// (exactly cvt.rzi.{u32,s32}.f32).
__device__ __forceinline__ uint32_t f2u(float f) { return __float2uint_rz(f); }
__device__ __forceinline__ int32_t f2i(float f) { return __float2int_rz(f); }

// ---- Bbox.rayIntersection (src/linalg.zig:324-349) ------------------------------------
__device__ __forceinline__ bool bbox_ray(v3 bmin, v3 bmax, const Ray& r, float& t, float& t_far)
{
    const bool sx = r.d.x < 0.0f, sy = r.d.y < 0.0f, sz = r.d.z < 0.0f;
    const v3 nr = mk(sx ? bmax.x : bmin.x, sy ? bmax.y : bmin.y, sz ? bmax.z : bmin.z);
    const v3 fr = mk(sx ? bmin.x : bmax.x, sy ? bmin.y : bmax.y, sz ? bmin.z : bmax.z);
    const v3 mn = (nr - r.o) / r.d;
    const v3 mx = (fr - r.o) / r.d;
    float tmin = mn.x, tmax = mx.x;
    if ((tmin > mx.y) || (tmax < mn.y)) return false;
    tmin = fmaxf(tmin, mn.y);
    tmax = fminf(tmax, mx.y);
    if ((tmin > mx.z) || (tmax < mn.z)) return false;
    tmin = fmaxf(tmin, mn.z);
    tmax = fminf(tmax, mx.z);
    t = tmin;
    t_far = tmax;
    return true;
}

// ---- Grid.traceRay / Iterator (src/linalg.zig:443-497) --------------------------------
struct GridIter {
    uint32_t cx, cy, cz;       // cell
    uint32_t ex, ey, ez;       // exit
    uint32_t sx, sy, sz;       // step (1 or 0xFFFFFFFF, wrapping add)
    float dx, dy, dz;          // t_delta
    float tx, ty, tz;          // t_next_crossing
};

__device__ __forceinline__ bool grid_trace_ray(const DevScene& g, const Ray& r, GridIter& it)
{
    float t_hit, t_far;
    const v3 bmin = from(g.bmin), bmax = from(g.bmax);
    if (!bbox_ray(bmin, bmax, r, t_hit, t_far)) return false;
    if (t_far < 0.0f) return false;            // DEFINED corner (oracle/rt_oracle.h): grid behind the ray
    t_hit = fmaxf(0.0f, t_hit);
    const v3 cs = from(g.cell);
    const v3 td = cs / r.d;
    const v3 local = ray_at(r, t_hit) - bmin;
    const v3 cf = local / cs;
    const bool nx = r.d.x < 0.0f, ny = r.d.y < 0.0f, nz = r.d.z < 0.0f;
    const uint32_t lx = g.res[0] - 1u, ly = g.res[1] - 1u, lz = g.res[2] - 1u;
    it.cx = min(f2u(cf.x), lx); it.cy = min(f2u(cf.y), ly); it.cz = min(f2u(cf.z), lz);
    it.sx = nx ? 0xFFFFFFFFu : 1u; it.sy = ny ? 0xFFFFFFFFu : 1u; it.sz = nz ? 0xFFFFFFFFu : 1u;
    it.ex = nx ? 0u : lx; it.ey = ny ? 0u : ly; it.ez = nz ? 0u : lz;
    it.dx = fabsf(td.x); it.dy = fabsf(td.y); it.dz = fabsf(td.z);
    const v3 nc = mk((float)(it.cx + (nx ? 0u : 1u)), (float)(it.cy + (ny ? 0u : 1u)), (float)(it.cz + (nz ? 0u : 1u)));
    const v3 tn = splat(t_hit) + ((nc * cs) - local) / r.d;
    it.tx = tn.x; it.ty = tn.y; it.tz = tn.z;
    return true;
}

// Iterator.next. The reference's table {2,1,2,1,2,2,0,0}[(tx<ty)<<2|(tx<tz)<<1|(ty<tz)]
// (linalg.zig:479-484) is this decision tree (all 8 keys checked in tests/test_oracle_golden.py
// through the oracle's table form): ties prefer z, then y.
__device__ __forceinline__ float grid_next(GridIter& it)
{
    const float inf = __int_as_float(0x7f800000);
    if (it.tx < it.ty) {
        if (it.tx < it.tz) {
            if (it.cx == it.ex) return inf;
            const float t = it.tx; it.cx += it.sx; it.tx += it.dx; return t;
        }
    } else if (it.ty < it.tz) {
        if (it.cy == it.ey) return inf;
        const float t = it.ty; it.cy += it.sy; it.ty += it.dy; return t;
    }
    if (it.cz == it.ez) return inf;
    const float t = it.tz; it.cz += it.sz; it.tz += it.dz; return t;
}

// ---- Triangle.rayIntersection (src/linalg.zig:696-722) --------------------------------
__host__ __device__ __forceinline__ bool tri_ray(v3 v0, v3 e1, v3 e2, const Ray& r, float& t, float& u_, float& v_)
{
    const v3 p = cross(r.d, e2);
    const float det = dot(e1, p);
    if (det < 0.00000001f) return false;
    const float inv = 1.0f / det;
    const v3 tv = r.o - v0;
    const float u = dot(tv, p) * inv;
    if (u < 0.0f || u > 1.0f) return false;
    const v3 q = cross(tv, e1);
    const float v = dot(r.d, q) * inv;
    if (v < 0.0f || u + v > 1.0f) return false;
    t = dot(e2, q) * inv;
    u_ = u;
    v_ = v;
    return true;
}

// ---- Texture(T).sample (src/stage3.zig:82-123) ----------------------------------------
__device__ __forceinline__ int clampi(int v, int lo, int hi) { const int m = v < hi ? v : hi; return lo > m ? lo : m; }
__device__ __forceinline__ int fmod_i(int a, int n) { const int r = a % n; return r < 0 ? r + n : r; }
__device__ __forceinline__ float lerpf(float a, float b, float t) { return __fmaf_rn(b - a, t, a); }   // std.math.lerp
__device__ __forceinline__ float fracf_(float v) { return fabsf(v - truncf(v)); }                      // stage3.zig:93-95

struct TexTaps { int i11, i21, i12, i22; float fu, fv; };
__device__ __forceinline__ TexTaps tex_taps(const DevTexture& tx, float u, float v)
{
    const int ui = f2i(floorf(tx.wf * u));
    const int vi = f2i(floorf(tx.hf * v));
    const int x1 = fmod_i(clampi(ui, tx.u_min, tx.u_max), tx.w);
    const int y1 = fmod_i(clampi(vi, tx.v_min, tx.v_max), tx.h);
    const int x2 = fmod_i(clampi((int)((uint32_t)ui + 1u), tx.u_min, tx.u_max), tx.w);
    const int y2 = fmod_i(clampi((int)((uint32_t)vi + 1u), tx.v_min, tx.v_max), tx.h);
    TexTaps t;
    t.i11 = y1 * tx.w + x1; t.i21 = y1 * tx.w + x2; t.i12 = y2 * tx.w + x1; t.i22 = y2 * tx.w + x2;
    t.fu = fracf_(u); t.fv = fracf_(v);
    return t;
}
// A 1x1 texture (the dummy textures of materials without an image, src/stage1.zig:411-425 — most emissive and
// transparency textures): all four taps are the one texel a, and with finite a, u, v both lerps are fma(+0, f, a) with
// f = |frac| >= 0, i.e. a + (+0) — the value returned here, bit for bit (-0 becomes +0 on both routes). Anything
// non-finite takes the general route, which makes the reference's NaN.
__device__ __forceinline__ bool tex_is_const(const DevTexture& tx, float u, float v)
{
    return (tx.w | tx.h) == 1 && fabsf(u) < __int_as_float(0x7f800000) && fabsf(v) < __int_as_float(0x7f800000);
}
__device__ __forceinline__ v3 tex_sample3(const DevTexture& tx, float u, float v)
{
    if (tex_is_const(tx, u, v)) {
        const float4 a = __ldg((const float4*)tx.texels);
        const float big = __int_as_float(0x7f800000);
        if (fabsf(a.x) < big && fabsf(a.y) < big && fabsf(a.z) < big) return mk(a.x + 0.0f, a.y + 0.0f, a.z + 0.0f);
    }
    const TexTaps t = tex_taps(tx, u, v);
    const float4* p = (const float4*)tx.texels;
    const float4 a = __ldg(p + t.i11), b = __ldg(p + t.i21), c = __ldg(p + t.i12), d = __ldg(p + t.i22);
    const v3 r1 = mk(lerpf(a.x, b.x, t.fu), lerpf(a.y, b.y, t.fu), lerpf(a.z, b.z, t.fu));
    const v3 r2 = mk(lerpf(c.x, d.x, t.fu), lerpf(c.y, d.y, t.fu), lerpf(c.z, d.z, t.fu));
    return mk(lerpf(r1.x, r2.x, t.fv), lerpf(r1.y, r2.y, t.fv), lerpf(r1.z, r2.z, t.fv));
}
__device__ __forceinline__ float tex_sample1(const DevTexture& tx, float u, float v)
{
    if (tex_is_const(tx, u, v)) {
        const float a = __ldg((const float*)tx.texels);
        if (fabsf(a) < __int_as_float(0x7f800000)) return a + 0.0f;
    }
    const TexTaps t = tex_taps(tx, u, v);
    const float* p = (const float*)tx.texels;
    const float a = __ldg(p + t.i11), b = __ldg(p + t.i21), c = __ldg(p + t.i12), d = __ldg(p + t.i22);
    return lerpf(lerpf(a, b, t.fu), lerpf(c, d, t.fu), t.fv);
}

// ---- sky, camera, tonemap --------------------------------------------------------------
// Scene.getEnvColor (src/stage3.zig:144-150)
__device__ __forceinline__ v3 env_color(v3 d)
{
    const float t = 0.5f * (d.y + 1.0f);
    return scale(splat(1.0f), 1.0f - t) + scale(mk(0.5f, 0.7f, 1.0f), t);
}
// Camera.getRay (src/stage3.zig:27-35)
__device__ __forceinline__ Ray camera_ray(const DevCamera& c, float x, float y)
{
    Ray r;
    r.o = from(c.origin);
    r.d = normalize((from(c.llc) + scale(from(c.right), x)) + scale(from(c.up), y));
    return r;
}
// Vec3.toRGB (src/linalg.zig:150-159): pow(1/2.2), upper clamp only, *256, truncate.
__device__ __forceinline__ uint32_t to_rgb8(float c)
{
    const float g = fminf(powf(c, (float)(1.0 / 2.2)), 0.999999f) * 256.0f;
    const uint32_t q = f2u(g);
    return q > 255u ? 255u : q;
}

// ---- RNG: Philox4x32-10, counter (pixel, sample, block, 0), key = seed; top 24 bits -> [0,1).
// Shared definition with oracle/rt_oracle.c (orc_rng_block). Replaces std.rand.DefaultPrng
// (src/stage3.zig:225-226), whose stream depends on the worker-thread count.
__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                              uint32_t k0, uint32_t k1, uint32_t out[4])
{
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t h0 = __umulhi(0xD2511F53u, c0), l0 = 0xD2511F53u * c0;
        const uint32_t h1 = __umulhi(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
        const uint32_t n0 = h1 ^ c1 ^ k0, n2 = h0 ^ c3 ^ k1;
        c0 = n0; c1 = l1; c2 = n2; c3 = l0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
__device__ __forceinline__ void rng_block(uint32_t seed_lo, uint32_t seed_hi, uint32_t pixel, uint32_t sample, uint32_t block, float out[4])
{
    uint32_t r[4];
    philox4x32_10(pixel, sample, block, 0u, seed_lo, seed_hi, r);
#pragma unroll
    for (int i = 0; i < 4; ++i) out[i] = (float)(r[i] >> 8) * (1.0f / 16777216.0f);
}
// Uniform direction on the sphere from two uniforms using only correctly-rounded operations
// (orc_unit_vector). Same distribution as Vec3.randomUnitVector (src/linalg.zig:140-148).
__device__ __forceinline__ v3 unit_vector(float u1, float u2)
{
    const float z = 0.99999994f - 2.0f * u1;       // bin centre of u1: the poles z = +-1 never occur (see orc_unit_vector)
    const float r = sqrtf(fmaxf(0.0f, 1.0f - z * z));
    const float t4 = u2 * 4.0f;
    const float kf = floorf(t4 + 0.5f);
    const float th = (t4 - kf) * 1.57079632679489661923f;
    const float s2 = th * th;
    float sp = __fmaf_rn(s2, -1.9515295891e-4f, 8.3321608736e-3f);
    sp = __fmaf_rn(sp, s2, -1.6666654611e-1f);
    const float sn = __fmaf_rn(sp * s2, th, th);
    float cp = __fmaf_rn(s2, 2.443315711809948e-5f, -1.388731625493765e-3f);
    cp = __fmaf_rn(cp, s2, 4.166664568298827e-2f);
    const float cs = __fmaf_rn(cp, s2 * s2, __fmaf_rn(s2, -0.5f, 1.0f));
    const int q = (int)kf & 3;
    float c, s;
    if (q == 0) { c = cs; s = sn; }
    else if (q == 1) { c = -sn; s = cs; }
    else if (q == 2) { c = -cs; s = -sn; }
    else { c = sn; s = -cs; }
    return mk(r * c, r * s, z);
}

}  // namespace rtc
