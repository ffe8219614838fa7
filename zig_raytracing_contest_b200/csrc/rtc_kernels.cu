// rtc_kernels.cu — sm_100a render kernels for the reference's per-pixel path-tracing loop
// (src/stage3.zig:152-245). Compiled twice: -DRTC_NS=exact -fmad=false and -DRTC_NS=fast -fmad=true.
//
// Mapping. One lane = one (pixel, sample) path; a warp takes a group of 32/min(S,32) consecutive
// pixels from a persistent work counter. Scene.traceRay is WARP-COOPERATIVE (trace32 below): every
// lane walks its own ray's DDA, but the triangle tests of all 32 rays are pooled and dealt out one
// (ray, triangle) pair per lane, so the Möller–Trumbore loop — the reference's innermost loop —
// runs with (nearly) full lanes whatever the per-cell triangle counts are, and 32 consecutive
// triangles of a cell are fetched with three coalesced LDG.128 (SoA float4 arrays).
// The per-lane version this replaced ran at 5.2 of 32 active lanes (profiles/r1a_*).
#include "rtc_kernels.h"

#ifndef RTC_NS
#error "compile with -DRTC_NS=exact or -DRTC_NS=fast"
#endif

namespace rtc {
namespace RTC_NS {

static constexpr int kThreads = 128;
static constexpr int kWarps = kThreads / 32;
static constexpr uint32_t kMiss = 0xFFFFFFFFu;
static constexpr uint32_t kFull = 0xFFFFFFFFu;
static constexpr int kMaxStep = 8;      // DDA steps a lane may take per round before the pooled triangle pass

struct Hit { float t, u, v; uint32_t tri; };
struct Counters { uint32_t rays, cells, tris, hits, paths; };

// Per-warp staging for trace32.
struct WarpShared {
    float4 item[32][2];                 // compacted pending (ray, cell): {o.xyz, begin-excl bits} {d.xyz, owner lane bits}
    unsigned long long best[32];        // per owner lane: min over (t bits << 32 | baked triangle index)
    float col[32][3];                   // per-sample radiance for the in-order pixel sum
};

// Grid.Iterator (src/linalg.zig:471-497) in a register-lean, branch-free form. The cell travels packed
// (x | y<<10 | z<<20) next to the packed exit cell; t_delta carries the step direction in its sign bit.
// The t_next_crossing / t_delta arithmetic is the reference's, add for add. `fine` is the occupancy mask
// of the 4x4x4 block the cell lies in, `bi` that block's index (rtc_device.cuh).
struct Trav {
    float tx, ty, tz;                   // t_next_crossing
    float dx, dy, dz;                   // |.| = t_delta; sign bit set = stepping towards 0 on that axis
    uint32_t c, ec;                     // packed cell, packed exit cell
    uint32_t bi;
    unsigned long long fine;
};

static constexpr uint32_t kBlockBits = 0x3FCFF3FCu;    // bits 2..9 of each 10-bit field: the block coordinates

template <bool CSM>
__device__ __forceinline__ void trav_load_block(const DevScene& sc, const uint32_t* __restrict__ coarse, Trav& s)
{
    const uint32_t bx = (s.c >> 2) & 0xFFu, by = (s.c >> 12) & 0xFFu, bz = (s.c >> 22) & 0xFFu;
    s.bi = bx + sc.nbx * (by + sc.nby * bz);
    const uint32_t w = CSM ? coarse[s.bi >> 5] : __ldg(sc.coarse + (s.bi >> 5));
    s.fine = ((w >> (s.bi & 31u)) & 1u) ? __ldg(sc.fine + s.bi) : 0ull;
}
// bit of the cell inside its block: x&3 | (y&3)<<2 | (z&3)<<4
__device__ __forceinline__ uint32_t trav_cell_bit(uint32_t c) { return (c & 3u) | ((c >> 8) & 0xCu) | ((c >> 16) & 0x30u); }

template <bool CSM>
__device__ __forceinline__ bool trav_init(const DevScene& sc, const uint32_t* __restrict__ coarse, const Ray& r, Trav& s)
{
    GridIter it;
    if (!grid_trace_ray(sc, r, it)) return false;
    s.tx = it.tx; s.ty = it.ty; s.tz = it.tz;
    s.dx = __uint_as_float(__float_as_uint(it.dx) | (it.sx == 1u ? 0u : 0x80000000u));
    s.dy = __uint_as_float(__float_as_uint(it.dy) | (it.sy == 1u ? 0u : 0x80000000u));
    s.dz = __uint_as_float(__float_as_uint(it.dz) | (it.sz == 1u ? 0u : 0x80000000u));
    s.c = it.cx | (it.cy << 10) | (it.cz << 20);
    s.ec = it.ex | (it.ey << 10) | (it.ez << 20);
    trav_load_block<CSM>(sc, coarse, s);
    return true;
}

// Iterator.next: axis = {2,1,2,1,2,2,0,0}[(tx<ty)<<2 | (tx<tz)<<1 | (ty<tz)] (linalg.zig:479-484):
// x iff tx<ty && tx<tz; y iff !(tx<ty) && ty<tz; otherwise z (ties prefer z, then y). Returns +inf without
// moving when the cell is the exit cell on that axis (linalg.zig:486-488).
__device__ __forceinline__ float trav_next(Trav& s)
{
    const bool xy = s.tx < s.ty, xz = s.tx < s.tz, yz = s.ty < s.tz;
    const bool ax = xy && xz, ay = !xy && yz, az = !ax && !ay;
    const float t = ax ? s.tx : (ay ? s.ty : s.tz);
    const float d = ax ? s.dx : (ay ? s.dy : s.dz);
    const uint32_t m = ax ? 0x3FFu : (ay ? 0xFFC00u : 0x3FF00000u);
    if (((s.c ^ s.ec) & m) == 0u) return __int_as_float(0x7f800000);
    const uint32_t one = m & 0x00100401u;
    s.c = (__float_as_uint(d) >> 31) ? s.c - one : s.c + one;
    const float tn = t + fabsf(d);
    s.tx = ax ? tn : s.tx; s.ty = ay ? tn : s.ty; s.tz = az ? tn : s.tz;
    return t;
}

// Scene.traceRay (src/stage3.zig:152-186) for the 32 rays of a warp. Must be called by all 32 lanes
// (inactive lanes pass active = false). Semantics per ray are the reference's: cells in DDA order;
// within a cell the winner is the smallest (t, baked index) with t > 0, and it replaces the running
// best only if strictly nearer — which is what the sequential `nearest.t > hit.t and hit.t > 0`
// (stage3.zig:174) yields when triangles are tested in index order; the loop ends at the first cell
// whose exit crossing is >= best.t (stage3.zig:179-182). The block occupancy masks only skip the cells[] load
// of empty cells. u and v of the winner are recomputed once at the end with the same arithmetic.
template <bool STATS, bool CSM>
__device__ __forceinline__ Hit trace32(const DevScene& sc, const uint32_t* __restrict__ coarse, const Ray& r, bool active,
                                       WarpShared& ws, Counters& cn)
{
    const float inf = __int_as_float(0x7f800000);
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lt_mask = (1u << lane) - 1u, le_mask = lt_mask | (1u << lane);
    Hit best; best.t = inf; best.u = 0.0f; best.v = 0.0f; best.tri = kMiss;
    Trav s;
    bool alive = false;
    if (active) { cn.rays += 1; alive = trav_init<CSM>(sc, coarse, r, s); }
    while (__ballot_sync(kFull, alive)) {
        // ---- phase A: every live lane walks to its next non-empty cell (at most kMaxStep cells) ----
        uint32_t begin = 0, cnt = 0;
        if (alive) {
#pragma unroll 1
            for (int k = 0; k < kMaxStep; ++k) {
                if (STATS) cn.cells += 1;
                const uint32_t b = trav_cell_bit(s.c);
                if ((uint32_t)(s.fine >> b) & 1u) {
                    const uint2 c = __ldg(sc.cells + (s.bi * 64u + b));
                    begin = c.x; cnt = c.y - c.x;
                    break;
                }
                const uint32_t c0 = s.c;
                const float tn = trav_next(s);
                if (best.t <= tn) { alive = false; break; }
                if ((s.c ^ c0) & kBlockBits) trav_load_block<CSM>(sc, coarse, s);
            }
        }
        if (STATS) cn.tris += cnt;
        const uint32_t pend = __ballot_sync(kFull, cnt != 0u);
        if (pend == 0u) continue;
        // ---- phase B: pooled (ray, triangle) pairs, one per lane per pass ----
        uint32_t incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t v = __shfl_up_sync(kFull, incl, d); if (lane >= (uint32_t)d) incl += v; }
        const uint32_t excl = incl - cnt;
        const uint32_t total = __shfl_sync(kFull, incl, 31);
        if (cnt != 0u) {
            const uint32_t rank = __popc(pend & lt_mask);
            ws.item[rank][0] = make_float4(r.o.x, r.o.y, r.o.z, __uint_as_float(begin - excl));
            ws.item[rank][1] = make_float4(r.d.x, r.d.y, r.d.z, __uint_as_float(lane));
            ws.best[lane] = ~0ull;
        }
        __syncwarp();
        for (uint32_t base = 0; base < total; base += 32u) {
            const bool mine = cnt != 0u && excl >= base && excl < base + 32u;
            const uint32_t starts = __reduce_or_sync(kFull, mine ? 1u << (excl - base) : 0u);
            const uint32_t before = __popc(__ballot_sync(kFull, cnt != 0u && excl < base));
            const uint32_t j = base + lane;
            if (j < total) {
                const uint32_t rank = before + __popc(starts & le_mask) - 1u;
                const float4 i0 = ws.item[rank][0], i1 = ws.item[rank][1];
                const uint32_t ti = j + __float_as_uint(i0.w);
                const float4 a = __ldg(sc.tri_v0 + ti), b = __ldg(sc.tri_e1 + ti), c = __ldg(sc.tri_e2 + ti);
                Ray q; q.o = mk(i0.x, i0.y, i0.z); q.d = mk(i1.x, i1.y, i1.z);
                float t, u, v;
                if (tri_ray(mk(a.x, a.y, a.z), mk(b.x, b.y, b.z), mk(c.x, c.y, c.z), q, t, u, v) && t > 0.0f)
                    atomicMin(&ws.best[__float_as_uint(i1.w)], ((unsigned long long)__float_as_uint(t) << 32) | ti);
            }
        }
        __syncwarp();
        // ---- phase C: owners fold the cell's winner in and take the cell's exit crossing ----
        if (cnt != 0u) {
            const unsigned long long k = ws.best[lane];
            const float t = __uint_as_float((uint32_t)(k >> 32));
            if (k != ~0ull && best.t > t) { best.t = t; best.tri = (uint32_t)k; }
            const uint32_t c0 = s.c;
            const float tn = trav_next(s);
            if (best.t <= tn) alive = false;
            else if ((s.c ^ c0) & kBlockBits) trav_load_block<CSM>(sc, coarse, s);
        }
        __syncwarp();
    }
    if (best.tri != kMiss) {
        const float4 a = __ldg(sc.tri_v0 + best.tri), b = __ldg(sc.tri_e1 + best.tri), c = __ldg(sc.tri_e2 + best.tri);
        float t;
        tri_ray(mk(a.x, a.y, a.z), mk(b.x, b.y, b.z), mk(c.x, c.y, c.z), r, t, best.u, best.v);
    }
    return best;
}

// Triangle.Data (src/stage3.zig:42-71) as stored on the device: 5 float4 per SOURCE triangle.
struct TriData { v3 n0, n1, n2; float uv0[2], uv1[2], uv2[2]; uint32_t material; };
__device__ __forceinline__ uint32_t src_of(const DevScene& sc, uint32_t baked)
{
    return __float_as_uint(__ldg(&sc.tri_v0[baked].w));
}
__device__ __forceinline__ TriData load_tri_data(const DevScene& sc, uint32_t src)
{
    const float4* p = sc.tri_data + 5u * (size_t)src;
    const float4 q0 = __ldg(p), q1 = __ldg(p + 1), q2 = __ldg(p + 2), q3 = __ldg(p + 3);
    TriData d;
    d.n0 = mk(q0.x, q0.y, q0.z); d.n1 = mk(q0.w, q1.x, q1.y); d.n2 = mk(q1.z, q1.w, q2.x);
    d.uv0[0] = q2.y; d.uv0[1] = q2.z; d.uv1[0] = q2.w; d.uv1[1] = q3.x; d.uv2[0] = q3.y; d.uv2[1] = q3.z;
    d.material = __float_as_uint(q3.w);
    return d;
}
// Triangle.Data.interpolate (src/stage3.zig:53-71), evaluation order as written.
__device__ __forceinline__ v3 interp_normal(const TriData& d, float u, float v)
{
    const float w = 1.0f - u - v;
    return (scale(d.n0, w) + scale(d.n1, u)) + scale(d.n2, v);
}
__device__ __forceinline__ void interp_uv(const TriData& d, float u, float v, float& tu, float& tv)
{
    const float w = 1.0f - u - v;
    tu = d.uv0[0] * w + d.uv1[0] * u + d.uv2[0] * v;
    tv = d.uv0[1] * w + d.uv1[1] * u + d.uv2[1] * v;
}


// traceRayRecursive (src/stage3.zig:188-220) for the 32 paths of a warp, in iterative form. The
// reference returns `emissive + albedo * recurse(...)` per scattering level; the exact build keeps
// the (emissive, albedo) pair of every scattering level and unwinds innermost-first so the float
// evaluation order is the recursion's. Alpha pass-through levels (stage3.zig:207-213) consume a
// depth level and an RNG block but neither attenuate nor push. The bounce loop is warp-uniform
// (trace32 needs all lanes); finished lanes ride along inactive.
template <bool STATS, int MAXB, bool CSM>
__device__ __forceinline__ v3 trace_path32(const DevScene& sc, const uint32_t* __restrict__ coarse, const DevParams& prm, Ray ray,
                                           bool on, uint32_t pixel, uint32_t sample, WarpShared& ws, Counters& cn)
{
#ifdef RTC_MATH_FAST_BUILD
    v3 radiance = splat(0.0f), throughput = splat(1.0f);
#else
    float st[MAXB][6];
    int top = 0;
#endif
    const float eps = 1.1920928955078125e-07f;          // std.math.floatEps(f32)
    v3 value = splat(0.0f);                              // depth == 0 -> zeroes (stage3.zig:189-191)
    uint32_t depth = prm.max_bounce;
    on = on && depth > 0u;
    while (__ballot_sync(kFull, on)) {
        const Hit hit = trace32<STATS, CSM>(sc, coarse, ray, on, ws, cn);
        if (!on) continue;
        if (hit.tri == kMiss) { value = env_color(ray.d); on = false; continue; }
        cn.hits += 1;
        const TriData td = load_tri_data(sc, src_of(sc, hit.tri));
        const DevMaterial* m = sc.materials + td.material;
        float tu, tv;
        interp_uv(td, hit.u, hit.v, tu, tv);
        const v3 albedo = tex_sample3(m->base, tu, tv);
        const v3 emissive = tex_sample3(m->emissive, tu, tv);
        const float transparency = tex_sample1(m->alpha, tu, tv);
        const v3 normal = interp_normal(td, hit.u, hit.v);
        float rnd[4];
        rng_block(prm, pixel, sample, 1u + (prm.max_bounce - depth), rnd);
        ray.o = ray_at(ray, hit.t + eps);
        depth -= 1u;
        if (depth == 0u) on = false;
        if (rnd[0] > transparency) continue;             // same direction, no attenuation
        ray.d = normalize(normal + unit_vector(rnd[1], rnd[2]));
#ifdef RTC_MATH_FAST_BUILD
        radiance = radiance + throughput * emissive;
        throughput = throughput * albedo;
#else
        st[top][0] = emissive.x; st[top][1] = emissive.y; st[top][2] = emissive.z;
        st[top][3] = albedo.x; st[top][4] = albedo.y; st[top][5] = albedo.z;
        ++top;
#endif
    }
#ifdef RTC_MATH_FAST_BUILD
    return radiance + throughput * value;
#else
    while (top > 0) {
        --top;
        value = mk(st[top][0], st[top][1], st[top][2]) + mk(st[top][3], st[top][4], st[top][5]) * value;
    }
    return value;
#endif
}

__device__ __forceinline__ void flush_counters(const Counters& cn, unsigned long long* counters, bool stats)
{
    uint32_t v[5] = { cn.rays, cn.cells, cn.tris, cn.hits, cn.paths };
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        unsigned long long x = v[i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(kFull, x, o);
        if ((threadIdx.x & 31) == 0 && x != 0 && (stats || (i != CNT_CELLS && i != CNT_TRIS)))
            atomicAdd(counters + i, x);
    }
}

// The coarse block bitmap staged in (dynamic) shared memory once per CTA; nullptr when it does not fit
// (CSM = false: probed in global memory instead).
template <bool CSM>
__device__ __forceinline__ const uint32_t* stage_coarse(const DevScene& sc)
{
    extern __shared__ uint32_t s_coarse[];
    if (!CSM) return nullptr;
    for (uint32_t i = threadIdx.x; i < sc.coarse_words; i += kThreads) s_coarse[i] = __ldg(sc.coarse + i);
    __syncthreads();
    return s_coarse;
}

template <bool STATS, int MAXB, bool CSM>
__global__ void __launch_bounds__(kThreads)
k_path(const __grid_constant__ DevScene sc, const __grid_constant__ DevCamera cam,
       const __grid_constant__ DevParams prm, float4* __restrict__ accum,
       unsigned long long* __restrict__ counters, uint32_t* __restrict__ work)
{
    __shared__ WarpShared s_ws[kWarps];
    const uint32_t* coarse = stage_coarse<CSM>(sc);
    const uint32_t lane = threadIdx.x & 31u;
    WarpShared& ws = s_ws[threadIdx.x >> 5];
    const uint32_t S = prm.num_samples;
    const uint32_t SL = S < 32u ? S : 32u;              // lanes per pixel
    const uint32_t ppw = 32u / SL;                      // pixels per warp pass
    const uint32_t passes = (S + 31u) / 32u;
    const uint32_t groups = (prm.num_pixels + ppw - 1u) / ppw;
    const uint32_t pj = lane / SL, sj = lane - pj * SL;
    Counters cn = { 0, 0, 0, 0, 0 };
    for (;;) {
        uint32_t g = 0;
        if (lane == 0) g = atomicAdd(work, 1u);
        g = __shfl_sync(kFull, g, 0);
        if (g >= groups) break;
        const uint32_t local = g * ppw + pj;
        const bool lane_on = (pj < ppw) && (local < prm.num_pixels);
        const uint32_t pixel = prm.first_pixel + local;
        const float x = (float)(pixel % cam.w), y = (float)(pixel / cam.w);
        v3 running = splat(0.0f);
        for (uint32_t pass = 0; pass < passes; ++pass) {
            const uint32_t s = pass * 32u + sj;
            const bool on = lane_on && s < S;
            const uint32_t sample = prm.sample_offset + s;
            Ray ray; ray.o = splat(0.0f); ray.d = splat(0.0f);
            if (on) {
                float jit[4];
                rng_block(prm, pixel, sample, 0u, jit);
                ray = camera_ray(cam, x + jit[0], y + jit[1]);
                cn.paths += 1;
            }
            const v3 col = trace_path32<STATS, MAXB, CSM>(sc, coarse, prm, ray, on, pixel, sample, ws, cn);
            ws.col[lane][0] = col.x; ws.col[lane][1] = col.y; ws.col[lane][2] = col.z;
            __syncwarp();
            if (lane_on && sj == 0) {                    // pixel = pixel.add(ray_color), in sample order (stage3.zig:237-241)
                const uint32_t left = S - pass * 32u;
                const uint32_t n = left < SL ? left : SL;
                for (uint32_t k = 0; k < n; ++k)
                    running = running + mk(ws.col[lane + k][0], ws.col[lane + k][1], ws.col[lane + k][2]);
            }
            __syncwarp();
        }
        if (lane_on && sj == 0) {
            float4 a = accum[pixel];
            a.x = a.x + running.x; a.y = a.y + running.y; a.z = a.z + running.z; a.w = a.w + (float)S;
            accum[pixel] = a;
        }
    }
    flush_counters(cn, counters, STATS);
}

// Deterministic slice (RTC_MODE_PRIMARY): Camera.getRay at the pixel centre + Scene.traceRay, then
// traceRayRecursive with depth 1 and no RNG: hit -> emissive sample, miss -> sky. One lane per
// pixel; a warp covers 32 consecutive pixels of a row.
template <bool STATS, bool CSM>
__global__ void __launch_bounds__(kThreads)
k_primary(const __grid_constant__ DevScene sc, const __grid_constant__ DevCamera cam,
          const __grid_constant__ DevParams prm, float4* __restrict__ accum, PrimaryOut out,
          unsigned long long* __restrict__ counters)
{
    __shared__ WarpShared s_ws[kWarps];
    const uint32_t* coarse = stage_coarse<CSM>(sc);
    WarpShared& ws = s_ws[threadIdx.x >> 5];
    Counters cn = { 0, 0, 0, 0, 0 };
    const uint32_t stride = gridDim.x * kThreads;
    const uint32_t rounds = (prm.num_pixels + stride - 1u) / stride;
    for (uint32_t it = 0; it < rounds; ++it) {
        const uint32_t local = it * stride + blockIdx.x * kThreads + threadIdx.x;
        const bool on = local < prm.num_pixels;
        const uint32_t pixel = prm.first_pixel + local;
        Ray ray; ray.o = splat(0.0f); ray.d = splat(0.0f);
        if (on) {
            const float x = (float)(pixel % cam.w), y = (float)(pixel / cam.w);
            ray = camera_ray(cam, x + 0.5f, y + 0.5f);
            cn.paths += 1;
        }
        const Hit hit = trace32<STATS, CSM>(sc, coarse, ray, on, ws, cn);
        if (!on) continue;
        v3 col;
        uint32_t id = kMiss;
        if (hit.tri == kMiss) {
            col = env_color(ray.d);
        } else {
            cn.hits += 1;
            id = src_of(sc, hit.tri);
            const TriData td = load_tri_data(sc, id);
            float tu, tv;
            interp_uv(td, hit.u, hit.v, tu, tv);
            col = tex_sample3(sc.materials[td.material].emissive, tu, tv);
        }
        if (out.hit_id) out.hit_id[pixel] = id;
        if (out.hit_tuv) {
            out.hit_tuv[3 * (size_t)pixel] = hit.t;
            out.hit_tuv[3 * (size_t)pixel + 1] = hit.u;
            out.hit_tuv[3 * (size_t)pixel + 2] = hit.v;
        }
        float4 a = accum[pixel];
        a.x = a.x + col.x; a.y = a.y + col.y; a.z = a.z + col.z; a.w = a.w + 1.0f;
        accum[pixel] = a;
    }
    flush_counters(cn, counters, STATS);
}

// img[i] = pixel.mul(inv_num_samples).toRGB() (src/stage3.zig:223,242):
// inv = 1/(f32)S computed once, then multiplied.
__global__ void __launch_bounds__(kThreads)
k_resolve(const float4* __restrict__ accum, uint32_t num_pixels, uint32_t total_samples,
          uint8_t* __restrict__ rgb, float* __restrict__ radiance)
{
    const uint32_t i = blockIdx.x * kThreads + threadIdx.x;
    if (i >= num_pixels) return;
    const float inv = 1.0f / (float)total_samples;
    const float4 a = accum[i];
    const float r = a.x * inv, g = a.y * inv, b = a.z * inv;
    if (radiance) { radiance[3 * (size_t)i] = r; radiance[3 * (size_t)i + 1] = g; radiance[3 * (size_t)i + 2] = b; }
    if (rgb) {
        rgb[3 * (size_t)i] = (uint8_t)to_rgb8(r);
        rgb[3 * (size_t)i + 1] = (uint8_t)to_rgb8(g);
        rgb[3 * (size_t)i + 2] = (uint8_t)to_rgb8(b);
    }
}

template <typename K>
static int blocks_per_sm(K kernel, size_t smem)
{
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, kThreads, smem) != cudaSuccess || n < 1) n = 1;
    return n;
}

static bool coarse_in_smem(const DevScene& sc) { return sc.coarse_words <= kCoarseSmemMaxWords; }

template <bool STATS, int MAXB, bool CSM>
static cudaError_t run_path(const DevScene& sc, const DevCamera& cam, const DevParams& prm, float4* accum,
                            unsigned long long* counters, uint32_t* work, const LaunchCfg& cfg)
{
    auto kern = k_path<STATS, MAXB, CSM>;
    const size_t smem = CSM ? sc.coarse_words * sizeof(uint32_t) : 0;
    const int grid = cfg.sm_count * blocks_per_sm(kern, smem);   // persistent: one wave, dynamic work counter
    cudaError_t e = cudaMemsetAsync(work, 0, sizeof(uint32_t), cfg.stream);
    if (e != cudaSuccess) return e;
    kern<<<grid, kThreads, smem, cfg.stream>>>(sc, cam, prm, accum, counters, work);
    return cudaGetLastError();
}

template <bool STATS, int MAXB>
static cudaError_t run_path_csm(const DevScene& sc, const DevCamera& cam, const DevParams& prm, float4* accum,
                                unsigned long long* counters, uint32_t* work, const LaunchCfg& cfg)
{
    return coarse_in_smem(sc) ? run_path<STATS, MAXB, true>(sc, cam, prm, accum, counters, work, cfg)
                              : run_path<STATS, MAXB, false>(sc, cam, prm, accum, counters, work, cfg);
}

cudaError_t launch_path(const DevScene& sc, const DevCamera& cam, const DevParams& prm, float4* accum,
                        unsigned long long* counters, uint32_t* work, const LaunchCfg& cfg)
{
    if (prm.num_pixels == 0 || prm.num_samples == 0) return cudaSuccess;
    if (prm.max_bounce <= 8)
        return cfg.stats ? run_path_csm<true, 8>(sc, cam, prm, accum, counters, work, cfg)
                         : run_path_csm<false, 8>(sc, cam, prm, accum, counters, work, cfg);
    if (prm.max_bounce <= 64)
        return cfg.stats ? run_path_csm<true, 64>(sc, cam, prm, accum, counters, work, cfg)
                         : run_path_csm<false, 64>(sc, cam, prm, accum, counters, work, cfg);
    return cudaErrorInvalidValue;
}

template <bool STATS, bool CSM>
static cudaError_t run_primary(const DevScene& sc, const DevCamera& cam, const DevParams& prm, float4* accum,
                               PrimaryOut out, unsigned long long* counters, const LaunchCfg& cfg)
{
    auto kern = k_primary<STATS, CSM>;
    const size_t smem = CSM ? sc.coarse_words * sizeof(uint32_t) : 0;
    const uint32_t need = (prm.num_pixels + kThreads - 1) / kThreads;
    const uint32_t cap = (uint32_t)(cfg.sm_count * blocks_per_sm(kern, smem));
    kern<<<need < cap ? need : cap, kThreads, smem, cfg.stream>>>(sc, cam, prm, accum, out, counters);
    return cudaGetLastError();
}

cudaError_t launch_primary(const DevScene& sc, const DevCamera& cam, const DevParams& prm, float4* accum,
                           PrimaryOut out, unsigned long long* counters, const LaunchCfg& cfg)
{
    if (prm.num_pixels == 0) return cudaSuccess;
    if (coarse_in_smem(sc))
        return cfg.stats ? run_primary<true, true>(sc, cam, prm, accum, out, counters, cfg)
                         : run_primary<false, true>(sc, cam, prm, accum, out, counters, cfg);
    return cfg.stats ? run_primary<true, false>(sc, cam, prm, accum, out, counters, cfg)
                     : run_primary<false, false>(sc, cam, prm, accum, out, counters, cfg);
}

cudaError_t launch_resolve(const float4* accum, uint32_t num_pixels, uint32_t total_samples,
                           uint8_t* rgb, float* radiance, cudaStream_t stream)
{
    if (num_pixels == 0) return cudaSuccess;
    k_resolve<<<(num_pixels + kThreads - 1) / kThreads, kThreads, 0, stream>>>(accum, num_pixels, total_samples, rgb, radiance);
    return cudaGetLastError();
}

}  // namespace RTC_NS
}  // namespace rtc
