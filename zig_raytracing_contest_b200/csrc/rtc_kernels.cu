// rtc_kernels.cu — sm_100a render kernels for the reference's per-pixel path-tracing loop
// (src/stage3.zig:152-245). Compiled twice: -DRTC_NS=exact -fmad=false and -DRTC_NS=fast -fmad=true.
//
// Mapping. One lane = one (pixel, sample) path. A warp takes a group of 32/min(S,32) consecutive
// pixels from a persistent work counter, so the primary rays of a warp are coherent (same or
// neighbouring pixels -> same cells, same triangle records, broadcast loads). Per-sample radiance
// is staged in shared memory and summed by the pixel's first lane IN SAMPLE ORDER, which
// reproduces renderWorker's sequential `pixel = pixel.add(ray_color)` (stage3.zig:237-241) bit
// for bit in the exact build.
#include "rtc_kernels.h"

#ifndef RTC_NS
#error "compile with -DRTC_NS=exact or -DRTC_NS=fast"
#endif

namespace rtc {
namespace RTC_NS {

static constexpr int kThreads = 256;
static constexpr int kWarps = kThreads / 32;
static constexpr uint32_t kMiss = 0xFFFFFFFFu;

struct Hit { float t, u, v; uint32_t tri; };

struct Counters { uint32_t rays, cells, tris, hits, paths; };

// Scene.traceRay (src/stage3.zig:152-186): nearest hit over the grid cells along the ray, cells in
// DDA order, triangles of a cell in baked order, strict `best.t > t && t > 0` (first tested wins).
// occ[] only skips the cells[] load of empty cells; the DDA state advances exactly as in the
// reference so every t_next_crossing is the reference's.
template <bool STATS>
__device__ __forceinline__ Hit trace_ray(const DevScene& sc, const Ray& r, Counters& cn)
{
    const float inf = __int_as_float(0x7f800000);
    Hit best; best.t = inf; best.u = 0.0f; best.v = 0.0f; best.tri = kMiss;
    cn.rays += 1;
    GridIter it;
    if (!grid_trace_ray(sc, r, it)) return best;
    const uint32_t rx = sc.res[0], rxy = sc.res[0] * sc.res[1];
    for (;;) {
        const uint32_t ci = it.cz * rxy + it.cy * rx + it.cx;
        if (STATS) cn.cells += 1;
        if ((__ldg(sc.occ + (ci >> 5)) >> (ci & 31u)) & 1u) {
            const uint2 c = __ldg(sc.cells + ci);
            if (STATS) cn.tris += c.y - c.x;
            for (uint32_t ti = c.x; ti < c.y; ++ti) {
                const float4 a = __ldg(sc.tri_pos + 3u * (size_t)ti);
                const float4 b = __ldg(sc.tri_pos + 3u * (size_t)ti + 1);
                const float4 c2 = __ldg(sc.tri_pos + 3u * (size_t)ti + 2);
                float t, u, v;
                if (tri_ray(mk(a.x, a.y, a.z), mk(b.x, b.y, b.z), mk(c2.x, c2.y, c2.z), r, t, u, v)) {
                    if (best.t > t && t > 0.0f) { best.t = t; best.u = u; best.v = v; best.tri = ti; }
                }
            }
        }
        const float tn = grid_next(it);
        if (best.t <= tn) break;
    }
    return best;
}

// Triangle.Data (src/stage3.zig:42-71) as stored on the device: 5 float4 per SOURCE triangle.
struct TriData { v3 n0, n1, n2; float uv0[2], uv1[2], uv2[2]; uint32_t material; };
__device__ __forceinline__ uint32_t src_of(const DevScene& sc, uint32_t baked)
{
    return __float_as_uint(__ldg(&sc.tri_pos[3u * (size_t)baked].w));
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

// traceRayRecursive (src/stage3.zig:188-220) in iterative form. The reference returns
// `emissive + albedo * recurse(...)` per scattering level; the exact build keeps the (emissive,
// albedo) pair of every scattering level and unwinds innermost-first so the float evaluation
// order is the recursion's. Alpha pass-through levels (stage3.zig:207-213) consume a depth level
// and an RNG block but neither attenuate nor push.
template <bool STATS, int MAXB>
__device__ __forceinline__ v3 trace_path(const DevScene& sc, const DevParams& prm, Ray ray,
                                         uint32_t pixel, uint32_t sample, Counters& cn)
{
    const float inf = __int_as_float(0x7f800000);
    const float eps = 1.1920928955078125e-07f;          // std.math.floatEps(f32)
#ifdef RTC_MATH_FAST_BUILD
    v3 radiance = splat(0.0f), throughput = splat(1.0f);
#else
    float st[MAXB][6];
    int top = 0;
#endif
    v3 value = splat(0.0f);                              // depth == 0 -> zeroes (stage3.zig:189-191)
    for (uint32_t depth = prm.max_bounce; depth > 0; --depth) {
        const Hit hit = trace_ray<STATS>(sc, ray, cn);
        if (hit.t == inf) { value = env_color(ray.d); break; }
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
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xFFFFFFFFu, x, o);
        if ((threadIdx.x & 31) == 0 && x != 0 && (stats || (i != CNT_CELLS && i != CNT_TRIS)))
            atomicAdd(counters + i, x);
    }
}

template <bool STATS, int MAXB>
__global__ void __launch_bounds__(kThreads)
k_path(const __grid_constant__ DevScene sc, const __grid_constant__ DevCamera cam,
       const __grid_constant__ DevParams prm, float4* __restrict__ accum,
       unsigned long long* __restrict__ counters, uint32_t* __restrict__ work)
{
    __shared__ float s_col[kWarps][32][3];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
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
        g = __shfl_sync(0xFFFFFFFFu, g, 0);
        if (g >= groups) break;
        const uint32_t local = g * ppw + pj;
        const bool lane_on = (pj < ppw) && (local < prm.num_pixels);
        const uint32_t pixel = prm.first_pixel + local;
        const float x = (float)(pixel % cam.w), y = (float)(pixel / cam.w);
        v3 running = splat(0.0f);
        for (uint32_t pass = 0; pass < passes; ++pass) {
            const uint32_t s = pass * 32u + sj;
            v3 col = splat(0.0f);
            if (lane_on && s < S) {
                const uint32_t sample = prm.sample_offset + s;
                float jit[4];
                rng_block(prm, pixel, sample, 0u, jit);
                const Ray ray = camera_ray(cam, x + jit[0], y + jit[1]);
                cn.paths += 1;
                col = trace_path<STATS, MAXB>(sc, prm, ray, pixel, sample, cn);
            }
            s_col[warp][lane][0] = col.x; s_col[warp][lane][1] = col.y; s_col[warp][lane][2] = col.z;
            __syncwarp();
            if (lane_on && sj == 0) {
                const uint32_t left = S - pass * 32u;
                const uint32_t n = left < SL ? left : SL;
                for (uint32_t k = 0; k < n; ++k)
                    running = running + mk(s_col[warp][lane + k][0], s_col[warp][lane + k][1], s_col[warp][lane + k][2]);
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
template <bool STATS>
__global__ void __launch_bounds__(kThreads)
k_primary(const __grid_constant__ DevScene sc, const __grid_constant__ DevCamera cam,
          const __grid_constant__ DevParams prm, float4* __restrict__ accum, PrimaryOut out,
          unsigned long long* __restrict__ counters)
{
    const float inf = __int_as_float(0x7f800000);
    Counters cn = { 0, 0, 0, 0, 0 };
    const uint32_t stride = gridDim.x * kThreads;
    for (uint32_t local = blockIdx.x * kThreads + threadIdx.x; local < prm.num_pixels; local += stride) {
        const uint32_t pixel = prm.first_pixel + local;
        const float x = (float)(pixel % cam.w), y = (float)(pixel / cam.w);
        const Ray ray = camera_ray(cam, x + 0.5f, y + 0.5f);
        cn.paths += 1;
        const Hit hit = trace_ray<STATS>(sc, ray, cn);
        v3 col;
        uint32_t id = kMiss;
        if (hit.t == inf) {
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
static int blocks_per_sm(K kernel)
{
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, kThreads, 0) != cudaSuccess || n < 1) n = 1;
    return n;
}

template <bool STATS, int MAXB>
static cudaError_t run_path(const DevScene& sc, const DevCamera& cam, const DevParams& prm, float4* accum,
                            unsigned long long* counters, uint32_t* work, const LaunchCfg& cfg)
{
    auto kern = k_path<STATS, MAXB>;
    const int grid = cfg.sm_count * blocks_per_sm(kern);   // persistent: one wave, dynamic work counter
    cudaError_t e = cudaMemsetAsync(work, 0, sizeof(uint32_t), cfg.stream);
    if (e != cudaSuccess) return e;
    kern<<<grid, kThreads, 0, cfg.stream>>>(sc, cam, prm, accum, counters, work);
    return cudaGetLastError();
}

cudaError_t launch_path(const DevScene& sc, const DevCamera& cam, const DevParams& prm, float4* accum,
                        unsigned long long* counters, uint32_t* work, const LaunchCfg& cfg)
{
    if (prm.num_pixels == 0 || prm.num_samples == 0) return cudaSuccess;
    if (prm.max_bounce <= 8)
        return cfg.stats ? run_path<true, 8>(sc, cam, prm, accum, counters, work, cfg)
                         : run_path<false, 8>(sc, cam, prm, accum, counters, work, cfg);
    if (prm.max_bounce <= 64)
        return cfg.stats ? run_path<true, 64>(sc, cam, prm, accum, counters, work, cfg)
                         : run_path<false, 64>(sc, cam, prm, accum, counters, work, cfg);
    return cudaErrorInvalidValue;
}

cudaError_t launch_primary(const DevScene& sc, const DevCamera& cam, const DevParams& prm, float4* accum,
                           PrimaryOut out, unsigned long long* counters, const LaunchCfg& cfg)
{
    if (prm.num_pixels == 0) return cudaSuccess;
    const uint32_t need = (prm.num_pixels + kThreads - 1) / kThreads;
    if (cfg.stats) {
        const uint32_t cap = (uint32_t)(cfg.sm_count * blocks_per_sm(k_primary<true>));
        k_primary<true><<<need < cap ? need : cap, kThreads, 0, cfg.stream>>>(sc, cam, prm, accum, out, counters);
    } else {
        const uint32_t cap = (uint32_t)(cfg.sm_count * blocks_per_sm(k_primary<false>));
        k_primary<false><<<need < cap ? need : cap, kThreads, 0, cfg.stream>>>(sc, cam, prm, accum, out, counters);
    }
    return cudaGetLastError();
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
