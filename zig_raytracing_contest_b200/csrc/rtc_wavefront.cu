// rtc_wavefront.cu — sm_100a render kernels for the reference's per-pixel path-tracing loop
// (src/stage3.zig:152-245), as a WAVEFRONT. Compiled twice: -DRTC_NS=exact -fmad=false and
// -DRTC_NS=fast -fmad=true.
//
// A batch is a pixel range x a sample range = up to `cap` path slots. Per batch:
//   k_gen     renderWorker's `camera.getRay(x + U, y + U)` (stage3.zig:238) + Grid.traceRay's DDA set-up
//             (linalg.zig:443-469) for every slot -> ray queue 0
//   per bounce d = 0 .. max_bounce-1:
//     k_trace Scene.traceRay (stage3.zig:152-186) for every queued ray. Persistent warps; a lane whose ray
//             ends writes its hit and IMMEDIATELY takes the next ray of the queue (no lane waits for the
//             longest ray of its warp). Every lane walks its own DDA; the triangle tests of the 32 rays in
//             flight are pooled and dealt out one (ray, triangle) pair per lane.
//     k_shade the body of traceRayRecursive (stage3.zig:193-219) for every traced ray: sky on a miss, else
//             attributes + textures + RNG; alpha pass-through or Lambertian scatter; survivors are
//             ballot-compacted into the other ray queue (with their DDA set-up done here, at full lanes)
//   k_accum   unwinds `emissive + albedo * recurse` innermost-first per path (the recursion's evaluation
//             order) and adds the samples to the pixel in sample order (stage3.zig:237-241)
// The first version of this path was one persistent megakernel (one lane = one path); its profile
// (profiles/r1b_*) showed 16 of 32 lanes active and 25 % occupancy at 128 registers.
#include "rtc_kernels.h"
#include "rtc_trav.cuh"
#include "rtc_bvh.cuh"

#ifndef RTC_NS
#error "compile with -DRTC_NS=exact or -DRTC_NS=fast"
#endif

namespace rtc {
namespace RTC_NS {

static constexpr int kTraceThreads = 256;
static constexpr int kTraceWarps = kTraceThreads / 32;
static constexpr int kThreads = 256;                 // gen / shade / accum / resolve
static constexpr uint32_t kMiss = 0xFFFFFFFFu;
static constexpr uint32_t kFull = 0xFFFFFFFFu;
#ifndef RTC_TRACE_MINB
#define RTC_TRACE_MINB 4
#endif
#ifndef RTC_KMAXSTEP
#define RTC_KMAXSTEP 4
#endif
#ifndef RTC_TWO_LEVEL
#define RTC_TWO_LEVEL 1
#endif

static constexpr int kMaxStep = RTC_KMAXSTEP;        // DDA steps a lane may take per round before the pooled triangle pass
#ifndef RTC_CHUNK
#define RTC_CHUNK 64
#endif
static constexpr uint32_t kChunk = RTC_CHUNK;        // queue entries a warp reserves per atomic
// A warp that reserves a chunk of its input queue asks L2 for the chunk's lines: the queue was written by the previous
// kernel (6 GB per bounce of cfg3) and comes from DRAM, and the chunk is consumed a few entries at a time over many
// rounds — every refill but the first then finds its entries in L2. +0.4 % on cfg3 (profiles/r2z_ab_stack.txt).
#ifndef RTC_QUEUE_PREFETCH
#define RTC_QUEUE_PREFETCH 1
#endif
template <uint32_t PLANES>
__device__ __forceinline__ void queue_prefetch_chunk(const RayQueue& q, uint32_t base, uint32_t end, uint32_t lane)
{
#if RTC_QUEUE_PREFETCH
    // 8 entries of 16 B per 128-byte line: lane -> (plane, line of the chunk)
    constexpr uint32_t kLines = kChunk / 8u;
    static_assert(PLANES * kLines <= 32u, "one prefetch per lane covers the chunk");
    if (lane < PLANES * kLines) {
        const uint32_t plane = lane / kLines, e = base + 8u * (lane % kLines);
        const float4* a = plane == 0u ? q.a0 : plane == 1u ? q.a1 : plane == 2u ? q.a2 : q.a3;
        if (e < end) asm volatile("prefetch.global.L2 [%0];" :: "l"(a + e));
    }
#endif
}

// ---------------------------------------------------------------------------------------------------------
// Ray queue entry (64 B, SoA over four float4 arrays): the ray and its DDA set-up.
//   a0 = o.xyz | slot      a1 = d.xyz | packed cell      a2 = t_next.xyz | packed exit cell
//   a3 = signed t_delta.xyz | scatter levels pushed so far
struct QueuedRay {
    Ray r;
    uint32_t slot, top;
    float tx, ty, tz, dx, dy, dz;
    uint32_t c, ec;
};
__device__ __forceinline__ bool queue_setup(const DevScene& sc, const Ray& r, QueuedRay& q)
{
    GridIter it;
    if (!grid_trace_ray(sc, r, it)) return false;
    q.r = r;
    q.tx = it.tx; q.ty = it.ty; q.tz = it.tz;
    q.dx = __uint_as_float(__float_as_uint(it.dx) | (it.sx == 1u ? 0u : 0x80000000u));
    q.dy = __uint_as_float(__float_as_uint(it.dy) | (it.sy == 1u ? 0u : 0x80000000u));
    q.dz = __uint_as_float(__float_as_uint(it.dz) | (it.sz == 1u ? 0u : 0x80000000u));
    q.c = pack_cell(sc.shift, it.cx, it.cy, it.cz);
    q.ec = pack_cell(sc.shift, it.ex, it.ey, it.ez);
    // empty blocks are taken in one step (rtc_trav.cuh) unless the grid's boundary can lie inside a block or a
    // crossing value is -inf / NaN (a -0.0 direction component): those rays walk cell by cell
    const bool two_level = ((sc.res[0] | sc.res[1] | sc.res[2]) & 3u) == 0u && it.tx > -f_inf() && it.ty > -f_inf() && it.tz > -f_inf();
    if (!two_level) q.ec |= kTravCellOnly;
    return true;
}
template <bool BVH>
__device__ __forceinline__ void queue_store(const RayQueue& q, uint32_t i, const QueuedRay& e)
{
    // a0 / a1 carry what k_shade reads back (origin + slot, direction + top), a2 / a3 the iterator state only k_trace
    // needs. Queues, hit records and the per-slot shading stack are written once and read once, gigabytes per launch:
    // streaming (evict-first) accesses keep them from pushing the scene out of L2.
    __stcs(q.a0 + i, make_float4(e.r.o.x, e.r.o.y, e.r.o.z, __uint_as_float(e.slot)));
    __stcs(q.a1 + i, make_float4(e.r.d.x, e.r.d.y, e.r.d.z, __uint_as_float(e.top)));
    if (BVH) {                                          // RTC_ACCEL_BVH: 1/d and the octant word instead of the iterator, 48 B / ray
        __stcs(q.a2 + i, make_float4(e.tx, e.ty, e.tz, __uint_as_float(e.ec)));
        return;
    }
    __stcs(q.a2 + i, make_float4(e.tx, e.ty, e.tz, __uint_as_float(e.ec)));
    __stcs(q.a3 + i, make_float4(e.dx, e.dy, e.dz, __uint_as_float(e.c)));
}
// Appends the rays of the threads with `push` to the queue. Every thread of the CTA calls this (the callers loop a
// CTA-uniform number of rounds). ONE atomic per CTA and round: the per-warp counts go through shared memory, thread 0
// reserves the CTA's range. With one atomic per warp the 6.5e5 same-address atomics of a 20 M-ray launch were what
// k_gen and k_shade waited on (profiles/r2d_*: k_gen at 24 % issue-active while writing 1 GB).
#ifndef RTC_QSORT
#define RTC_QSORT 0
#endif
template <bool BVH>
__device__ __forceinline__ void queue_push(const RayQueue& q, uint32_t* count, bool push, const QueuedRay& e)
{
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    constexpr int kWarps = kThreads / 32;
#if RTC_QSORT
    // EXPERIMENT, off by default (profiles/r2j_ab_*: the traversal kernels gain 0.2 % from it, k_gen / k_shade pay 3.5 ms
    // per frame for the sort: 2974 vs 3041 Mrays/s with the BVH, 1778 vs 1809 with the grid).
    // The CTA's rays of this round go out GROUPED BY DIRECTION OCTANT (a counting sort over 8 buckets in shared
    // memory): consecutive queue entries — the rays one traversal warp picks up together — then share the sign of every
    // direction component, i.e. the DDA's step directions / the BVH's child order, and they come from neighbouring
    // pixels. The order of the queue has no influence on the results (every ray carries its slot).
    __shared__ uint32_t s_cnt[8 * kWarps];              // [bucket][warp] -> count, then exclusive offset inside the CTA's range
    __shared__ uint32_t s_base;
    if (threadIdx.x < 8 * kWarps) s_cnt[threadIdx.x] = 0u;
    __syncthreads();
    const uint32_t bkt = push ? ((e.r.d.x < 0.0f ? 1u : 0u) | (e.r.d.y < 0.0f ? 2u : 0u) | (e.r.d.z < 0.0f ? 4u : 0u)) : 8u;
    const uint32_t peers = __match_any_sync(kFull, bkt);
    const uint32_t rank = (uint32_t)__popc(peers & ((1u << lane) - 1u));
    if (push && rank == 0u) s_cnt[bkt * kWarps + warp] = (uint32_t)__popc(peers);
    __syncthreads();
    if (warp == 0) {
        // exclusive scan of the 64 counts in bucket-major order: two per lane
        const uint32_t c0 = s_cnt[2u * lane], c1 = s_cnt[2u * lane + 1u];
        uint32_t incl = c0 + c1;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t v = __shfl_up_sync(kFull, incl, d); if (lane >= (uint32_t)d) incl += v; }
        const uint32_t excl = incl - (c0 + c1);
        s_cnt[2u * lane] = excl; s_cnt[2u * lane + 1u] = excl + c0;
        if (lane == 31u) s_base = incl ? atomicAdd(count, incl) : 0u;
    }
    __syncthreads();
    if (push) queue_store<BVH>(q, s_base + s_cnt[bkt * kWarps + warp] + rank, e);
#else
    __shared__ uint32_t s_cnt[kWarps];
    __shared__ uint32_t s_base;
    const uint32_t m = __ballot_sync(kFull, push);
    if (lane == 0) s_cnt[warp] = (uint32_t)__popc(m);
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t total = 0;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) { const uint32_t c = s_cnt[w]; s_cnt[w] = total; total += c; }
        s_base = total ? atomicAdd(count, total) : 0u;
    }
    __syncthreads();
    if (push) queue_store<BVH>(q, s_base + s_cnt[warp] + __popc(m & ((1u << lane) - 1u)), e);
#endif
}

// What a new ray needs before it is queued. Grid: Grid.traceRay's iterator (a ray that misses the grid's box is a
// miss right away, stage3.zig:159). BVH: 1/d and the octant word (the root's slab test plays the part of the grid's
// box test in k_trace_bvh).
template <bool BVH>
__device__ __forceinline__ bool ray_setup(const DevScene& sc, const Ray& r, QueuedRay& q)
{
    if (BVH) {
        BvhRay br;
        bvh_ray_init(br, r.o, r.d);                     // three divisions, here at full lanes instead of at k_trace_bvh's refill
        q.r = r; q.tx = br.ix; q.ty = br.iy; q.tz = br.iz; q.ec = br.octinv4;
        return true;
    }
    return queue_setup(sc, r, q);
}

struct Counters { uint32_t rays, cells, tris, hits, paths; };
__device__ __forceinline__ void flush_counters(const Counters& cn, unsigned long long* counters)
{
    uint32_t v[5] = { cn.rays, cn.cells, cn.tris, cn.hits, cn.paths };
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        unsigned long long x = v[i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(kFull, x, o);
        if ((threadIdx.x & 31) == 0 && x != 0) atomicAdd(counters + i, x);
    }
}

// The coarse block bitmap staged in (dynamic) shared memory once per CTA; nullptr when it does not fit
// (CSM = false: probed in global memory instead).
template <bool CSM>
__device__ __forceinline__ const uint32_t* stage_coarse(const DevScene& sc)
{
    extern __shared__ uint32_t s_coarse[];
    if (!CSM) return nullptr;
    for (uint32_t i = threadIdx.x; i < sc.coarse_words; i += blockDim.x) s_coarse[i] = __ldg(sc.coarse + i);
    __syncthreads();
    return s_coarse;
}

// ---------------------------------------------------------------------------------------------------------
// k_gen: one thread per path slot. slot = sample_local * num_pixels + pixel_local, so consecutive lanes are
// consecutive pixels of a row (coherent primary rays). RNG block 0 of the path = the two jitter draws
// (stage3.zig:238). PRIMARY: pixel centre, no RNG.
template <bool PRIMARY, bool BVH>
__global__ void __launch_bounds__(kThreads)
k_gen(const __grid_constant__ DevScene sc, const __grid_constant__ DevCamera cam, const __grid_constant__ WaveParams wp,
      const __grid_constant__ WaveScratch ws, PrimaryOut out, unsigned long long* __restrict__ counters)
{
    Counters cn = { 0, 0, 0, 0, 0 };
    const uint32_t nslots = wp.num_pixels * wp.ns;
    const uint32_t stride = gridDim.x * kThreads;
    const uint32_t rounds = (nslots + stride - 1u) / stride;
    for (uint32_t it = 0; it < rounds; ++it) {
        const uint32_t slot = it * stride + blockIdx.x * kThreads + threadIdx.x;
        const bool on = slot < nslots;
        bool push = false;
        QueuedRay e;
        if (on) {
            const uint32_t sl = slot / wp.num_pixels, pl = slot - sl * wp.num_pixels;
            const uint32_t pixel = slot_pixel(wp, pl);
            const float x = (float)(pixel % cam.w), y = (float)(pixel / cam.w);
            cn.paths += 1;
            float4 leaf = make_float4(0.0f, 0.0f, 0.0f, __uint_as_float(0u));     // depth == 0 -> zeroes (stage3.zig:189-191)
            if (wp.max_bounce > 0u) {
                Ray ray;
                if (PRIMARY) {
                    ray = camera_ray(cam, x + 0.5f, y + 0.5f);
                } else {
                    float jit[4];
                    rng_block(wp.seed_lo, wp.seed_hi, pixel, wp.s0 + sl, 0u, jit);
                    ray = camera_ray(cam, x + jit[0], y + jit[1]);
                }
                cn.rays += 1;
                e.slot = slot; e.top = 0u;
                push = ray_setup<BVH>(sc, ray, e);
                if (!push) {                                    // Grid.traceRay -> null: a miss (stage3.zig:159)
                    const v3 sky = env_color(ray.d);
                    leaf = make_float4(sky.x, sky.y, sky.z, __uint_as_float(0u));
                    if (PRIMARY) {
                        if (out.hit_id) out.hit_id[pixel] = kMiss;
                        if (out.hit_tuv) {
                            out.hit_tuv[3 * (size_t)pixel] = __int_as_float(0x7f800000);
                            out.hit_tuv[3 * (size_t)pixel + 1] = 0.0f;
                            out.hit_tuv[3 * (size_t)pixel + 2] = 0.0f;
                        }
                    }
                }
            }
            if (!push) __stcs(ws.leaf + slot, leaf);
        }
        queue_push<BVH>(ws.q[0], ws.counts, push, e);
    }
    flush_counters(cn, counters);
}

// ---------------------------------------------------------------------------------------------------------
// k_trace: Scene.traceRay (src/stage3.zig:152-186) over a ray queue. Semantics per ray are the reference's:
// cells in DDA order; within a cell the winner is the smallest (t, baked index) with t > 0, and it replaces
// the running best only if strictly nearer — which is what the sequential `nearest.t > hit.t and hit.t > 0`
// (stage3.zig:174) yields when triangles are tested in index order; the ray ends at the first cell whose
// exit crossing is >= best.t (stage3.zig:179-182). The block occupancy masks only skip the cells[] load of
// empty cells.
struct TraceShared {
    float4 ro[32];                      // per lane: ray origin (.w unused)
    float4 rd[32];                      // per lane: ray direction
    unsigned long long best[32];        // per lane: min over the cell's (t bits << 32 | baked triangle index)
    uint2 item[32];                     // compacted pending list: {owner lane, begin - exclusive prefix}
};

template <bool STATS, bool CSM>
__global__ void __launch_bounds__(kTraceThreads, RTC_TRACE_MINB)
k_trace(const __grid_constant__ DevScene sc, const RayQueue q, const uint32_t* __restrict__ count_ptr,
        uint32_t* __restrict__ head_ptr, uint2* __restrict__ hits, unsigned long long* __restrict__ counters)
{
    // The instrumented build counts every cell the reference visits, so it walks cell by cell; the production
    // build takes empty blocks in one step (same triangle-holding cells, same t values: tests/test_trav_twolevel.py).
    constexpr bool TWO = !STATS && RTC_TWO_LEVEL;
    __shared__ TraceShared s_ws[kTraceWarps];
    const uint32_t* coarse = stage_coarse<CSM>(sc);
    TraceShared& ws = s_ws[threadIdx.x >> 5];
    const float inf = __int_as_float(0x7f800000);
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lt_mask = (1u << lane) - 1u, le_mask = lt_mask | (1u << lane);
    const uint32_t n = __ldg(count_ptr);
    Counters cn = { 0, 0, 0, 0, 0 };
    uint32_t next = 0, end = 0;         // the warp's reserved chunk of queue entries (uniform)
    bool exhausted = false;             // uniform
    bool alive = false;
    Trav s;
    float best_t = inf;
    uint32_t best_tri = kMiss, qi = 0;
    s.tx = s.ty = s.tz = s.dx = s.dy = s.dz = 0.0f; s.c = s.ec = s.fine_lo = s.fine_hi = 0u;
    for (;;) {
        // ---- refill: idle lanes take the next rays of the queue ----
        const uint32_t idle = __ballot_sync(kFull, !alive);
        if (idle != 0u && !exhausted) {
            if (next == end) {
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd(head_ptr, kChunk);
                base = __shfl_sync(kFull, base, 0);
                if (base >= n) { exhausted = true; next = end = 0; }
                else { next = base; end = base + kChunk < n ? base + kChunk : n; queue_prefetch_chunk<4u>(q, next, end, lane); }
            }
            const uint32_t avail = end - next, want = __popc(idle);
            const uint32_t take = want < avail ? want : avail;
            if (!alive && (uint32_t)__popc(idle & lt_mask) < take) {
                qi = next + __popc(idle & lt_mask);
                const float4 a0 = __ldcs(q.a0 + qi), a1 = __ldcs(q.a1 + qi), a2 = __ldcs(q.a2 + qi), a3 = __ldcs(q.a3 + qi);
                ws.ro[lane] = a0; ws.rd[lane] = a1;
                s.tx = a2.x; s.ty = a2.y; s.tz = a2.z; s.dx = a3.x; s.dy = a3.y; s.dz = a3.z;
                s.c = __float_as_uint(a3.w); s.ec = __float_as_uint(a2.w);
                trav_load_block<CSM>(sc, coarse, s);
                best_t = inf; best_tri = kMiss;
                alive = true;
            }
            next += take;
        }
        if (__ballot_sync(kFull, alive) == 0u) {
            if (exhausted) break;
            continue;
        }
        // ---- phase A: every live lane walks to its next non-empty cell (at most kMaxStep cells) ----
        uint32_t begin = 0, cnt = 0;
        float enter_t = 0.0f;
        bool done = false;
        if (alive) {
#pragma unroll 1
            for (int k = 0; k < kMaxStep; ++k) {
                if (STATS) cn.cells += 1;
                if (trav_occupied(s)) {
                    const uint2 c = __ldg(sc.cells + s.c);
                    begin = c.x; cnt = c.y - c.x;
                    break;
                }
                const uint32_t w = trav_advance<CSM, TWO>(sc, coarse, s, best_t, enter_t);
                if (w == kWalkDone) { done = true; break; }
                if (TWO && w >= kWalkEnter) trav_enter(s, enter_t, w - kWalkEnter);
            }
        }
        if (STATS) cn.tris += cnt;
        const uint32_t pend = __ballot_sync(kFull, cnt != 0u);
        if (pend != 0u) {
            // ---- phase B: pooled (ray, triangle) pairs, one per lane per pass ----
            uint32_t incl = cnt;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const uint32_t v = __shfl_up_sync(kFull, incl, d); if (lane >= (uint32_t)d) incl += v; }
            const uint32_t excl = incl - cnt;
            const uint32_t total = __shfl_sync(kFull, incl, 31);
            if (cnt != 0u) {
                ws.item[__popc(pend & lt_mask)] = make_uint2(lane, begin - excl);
                ws.best[lane] = ~0ull;
            }
            __syncwarp();
            for (uint32_t base = 0; base < total; base += 32u) {
                const bool mine = cnt != 0u && excl >= base && excl < base + 32u;
                const uint32_t starts = __reduce_or_sync(kFull, mine ? 1u << (excl - base) : 0u);
                const uint32_t before = __popc(__ballot_sync(kFull, cnt != 0u && excl < base));
                const uint32_t j = base + lane;
                if (j < total) {
                    const uint2 it = ws.item[before + __popc(starts & le_mask) - 1u];
                    const uint32_t ti = j + it.y;
                    const float4 a = __ldg(sc.tri_v0 + ti), b = __ldg(sc.tri_e1 + ti), c = __ldg(sc.tri_e2 + ti);
                    const float4 o4 = ws.ro[it.x], d4 = ws.rd[it.x];
                    Ray r; r.o = mk(o4.x, o4.y, o4.z); r.d = mk(d4.x, d4.y, d4.z);
                    float t, u, v;
                    if (tri_ray(mk(a.x, a.y, a.z), mk(b.x, b.y, b.z), mk(c.x, c.y, c.z), r, t, u, v) && t > 0.0f)
                        atomicMin(&ws.best[it.x], ((unsigned long long)__float_as_uint(t) << 32) | ti);
                }
            }
            __syncwarp();
            // ---- phase C: owners fold the cell's winner in and take the cell's exit crossing ----
            if (cnt != 0u) {
                const unsigned long long k = ws.best[lane];
                const float t = __uint_as_float((uint32_t)(k >> 32));
                if (k != ~0ull && best_t > t) { best_t = t; best_tri = (uint32_t)k; }
                float unused_t;
                if (trav_advance<CSM, TWO>(sc, coarse, s, best_t, unused_t) == kWalkDone) done = true;
            }
            __syncwarp();
        }
        if (done) {
            __stcs(hits + qi, make_uint2(__float_as_uint(best_t), best_tri));
            alive = false;
        }
    }
    if (STATS) flush_counters(cn, counters);
}

// ---------------------------------------------------------------------------------------------------------
// k_trace_bvh: Scene.traceRay (src/stage3.zig:152-186) over a ray queue with the 8-wide BVH of rtc_bvh.cuh in
// place of the grid (RTC_ACCEL_BVH). Persistent warps with the same per-lane refill as k_trace. One round:
//   A  every live lane pops the nearest pending child of its own ray and tests that node's eight quantised boxes
//      (one 80-byte node at a 96-byte stride = three sector-aligned 256-bit loads; the lanes run this part converged);
//   B  the triangles of the leaves that were hit — a few per lane, none for most — are POOLED over the warp and
//      dealt out one (ray, triangle) pair per lane per pass (48-byte records at a 64-byte stride, two 256-bit loads, the reference's
//      Triangle.rayIntersection). Tested per lane they ran at 3.5 of 32 lanes (profiles/r2b_*). Winner per ray =
//      atomicMin on (t bits << 32 | source id) in shared memory: smallest t > 0, smaller source id on exact ties;
//   C  owners fold their winner in (strictly nearer, or equal and smaller source id) and pop their stack.
// The hit leaves as {t, u, v, source id} (16 B, in the a3 plane of the input queue, which BVH rays do not use), so
// the shading kernel neither fetches the triangle again nor repeats its test.
// The traversal stack is per lane, in local memory (L1-resident: a few entries deep in practice).
#ifndef RTC_BVH_MINB
#define RTC_BVH_MINB 4
#endif
#ifndef RTC_BVH_REFILL
#define RTC_BVH_REFILL 4
#endif
static constexpr uint32_t kBvhRefillMin = RTC_BVH_REFILL;     // refill when at least this many lanes of the warp are idle
static constexpr uint32_t kBvhPool = 32u * 24u;               // a lane's triangle group holds at most 24 triangles
#ifndef RTC_BVH_TOP
#define RTC_BVH_TOP 0
#endif
// Nodes staged in shared memory (breadth-first order = the top levels). EXPERIMENT, off: 9 nodes +0.4 %, 73 or 256 nodes
// -0.9 % on the atrium, -1.9 % on the 5 M-triangle scene (profiles/r2l_ab_*): the per-node branch between the two
// fetch paths costs what the saved L1 lookups give.
static constexpr uint32_t kBvhTop = RTC_BVH_TOP;
#ifndef RTC_BVH_DIRECT
#define RTC_BVH_DIRECT 3
#endif
static constexpr uint32_t kBvhDirect = RTC_BVH_DIRECT;        // rounds whose busiest lane has at most this many triangles skip the pool

// EXPERIMENT, off. RTC_BVH_SSTACK = K > 0: the first K stack entries of every lane in shared memory (lane-contiguous,
// conflict-free), deeper ones in local memory — the pops are long-scoreboard stalls (4.7 % of the samples: their L1 lines
// are evicted between push and pop), but K = 2 / 4 / 8 measured -1.1 % / -0.25 % / -0.55 % (profiles/r2z_ab_stack.txt):
// the branch per access and the L1 capacity given to shared memory cost what the faster pops give.
#ifndef RTC_BVH_SSTACK
#define RTC_BVH_SSTACK 0
#endif
// Measured and dropped: asking L1 (prefetch.global.L1) for the node a lane visits next as soon as it is known (after the node test /
// after a pop), or for the first triangle of a group ahead of the pooled tests: -23 % / -55 % / -2 % (profiles/r2z_ab_node_prefetch.txt).
// Every lane's prefetch is one more L1 tag lookup per sector, and the L1 tag rate is the kernel's second roof (l1tex 0.73).
static constexpr uint32_t kBvhSStack = RTC_BVH_SSTACK;
struct BvhHybridStack {
    uint2* sh;                 // entry k of this lane at sh[k * kTraceThreads]
    uint2* lo;                 // local memory, entries kBvhSStack..
    __device__ __forceinline__ void put(uint32_t k, uint2 v) { if (k < kBvhSStack) sh[k * kTraceThreads] = v; else lo[k - kBvhSStack] = v; }
    __device__ __forceinline__ uint2 get(uint32_t k) const { return k < kBvhSStack ? sh[k * kTraceThreads] : lo[k - kBvhSStack]; }
};

struct BvhShared {
    float4 ro[32];                      // per lane: ray origin
    float4 rd[32];                      // per lane: ray direction
    unsigned long long best[32];        // per lane: min over this round's (t bits << 32 | source id)
    float2 wuv[32];                     // per lane: u, v of the test holding best (this round)
    float2 buv[32];                     // per lane: u, v of the ray's nearest hit so far
    uint32_t base[32];                  // per lane: tri_base of its triangle group
    unsigned short pool[kBvhPool];      // (owner lane << 8 | triangle offset) per pooled pair
};

template <bool STATS>
__global__ void __launch_bounds__(kTraceThreads, RTC_BVH_MINB)
k_trace_bvh(const BvhView bv, const RayQueue q, const uint32_t* __restrict__ count_ptr, uint32_t* __restrict__ head_ptr,
            float4* __restrict__ hits, unsigned long long* __restrict__ counters)
{
    __shared__ BvhShared s_ws[kTraceWarps];
    __shared__ float4 s_top[kBvhTop > 0u ? 5u * kBvhTop : 1u];
    BvhShared& ws = s_ws[threadIdx.x >> 5];
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t n = __ldg(count_ptr);
    // the top of the tree (breadth-first node order: the first nodes are the first levels), once per CTA
    const uint32_t ntop = bv.num_nodes < kBvhTop ? bv.num_nodes : kBvhTop;
    for (uint32_t k = threadIdx.x; k < 5u * ntop; k += kTraceThreads) s_top[k] = __ldg(bv.nodes + kBvhNodeQuads * (k / 5u) + (k % 5u));
    __syncthreads();
    Counters cn = { 0, 0, 0, 0, 0 };
    uint32_t next = 0, end = 0;
    bool exhausted = false, alive = false;
    uint32_t qi = 0;
    BvhRay br; BvhTrav s;
    br.ox = br.oy = br.oz = 0.0f; br.ix = br.iy = br.iz = 1.0f; br.octinv4 = 0u;
    bvh_trav_begin(s);
    uint2 stack_mem[kBvhStack];
#if RTC_BVH_SSTACK
    __shared__ uint2 s_stack[kBvhSStack * kTraceThreads];
    BvhHybridStack stack{ s_stack + threadIdx.x, stack_mem };
#else
    BvhArrayStack stack{ stack_mem, 1u };
#endif
    for (;;) {
        // ---- refill: idle lanes take the next rays of the queue ----
        const uint32_t idle = __ballot_sync(kFull, !alive);
        if (idle != 0u && !exhausted && ((uint32_t)__popc(idle) >= kBvhRefillMin || idle == kFull)) {
            if (next == end) {
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd(head_ptr, kChunk);
                base = __shfl_sync(kFull, base, 0);
                if (base >= n) { exhausted = true; next = end = 0; }
                else {
                    next = base; end = base + kChunk < n ? base + kChunk : n;
                    queue_prefetch_chunk<3u>(q, next, end, lane);
                }
            }
            const uint32_t avail = end - next, want = __popc(idle);
            const uint32_t take = want < avail ? want : avail;
            if (!alive && (uint32_t)__popc(idle & lt_mask) < take) {
                qi = next + __popc(idle & lt_mask);
                const float4 a0 = __ldcs(q.a0 + qi), a1 = __ldcs(q.a1 + qi), a2 = __ldcs(q.a2 + qi);
                ws.ro[lane] = a0; ws.rd[lane] = a1;
                br.ox = a0.x; br.oy = a0.y; br.oz = a0.z; br.ix = a2.x; br.iy = a2.y; br.iz = a2.z;
                br.octinv4 = __float_as_uint(a2.w);
                bvh_trav_begin(s);
                alive = true;
            }
            next += take;
        }
        if (__ballot_sync(kFull, alive) == 0u) {
            if (exhausted) break;
            continue;
        }
        // ---- A: one node per live lane ----
        uint2 tg = make_uint2(0u, 0u);
        if (alive) tg = bvh_step_nodes<STATS>(bv, br, s, stack, cn.cells, s_top, ntop);
        // ---- B: pooled triangle tests ----
        const uint32_t cnt = __popc(tg.y);
        if (STATS) cn.tris += cnt;
        const uint32_t busiest = __reduce_max_sync(kFull, cnt);
        if (busiest != 0u && busiest <= kBvhDirect) {
            // few triangles per lane (the incoherent case): pooling would cost more (scan, pool fill, atomics, two
            // warp barriers per pass) than the passes it saves — every lane tests its own, `busiest` passes
            if (cnt != 0u) {
                const float4 o4 = ws.ro[lane], d4 = ws.rd[lane];
                Ray r; r.o = mk(o4.x, o4.y, o4.z); r.d = mk(d4.x, d4.y, d4.z);
                uint32_t m = tg.y;
                do {
                    const uint32_t b = 31u - (uint32_t)__clz((int)m);
                    m &= ~(1u << b);
                    const float4* p = bv.tris + kBvhTriQuads * (size_t)(tg.x + b);
                    const Quad2 t01 = bvh_ld256(p), t23 = bvh_ld256(p + 2);
                    const float4 a = t01.lo, e1 = t01.hi, e2 = t23.lo;
                    float t, u, v;
                    if (tri_ray(mk(a.x, a.y, a.z), mk(e1.x, e1.y, e1.z), mk(e2.x, e2.y, e2.z), r, t, u, v) && t > 0.0f) {
                        const uint32_t src = __float_as_uint(a.w);
                        if (s.best_t > t || (s.best_t == t && src < s.best_tri)) { s.best_t = t; s.best_tri = src; ws.buv[lane] = make_float2(u, v); }
                    }
                } while (m != 0u);
            }
        } else if (busiest != 0u) {
            uint32_t incl = cnt;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const uint32_t v = __shfl_up_sync(kFull, incl, d); if (lane >= (uint32_t)d) incl += v; }
            const uint32_t total = __shfl_sync(kFull, incl, 31);
            if (cnt != 0u) {
                uint32_t o = incl - cnt, m = tg.y;
                ws.base[lane] = tg.x;
                ws.best[lane] = ~0ull;
                do {
                    const uint32_t b = 31u - (uint32_t)__clz((int)m);
                    m &= ~(1u << b);
                    ws.pool[o++] = (unsigned short)((lane << 8) | b);
                } while (m != 0u);
            }
            __syncwarp();
            for (uint32_t base = 0; base < total; base += 32u) {
                const uint32_t j = base + lane;
                unsigned long long key = ~0ull;
                uint32_t owner = 0u;
                float u = 0.0f, v = 0.0f;
                if (j < total) {
                    const uint32_t e = ws.pool[j];
                    owner = e >> 8;
                    const uint32_t ti = ws.base[owner] + (e & 0xFFu);
                    const float4* p = bv.tris + kBvhTriQuads * (size_t)ti;
                    const Quad2 t01 = bvh_ld256(p), t23 = bvh_ld256(p + 2);
                    const float4 a = t01.lo, e1 = t01.hi, e2 = t23.lo;
                    const float4 o4 = ws.ro[owner], d4 = ws.rd[owner];
                    Ray r; r.o = mk(o4.x, o4.y, o4.z); r.d = mk(d4.x, d4.y, d4.z);
                    float t;
                    if (tri_ray(mk(a.x, a.y, a.z), mk(e1.x, e1.y, e1.z), mk(e2.x, e2.y, e2.z), r, t, u, v) && t > 0.0f) {
                        key = ((unsigned long long)__float_as_uint(t) << 32) | __float_as_uint(a.w);
                        atomicMin(&ws.best[owner], key);
                    }
                }
                __syncwarp();
                if (key != ~0ull && ws.best[owner] == key) ws.wuv[owner] = make_float2(u, v);     // source ids are unique: one winner
                __syncwarp();
            }
            // ---- C (first half): owners fold the round's winner in ----
            if (cnt != 0u) {
                const unsigned long long k = ws.best[lane];
                if (k != ~0ull) {
                    const float t = __uint_as_float((uint32_t)(k >> 32));
                    // s.best_tri holds the SOURCE id here (the shading kernel needs nothing else of the triangle)
                    if (s.best_t > t || (s.best_t == t && (uint32_t)k < s.best_tri)) {
                        s.best_t = t; s.best_tri = (uint32_t)k; ws.buv[lane] = ws.wuv[lane];
                    }
                }
            }
            __syncwarp();
        }
        // ---- C (second half): next group, or the ray is done ----
        if (alive && bvh_step_pop(s, stack)) {
            // the whole hit travels to the shading kernel: t, u, v and the source triangle (16 B; no triangle fetch there)
            const float2 uv = s.best_tri != kMiss ? ws.buv[lane] : make_float2(0.0f, 0.0f);
            __stcs(hits + qi, make_float4(s.best_t, uv.x, uv.y, __uint_as_float(s.best_tri)));
            alive = false;
        }
    }
    if (STATS) flush_counters(cn, counters);
}

// ---------------------------------------------------------------------------------------------------------
// Triangle.Data (src/stage3.zig:42-71) as stored on the device: 4 float4 (64 B) per SOURCE triangle.
struct TriData { v3 n0, n1, n2; float uv0[2], uv1[2], uv2[2]; uint32_t material; };
__device__ __forceinline__ TriData load_tri_data(const DevScene& sc, uint32_t src)
{
    const float4* p = sc.tri_data + 4u * (size_t)src;
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

// k_shade: the body of traceRayRecursive (src/stage3.zig:193-219) for every ray of bounce `bounce`. The hit's
// u, v are recomputed from the winning triangle with the arithmetic of the test (same inputs, same bits).
// RNG block 1 + bounce of the path: draw 0 = the alpha test (always drawn, stage3.zig:207), draws 1, 2 = the
// scatter direction. A scattering level pushes (emissive, albedo) for k_accum; an alpha pass-through level
// (stage3.zig:207-213) consumes a depth level and neither attenuates nor pushes.
#ifndef RTC_SHADE_MINB
#define RTC_SHADE_MINB 4
#endif
static constexpr uint32_t kMatSmem = 32;             // material tables up to this size are staged in shared memory
// Rounds ahead for which a thread asks L2 for its queue entry and hit record (0 = off): the chain of dependent loads of one
// ray starts with three DRAM reads (the queue was streamed out by the previous kernels) and nothing overlaps them but the
// other warps of the SM. Distance 1 / 2 / 4: +1.0 % / +1.1 % / +0.8 % on cfg3 (profiles/r2z_ab_shade_prefetch.txt).
#ifndef RTC_SHADE_PREFETCH
#define RTC_SHADE_PREFETCH 2
#endif
// Measured and dropped: reading the NEXT round's hit id one round ahead and asking its triangle record into L1 at the end of
// the round (the second link of the chain): -0.6 % with the BVH, -1.2 % with the grid (profiles/r2z_ab_shade_tri.txt): the extra
// live register spills and the extra load costs more than the L2 latency it hides.

template <bool PRIMARY, bool BVH>
__global__ void __launch_bounds__(kThreads, RTC_SHADE_MINB)
k_shade(const __grid_constant__ DevScene sc, const __grid_constant__ DevCamera cam, const __grid_constant__ WaveParams wp,
        const __grid_constant__ WaveScratch ws, uint32_t bounce, PrimaryOut out, unsigned long long* __restrict__ counters)
{
    Counters cn = { 0, 0, 0, 0, 0 };
    const RayQueue qin = ws.q[bounce & 1u], qout = ws.q[(bounce + 1u) & 1u];
    const uint32_t n = __ldg(ws.counts + bounce);
    const uint32_t stride = gridDim.x * kThreads;
    const uint32_t rounds = (n + stride - 1u) / stride;
    const float eps = 1.1920928955078125e-07f;           // std.math.floatEps(f32)
    // the material table (three texture headers per material) sits in shared memory: one level less in the chain of
    // dependent global loads hit -> triangle data -> material -> texels
    __shared__ DevMaterial s_mats[kMatSmem];
    const bool msm = sc.num_materials <= kMatSmem;
    if (msm) {
        const uint32_t words = sc.num_materials * (uint32_t)(sizeof(DevMaterial) / 4u);
        for (uint32_t k = threadIdx.x; k < words; k += kThreads)
            reinterpret_cast<uint32_t*>(s_mats)[k] = __ldg(reinterpret_cast<const uint32_t*>(sc.materials) + k);
        __syncthreads();
    }
    for (uint32_t it = 0; it < rounds; ++it) {
        const uint32_t i = it * stride + blockIdx.x * kThreads + threadIdx.x;
        bool push = false;
        QueuedRay e;
#if RTC_SHADE_PREFETCH
        {
            const uint32_t ip = i + RTC_SHADE_PREFETCH * stride;
            if (ip < n) {
                asm volatile("prefetch.global.L2 [%0];" :: "l"(qin.a0 + ip));
                asm volatile("prefetch.global.L2 [%0];" :: "l"(qin.a1 + ip));
                if (BVH) asm volatile("prefetch.global.L2 [%0];" :: "l"(qin.a3 + ip));
                else asm volatile("prefetch.global.L2 [%0];" :: "l"(ws.hits + ip));
            }
        }
#endif
        if (i < n) {
            const float4 a0 = __ldcs(qin.a0 + i), a1 = __ldcs(qin.a1 + i);
            // grid: {t bits, baked triangle}; BVH: {t, u, v, source id} in the a3 plane of the input queue
            uint2 h;
            float t = 0.0f, u = 0.0f, v = 0.0f;
            uint32_t src = kMiss;
            if (BVH) {
                const float4 h4 = __ldcs(qin.a3 + i);
                t = h4.x; u = h4.y; v = h4.z; src = __float_as_uint(h4.w);
                h = make_uint2(__float_as_uint(h4.x), src);
            } else {
                h = __ldcs(ws.hits + i);
            }
            const uint32_t slot = __float_as_uint(a0.w);
            uint32_t top = __float_as_uint(a1.w);
            Ray ray; ray.o = mk(a0.x, a0.y, a0.z); ray.d = mk(a1.x, a1.y, a1.z);
            const uint32_t sl = slot / wp.num_pixels, pl = slot - sl * wp.num_pixels;
            const uint32_t pixel = slot_pixel(wp, pl);
            v3 leaf = splat(0.0f);
            if (h.y == kMiss) {
                leaf = env_color(ray.d);
                if (PRIMARY) {
                    if (out.hit_id) out.hit_id[pixel] = kMiss;
                    if (out.hit_tuv) {
                        out.hit_tuv[3 * (size_t)pixel] = __uint_as_float(h.x);
                        out.hit_tuv[3 * (size_t)pixel + 1] = 0.0f;
                        out.hit_tuv[3 * (size_t)pixel + 2] = 0.0f;
                    }
                }
            } else {
                cn.hits += 1;
                if (!BVH) {
                    // the hit's u, v are recomputed from the winning triangle with the arithmetic of the test (same inputs, same bits)
                    const float4 ta = __ldg(sc.tri_v0 + h.y), tb = __ldg(sc.tri_e1 + h.y), tc = __ldg(sc.tri_e2 + h.y);
                    tri_ray(mk(ta.x, ta.y, ta.z), mk(tb.x, tb.y, tb.z), mk(tc.x, tc.y, tc.z), ray, t, u, v);
                    t = __uint_as_float(h.x);
                    src = __float_as_uint(ta.w);
                }
                const TriData td = load_tri_data(sc, src);
                const DevMaterial* m = (msm ? s_mats : sc.materials) + td.material;
                float tu, tv;
                interp_uv(td, u, v, tu, tv);
                if (PRIMARY) {
                    // depth-1 slice: traceRayRecursive(ray, 1) = emissive + albedo * zeroes
                    leaf = tex_sample3(m->emissive, tu, tv);
                    if (out.hit_id) out.hit_id[pixel] = src;
                    if (out.hit_tuv) {
                        out.hit_tuv[3 * (size_t)pixel] = t;
                        out.hit_tuv[3 * (size_t)pixel + 1] = u;
                        out.hit_tuv[3 * (size_t)pixel + 2] = v;
                    }
                } else {
                    const v3 albedo = tex_sample3(m->base, tu, tv);
                    const v3 emissive = tex_sample3(m->emissive, tu, tv);
                    const float transparency = tex_sample1(m->alpha, tu, tv);
                    const v3 normal = interp_normal(td, u, v);
                    float rnd[4];
                    rng_block(wp.seed_lo, wp.seed_hi, pixel, wp.s0 + sl, 1u + bounce, rnd);
                    ray.o = ray_at(ray, t + eps);
                    if (!(rnd[0] > transparency)) {          // scatter (stage3.zig:214-219)
                        ray.d = normalize(normal + unit_vector(rnd[1], rnd[2]));
                        const size_t k = (size_t)top * ws.cap + slot;
                        __stcs(ws.stack_e + k, make_float4(emissive.x, emissive.y, emissive.z, albedo.x));
                        __stcs(ws.stack_a + k, make_float2(albedo.y, albedo.z));
                        top += 1u;
                    }
                    if (bounce + 1u < wp.max_bounce) {       // depth - 1 > 0: the recursion goes on
                        cn.rays += 1;
                        e.slot = slot; e.top = top;
                        push = ray_setup<BVH>(sc, ray, e);
                        if (!push) leaf = env_color(ray.d);
                    }
                }
            }
            if (!push) __stcs(ws.leaf + slot, make_float4(leaf.x, leaf.y, leaf.z, __uint_as_float(top)));
        }
        queue_push<BVH>(qout, ws.counts + bounce + 1u, push, e);
    }
    flush_counters(cn, counters);
}

// ---------------------------------------------------------------------------------------------------------
// k_accum: one thread per pixel of the batch. For every sample in order: value = leaf; then
// value = emissive + albedo * value from the innermost scattering level outwards (the float evaluation
// order of `emissive.add(albedo.mul(traceRayRecursive(...)))`, stage3.zig:219); pixel = pixel.add(value)
// (stage3.zig:239-240). The running sum continues from accum[], so sample ranges rendered by successive
// batches or calls add up in the reference's order.
__global__ void __launch_bounds__(kThreads)
k_accum(const __grid_constant__ WaveParams wp, const __grid_constant__ WaveScratch ws, float4* __restrict__ accum)
{
    const uint32_t pl = blockIdx.x * kThreads + threadIdx.x;
    if (pl >= wp.num_pixels) return;
    const uint32_t pixel = slot_pixel(wp, pl);
    float4 a = accum[pixel];
    for (uint32_t sl = 0; sl < wp.ns; ++sl) {
        const uint32_t slot = sl * wp.num_pixels + pl;
        const float4 lf = __ldcs(ws.leaf + slot);
        v3 value = mk(lf.x, lf.y, lf.z);
        for (uint32_t l = __float_as_uint(lf.w); l > 0u; --l) {
            const size_t k = (size_t)(l - 1u) * ws.cap + slot;
            const float4 ea = __ldcs(ws.stack_e + k);
            const float2 a2 = __ldcs(ws.stack_a + k);
            value = mk(ea.x, ea.y, ea.z) + mk(ea.w, a2.x, a2.y) * value;
        }
        a.x = a.x + value.x; a.y = a.y + value.y; a.z = a.z + value.z;
    }
    a.w = a.w + (float)wp.ns;
    accum[pixel] = a;
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

// Sample sharding inside one process (rtc_params.num_gpus > 1, RTC_SHARD_SAMPLES): the reduce of the per-device
// accumulators and the resolve in ONE kernel over NVLink peer memory. Device g runs it on ITS slice of the pixels and
// reads that slice of every device's accumulator directly (P2P loads, 16 B per pixel and peer), adds them in device
// order — a fixed order, so the frame does not depend on timing — scales by 1/S and writes RGB8 (+ radiance): a
// reduce-scatter fused with stage3.zig:242, after which every device copies its own slice to the host frame.
__global__ void __launch_bounds__(kThreads)
k_combine_resolve(const PeerAccums pa, uint32_t first, uint32_t count, uint32_t total_samples,
                  uint8_t* __restrict__ rgb, float* __restrict__ radiance)
{
    const uint32_t k = blockIdx.x * kThreads + threadIdx.x;
    if (k >= count) return;
    const uint32_t i = first + k;
    float4 a = pa.p[0][i];
    for (uint32_t j = 1; j < pa.n; ++j) {
        const float4 b = __ldcs(pa.p[j] + i);
        a.x = a.x + b.x; a.y = a.y + b.y; a.z = a.z + b.z;
    }
    const float inv = 1.0f / (float)total_samples;
    const float r = a.x * inv, g = a.y * inv, b = a.z * inv;
    if (radiance) { radiance[3 * (size_t)i] = r; radiance[3 * (size_t)i + 1] = g; radiance[3 * (size_t)i + 2] = b; }
    if (rgb) {
        rgb[3 * (size_t)i] = (uint8_t)to_rgb8(r);
        rgb[3 * (size_t)i + 1] = (uint8_t)to_rgb8(g);
        rgb[3 * (size_t)i + 2] = (uint8_t)to_rgb8(b);
    }
}

// One process per GPU (rtc_render_sharded, RTC_SHARD_SAMPLES): the accumulators travel as rgb only — the sample count
// in .w is the same on every rank's frame and a quarter of the bytes. k_pack_rgb: float4 -> 3 floats; k_resolve3: the
// resolve from the summed 3-float rows.
__global__ void __launch_bounds__(kThreads)
k_pack_rgb(const float4* __restrict__ accum, uint32_t num_pixels, float* __restrict__ rgb3)
{
    const uint32_t i = blockIdx.x * kThreads + threadIdx.x;
    if (i >= num_pixels) return;
    const float4 a = accum[i];
    rgb3[3 * (size_t)i] = a.x; rgb3[3 * (size_t)i + 1] = a.y; rgb3[3 * (size_t)i + 2] = a.z;
}
__global__ void __launch_bounds__(kThreads)
k_resolve3(const float* __restrict__ sum3, uint32_t num_pixels, uint32_t total_samples, uint8_t* __restrict__ rgb)
{
    const uint32_t i = blockIdx.x * kThreads + threadIdx.x;
    if (i >= num_pixels) return;
    const float inv = 1.0f / (float)total_samples;
    rgb[3 * (size_t)i] = (uint8_t)to_rgb8(sum3[3 * (size_t)i] * inv);
    rgb[3 * (size_t)i + 1] = (uint8_t)to_rgb8(sum3[3 * (size_t)i + 1] * inv);
    rgb[3 * (size_t)i + 2] = (uint8_t)to_rgb8(sum3[3 * (size_t)i + 2] * inv);
}

// ---------------------------------------------------------------------------------------------------------
template <typename K>
static int blocks_per_sm(K kernel, int threads, size_t smem)
{
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, threads, smem) != cudaSuccess || n < 1) n = 1;
    return n;
}

template <bool STATS, bool CSM>
static void run_trace(const DevScene& sc, const WaveScratch& ws, uint32_t bounce, unsigned long long* counters, const LaunchCfg& cfg)
{
    auto kern = k_trace<STATS, CSM>;
    const size_t smem = CSM ? sc.coarse_words * sizeof(uint32_t) : 0;
    // occupancy and the shared-memory carveout are per DEVICE state (cudaFuncSetAttribute applies to the current
    // device): cached in the scene's LaunchCache, not in process-wide statics
    LaunchCache& lc = *cfg.cache;
    int* per_sm = lc.trace_per_sm;
    if (lc.trace_smem[STATS] != smem) {
        per_sm[STATS] = blocks_per_sm(kern, kTraceThreads, smem); lc.trace_smem[STATS] = smem;
        // Ask for no more shared memory than the resident CTAs use (static + coarse bitmap + 1 KiB the system reserves
        // per CTA), so that the rest of the 256 KiB stays L1: the triangle and cell gathers are what the kernel waits
        // on. +1.2 % on the 5 M-triangle scene, +0.9 % on the atrium; a carveout below the need costs occupancy (-14 %),
        // hence the margin. RTC_CARVEOUT=<percent> overrides.
        cudaFuncAttributes fa;
        cudaDeviceProp prop;
        int dev = 0;
        int pct = -1;
        if (cudaGetDevice(&dev) == cudaSuccess && cudaGetDeviceProperties(&prop, dev) == cudaSuccess &&
            cudaFuncGetAttributes(&fa, kern) == cudaSuccess && prop.sharedMemPerMultiprocessor > 0) {
            const double need = (double)per_sm[STATS] * ((double)fa.sharedSizeBytes + (double)smem + 1024.0);
            pct = (int)(100.0 * need / (double)prop.sharedMemPerMultiprocessor) + 3;
            if (pct > 100) pct = 100;
        }
        if (const char* env = getenv("RTC_CARVEOUT")) pct = atoi(env);
        if (pct >= 0) cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
        cudaGetLastError();
    }
    TraceTimer* tm = cfg.timer;
    const bool timed = tm && tm->used < tm->capacity;
    if (tm && !timed) tm->dropped += 1;
    if (timed) cudaEventRecord(tm->ev[2 * tm->used], cfg.stream);
    kern<<<cfg.sm_count * per_sm[STATS], kTraceThreads, smem, cfg.stream>>>(sc, ws.q[bounce & 1u], ws.counts + bounce,
                                                                             ws.counts + kMaxBounce + 1 + bounce, ws.hits, counters);
    if (timed) { cudaEventRecord(tm->ev[2 * tm->used + 1], cfg.stream); tm->used += 1; }
}

template <bool STATS>
static void run_trace_bvh(const BvhView& bv, const WaveScratch& ws, uint32_t bounce, unsigned long long* counters, const LaunchCfg& cfg)
{
    auto kern = k_trace_bvh<STATS>;
    int& per_sm = cfg.cache->bvh_per_sm[STATS];
    if (per_sm == 0) per_sm = blocks_per_sm(kern, kTraceThreads, 0);
    TraceTimer* tm = cfg.timer;
    const bool timed = tm && tm->used < tm->capacity;
    if (tm && !timed) tm->dropped += 1;
    if (timed) cudaEventRecord(tm->ev[2 * tm->used], cfg.stream);
    kern<<<cfg.sm_count * per_sm, kTraceThreads, 0, cfg.stream>>>(bv, ws.q[bounce & 1u], ws.counts + bounce,
                                                                   ws.counts + kMaxBounce + 1 + bounce, ws.q[bounce & 1u].a3, counters);
    if (timed) { cudaEventRecord(tm->ev[2 * tm->used + 1], cfg.stream); tm->used += 1; }
}

template <bool PRIMARY, bool BVH>
static void run_gen(uint32_t g, const DevScene& sc, const DevCamera& cam, const WaveParams& wp, const WaveScratch& ws,
                    PrimaryOut out, unsigned long long* counters, cudaStream_t st)
{
    k_gen<PRIMARY, BVH><<<g, kThreads, 0, st>>>(sc, cam, wp, ws, out, counters);
}
template <bool PRIMARY, bool BVH>
static void run_shade(uint32_t g, const DevScene& sc, const DevCamera& cam, const WaveParams& wp, const WaveScratch& ws,
                      uint32_t b, PrimaryOut out, unsigned long long* counters, cudaStream_t st)
{
    k_shade<PRIMARY, BVH><<<g, kThreads, 0, st>>>(sc, cam, wp, ws, b, out, counters);
}

cudaError_t launch_render(const DevScene& sc, const DevCamera& cam, const DevParams& prm, bool primary, float4* accum,
                          PrimaryOut out, unsigned long long* counters, const WaveScratch& ws, const LaunchCfg& cfg,
                          uint32_t* launches, const BvhView* bvh)
{
    if (prm.num_pixels == 0 || (!primary && prm.num_samples == 0)) return cudaSuccess;
    const uint32_t S = primary ? 1u : prm.num_samples;
    const uint32_t B = primary ? 1u : prm.max_bounce;
    if (B > kMaxBounce || B > ws.max_levels || ws.cap == 0) return cudaErrorInvalidValue;
    const bool csm = sc.coarse_words <= kCoarseSmemMaxWords;
    const int wide = cfg.sm_count * 8;                // grid-stride kernels: 8 CTAs of 256 threads per SM
    // batches: pixel chunks of at most cap pixels x as many samples as fit into cap slots
    const uint32_t px_chunk = prm.num_pixels < ws.cap ? prm.num_pixels : ws.cap;
    for (uint32_t p0 = 0; p0 < prm.num_pixels; p0 += px_chunk) {
        const uint32_t np = prm.num_pixels - p0 < px_chunk ? prm.num_pixels - p0 : px_chunk;
        const uint32_t s_chunk = ws.cap / np < S ? ws.cap / np : S;
        for (uint32_t s0 = 0; s0 < S; s0 += s_chunk) {
            WaveParams wp;
            wp.first_pixel = prm.first_pixel; wp.num_pixels = np;
            wp.pl0 = p0; wp.strip_px = prm.strip_px; wp.stride_px = prm.stride_px;
            wp.s0 = prm.sample_offset + s0; wp.ns = S - s0 < s_chunk ? S - s0 : s_chunk;
            wp.max_bounce = B; wp.seed_lo = prm.seed_lo; wp.seed_hi = prm.seed_hi;
            cudaError_t e = cudaMemsetAsync(ws.counts, 0, 2 * (kMaxBounce + 1) * sizeof(uint32_t), cfg.stream);
            if (e != cudaSuccess) return e;
            const uint32_t nslots = wp.num_pixels * wp.ns;
            const uint32_t need = (nslots + kThreads - 1) / kThreads;
            const uint32_t g = need < (uint32_t)wide ? need : (uint32_t)wide;
            if (bvh) { if (primary) run_gen<true, true>(g, sc, cam, wp, ws, out, counters, cfg.stream); else run_gen<false, true>(g, sc, cam, wp, ws, out, counters, cfg.stream); }
            else     { if (primary) run_gen<true, false>(g, sc, cam, wp, ws, out, counters, cfg.stream); else run_gen<false, false>(g, sc, cam, wp, ws, out, counters, cfg.stream); }
            *launches += 1;
            for (uint32_t b = 0; b < B; ++b) {
                if (bvh) { if (cfg.stats) run_trace_bvh<true>(*bvh, ws, b, counters, cfg); else run_trace_bvh<false>(*bvh, ws, b, counters, cfg); }
                else if (csm) { if (cfg.stats) run_trace<true, true>(sc, ws, b, counters, cfg); else run_trace<false, true>(sc, ws, b, counters, cfg); }
                else     { if (cfg.stats) run_trace<true, false>(sc, ws, b, counters, cfg); else run_trace<false, false>(sc, ws, b, counters, cfg); }
                if (bvh) { if (primary) run_shade<true, true>(g, sc, cam, wp, ws, b, out, counters, cfg.stream); else run_shade<false, true>(g, sc, cam, wp, ws, b, out, counters, cfg.stream); }
                else     { if (primary) run_shade<true, false>(g, sc, cam, wp, ws, b, out, counters, cfg.stream); else run_shade<false, false>(g, sc, cam, wp, ws, b, out, counters, cfg.stream); }
                *launches += 2;
            }
            k_accum<<<(wp.num_pixels + kThreads - 1) / kThreads, kThreads, 0, cfg.stream>>>(wp, ws, accum);
            *launches += 1;
            e = cudaGetLastError();
            if (e != cudaSuccess) return e;
        }
    }
    return cudaSuccess;
}

cudaError_t launch_resolve(const float4* accum, uint32_t num_pixels, uint32_t total_samples,
                           uint8_t* rgb, float* radiance, cudaStream_t stream)
{
    if (num_pixels == 0) return cudaSuccess;
    k_resolve<<<(num_pixels + kThreads - 1) / kThreads, kThreads, 0, stream>>>(accum, num_pixels, total_samples, rgb, radiance);
    return cudaGetLastError();
}

cudaError_t launch_combine_resolve(const PeerAccums& pa, uint32_t first, uint32_t count, uint32_t total_samples,
                                   uint8_t* rgb, float* radiance, cudaStream_t stream)
{
    if (count == 0) return cudaSuccess;
    k_combine_resolve<<<(count + kThreads - 1) / kThreads, kThreads, 0, stream>>>(pa, first, count, total_samples, rgb, radiance);
    return cudaGetLastError();
}
cudaError_t launch_pack_rgb(const float4* accum, uint32_t num_pixels, float* rgb3, cudaStream_t stream)
{
    if (num_pixels == 0) return cudaSuccess;
    k_pack_rgb<<<(num_pixels + kThreads - 1) / kThreads, kThreads, 0, stream>>>(accum, num_pixels, rgb3);
    return cudaGetLastError();
}
cudaError_t launch_resolve3(const float* sum3, uint32_t num_pixels, uint32_t total_samples, uint8_t* rgb, cudaStream_t stream)
{
    if (num_pixels == 0) return cudaSuccess;
    k_resolve3<<<(num_pixels + kThreads - 1) / kThreads, kThreads, 0, stream>>>(sum3, num_pixels, total_samples, rgb);
    return cudaGetLastError();
}

}  // namespace RTC_NS
}  // namespace rtc
