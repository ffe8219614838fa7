// rtc_build.cu — Geometry.build + bakeInto (src/stage2.zig:44-164) on the device, from a triangle soup that
// was uploaded once. SURVEY §8(f) #1: the step immediately before the render path, which the reference runs
// serially on one host thread.
//
//   k_bbox      initGrid (stage2.zig:44-57): min/max over all vertices            -> Grid.init (linalg.zig:412-418)
//   k_bin       initCells / initIndices (stage2.zig:59-81, 104-125): per triangle, the cells of its bounding
//               range that pass intersectsTriangleAabb (linalg.zig:500-563). COUNT adds to the cell's counter,
//               EMIT appends the key (cell << 32 | triangle). Triangles whose range exceeds kInlineCells
//               are deferred to k_bin_big, one CTA per triangle.
//   scan        exclusive sum of the counters = Cell.begin (stage2.zig:82-96)      (cub::DeviceScan)
//   sort        the keys, ascending: cell-major, source-triangle order inside the cell — the order the
//               reference's serial scatter produces and Scene.traceRay's tie-break depends on (cub::DeviceRadixSort)
//   k_cells     Cell{begin,end} in the reference's x-fastest order + the empty/min/max statistics (stage2.zig:76-100)
//   k_layout    the same cells by packed index + the two occupancy levels (rtc_device.cuh)
//   k_bake_soa  bakeInto (stage2.zig:137-164) + Triangle.init (linalg.zig:688-694) into the SoA float4 arrays
//   k_pack_data Triangle.Data per SOURCE triangle in the 4-float4 (64-byte, sector-aligned) layout k_shade reads
//   k_bake_aos  the same bake in the C ABI's AoS layout, for rtc_geometry_build_device (parity with the host build)
//
// Compiled with -fmad=false, IEEE div/sqrt: the overlap test decides which cells hold a triangle and must
// evaluate exactly like the reference's fp32 code. CUB (scan, radix sort) is library plumbing of the build
// step, not part of the render path.
#include "rtc_gltf_plan.h"
#include "rtc_internal.h"

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstring>
#include <memory>
#include <new>

namespace rtc {
namespace build {

static constexpr int kThreads = 256;
static constexpr uint32_t kInlineCells = 64;       // ranges up to this many cells are tested by the triangle's own thread

struct GridP {
    float bmin[3], cell[3];
    uint32_t res[3];
};

// ---- initGrid: order-independent min/max through the monotone float -> int map (-0 < +0) ----
__device__ __forceinline__ int f2ord(float f) { const int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7FFFFFFF; }
__host__ __device__ __forceinline__ float ord2f_bits(int i, bool)
{
    const int b = i >= 0 ? i : i ^ 0x7FFFFFFF;
    float f;
    memcpy(&f, &b, sizeof f);
    return f;
}

__global__ void __launch_bounds__(kThreads) k_bbox(const float* __restrict__ pos, uint64_t num_vertices, int* __restrict__ out /* min xyz, max xyz */)
{
    int lo[3] = { 0x7F800000, 0x7F800000, 0x7F800000 };                 // +inf
    int hi[3] = { (int)0x807FFFFF, (int)0x807FFFFF, (int)0x807FFFFF };  // ord(-inf)
    for (uint64_t v = (uint64_t)blockIdx.x * kThreads + threadIdx.x; v < num_vertices; v += (uint64_t)gridDim.x * kThreads) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float p = __ldg(pos + 3 * v + a);
            if (p != p) continue;                                       // @min/@max skip NaN (fminf/fmaxf)
            const int o = f2ord(p);
            lo[a] = min(lo[a], o); hi[a] = max(hi[a], o);
        }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        lo[a] = __reduce_min_sync(0xFFFFFFFFu, lo[a]);
        hi[a] = __reduce_max_sync(0xFFFFFFFFu, hi[a]);
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) { atomicMin(out + a, lo[a]); atomicMax(out + 3 + a, hi[a]); }
    }
}

// ---- intersectsTriangleAabbSat / intersectsTriangleAabb (src/linalg.zig:500-563), expression for expression
// like the host build (rtc_geometry.cpp) and the oracle ----
__device__ __forceinline__ bool sat(v3 a, v3 b, v3 c, v3 e, v3 axis)
{
    const float p0 = dot(a, axis), p1 = dot(b, axis), p2 = dot(c, axis);
    const float r = e.x * fabsf(dot(mk(1.0f, 0.0f, 0.0f), axis)) + e.y * fabsf(dot(mk(0.0f, 1.0f, 0.0f), axis)) +
                    e.z * fabsf(dot(mk(0.0f, 0.0f, 1.0f), axis));
    const float hi = fmaxf(p0, fmaxf(p1, p2)), lo = fminf(p0, fminf(p1, p2));
    return !(fmaxf(-hi, lo) > r);
}

__device__ __forceinline__ bool tri_box(v3 t0, v3 t1, v3 t2, v3 bmin, v3 bmax)
{
    const v3 center = scale(bmax + bmin, 0.5f), e = scale(bmax - bmin, 0.5f);
    const v3 a = t0 - center, b = t1 - center, c = t2 - center;
    struct { v3 ab, bc, ca, n; } x;
    x.ab = normalize(b - a); x.bc = normalize(c - b); x.ca = normalize(a - c);
    x.n = cross(x.ab, x.bc);
    if (!sat(a, b, c, e, mk(0.0f, -x.ab.z, x.ab.y))) return false;
    if (!sat(a, b, c, e, mk(0.0f, -x.bc.z, x.bc.y))) return false;
    if (!sat(a, b, c, e, mk(0.0f, -x.ca.z, x.ca.y))) return false;
    if (!sat(a, b, c, e, mk(x.ab.z, 0.0f, -x.ab.x))) return false;
    if (!sat(a, b, c, e, mk(x.bc.z, 0.0f, -x.bc.x))) return false;
    if (!sat(a, b, c, e, mk(x.ca.z, 0.0f, -x.ca.x))) return false;
    if (!sat(a, b, c, e, mk(-x.ab.y, x.ab.x, 0.0f))) return false;
    if (!sat(a, b, c, e, mk(-x.bc.y, x.bc.x, 0.0f))) return false;
    if (!sat(a, b, c, e, mk(-x.ca.y, x.ca.x, 0.0f))) return false;
    if (!sat(a, b, c, e, mk(1.0f, 0.0f, 0.0f))) return false;
    if (!sat(a, b, c, e, mk(0.0f, 1.0f, 0.0f))) return false;
    if (!sat(a, b, c, e, mk(0.0f, 0.0f, 1.0f))) return false;
    return sat(a, b, c, e, x.n);
}

struct Tri { v3 p0, p1, p2; };
__device__ __forceinline__ Tri load_tri(const float* __restrict__ pos, uint32_t t)
{
    const float* p = pos + 9 * (size_t)t;
    Tri r;
    r.p0 = mk(__ldg(p), __ldg(p + 1), __ldg(p + 2));
    r.p1 = mk(__ldg(p + 3), __ldg(p + 4), __ldg(p + 5));
    r.p2 = mk(__ldg(p + 6), __ldg(p + 7), __ldg(p + 8));
    return r;
}
// Grid.getCellIdx (src/linalg.zig:424-427)
__device__ __forceinline__ void cell_idx(const GridP& g, v3 p, uint32_t out[3])
{
    out[0] = min(f2u((p.x - g.bmin[0]) / g.cell[0]), g.res[0] - 1u);
    out[1] = min(f2u((p.y - g.bmin[1]) / g.cell[1]), g.res[1] - 1u);
    out[2] = min(f2u((p.z - g.bmin[2]) / g.cell[2]), g.res[2] - 1u);
}
__device__ __forceinline__ void tri_range(const GridP& g, const Tri& t, uint32_t lo[3], uint32_t hi[3])
{
    const v3 mn = mk(fminf(t.p0.x, fminf(t.p1.x, t.p2.x)), fminf(t.p0.y, fminf(t.p1.y, t.p2.y)), fminf(t.p0.z, fminf(t.p1.z, t.p2.z)));
    const v3 mx = mk(fmaxf(t.p0.x, fmaxf(t.p1.x, t.p2.x)), fmaxf(t.p0.y, fmaxf(t.p1.y, t.p2.y)), fmaxf(t.p0.z, fmaxf(t.p1.z, t.p2.z)));
    cell_idx(g, mn, lo);
    cell_idx(g, mx, hi);
}
// Grid.getCellBbox (src/linalg.zig:433-441) + the overlap test for one cell of the range
__device__ __forceinline__ bool cell_overlaps(const GridP& g, const Tri& t, uint32_t x, uint32_t y, uint32_t z)
{
    const v3 bmin = mk(g.bmin[0] + g.cell[0] * (float)x, g.bmin[1] + g.cell[1] * (float)y, g.bmin[2] + g.cell[2] * (float)z);
    const v3 bmax = bmin + mk(g.cell[0], g.cell[1], g.cell[2]);
    return tri_box(t.p0, t.p1, t.p2, bmin, bmax);
}

struct BinOut {
    uint32_t* count;                    // per cell (COUNT)
    unsigned long long* keys;           // EMIT: cell << 32 | triangle
    unsigned long long* cursor;         // [0] keys written (EMIT) / pairs counted (COUNT), [1] deferred triangles
    uint32_t* big;                      // deferred triangle ids
};

template <bool EMIT>
__device__ __forceinline__ void visit(const BinOut& o, uint32_t cell, uint32_t tri)
{
    if (EMIT) {
        // one atomic per converged group of lanes, order inside the group is irrelevant (the keys get sorted)
        const uint32_t m = __activemask();
        const uint32_t lane = threadIdx.x & 31u;
        const int leader = __ffs(m) - 1;
        unsigned long long base = 0;
        if ((int)lane == leader) base = atomicAdd(o.cursor, (unsigned long long)__popc(m));
        base = __shfl_sync(m, base, leader);
        o.keys[base + __popc(m & ((1u << lane) - 1u))] = ((unsigned long long)cell << 32) | tri;
    } else {
        atomicAdd(o.count + cell, 1u);
    }
}

template <bool EMIT>
__global__ void __launch_bounds__(kThreads) k_bin(const float* __restrict__ pos, uint32_t num_tris, const __grid_constant__ GridP g, BinOut o)
{
    const uint32_t t = blockIdx.x * kThreads + threadIdx.x;
    uint32_t mine = 0;
    if (t < num_tris) {
        const Tri tri = load_tri(pos, t);
        uint32_t lo[3], hi[3];
        tri_range(g, tri, lo, hi);
        const uint64_t n = (uint64_t)(hi[0] - lo[0] + 1u) * (hi[1] - lo[1] + 1u) * (hi[2] - lo[2] + 1u);
        if (n > kInlineCells) {
            o.big[atomicAdd(o.cursor + 1, 1ull)] = t;
        } else {
            for (uint32_t z = lo[2]; z <= hi[2]; ++z)
                for (uint32_t y = lo[1]; y <= hi[1]; ++y)
                    for (uint32_t x = lo[0]; x <= hi[0]; ++x)
                        if (cell_overlaps(g, tri, x, y, z)) { visit<EMIT>(o, (z * g.res[1] + y) * g.res[0] + x, t); mine += 1; }
        }
    }
    if (!EMIT) {
        mine = __reduce_add_sync(0xFFFFFFFFu, mine);
        if ((threadIdx.x & 31) == 0 && mine) atomicAdd(o.cursor, (unsigned long long)mine);
    }
}

// Deferred triangles (walls, floors: ranges of thousands of cells): one CTA per triangle, lanes stride the range.
template <bool EMIT>
__global__ void __launch_bounds__(kThreads) k_bin_big(const float* __restrict__ pos, const __grid_constant__ GridP g, BinOut o)
{
    const uint32_t nbig = (uint32_t)o.cursor[1];
    uint32_t mine = 0;
    for (uint32_t b = blockIdx.x; b < nbig; b += gridDim.x) {
        const uint32_t t = o.big[b];
        const Tri tri = load_tri(pos, t);
        uint32_t lo[3], hi[3];
        tri_range(g, tri, lo, hi);
        const uint32_t nx = hi[0] - lo[0] + 1u, ny = hi[1] - lo[1] + 1u, nz = hi[2] - lo[2] + 1u;
        const uint64_t n = (uint64_t)nx * ny * nz;
        for (uint64_t i = threadIdx.x; i < n; i += kThreads) {
            const uint32_t x = lo[0] + (uint32_t)(i % nx), y = lo[1] + (uint32_t)((i / nx) % ny), z = lo[2] + (uint32_t)(i / ((uint64_t)nx * ny));
            if (cell_overlaps(g, tri, x, y, z)) { visit<EMIT>(o, (z * g.res[1] + y) * g.res[0] + x, t); mine += 1; }
        }
    }
    if (!EMIT) {
        mine = __reduce_add_sync(0xFFFFFFFFu, mine);
        if ((threadIdx.x & 31) == 0 && mine) atomicAdd(o.cursor, (unsigned long long)mine);
    }
}

// Cell{begin,end} (x-fastest order) + stage2's statistics: stats = {empty cells, min, max over non-empty cells}
__global__ void __launch_bounds__(kThreads) k_cells(const uint32_t* __restrict__ count, const uint32_t* __restrict__ begin, uint32_t num_cells,
                                                    uint2* __restrict__ cells, uint32_t* __restrict__ stats)
{
    const uint32_t c = blockIdx.x * kThreads + threadIdx.x;
    uint32_t n = 0;
    const bool on = c < num_cells;
    if (on) { n = count[c]; const uint32_t b = begin[c]; cells[c] = make_uint2(b, b + n); }
    const uint32_t empty = __popc(__ballot_sync(0xFFFFFFFFu, on && n == 0u));
    const uint32_t mn = __reduce_min_sync(0xFFFFFFFFu, on && n ? n : 0xFFFFFFFFu);
    const uint32_t mx = __reduce_max_sync(0xFFFFFFFFu, on ? n : 0u);
    if ((threadIdx.x & 31) == 0) {
        if (empty) atomicAdd(stats, empty);
        if (mn != 0xFFFFFFFFu) atomicMin(stats + 1, mn);
        if (mx) atomicMax(stats + 2, mx);
    }
}

// The cells by packed index (rtc_device.cuh) and the two occupancy levels. cells_b / fine / coarse are zeroed before.
struct Packing { uint32_t shift[3]; };
__global__ void __launch_bounds__(kThreads) k_layout(const uint2* __restrict__ cells, const __grid_constant__ GridP g, const __grid_constant__ Packing pk,
                                                     uint2* __restrict__ cells_b, unsigned long long* __restrict__ fine, uint32_t* __restrict__ coarse)
{
    const uint32_t c = blockIdx.x * kThreads + threadIdx.x;
    const uint32_t num_cells = g.res[0] * g.res[1] * g.res[2];
    if (c >= num_cells) return;
    const uint32_t x = c % g.res[0], y = (c / g.res[0]) % g.res[1], z = c / (g.res[0] * g.res[1]);
    const uint2 cell = cells[c];
    const uint32_t pc = pack_cell(pk.shift, x, y, z);
    cells_b[pc] = cell;
    if (cell.y > cell.x) {
        const uint32_t bi = pc >> 6;
        atomicOr(fine + bi, 1ull << (pc & 63u));
        atomicOr(coarse + (bi >> 5), 1u << (bi & 31u));
    }
}

// bakeInto (src/stage2.zig:137-164): duplicate the triangles in cell order; Triangle.init (src/linalg.zig:688-694):
// {v0, e1 = v1 - v0, e2 = v2 - v0}. v0.w carries the source-triangle id.
__global__ void __launch_bounds__(kThreads) k_bake_soa(const unsigned long long* __restrict__ keys, uint64_t num_baked, const float* __restrict__ pos,
                                                       float4* __restrict__ v0, float4* __restrict__ e1, float4* __restrict__ e2)
{
    const uint64_t i = (uint64_t)blockIdx.x * kThreads + threadIdx.x;
    if (i >= num_baked) return;
    const uint32_t s = (uint32_t)keys[i];
    const Tri t = load_tri(pos, s);
    const v3 a = t.p1 - t.p0, b = t.p2 - t.p0;
    v0[i] = make_float4(t.p0.x, t.p0.y, t.p0.z, __uint_as_float(s));
    e1[i] = make_float4(a.x, a.y, a.z, 0.0f);
    e2[i] = make_float4(b.x, b.y, b.z, 0.0f);
}

// rtc_tri_data (17 words, AoS) -> 4 float4 = 64 B per source triangle: 9 normal + 6 uv + material (the record's index
// IS the source id). flag[0] is set when a material index is out of range.
__global__ void __launch_bounds__(kThreads) k_pack_data(const uint32_t* __restrict__ data, uint32_t num_tris, uint32_t num_materials,
                                                        uint32_t* __restrict__ out, uint32_t* __restrict__ flag)
{
    const uint64_t i = (uint64_t)blockIdx.x * kThreads + threadIdx.x;
    if (i >= (uint64_t)num_tris * 16u) return;
    const uint32_t t = (uint32_t)(i / 16u), w = (uint32_t)(i % 16u);
    const uint32_t v = data[17 * (size_t)t + w];
    if (w == 15u && v >= num_materials) atomicOr(flag, 1u);
    out[i] = v;
}

// The bake in the C ABI's layout: rtc_tri_pos (12 floats) and rtc_tri_data (17 words, src_tri = source index).
__global__ void __launch_bounds__(kThreads) k_bake_aos(const unsigned long long* __restrict__ keys, uint64_t num_baked, const float* __restrict__ pos,
                                                       const uint32_t* __restrict__ data, float4* __restrict__ tri_pos, uint32_t* __restrict__ tri_data)
{
    const uint64_t i = (uint64_t)blockIdx.x * kThreads + threadIdx.x;
    if (i >= num_baked) return;
    const uint32_t s = (uint32_t)keys[i];
    const Tri t = load_tri(pos, s);
    const v3 a = t.p1 - t.p0, b = t.p2 - t.p0;
    tri_pos[3 * i] = make_float4(t.p0.x, t.p0.y, t.p0.z, 0.0f);
    tri_pos[3 * i + 1] = make_float4(a.x, a.y, a.z, 0.0f);
    tri_pos[3 * i + 2] = make_float4(b.x, b.y, b.z, 0.0f);
    for (int w = 0; w < 16; ++w) tri_data[17 * i + w] = data[17 * (size_t)s + w];
    tri_data[17 * i + 16] = s;
}

// ---------------------------------------------------------------------------------------------------------
// loadTriangles (src/stage1.zig:217-259) on the device: one thread per triangle of the soup. The primitive the
// triangle belongs to is found by bisection over the table's first_tri column; the three indices are read in
// their component type, range-checked against the accessors (flag[0] = a vertex index is out of range, what the
// host loader reports as an error), the positions go through the node's global transform and the normals through
// its direction part + normalize — the reference's fp32 operations in the reference's order (linalg.zig:263-278,
// :119-125), this file being compiled without FMA contraction. Accessor data may sit at any byte offset (glTF
// asks for 4-byte alignment, files in the wild do not always comply): unaligned words are put together from bytes.
__device__ __forceinline__ uint32_t ld_u32_any(const uint8_t* p)
{
    if ((reinterpret_cast<uintptr_t>(p) & 3u) == 0u) return __ldg(reinterpret_cast<const uint32_t*>(p));
    return (uint32_t)__ldg(p) | ((uint32_t)__ldg(p + 1) << 8) | ((uint32_t)__ldg(p + 2) << 16) | ((uint32_t)__ldg(p + 3) << 24);
}
__device__ __forceinline__ v3 ld_v3_any(const uint8_t* p)
{
    return mk(__uint_as_float(ld_u32_any(p)), __uint_as_float(ld_u32_any(p + 4)), __uint_as_float(ld_u32_any(p + 8)));
}
__device__ __forceinline__ uint32_t ld_index(const uint8_t* p, uint32_t component)
{
    if (component == 5125u) return ld_u32_any(p);
    if (component == 5123u) {
        if ((reinterpret_cast<uintptr_t>(p) & 1u) == 0u) return __ldg(reinterpret_cast<const uint16_t*>(p));
        return (uint32_t)__ldg(p) | ((uint32_t)__ldg(p + 1) << 8);
    }
    return __ldg(p);
}

__global__ void __launch_bounds__(kThreads) k_assemble(const uint8_t* __restrict__ blob, const RtcPrimRef* __restrict__ prims, uint32_t num_prims,
                                                       uint32_t num_tris, float* __restrict__ pos_out, uint32_t* __restrict__ data_out,
                                                       uint32_t* __restrict__ flag)
{
    __shared__ __align__(16) float sh_pos[9 * kThreads];
    __shared__ __align__(16) uint32_t sh_data[17 * kThreads];
    const uint32_t g0 = blockIdx.x * kThreads;
    const uint32_t g = min(g0 + threadIdx.x, num_tris - 1u);      // threads past the end redo the last triangle (not stored)
    uint32_t lo = 0u, hi = num_prims;                       // last primitive with first_tri <= g
    while (hi - lo > 1u) {
        const uint32_t mid = (lo + hi) >> 1;
        if (__ldg(&prims[mid].first_tri) <= g) lo = mid; else hi = mid;
    }
    const RtcPrimRef& pr = prims[lo];
    const uint32_t t = g - __ldg(&pr.first_tri);
    const v3 c0 = mk(__ldg(&pr.m[0]), __ldg(&pr.m[1]), __ldg(&pr.m[2]));
    const v3 c1 = mk(__ldg(&pr.m[4]), __ldg(&pr.m[5]), __ldg(&pr.m[6]));
    const v3 c2 = mk(__ldg(&pr.m[8]), __ldg(&pr.m[9]), __ldg(&pr.m[10]));
    const v3 c3 = mk(__ldg(&pr.m[12]), __ldg(&pr.m[13]), __ldg(&pr.m[14]));
    const uint32_t comp = __ldg(&pr.index_component), has_uv = __ldg(&pr.has_uv);
    const uint32_t s_ind = __ldg(&pr.stride_ind), s_pos = __ldg(&pr.stride_pos), s_nrm = __ldg(&pr.stride_nrm), s_uv = __ldg(&pr.stride_uv);
    const uint8_t* b_ind = blob + __ldg(&pr.off_ind);
    const uint8_t* b_pos = blob + __ldg(&pr.off_pos);
    const uint8_t* b_nrm = blob + __ldg(&pr.off_nrm);
    const uint8_t* b_uv = blob + __ldg(&pr.off_uv);
    const uint32_t n_pos = __ldg(&pr.count_pos), n_nrm = __ldg(&pr.count_nrm), n_uv = __ldg(&pr.count_uv);
    // records are staged in shared memory (strides 9 and 17 words: conflict-free) and leave the CTA as one contiguous,
    // 16-byte-vectorised run each — per-thread 36 B / 68 B records written directly touch every sector 9 / 17 times
    float* po = sh_pos + 9u * threadIdx.x;
    uint32_t* dout = sh_data + 17u * threadIdx.x;
    bool bad = false;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const uint32_t v = ld_index(b_ind + (size_t)(3u * t + (uint32_t)i) * s_ind, comp);
        v3 p = mk(0.0f, 0.0f, 0.0f), n = p;
        uint32_t uv0 = 0u, uv1 = 0u;
        if (v >= n_pos || v >= n_nrm || (has_uv && v >= n_uv)) bad = true;
        else {
            const v3 vp = ld_v3_any(b_pos + (size_t)v * s_pos);
            const v3 vn = ld_v3_any(b_nrm + (size_t)v * s_nrm);
            p = ((scale(c0, vp.x) + scale(c1, vp.y)) + scale(c2, vp.z)) + c3;
            n = normalize((scale(c0, vn.x) + scale(c1, vn.y)) + scale(c2, vn.z));   // not the inverse transpose (TODO in the reference)
            if (has_uv) { uv0 = ld_u32_any(b_uv + (size_t)v * s_uv); uv1 = ld_u32_any(b_uv + (size_t)v * s_uv + 4); }
        }
        po[3 * i] = p.x; po[3 * i + 1] = p.y; po[3 * i + 2] = p.z;
        dout[3 * i] = __float_as_uint(n.x); dout[3 * i + 1] = __float_as_uint(n.y); dout[3 * i + 2] = __float_as_uint(n.z);
        dout[9 + 2 * i] = uv0; dout[10 + 2 * i] = uv1;
    }
    dout[15] = __ldg(&pr.material);
    dout[16] = g;
    if (bad) atomicOr(flag, 1u);
    __syncthreads();
    const uint32_t live = min((uint32_t)kThreads, num_tris - g0);
    float4* gp = reinterpret_cast<float4*>(pos_out + 9 * (size_t)g0);            // g0 * 36 B and g0 * 68 B are multiples of 16
    uint4* gd = reinterpret_cast<uint4*>(data_out + 17 * (size_t)g0);
    const uint32_t np4 = (9u * live) / 4u, nd4 = (17u * live) / 4u;
    for (uint32_t i = threadIdx.x; i < np4; i += kThreads) gp[i] = reinterpret_cast<const float4*>(sh_pos)[i];
    for (uint32_t i = threadIdx.x; i < nd4; i += kThreads) gd[i] = reinterpret_cast<const uint4*>(sh_data)[i];
    for (uint32_t i = 4u * np4 + threadIdx.x; i < 9u * live; i += kThreads) pos_out[9 * (size_t)g0 + i] = sh_pos[i];
    for (uint32_t i = 4u * nd4 + threadIdx.x; i < 17u * live; i += kThreads) data_out[17 * (size_t)g0 + i] = sh_data[i];
}

// ---------------------------------------------------------------------------------------------------------
// Texel linearisation (stb_image.h:1856-1877 + src/stage1.zig:433-440, 455-468) on the device: the decoded
// 8-bit RGBA image goes up once (4 B/texel instead of 12-16 B of floats), every texture made from it is one
// launch. lut = pow(c/255.0f, 2.2f) evaluated in double on the host and rounded to float (256 entries); the
// per-texel arithmetic is one fp32 multiply by the material's factor (RGB), c/255 in fp32 (BLEND) or a compare
// of that against alphaCutoff (MASK) — bit-identical to the host loader's.
__global__ void __launch_bounds__(kThreads) k_texels_rgb(const uchar4* __restrict__ img, uint32_t n, const float* __restrict__ lut,
                                                         float k0, float k1, float k2, float4* __restrict__ out)
{
    __shared__ float s_lut[256];
    s_lut[threadIdx.x] = lut[threadIdx.x];          // kThreads == 256
    __syncthreads();
    for (uint32_t i = blockIdx.x * kThreads + threadIdx.x; i < n; i += gridDim.x * kThreads) {
        const uchar4 c = __ldg(img + i);
        out[i] = make_float4(s_lut[c.x] * k0, s_lut[c.y] * k1, s_lut[c.z] * k2, 0.0f);
    }
}
__global__ void __launch_bounds__(kThreads) k_texels_alpha(const uchar4* __restrict__ img, uint32_t n, uint32_t mask, float cutoff,
                                                           float* __restrict__ out)
{
    for (uint32_t i = blockIdx.x * kThreads + threadIdx.x; i < n; i += gridDim.x * kThreads) {
        const float a = (float)__ldg(img + i).w / 255.0f;
        out[i] = mask ? (a > cutoff ? 1.0f : 0.0f) : a;
    }
}

// ---------------------------------------------------------------------------------------------------------
// Host side of the build. Everything is allocated with cudaMalloc and owned by `Built` until released.
struct Built {
    rtc_grid grid{};
    float* d_pos = nullptr;             // the soup (9 floats / triangle)
    uint32_t* d_data = nullptr;         // rtc_tri_data records (17 words / triangle)
    uint2* d_cells = nullptr;           // num_cells, x-fastest
    unsigned long long* d_keys = nullptr;   // num_baked sorted keys
    uint32_t num_src = 0;
    uint64_t num_cells = 0, num_baked = 0;
    uint32_t empty_cells = 0, min_tris = 0, max_tris = 0;
    float device_ms = 0.0f;
    float assemble_ms = 0.0f;           // k_assemble (glTF plan -> soup), 0 when the soup came from the host
    std::vector<void*> tmp;             // scratch freed at the end of the build
    ~Built()
    {
        for (void* p : tmp) cudaFree(p);
        cudaFree(d_pos); cudaFree(d_data); cudaFree(d_cells); cudaFree(d_keys);
    }
};

template <typename T>
static int dalloc(T** out, size_t count, std::vector<void*>* track)
{
    void* p = nullptr;
    const cudaError_t e = cudaMalloc(&p, (count ? count : 1) * sizeof(T));
    if (e != cudaSuccess) return rtc_fail_(RTC_ERR_OUT_OF_MEMORY, "cudaMalloc(%zu bytes) in the grid build: %s", count * sizeof(T), cudaGetErrorString(e));
    if (track) track->push_back(p);
    *out = static_cast<T*>(p);
    return RTC_OK;
}

// The soup from host memory: b->d_pos / b->d_data.
static int soup_upload(const float* positions, const rtc_tri_data* data, uint32_t num_tris, cudaStream_t st, Built* b)
{
    static_assert(sizeof(rtc_tri_data) == 17 * sizeof(uint32_t), "rtc_tri_data is 17 words");
    if (int rc = dalloc(&b->d_pos, (size_t)num_tris * 9, nullptr)) return rc;
    if (int rc = dalloc(&b->d_data, (size_t)num_tris * 17, nullptr)) return rc;
    RTC_CU(cudaMemcpyAsync(b->d_pos, positions, (size_t)num_tris * 9 * sizeof(float), cudaMemcpyHostToDevice, st));
    RTC_CU(cudaMemcpyAsync(b->d_data, data, (size_t)num_tris * sizeof(rtc_tri_data), cudaMemcpyHostToDevice, st));
    return RTC_OK;
}

// The soup from a glTF plan: the referenced buffers and the primitive table go up, k_assemble writes
// b->d_pos / b->d_data.
static int soup_assemble(const RtcGltfPlan& plan, cudaStream_t st, Built* b)
{
    static_assert(sizeof(rtc_tri_data) == 17 * sizeof(uint32_t), "rtc_tri_data is 17 words");
    const uint32_t num_tris = plan.num_tris;
    if (int rc = dalloc(&b->d_pos, (size_t)num_tris * 9, nullptr)) return rc;
    if (int rc = dalloc(&b->d_data, (size_t)num_tris * 17, nullptr)) return rc;
    if (!num_tris) return RTC_OK;
    uint8_t* d_blob = nullptr; RtcPrimRef* d_prims = nullptr; uint32_t* d_flag = nullptr;
    if (int rc = dalloc(&d_blob, (size_t)plan.blob_bytes + 16, &b->tmp)) return rc;
    if (int rc = dalloc(&d_prims, plan.prims.size(), &b->tmp)) return rc;
    if (int rc = dalloc(&d_flag, 1, &b->tmp)) return rc;
    for (size_t i = 0; i < plan.buffers.size(); ++i)
        if (plan.buffer_base[i] != ~(uint64_t)0 && !plan.buffers[i].empty())
            RTC_CU(cudaMemcpyAsync(d_blob + plan.buffer_base[i], plan.buffers[i].data(), plan.buffers[i].size(), cudaMemcpyHostToDevice, st));
    RTC_CU(cudaMemcpyAsync(d_prims, plan.prims.data(), plan.prims.size() * sizeof(RtcPrimRef), cudaMemcpyHostToDevice, st));
    RTC_CU(cudaMemsetAsync(d_flag, 0, sizeof(uint32_t), st));
    cudaEvent_t e0, e1;
    RTC_CU(cudaEventCreate(&e0));
    RTC_CU(cudaEventCreate(&e1));
    struct EvGuard { cudaEvent_t a, b; ~EvGuard() { cudaEventDestroy(a); cudaEventDestroy(b); } } evg{ e0, e1 };
    RTC_CU(cudaEventRecord(e0, st));
    k_assemble<<<(num_tris + kThreads - 1) / kThreads, kThreads, 0, st>>>(d_blob, d_prims, (uint32_t)plan.prims.size(), num_tris, b->d_pos, b->d_data, d_flag);
    RTC_CU(cudaEventRecord(e1, st));
    uint32_t flag = 0;
    RTC_CU(cudaMemcpyAsync(&flag, d_flag, sizeof flag, cudaMemcpyDeviceToHost, st));
    RTC_CU(cudaStreamSynchronize(st));
    RTC_CU(cudaGetLastError());
    RTC_CU(cudaEventElapsedTime(&b->assemble_ms, e0, e1));
    if (flag) return rtc_fail_(RTC_ERR_FORMAT, "glTF: vertex index out of range");
    cudaFree(d_blob);                                   // the largest scratch: gone before the grid build allocates
    b->tmp.erase(std::find(b->tmp.begin(), b->tmp.end(), (void*)d_blob));
    return RTC_OK;
}

// The grid over the device-resident soup b->d_pos (num_tris triangles).
static int grid_build(uint32_t num_tris, const uint32_t res[3], cudaStream_t st, Built* b)
{
    const uint64_t num_cells = (uint64_t)res[0] * res[1] * res[2];
    if (num_cells >= ((uint64_t)1 << 32)) return rtc_fail_(RTC_ERR_UNSUPPORTED, "grid has 2^32 cells or more");
    b->num_cells = num_cells; b->num_src = num_tris;
    // device time = three bracketed kernel groups (bbox | count | scan, cells, emit, sort); the cudaMallocs and the
    // two small D2H reads between them (the grid box, the baked-triangle total) are host-side and not counted
    cudaEvent_t ev[6];
    for (cudaEvent_t& e : ev) RTC_CU(cudaEventCreate(&e));
    struct EvGuard { cudaEvent_t* e; ~EvGuard() { for (int i = 0; i < 6; ++i) cudaEventDestroy(e[i]); } } evg{ ev };

    if (int rc = dalloc(&b->d_cells, (size_t)num_cells, nullptr)) return rc;

    // initGrid
    int* d_box = nullptr;
    if (int rc = dalloc(&d_box, 6, &b->tmp)) return rc;
    const int box_init[6] = { 0x7F800000, 0x7F800000, 0x7F800000, (int)0x807FFFFF, (int)0x807FFFFF, (int)0x807FFFFF };
    RTC_CU(cudaMemcpyAsync(d_box, box_init, sizeof box_init, cudaMemcpyHostToDevice, st));
    const uint64_t nv = (uint64_t)num_tris * 3;
    RTC_CU(cudaEventRecord(ev[0], st));
    if (nv) {
        const unsigned gridx = (unsigned)std::min<uint64_t>((nv + kThreads - 1) / kThreads, 148u * 8u);
        k_bbox<<<gridx, kThreads, 0, st>>>(b->d_pos, nv, d_box);
    }
    RTC_CU(cudaEventRecord(ev[1], st));
    int box[6];
    RTC_CU(cudaMemcpyAsync(box, d_box, sizeof box, cudaMemcpyDeviceToHost, st));
    RTC_CU(cudaStreamSynchronize(st));
    GridP g;
    for (int a = 0; a < 3; ++a) {
        const float lo = ord2f_bits(box[a], true), hi = ord2f_bits(box[3 + a], true);
        b->grid.bbox_min[a] = lo; b->grid.bbox_max[a] = hi; b->grid.resolution[a] = res[a];
        b->grid.cell_size[a] = (hi - lo) / (float)res[a];            // Grid.init (linalg.zig:412-418)
        g.bmin[a] = lo; g.cell[a] = b->grid.cell_size[a]; g.res[a] = res[a];
    }

    // initCells: count
    uint32_t* d_count = nullptr; uint32_t* d_begin = nullptr; unsigned long long* d_cursor = nullptr; uint32_t* d_big = nullptr; uint32_t* d_stats = nullptr;
    if (int rc = dalloc(&d_count, (size_t)num_cells, &b->tmp)) return rc;
    if (int rc = dalloc(&d_begin, (size_t)num_cells, &b->tmp)) return rc;
    if (int rc = dalloc(&d_cursor, 2, &b->tmp)) return rc;
    if (int rc = dalloc(&d_big, (size_t)num_tris, &b->tmp)) return rc;
    if (int rc = dalloc(&d_stats, 3, &b->tmp)) return rc;
    RTC_CU(cudaMemsetAsync(d_count, 0, (size_t)num_cells * sizeof(uint32_t), st));
    RTC_CU(cudaMemsetAsync(d_cursor, 0, 2 * sizeof(unsigned long long), st));
    BinOut o; o.count = d_count; o.keys = nullptr; o.cursor = d_cursor; o.big = d_big;
    const unsigned tri_blocks = (num_tris + kThreads - 1) / kThreads;
    RTC_CU(cudaEventRecord(ev[2], st));
    if (num_tris) {
        k_bin<false><<<tri_blocks, kThreads, 0, st>>>(b->d_pos, num_tris, g, o);
        k_bin_big<false><<<148 * 8, kThreads, 0, st>>>(b->d_pos, g, o);
    }
    RTC_CU(cudaEventRecord(ev[3], st));
    unsigned long long h_cursor[2] = { 0, 0 };
    RTC_CU(cudaMemcpyAsync(h_cursor, d_cursor, sizeof h_cursor, cudaMemcpyDeviceToHost, st));
    RTC_CU(cudaStreamSynchronize(st));
    const uint64_t M = h_cursor[0];
    if (M >= ((uint64_t)1 << 32)) return rtc_fail_(RTC_ERR_UNSUPPORTED, "more than 2^32 baked triangles");
    b->num_baked = M;

    // exclusive scan -> Cell.begin; cells + statistics
    size_t scan_bytes = 0, sort_bytes = 0;
    RTC_CU(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, d_count, d_begin, (int)num_cells, st));
    unsigned long long* d_keys_in = nullptr;
    if (int rc = dalloc(&d_keys_in, (size_t)M, &b->tmp)) return rc;
    if (int rc = dalloc(&b->d_keys, (size_t)M, nullptr)) return rc;
    int cell_bits = 0;
    while (((uint64_t)1 << cell_bits) < num_cells) ++cell_bits;
    RTC_CU(cub::DeviceRadixSort::SortKeys(nullptr, sort_bytes, d_keys_in, b->d_keys, (int)M, 0, 32 + cell_bits, st));
    uint8_t* d_temp = nullptr;
    if (int rc = dalloc(&d_temp, std::max(scan_bytes, sort_bytes), &b->tmp)) return rc;
    RTC_CU(cudaEventRecord(ev[4], st));
    RTC_CU(cub::DeviceScan::ExclusiveSum(d_temp, scan_bytes, d_count, d_begin, (int)num_cells, st));
    const uint32_t stats_init[3] = { 0u, 0xFFFFFFFFu, 0u };
    RTC_CU(cudaMemcpyAsync(d_stats, stats_init, sizeof stats_init, cudaMemcpyHostToDevice, st));
    k_cells<<<(unsigned)((num_cells + kThreads - 1) / kThreads), kThreads, 0, st>>>(d_count, d_begin, (uint32_t)num_cells, b->d_cells, d_stats);

    // initIndices: emit the (cell, triangle) keys, sort
    if (M) {
        RTC_CU(cudaMemsetAsync(d_cursor, 0, 2 * sizeof(unsigned long long), st));
        o.keys = d_keys_in;
        k_bin<true><<<tri_blocks, kThreads, 0, st>>>(b->d_pos, num_tris, g, o);
        k_bin_big<true><<<148 * 8, kThreads, 0, st>>>(b->d_pos, g, o);
        RTC_CU(cub::DeviceRadixSort::SortKeys(d_temp, sort_bytes, d_keys_in, b->d_keys, (int)M, 0, 32 + cell_bits, st));
    }
    RTC_CU(cudaEventRecord(ev[5], st));
    uint32_t h_stats[3];
    RTC_CU(cudaMemcpyAsync(h_stats, d_stats, sizeof h_stats, cudaMemcpyDeviceToHost, st));
    RTC_CU(cudaMemcpyAsync(h_cursor, d_cursor, sizeof h_cursor, cudaMemcpyDeviceToHost, st));
    RTC_CU(cudaStreamSynchronize(st));
    RTC_CU(cudaGetLastError());
    if (M && h_cursor[0] != M) return rtc_fail_(RTC_ERR_CUDA, "grid build: the emit pass produced %llu keys, the count pass %llu", h_cursor[0], (unsigned long long)M);
    // stage2.zig:76-100 initialises min to maxInt and max to 0; the host build reports the same values
    b->empty_cells = h_stats[0]; b->min_tris = h_stats[1]; b->max_tris = h_stats[2];
    b->device_ms = 0.0f;
    for (int i = 0; i < 3; ++i) { float ms = 0.0f; RTC_CU(cudaEventElapsedTime(&ms, ev[2 * i], ev[2 * i + 1])); b->device_ms += ms; }
    for (void* p : b->tmp) cudaFree(p);
    b->tmp.clear();
    return RTC_OK;
}

}  // namespace build
}  // namespace rtc

using rtc::build::kThreads;

static int check_soup_args(const float* positions, const rtc_tri_data* data, uint32_t num_tris, const uint32_t* res)
{
    if ((num_tris && (!positions || !data)) || !res || !res[0] || !res[1] || !res[2])
        return rtc_fail_(RTC_ERR_INVALID_ARGUMENT, "null triangle soup or zero grid resolution");
    return RTC_OK;
}

static int pick_device(int* device)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) { cudaGetLastError(); return rtc_fail_(RTC_ERR_NO_DEVICE, "no CUDA device (this library has no CPU fallback)"); }
    if (*device < 0) RTC_CU(cudaGetDevice(device));
    if (*device >= n) return rtc_fail_(RTC_ERR_INVALID_ARGUMENT, "device %d of %d", *device, n);
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, *device) != cudaSuccess || p.major != 10)
        return rtc_fail_(RTC_ERR_NO_DEVICE, "device %d is not sm_100 (kernels are built for sm_100a only)", *device);
    return RTC_OK;
}

extern "C" int rtc_geometry_build_device(const float* positions, const rtc_tri_data* data, uint32_t num_tris,
                                         const uint32_t res[3], int device, rtc_geometry* out)
{
    if (!out) return rtc_fail_(RTC_ERR_INVALID_ARGUMENT, "rtc_geometry_build_device: out is null");
    memset(out, 0, sizeof *out);
    if (int rc = check_soup_args(positions, data, num_tris, res)) return rc;
    if (int rc = pick_device(&device)) return rc;
    RtcDeviceGuard guard(device);
    try {
        rtc::build::Built b;
        if (int rc = rtc::build::soup_upload(positions, data, num_tris, nullptr, &b)) return rc;
        if (int rc = rtc::build::grid_build(num_tris, res, nullptr, &b)) return rc;
        std::unique_ptr<RtcGeometryOwner> own(new RtcGeometryOwner());
        own->cells.resize(b.num_cells);
        own->tri_pos.resize(b.num_baked);
        own->tri_data.resize(b.num_baked);
        static_assert(sizeof(rtc_cell) == sizeof(uint2) && sizeof(rtc_tri_pos) == 3 * sizeof(float4), "ABI layouts");
        RTC_CU(cudaMemcpy(own->cells.data(), b.d_cells, b.num_cells * sizeof(uint2), cudaMemcpyDeviceToHost));
        if (b.num_baked) {
            float4* d_tp = nullptr; uint32_t* d_td = nullptr;
            if (int rc = rtc::build::dalloc(&d_tp, 3 * (size_t)b.num_baked, &b.tmp)) return rc;
            if (int rc = rtc::build::dalloc(&d_td, 17 * (size_t)b.num_baked, &b.tmp)) return rc;
            rtc::build::k_bake_aos<<<(unsigned)((b.num_baked + kThreads - 1) / kThreads), kThreads>>>(b.d_keys, b.num_baked, b.d_pos, b.d_data, d_tp, d_td);
            RTC_CU(cudaGetLastError());
            RTC_CU(cudaMemcpy(own->tri_pos.data(), d_tp, b.num_baked * sizeof(rtc_tri_pos), cudaMemcpyDeviceToHost));
            RTC_CU(cudaMemcpy(own->tri_data.data(), d_td, b.num_baked * sizeof(rtc_tri_data), cudaMemcpyDeviceToHost));
        }
        out->grid = b.grid;
        out->cells = own->cells.data(); out->num_cells = b.num_cells;
        out->tri_pos = own->tri_pos.data(); out->tri_data = own->tri_data.data();
        out->num_tris = b.num_baked; out->num_src_tris = num_tris;
        out->empty_cells = b.empty_cells; out->min_tris = b.min_tris; out->max_tris = b.max_tris;
        out->owner = own.release();
        return RTC_OK;
    } catch (const std::bad_alloc&) {
        return rtc_fail_(RTC_ERR_OUT_OF_MEMORY, "rtc_geometry_build_device: host allocation failed");
    }
}

// The plan's image-backed textures on the device: every image is uploaded once as RGBA8, every recipe is one
// launch into `alloc`-ed memory (the scene's, or scratch). preset[3*material + slot] receives the device texels.
template <typename Alloc>
static int textures_on_device(const RtcGltfPlan& plan, Alloc&& alloc, std::vector<void*>* scratch, std::vector<const void*>* preset, float* ms)
{
    preset->assign(3 * plan.materials.size(), nullptr);
    if (plan.recipes.empty()) return RTC_OK;
    float* d_lut = nullptr;
    if (int rc = rtc::build::dalloc(&d_lut, 256, scratch)) return rc;
    RTC_CU(cudaMemcpy(d_lut, plan.lut, sizeof plan.lut, cudaMemcpyHostToDevice));
    std::vector<uchar4*> d_img(plan.images8.size(), nullptr);
    cudaEvent_t e0, e1;
    RTC_CU(cudaEventCreate(&e0));
    RTC_CU(cudaEventCreate(&e1));
    struct EvGuard { cudaEvent_t a, b; ~EvGuard() { cudaEventDestroy(a); cudaEventDestroy(b); } } evg{ e0, e1 };
    for (const RtcTexelRecipe& r : plan.recipes) {                     // uploads first, so that the kernels run back to back
        const RtcImage8& im = plan.images8[r.image];
        const size_t n = (size_t)im.w * im.h;
        if (n == 0 || n > ((size_t)1 << 30)) return rtc_fail_(RTC_ERR_UNSUPPORTED, "texture image %u is empty or larger than 2^30 texels", r.image);
        if (!d_img[r.image]) {
            if (int rc = rtc::build::dalloc(&d_img[r.image], n, scratch)) return rc;
            RTC_CU(cudaMemcpyAsync(d_img[r.image], im.rgba.data(), n * 4, cudaMemcpyHostToDevice, nullptr));
        }
    }
    std::vector<void*> outs(plan.recipes.size(), nullptr);
    for (size_t i = 0; i < plan.recipes.size(); ++i) {
        const RtcTexelRecipe& r = plan.recipes[i];
        const size_t n = (size_t)plan.images8[r.image].w * plan.images8[r.image].h;
        if (r.kind == RTC_TEXEL_RGB) { float4* o = nullptr; if (int rc = alloc(&o, n)) return rc; outs[i] = o; }
        else { float* o = nullptr; if (int rc = alloc(&o, n)) return rc; outs[i] = o; }
    }
    RTC_CU(cudaEventRecord(e0, nullptr));
    for (size_t i = 0; i < plan.recipes.size(); ++i) {
        const RtcTexelRecipe& r = plan.recipes[i];
        const uint32_t n = plan.images8[r.image].w * plan.images8[r.image].h;
        const unsigned blocks = (unsigned)std::min<uint64_t>(((uint64_t)n + kThreads - 1) / kThreads, 148u * 16u);
        if (r.kind == RTC_TEXEL_RGB)
            rtc::build::k_texels_rgb<<<blocks, kThreads>>>(d_img[r.image], n, d_lut, r.k[0], r.k[1], r.k[2], (float4*)outs[i]);
        else
            rtc::build::k_texels_alpha<<<blocks, kThreads>>>(d_img[r.image], n, r.kind == RTC_TEXEL_ALPHA_MASK ? 1u : 0u, r.k[0], (float*)outs[i]);
        (*preset)[3 * (size_t)r.material + r.slot] = outs[i];
    }
    RTC_CU(cudaEventRecord(e1, nullptr));
    RTC_CU(cudaEventSynchronize(e1));
    RTC_CU(cudaGetLastError());
    if (ms) RTC_CU(cudaEventElapsedTime(ms, e0, e1));
    return RTC_OK;
}

// The scene around a soup that is already on the device (b.d_pos / b.d_data): grid build, packed layout, bake,
// materials. Shared by the soup and the glTF entry points.
static int scene_from_built(rtc_scene* s, rtc::build::Built& b, uint32_t num_tris, const uint32_t res[3],
                            const rtc_material* materials, uint32_t num_materials, rtc_build_info* info,
                            const void* const* preset = nullptr)
{
    if (int rc = rtc::build::grid_build(num_tris, res, nullptr, &b)) return rc;
    rtc::DevScene& d = s->dev;
    for (int i = 0; i < 3; ++i) {
        d.bmin[i] = b.grid.bbox_min[i]; d.bmax[i] = b.grid.bbox_max[i];
        d.res[i] = b.grid.resolution[i]; d.cell[i] = b.grid.cell_size[i];
    }
    const size_t nb = rtc_scene_set_packing_(d, res);
    uint2* d_cells = nullptr; unsigned long long* d_fine = nullptr; uint32_t* d_coarse = nullptr;
    float4* d_pos = nullptr; uint32_t* d_data = nullptr; uint32_t* d_flag = nullptr;
    const size_t M = (size_t)b.num_baked;
    if (int rc = s->alloc(&d_cells, nb * 64)) return rc;
    if (int rc = s->alloc(&d_fine, nb)) return rc;
    if (int rc = s->alloc(&d_coarse, (size_t)d.coarse_words)) return rc;
    if (int rc = s->alloc(&d_pos, 3 * M)) return rc;
    if (int rc = s->alloc(&d_data, 16 * (size_t)num_tris)) return rc;
    if (int rc = rtc::build::dalloc(&d_flag, 1, &b.tmp)) return rc;
    RTC_CU(cudaMemset(d_cells, 0, nb * 64 * sizeof(uint2)));
    RTC_CU(cudaMemset(d_fine, 0, nb * sizeof(unsigned long long)));
    RTC_CU(cudaMemset(d_coarse, 0, (size_t)d.coarse_words * sizeof(uint32_t)));
    RTC_CU(cudaMemset(d_flag, 0, sizeof(uint32_t)));
    rtc::build::GridP g;
    rtc::build::Packing pk;
    for (int a = 0; a < 3; ++a) { g.bmin[a] = b.grid.bbox_min[a]; g.cell[a] = b.grid.cell_size[a]; g.res[a] = res[a]; pk.shift[a] = d.shift[a]; }
    rtc::build::k_layout<<<(unsigned)((b.num_cells + kThreads - 1) / kThreads), kThreads>>>(b.d_cells, g, pk, d_cells, d_fine, d_coarse);
    if (M) rtc::build::k_bake_soa<<<(unsigned)((M + kThreads - 1) / kThreads), kThreads>>>(b.d_keys, M, b.d_pos, d_pos, d_pos + M, d_pos + 2 * M);
    if (num_tris) {
        const uint64_t words = (uint64_t)num_tris * 16u;
        rtc::build::k_pack_data<<<(unsigned)((words + kThreads - 1) / kThreads), kThreads>>>(b.d_data, num_tris, num_materials, d_data, d_flag);
    }
    RTC_CU(cudaGetLastError());
    uint32_t flag = 0;
    RTC_CU(cudaMemcpy(&flag, d_flag, sizeof flag, cudaMemcpyDeviceToHost));
    if (flag) return rtc_fail_(RTC_ERR_INVALID_ARGUMENT, "a triangle references a material index >= %u", num_materials);
    d.cells = d_cells; d.fine = d_fine; d.coarse = d_coarse;
    d.tri_v0 = d_pos; d.tri_e1 = d_pos + M; d.tri_e2 = d_pos + 2 * M; d.tri_data = (const float4*)d_data;
    s->num_tris = b.num_baked; s->num_src_tris = num_tris;
    if (info) {
        info->grid = b.grid; info->num_cells = b.num_cells; info->num_tris = b.num_baked; info->num_src_tris = num_tris;
        info->empty_cells = b.empty_cells; info->min_tris = b.min_tris; info->max_tris = b.max_tris;
        info->device_ms = b.device_ms; info->assemble_ms = b.assemble_ms;
    }
    if (int rc = rtc_scene_upload_materials_(s, materials, num_materials, preset)) return rc;
    return rtc_scene_finish_(s);
}

static int check_res(const uint32_t* res)
{
    if (!res || !res[0] || !res[1] || !res[2]) return rtc_fail_(RTC_ERR_INVALID_ARGUMENT, "zero grid resolution");
    for (int a = 0; a < 3; ++a)
        if (res[a] > rtc::kMaxGridRes)
            return rtc_fail_(RTC_ERR_UNSUPPORTED, "grid resolution %u on axis %d exceeds %u (cell coordinates are packed 10 bits per axis)", res[a], a, rtc::kMaxGridRes);
    return RTC_OK;
}

extern "C" int rtc_scene_create_from_soup(const float* positions, const rtc_tri_data* data, uint32_t num_tris,
                                          const uint32_t res[3], const rtc_material* materials, uint32_t num_materials,
                                          int device, rtc_scene** out, rtc_build_info* info)
{
    if (!out) return rtc_fail_(RTC_ERR_INVALID_ARGUMENT, "rtc_scene_create_from_soup: out is null");
    *out = nullptr;
    if (info) memset(info, 0, sizeof *info);
    if (int rc = check_soup_args(positions, data, num_tris, res)) return rc;
    if (int rc = check_res(res)) return rc;
    rtc_scene* s = nullptr;
    if (int rc = rtc_scene_new_(device, &s)) return rc;
    RtcDeviceGuard guard(s->device);
    const auto t0 = std::chrono::steady_clock::now();
    auto body = [&]() -> int {
        rtc::build::Built b;
        if (int rc = rtc::build::soup_upload(positions, data, num_tris, nullptr, &b)) return rc;
        return scene_from_built(s, b, num_tris, res, materials, num_materials, info);
    };
    const int rc = body();
    if (rc != RTC_OK) { rtc_keep_error_free_(s); return rc; }
    if (info) info->total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    *out = s;
    return RTC_OK;
}

// ---- stage1's triangle assembly on the device ---------------------------------------------

extern "C" int rtc_scene_create_from_gltf(const char* path, const char* camera_name, uint32_t width, uint32_t height,
                                          int num_threads, const uint32_t res[3], int device,
                                          rtc_scene** out, rtc_camera* camera, rtc_build_info* info)
{
    if (!out || !path) return rtc_fail_(RTC_ERR_INVALID_ARGUMENT, "rtc_scene_create_from_gltf: null argument");
    *out = nullptr;
    if (info) memset(info, 0, sizeof *info);
    if (int rc = check_res(res)) return rc;
    const auto t0 = std::chrono::steady_clock::now();
    try {
        RtcGltfPlan plan;
        if (int rc = rtc_gltf_plan_(path, camera_name, width, height, num_threads, &plan)) return rc;
        rtc_scene* s = nullptr;
        if (int rc = rtc_scene_new_(device, &s)) return rc;
        RtcDeviceGuard guard(s->device);
        auto body = [&]() -> int {
            rtc::build::Built b;
            if (int rc = rtc::build::soup_assemble(plan, nullptr, &b)) return rc;
            std::vector<const void*> preset;
            float texel_ms = 0.0f;
            if (int rc = textures_on_device(plan, [&](auto** p, size_t n) { return s->alloc(p, n); }, &b.tmp, &preset, &texel_ms)) return rc;
            if (int rc = scene_from_built(s, b, plan.num_tris, res, plan.materials.data(), (uint32_t)plan.materials.size(), info, preset.data())) return rc;
            if (info) info->assemble_ms += texel_ms;
            return RTC_OK;
        };
        const int rc = body();
        if (rc != RTC_OK) { rtc_keep_error_free_(s); return rc; }
        if (camera) *camera = plan.camera;
        if (info) info->total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        *out = s;
        return RTC_OK;
    } catch (const std::bad_alloc&) {
        return rtc_fail_(RTC_ERR_OUT_OF_MEMORY, "rtc_scene_create_from_gltf: host allocation failed");
    }
}

extern "C" int rtc_gltf_load_device(const char* path, const char* camera_name, uint32_t width, uint32_t height,
                                    int num_threads, int device, rtc_host_scene* out)
{
    if (!out || !path) return rtc_fail_(RTC_ERR_INVALID_ARGUMENT, "rtc_gltf_load_device: null argument");
    memset(out, 0, sizeof *out);
    try {
        RtcGltfPlan plan;
        if (int rc = rtc_gltf_plan_(path, camera_name, width, height, num_threads, &plan)) return rc;
        if (int rc = pick_device(&device)) return rc;
        RtcDeviceGuard guard(device);
        rtc::build::Built b;
        if (int rc = rtc::build::soup_assemble(plan, nullptr, &b)) return rc;
        std::unique_ptr<RtcHostSceneOwner> own(new RtcHostSceneOwner());
        own->positions.resize((size_t)plan.num_tris * 9);
        own->tri_data.resize(plan.num_tris);
        if (plan.num_tris) {
            RTC_CU(cudaMemcpy(own->positions.data(), b.d_pos, own->positions.size() * sizeof(float), cudaMemcpyDeviceToHost));
            RTC_CU(cudaMemcpy(own->tri_data.data(), b.d_data, own->tri_data.size() * sizeof(rtc_tri_data), cudaMemcpyDeviceToHost));
        }
        std::vector<const void*> preset;
        if (int rc = textures_on_device(plan, [&](auto** p, size_t n) { return rtc::build::dalloc(p, n, &b.tmp); }, &b.tmp, &preset, nullptr)) return rc;
        for (const RtcTexelRecipe& r : plan.recipes) {                  // back to the host in rtc_texture's layout (3 or 1 floats per texel)
            const size_t n = (size_t)plan.images8[r.image].w * plan.images8[r.image].h;
            const void* d = preset[3 * (size_t)r.material + r.slot];
            std::vector<float> px;
            if (r.kind == RTC_TEXEL_RGB) {
                std::vector<float4> tmp4(n);
                RTC_CU(cudaMemcpy(tmp4.data(), d, n * sizeof(float4), cudaMemcpyDeviceToHost));
                px.resize(3 * n);
                for (size_t i = 0; i < n; ++i) { px[3 * i] = tmp4[i].x; px[3 * i + 1] = tmp4[i].y; px[3 * i + 2] = tmp4[i].z; }
            } else {
                px.resize(n);
                RTC_CU(cudaMemcpy(px.data(), d, n * sizeof(float), cudaMemcpyDeviceToHost));
            }
            plan.texels.push_back(std::move(px));
            rtc_material& m = plan.materials[r.material];
            (r.slot == 0 ? m.base_color : (r.slot == 1 ? m.emissive : m.transparency)).texels = plan.texels.back().data();
        }
        own->materials = std::move(plan.materials);
        own->texels = std::move(plan.texels);
        out->camera = plan.camera;
        out->positions = own->positions.data(); out->tri_data = own->tri_data.data(); out->num_tris = plan.num_tris;
        out->materials = own->materials.data(); out->num_materials = (uint32_t)own->materials.size();
        out->num_images = plan.num_images;
        out->owner = own.release();
        return RTC_OK;
    } catch (const std::bad_alloc&) {
        return rtc_fail_(RTC_ERR_OUT_OF_MEMORY, "rtc_gltf_load_device: host allocation failed");
    }
}
