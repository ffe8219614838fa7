// rtc_geometry.cpp — host producer of the render path's inputs: Geometry.build + bakeInto
// (src/stage2.zig:44-164) on a triangle soup. The reference runs this serially; here the two
// triangle passes run on all host threads. Cell contents are kept in ascending source-triangle
// order (what the reference's serial scatter produces, stage2.zig:107-124), which fixes the
// tie-breaking order of Scene.traceRay.
//
// Compiled with -ffp-contract=off: the overlap test decides which cells hold a triangle and must
// evaluate exactly like the reference's fp32 code (src/linalg.zig:500-563).
#include "../../include/rtc.h"
#include "rtc_geometry_owner.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <new>
#include <thread>
#include <vector>

int rtc_set_error_(int code, const char* msg);   // rtc_api.cu

namespace {

struct V { float x, y, z; };
inline V mk(float x, float y, float z) { return V{ x, y, z }; }
inline V ld(const float* p) { return V{ p[0], p[1], p[2] }; }
inline V operator+(V a, V b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V operator-(V a, V b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V scale(V a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
inline V vmin(V a, V b) { return mk(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)); }
inline V vmax(V a, V b) { return mk(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }
inline float dot(V a, V b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }          // linalg.zig:189-191
inline V cross(V a, V b) { return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline V normalize(V a) { return scale(a, 1.0f / sqrtf(dot(a, a))); }               // linalg.zig:119-125

// @intFromFloat with the defined semantics of the render kernels (cvt.rzi.u32.f32).
inline uint32_t f2u(float f)
{
    if (!(f > 0.0f)) return 0u;
    if (f >= 4294967296.0f) return 0xFFFFFFFFu;
    return (uint32_t)f;
}

// intersectsTriangleAabbSat (src/linalg.zig:500-515)
inline bool sat(V a, V b, V c, V e, V axis)
{
    const float p0 = dot(a, axis), p1 = dot(b, axis), p2 = dot(c, axis);
    const float r = e.x * fabsf(dot(mk(1, 0, 0), axis)) + e.y * fabsf(dot(mk(0, 1, 0), axis)) +
                    e.z * fabsf(dot(mk(0, 0, 1), axis));
    const float hi = fmaxf(p0, fmaxf(p1, p2)), lo = fminf(p0, fminf(p1, p2));
    return !(fmaxf(-hi, lo) > r);
}

// intersectsTriangleAabb (src/linalg.zig:517-563): 9 edge-cross axes, 3 box axes, the face normal.
bool tri_box(V t0, V t1, V t2, V bmin, V bmax)
{
    const V center = scale(bmax + bmin, 0.5f), e = scale(bmax - bmin, 0.5f);
    const V a = t0 - center, b = t1 - center, c = t2 - center;
    const V ab = normalize(b - a), bc = normalize(c - b), ca = normalize(a - c);
    const V axes[13] = {
        mk(0.0f, -ab.z, ab.y), mk(0.0f, -bc.z, bc.y), mk(0.0f, -ca.z, ca.y),
        mk(ab.z, 0.0f, -ab.x), mk(bc.z, 0.0f, -bc.x), mk(ca.z, 0.0f, -ca.x),
        mk(-ab.y, ab.x, 0.0f), mk(-bc.y, bc.x, 0.0f), mk(-ca.y, ca.x, 0.0f),
        mk(1, 0, 0), mk(0, 1, 0), mk(0, 0, 1), cross(ab, bc),
    };
    for (const V& ax : axes)
        if (!sat(a, b, c, e, ax)) return false;
    return true;
}

struct Builder {
    rtc_grid grid;
    const float* pos;
    uint32_t num_tris;

    // Grid.getCellIdx (src/linalg.zig:424-427)
    void cell_idx(V p, uint32_t out[3]) const
    {
        const float q[3] = { (p.x - grid.bbox_min[0]) / grid.cell_size[0], (p.y - grid.bbox_min[1]) / grid.cell_size[1],
                             (p.z - grid.bbox_min[2]) / grid.cell_size[2] };
        for (int i = 0; i < 3; ++i) out[i] = std::min(f2u(q[i]), grid.resolution[i] - 1u);
    }
    // The triple loop of initCells / initIndices (src/stage2.zig:63-80, 107-124) for one triangle.
    template <typename F>
    void for_cells(uint32_t t, F&& visit) const
    {
        const V p0 = ld(pos + 9 * (size_t)t), p1 = ld(pos + 9 * (size_t)t + 3), p2 = ld(pos + 9 * (size_t)t + 6);
        uint32_t lo[3], hi[3];
        cell_idx(vmin(p0, vmin(p1, p2)), lo);
        cell_idx(vmax(p0, vmax(p1, p2)), hi);
        const uint64_t rx = grid.resolution[0], ry = grid.resolution[1];
        for (uint32_t z = lo[2]; z < hi[2] + 1; ++z)
            for (uint32_t y = lo[1]; y < hi[1] + 1; ++y)
                for (uint32_t x = lo[0]; x < hi[0] + 1; ++x) {
                    // Grid.getCellBbox (src/linalg.zig:433-441)
                    const V bmin = mk(grid.bbox_min[0] + grid.cell_size[0] * (float)x, grid.bbox_min[1] + grid.cell_size[1] * (float)y,
                                      grid.bbox_min[2] + grid.cell_size[2] * (float)z);
                    const V bmax = bmin + mk(grid.cell_size[0], grid.cell_size[1], grid.cell_size[2]);
                    if (tri_box(p0, p1, p2, bmin, bmax)) visit(z * rx * ry + y * rx + x);
                }
    }
};

template <typename F>
void parallel_for(uint64_t n, uint64_t grain, F&& body)
{
    unsigned hw = std::thread::hardware_concurrency();
    if (hw == 0) hw = 1;
    const uint64_t chunks = (n + grain - 1) / grain;
    const unsigned nt = (unsigned)std::min<uint64_t>(hw, chunks ? chunks : 1);
    if (nt <= 1) { body(0, n); return; }
    std::atomic<uint64_t> next(0);
    std::vector<std::thread> th;
    for (unsigned i = 0; i < nt; ++i)
        th.emplace_back([&] {
            for (;;) {
                const uint64_t c = next.fetch_add(1);
                if (c >= chunks) break;
                body(c * grain, std::min(n, (c + 1) * grain));
            }
        });
    for (auto& t : th) t.join();
}

}  // namespace

extern "C" int rtc_geometry_build(const float* positions, const rtc_tri_data* data, uint32_t num_tris,
                                  const uint32_t res[3], rtc_geometry* out)
{
    if (!out) return rtc_set_error_(RTC_ERR_INVALID_ARGUMENT, "rtc_geometry_build: out is null");
    memset(out, 0, sizeof *out);
    if ((num_tris && (!positions || !data)) || !res || !res[0] || !res[1] || !res[2])
        return rtc_set_error_(RTC_ERR_INVALID_ARGUMENT, "rtc_geometry_build: null input or zero grid resolution");
    const uint64_t num_cells = (uint64_t)res[0] * res[1] * res[2];
    if (num_cells >= ((uint64_t)1 << 32)) return rtc_set_error_(RTC_ERR_UNSUPPORTED, "grid has 2^32 cells or more");
    try {
        Builder b;
        b.pos = positions; b.num_tris = num_tris;
        // initGrid (src/stage2.zig:44-57), Grid.init (src/linalg.zig:412-418)
        V bmin = mk(INFINITY, INFINITY, INFINITY), bmax = mk(-INFINITY, -INFINITY, -INFINITY);
        for (uint64_t i = 0; i < (uint64_t)num_tris * 3; ++i) { const V p = ld(positions + 3 * i); bmin = vmin(bmin, p); bmax = vmax(bmax, p); }
        const float lo[3] = { bmin.x, bmin.y, bmin.z }, hi[3] = { bmax.x, bmax.y, bmax.z };
        for (int i = 0; i < 3; ++i) {
            b.grid.bbox_min[i] = lo[i]; b.grid.bbox_max[i] = hi[i]; b.grid.resolution[i] = res[i];
            b.grid.cell_size[i] = (hi[i] - lo[i]) / (float)res[i];
        }
        // initCells: count (stage2.zig:59-81), then exclusive scan + stats (:82-96)
        std::vector<std::atomic<uint32_t>> count(num_cells);
        for (auto& c : count) c.store(0, std::memory_order_relaxed);
        parallel_for(num_tris, 4096, [&](uint64_t t0, uint64_t t1) {
            for (uint64_t t = t0; t < t1; ++t)
                b.for_cells((uint32_t)t, [&](uint64_t cell) { count[cell].fetch_add(1, std::memory_order_relaxed); });
        });
        std::unique_ptr<RtcGeometryOwner> own(new RtcGeometryOwner());
        own->cells.resize(num_cells);
        uint64_t total = 0;
        uint32_t min_t = 0xFFFFFFFFu, max_t = 0, empty = 0;
        for (uint64_t c = 0; c < num_cells; ++c) {
            const uint32_t n = count[c].load(std::memory_order_relaxed);
            if (n) { min_t = std::min(min_t, n); max_t = std::max(max_t, n); } else ++empty;
            own->cells[c].begin = (uint32_t)total;
            own->cells[c].end = (uint32_t)total;       // used as the fill cursor below
            total += n;
            if (total >= ((uint64_t)1 << 32)) { return rtc_set_error_(RTC_ERR_UNSUPPORTED, "more than 2^32 baked triangles"); }
        }
        // initIndices (stage2.zig:104-125): parallel scatter, then restore ascending source order per cell
        std::vector<uint32_t> indices(total);
        std::vector<std::atomic<uint32_t>> cursor(num_cells);
        for (uint64_t c = 0; c < num_cells; ++c) cursor[c].store(own->cells[c].begin, std::memory_order_relaxed);
        parallel_for(num_tris, 4096, [&](uint64_t t0, uint64_t t1) {
            for (uint64_t t = t0; t < t1; ++t)
                b.for_cells((uint32_t)t, [&](uint64_t cell) { indices[cursor[cell].fetch_add(1, std::memory_order_relaxed)] = (uint32_t)t; });
        });
        parallel_for(num_cells, 1 << 14, [&](uint64_t c0, uint64_t c1) {
            for (uint64_t c = c0; c < c1; ++c) {
                own->cells[c].end = cursor[c].load(std::memory_order_relaxed);
                if (own->cells[c].end - own->cells[c].begin > 1)
                    std::sort(indices.begin() + own->cells[c].begin, indices.begin() + own->cells[c].end);
            }
        });
        // bakeInto (stage2.zig:137-164): duplicate triangles in cell order, Triangle.init (linalg.zig:688-694)
        own->tri_pos.resize(total);
        own->tri_data.resize(total);
        parallel_for(total, 1 << 14, [&](uint64_t i0, uint64_t i1) {
            for (uint64_t i = i0; i < i1; ++i) {
                const uint32_t s = indices[i];
                const V p0 = ld(positions + 9 * (size_t)s), p1 = ld(positions + 9 * (size_t)s + 3), p2 = ld(positions + 9 * (size_t)s + 6);
                const V e1 = p1 - p0, e2 = p2 - p0;
                rtc_tri_pos& tp = own->tri_pos[i];
                tp.v0[0] = p0.x; tp.v0[1] = p0.y; tp.v0[2] = p0.z; tp.v0[3] = 0.0f;
                tp.e1[0] = e1.x; tp.e1[1] = e1.y; tp.e1[2] = e1.z; tp.e1[3] = 0.0f;
                tp.e2[0] = e2.x; tp.e2[1] = e2.y; tp.e2[2] = e2.z; tp.e2[3] = 0.0f;
                own->tri_data[i] = data[s];
                own->tri_data[i].src_tri = s;
            }
        });
        out->grid = b.grid;
        out->cells = own->cells.data(); out->num_cells = num_cells;
        out->tri_pos = own->tri_pos.data(); out->tri_data = own->tri_data.data();
        out->num_tris = total; out->num_src_tris = num_tris;
        out->empty_cells = empty; out->min_tris = min_t; out->max_tris = max_t;
        out->owner = own.release();
        return RTC_OK;
    } catch (const std::bad_alloc&) {
        return rtc_set_error_(RTC_ERR_OUT_OF_MEMORY, "rtc_geometry_build: host allocation failed");
    }
}

extern "C" void rtc_geometry_free(rtc_geometry* g)
{
    if (!g) return;
    delete static_cast<RtcGeometryOwner*>(g->owner);
    memset(g, 0, sizeof *g);
}
