// rtc_bvh_build.cpp — host builder of the 8-wide quantised BVH that RTC_ACCEL_BVH walks (layout: rtc_bvh.cuh).
//
// The reference has no counterpart: its accelerator is the uniform grid of src/stage2.zig:44-129. This structure is
// built over the same triangles in the form the reference's intersection test consumes, {v0, e1, e2}
// (Triangle.init, src/linalg.zig:688-694), so that its boxes bound exactly what the test sees.
//
//   1. per triangle: box of {v0, v0+e1, v0+e2} evaluated in double and rounded outwards, padded by a few fp32 ulps of
//      the scene size (the slab test in rtc_bvh.cuh works in fp32: (p - o) * (1/d) carries an absolute error of that
//      order), centroid of the box;
//   2. top-down, one 8-wide node at a time: the node's triangle range is split by binned SAH (16 bins, all three
//      axes, object splits; median split when the bins cannot separate) — always the pending child with the largest
//      surface area among those still too big for a leaf, then, while slots remain, the leaf-sized children worth
//      separating — until eight children exist; children with at most three triangles become leaves. Subtrees are
//      built on all host threads, and so is every split of a long range (2^17 triangles or more: the top of the tree):
//      binning, a stable out-of-place partition and the part boxes (taken from the bins) are chunked over the threads.
//      Everything that decides a split is a min / max / count, so the tree does not depend on the number of threads. From depth 28 on splits are medians, which bounds the depth under the traversal
//      stack (kBvhStack);
//   3. breadth-first flattening: the inner children of a node get consecutive node indices, the triangles of its
//      leaf children consecutive slots of the BVH-ordered triangle list; children are assigned to the eight slots
//      greedily by (child centre - node centre) . (+-1, +-1, +-1) so that slot ^ ray octant orders them front to
//      back; child boxes are quantised to 8 bits per bound in the node's frame p + q * 2^e, rounded outwards. Slot
//      assignment and quantisation run on all threads; only the numbering is serial. Triangles of a leaf are stored in
//      ascending source order. RTC_TRACE_BVH=1 prints the phases to stderr.
#include "rtc_bvh_host.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <thread>

namespace {

struct Box {
    float lo[3], hi[3];
    void reset() { for (int a = 0; a < 3; ++a) { lo[a] = INFINITY; hi[a] = -INFINITY; } }
    void grow(const Box& b) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], b.lo[a]); hi[a] = std::max(hi[a], b.hi[a]); } }
    void grow(const float p[3]) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], p[a]); hi[a] = std::max(hi[a], p[a]); } }
    float area() const
    {
        const float x = hi[0] - lo[0], y = hi[1] - lo[1], z = hi[2] - lo[2];
        return (x < 0.0f) ? 0.0f : x * y + y * z + z * x;
    }
};

struct TNode {
    Box box;
    uint32_t first, count;
    int32_t child[8];
    uint32_t nchild;          // 0 = leaf
};

constexpr int kBins = 16;
constexpr uint32_t kLeafMax = 3;
constexpr uint32_t kMedianDepth = 28;

constexpr uint32_t kParallelMin = 1u << 17;       // ranges at least this long are split by several threads

// f(chunk, a, b) over [0, count) cut into `workers` chunks; chunk 0 runs on the calling thread
template <class F>
void parallel_chunks(uint32_t count, int workers, F&& f)
{
    if (workers < 1) workers = 1;
    const uint32_t per = (count + (uint32_t)workers - 1) / (uint32_t)workers;
    std::vector<std::thread> th;
    for (int c = 1; c < workers; ++c) {
        const uint32_t a = std::min<uint64_t>((uint64_t)c * per, count), b = std::min<uint64_t>((uint64_t)a + per, count);
        if (a < b) th.emplace_back([&f, c, a, b] { f(c, a, b); });
    }
    f(0, 0u, std::min(per, count));
    for (auto& t : th) t.join();
}

struct Bins {
    Box bb[3][kBins];
    uint32_t bn[3][kBins];
    void reset() { for (int a = 0; a < 3; ++a) for (int k = 0; k < kBins; ++k) { bb[a][k].reset(); bn[a][k] = 0; } }
    void merge(const Bins& o) { for (int a = 0; a < 3; ++a) for (int k = 0; k < kBins; ++k) { bb[a][k].grow(o.bb[a][k]); bn[a][k] += o.bn[a][k]; } }
};

struct Builder {
    const Box* tbox = nullptr;
    const float (*tcen)[3] = nullptr;
    std::unique_ptr<uint32_t[]> idx;
    std::unique_ptr<uint32_t[]> scratch;      // the out-of-place partition of long ranges (same offsets as idx)
    std::unique_ptr<TNode[]> nodes;
    std::atomic<uint32_t> num_nodes{ 0 };
    std::atomic<int> spare_threads{ 0 };
    std::atomic<uint32_t> max_depth{ 0 };
    int max_threads = 1;
    bool trace = false;
    std::chrono::steady_clock::time_point t_start;

    int grab(int want)
    {
        int got = 0;
        while (got < want) {
            int cur = spare_threads.load();
            if (cur <= 0) break;
            if (spare_threads.compare_exchange_weak(cur, cur - 1)) got += 1;
        }
        return got;
    }

    Box range_box(uint32_t first, uint32_t count) const
    {
        Box b; b.reset();
        for (uint32_t i = first; i < first + count; ++i) b.grow(tbox[idx[i]]);
        return b;
    }

    // Splits idx[first, first+count) in place; returns the size of the left part (0 < left < count) and the boxes of the
    // two parts. Every quantity that decides the split is a min / max / count over the range, so the result does not
    // depend on how many threads worked on it; ranges of kParallelMin or more are partitioned stably (out of place),
    // shorter ones with std::partition — a property of the length alone.
    uint32_t split(uint32_t first, uint32_t count, bool median_only, Box& lbox, Box& rbox)
    {
        const bool big = count >= kParallelMin;
        const int extra = big ? grab(max_threads - 1) : 0;
        const int workers = extra + 1;
        struct Release { Builder* b; int k; ~Release() { if (k) b->spare_threads.fetch_add(k); } } release{ this, extra };
        const uint32_t* ix = idx.get() + first;

        Box cb; cb.reset();
        if (workers == 1) {
            for (uint32_t i = 0; i < count; ++i) cb.grow(tcen[ix[i]]);
        } else {
            std::vector<Box> part((size_t)workers);
            for (auto& b : part) b.reset();
            parallel_chunks(count, workers, [&](int c, uint32_t a, uint32_t b) { Box x; x.reset(); for (uint32_t i = a; i < b; ++i) x.grow(tcen[ix[i]]); part[(size_t)c] = x; });
            for (auto& b : part) cb.grow(b);
        }
        int best_axis = -1, best_plane = 0;
        float best_cost = INFINITY;
        float scale[3] = { 0, 0, 0 };
        bool valid[3] = { false, false, false };
        std::unique_ptr<Bins> bins;
        if (!median_only) {
            for (int a = 0; a < 3; ++a) {
                const float ext = cb.hi[a] - cb.lo[a];
                if (!(ext > 0.0f) || !std::isfinite(ext)) continue;
                scale[a] = (float)kBins / ext; valid[a] = true;
            }
            auto fill = [&](Bins& out, uint32_t a0, uint32_t b0) {
                out.reset();
                for (uint32_t i = a0; i < b0; ++i) {
                    const uint32_t t = ix[i];
                    const Box& tb = tbox[t];
                    for (int a = 0; a < 3; ++a) {
                        if (!valid[a]) continue;
                        int k = (int)((tcen[t][a] - cb.lo[a]) * scale[a]);
                        k = k < 0 ? 0 : (k >= kBins ? kBins - 1 : k);
                        out.bb[a][k].grow(tb); out.bn[a][k] += 1;
                    }
                }
            };
            bins.reset(new Bins);
            if (workers == 1) fill(*bins, 0u, count);
            else {
                std::vector<Bins> part((size_t)workers);
                for (auto& b : part) b.reset();
                parallel_chunks(count, workers, [&](int c, uint32_t a, uint32_t b) { fill(part[(size_t)c], a, b); });
                bins->reset();
                for (auto& b : part) bins->merge(b);
            }
            for (int a = 0; a < 3; ++a) {
                if (!valid[a]) continue;
                const Box* bb = bins->bb[a]; const uint32_t* bn = bins->bn[a];
                float right_area[kBins]; uint32_t right_n[kBins];
                Box acc; acc.reset(); uint32_t n = 0;
                for (int k = kBins - 1; k > 0; --k) { acc.grow(bb[k]); n += bn[k]; right_area[k] = acc.area(); right_n[k] = n; }
                acc.reset(); n = 0;
                for (int k = 0; k < kBins - 1; ++k) {
                    acc.grow(bb[k]); n += bn[k];
                    if (n == 0 || right_n[k + 1] == 0) continue;
                    const float cost = acc.area() * (float)n + right_area[k + 1] * (float)right_n[k + 1];
                    if (cost < best_cost) { best_cost = cost; best_axis = a; best_plane = k; }
                }
            }
        }
        if (best_axis >= 0) {
            const int a = best_axis;
            auto is_left = [&](uint32_t t) {
                int k = (int)((tcen[t][a] - cb.lo[a]) * scale[a]);
                k = k < 0 ? 0 : (k >= kBins ? kBins - 1 : k);
                return k <= best_plane;
            };
            uint32_t* b = idx.get() + first;
            uint32_t left = 0;
            if (!big) {
                left = (uint32_t)(std::partition(b, b + count, is_left) - b);
            } else {
                // stable, out of place: count per chunk, then every chunk writes its lefts and rights to their final places
                uint32_t* out = scratch.get() + first;
                std::vector<uint32_t> nl((size_t)workers, 0u), lo_at((size_t)workers, 0u), hi_at((size_t)workers, 0u), ca((size_t)workers, 0u), cbnd((size_t)workers, 0u);
                parallel_chunks(count, workers, [&](int c, uint32_t x, uint32_t y) {
                    uint32_t n = 0; for (uint32_t i = x; i < y; ++i) n += is_left(b[i]) ? 1u : 0u;
                    nl[(size_t)c] = n; ca[(size_t)c] = x; cbnd[(size_t)c] = y;
                });
                for (int c = 0; c < workers; ++c) left += nl[(size_t)c];
                uint32_t l = 0, r = left;
                for (int c = 0; c < workers; ++c) { lo_at[(size_t)c] = l; hi_at[(size_t)c] = r; l += nl[(size_t)c]; r += (cbnd[(size_t)c] - ca[(size_t)c]) - nl[(size_t)c]; }
                parallel_chunks(count, workers, [&](int c, uint32_t x, uint32_t y) {
                    uint32_t l2 = lo_at[(size_t)c], r2 = hi_at[(size_t)c];
                    for (uint32_t i = x; i < y; ++i) { const uint32_t t = b[i]; if (is_left(t)) out[l2++] = t; else out[r2++] = t; }
                });
                parallel_chunks(count, workers, [&](int, uint32_t x, uint32_t y) { memcpy(b + x, out + x, (size_t)(y - x) * sizeof(uint32_t)); });
            }
            if (left > 0 && left < count) {
                lbox.reset(); rbox.reset();
                for (int k = 0; k < kBins; ++k) (k <= best_plane ? lbox : rbox).grow(bins->bb[a][k]);
                return left;
            }
        }
        // median along the widest centroid axis (ties / identical centroids: by index)
        int a = 0;
        for (int k = 1; k < 3; ++k) if (cb.hi[k] - cb.lo[k] > cb.hi[a] - cb.lo[a]) a = k;
        uint32_t* b = idx.get() + first;
        std::nth_element(b, b + count / 2, b + count, [&](uint32_t x, uint32_t y) {
            return tcen[x][a] < tcen[y][a] || (tcen[x][a] == tcen[y][a] && x < y);
        });
        lbox = range_box(first, count / 2); rbox = range_box(first + count / 2, count - count / 2);
        return count / 2;
    }

    void build_node(uint32_t ni, uint32_t depth)
    {
        uint32_t d0 = max_depth.load(std::memory_order_relaxed);
        while (depth > d0 && !max_depth.compare_exchange_weak(d0, depth, std::memory_order_relaxed)) {}
        struct Part { uint32_t first, count; Box box; };
        Part parts[8];
        uint32_t np = 1;
        parts[0] = { nodes[ni].first, nodes[ni].count, nodes[ni].box };
        const bool median_only = depth >= kMedianDepth;
        while (np < 8) {
            // the pending part most worth splitting: parts too big for a leaf first (largest area), then leaf-sized ones
            // (at half weight: separating two triangles of a leaf saves a test, not a subtree)
            int pick = -1; float key = -1.0f; bool pick_big = false;
            for (uint32_t k = 0; k < np; ++k) {
                if (parts[k].count < 2u) continue;
                const bool big = parts[k].count > kLeafMax;
                const float v = parts[k].box.area() * (big ? 1.0f : 0.5f);
                if ((big && !pick_big) || (big == pick_big && v > key)) { pick = (int)k; key = v; pick_big = big; }
            }
            if (pick < 0) break;
            const Part p = parts[pick];
            Box lb, rb;
            const uint32_t left = split(p.first, p.count, median_only, lb, rb);
            parts[pick] = { p.first, left, lb };
            parts[np++] = { p.first + left, p.count - left, rb };
        }
        if (trace && depth <= 2)
            fprintf(stderr, "[bvh]   +%.0f ms: depth %u node of %u triangles split into %u parts\n",
                    std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_start).count(), depth, nodes[ni].count, np);
        TNode& n = nodes[ni];
        n.nchild = np;
        const uint32_t base = num_nodes.fetch_add(np);
        for (uint32_t k = 0; k < np; ++k) {
            TNode& c = nodes[base + k];
            c.box = parts[k].box; c.first = parts[k].first; c.count = parts[k].count; c.nchild = 0;
            n.child[k] = (int32_t)(base + k);
        }
        std::vector<std::thread> spawned;
        for (uint32_t k = 0; k < np; ++k) {
            if (parts[k].count <= kLeafMax) continue;
            const uint32_t ci = base + k;
            if (parts[k].count > 16384 && spare_threads.fetch_sub(1) > 0) {
                spawned.emplace_back([this, ci, depth] { build_node(ci, depth + 1); spare_threads.fetch_add(1); });
            } else {
                if (parts[k].count > 16384) spare_threads.fetch_add(1);
                build_node(ci, depth + 1);
            }
        }
        for (auto& t : spawned) t.join();
    }
};

inline float next_down(float f) { return std::nextafter(f, -INFINITY); }
inline float next_up(float f) { return std::nextafter(f, INFINITY); }

}  // namespace

int rtc_bvh_build_host(const float* tris, uint32_t n, int threads, RtcBvhHost& out)
{
    out.nodes.clear(); out.order.clear(); out.depth = 0; out.num_nodes = 0;
    if (n == 0) {
        // one empty root: every slot empty, nothing is ever hit
        out.nodes.assign(24, 0u);
        const float one = 1.0f; uint32_t e; memcpy(&e, &one, 4);
        out.nodes[3] = (e >> 23) * 0x00010101u;
        out.num_nodes = 1;
        return 0;
    }
    if (threads < 1) threads = (int)std::thread::hardware_concurrency();
    if (threads < 1) threads = 1;
    const bool trace = getenv("RTC_TRACE_BVH") != nullptr;        // host phases to stderr
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto ms_since = [&](std::chrono::steady_clock::time_point t) { return std::chrono::duration<double, std::milli>(now() - t).count(); };
    auto t_phase = now();

    // 1. boxes and centroids
    // (no value-initialisation: the arrays are first touched by the threads that fill them)
    std::unique_ptr<Box[]> tbox(new Box[n]);
    std::unique_ptr<float[]> tcen(new float[(size_t)n * 3]);
    Box scene; scene.reset();
    {
        auto work = [&](uint32_t a, uint32_t b) {
            for (uint32_t i = a; i < b; ++i) {
                const float* t = tris + 12 * (size_t)i;
                Box bx;
                for (int k = 0; k < 3; ++k) {
                    const double v0 = t[k], v1 = (double)t[k] + (double)t[4 + k], v2 = (double)t[k] + (double)t[8 + k];
                    double lo = std::min(v0, std::min(v1, v2)), hi = std::max(v0, std::max(v1, v2));
                    if (!(lo == lo) || !(hi == hi)) { lo = 0.0; hi = 0.0; }       // NaN edges (degenerate input): never hit anyway
                    float flo = (float)lo, fhi = (float)hi;
                    if ((double)flo > lo) flo = next_down(flo);
                    if ((double)fhi < hi) fhi = next_up(fhi);
                    bx.lo[k] = std::max(flo, -3.0e38f); bx.hi[k] = std::min(fhi, 3.0e38f);
                }
                tbox[i] = bx;
            }
        };
        std::vector<Box> part((size_t)threads);
        for (auto& b : part) b.reset();
        parallel_chunks(n, threads, [&](int c, uint32_t a, uint32_t b) {
            work(a, b);
            Box x; x.reset();
            for (uint32_t i = a; i < b; ++i) x.grow(tbox[i]);
            part[(size_t)c] = x;
        });
        for (auto& b : part) scene.grow(b);
        if (trace) fprintf(stderr, "[bvh]   boxes %.1f ms\n", ms_since(t_phase));
        float size = 0.0f;
        for (int k = 0; k < 3; ++k) size = std::max(size, std::max(std::fabs(scene.lo[k]), std::fabs(scene.hi[k])));
        const float pad = size * 9.5367431640625e-07f;     // 2^-20 of the largest coordinate: ~8 fp32 ulps there
        parallel_chunks(n, threads, [&](int, uint32_t a, uint32_t b) {
            for (uint32_t i = a; i < b; ++i)
                for (int k = 0; k < 3; ++k) {
                    tbox[i].lo[k] -= pad; tbox[i].hi[k] += pad;
                    tcen[3 * (size_t)i + k] = 0.5f * tbox[i].lo[k] + 0.5f * tbox[i].hi[k];
                }
        });
        for (int k = 0; k < 3; ++k) { scene.lo[k] -= pad; scene.hi[k] += pad; }
    }

    if (trace) { fprintf(stderr, "[bvh] boxes + centroids %.1f ms\n", ms_since(t_phase)); t_phase = now(); }

    // 2. the tree
    Builder B;
    B.tbox = tbox.get();
    B.tcen = reinterpret_cast<const float (*)[3]>(tcen.get());
    B.idx.reset(new uint32_t[n]);
    if (n >= kParallelMin) B.scratch.reset(new uint32_t[n]);
    B.max_threads = threads;
    parallel_chunks(n, threads, [&](int, uint32_t a, uint32_t b) { for (uint32_t i = a; i < b; ++i) B.idx[i] = i; });
    B.nodes.reset(new TNode[(size_t)2 * n + 16]);
    B.num_nodes = 1;
    B.spare_threads = threads - 1;
    B.trace = trace; B.t_start = now();
    B.nodes[0].box = scene; B.nodes[0].first = 0; B.nodes[0].count = n; B.nodes[0].nchild = 0;
    // the root is always an inner node (a scene of <= 3 triangles: one inner node with one leaf child)
    if (n <= kLeafMax) {
        B.nodes[0].nchild = 1; B.nodes[0].child[0] = 1;
        B.nodes[1] = B.nodes[0]; B.nodes[1].nchild = 0;
        B.num_nodes = 2;
    } else {
        B.build_node(0, 1);
    }
    out.depth = std::max(1u, B.max_depth.load());
    if (trace) { fprintf(stderr, "[bvh] tree (%d threads) %.1f ms\n", threads, ms_since(t_phase)); t_phase = now(); }

    // 3. flatten. (a) slot assignment per inner node, on all threads; (b) breadth-first numbering, serial and light: the inner
    // children of a node get consecutive indices in slot order, its leaves' triangles consecutive places in the triangle order
    // (ascending source index inside a leaf: independent of how the ranges were permuted); (c) quantisation, on all threads.
    const uint32_t tn_count = B.num_nodes.load();
    // per tree node and slot: index into child[] (0xFF = empty), and what sits there (0 = empty, 1..3 = leaf of so many triangles, 0xFF = inner)
    std::unique_ptr<uint8_t[]> child_in_all(new uint8_t[(size_t)tn_count * 8]), kind_all(new uint8_t[(size_t)tn_count * 8]);
    std::atomic<int> err{ 0 };
    parallel_chunks(tn_count, threads, [&](int, uint32_t a0, uint32_t b0) {
        for (uint32_t ti = a0; ti < b0; ++ti) {
            const TNode& tn = B.nodes[ti];
            memset(&child_in_all[(size_t)ti * 8], 0xFF, 8); memset(&kind_all[(size_t)ti * 8], 0, 8);
            if (tn.nchild == 0) continue;
            float cen[3];
            for (int k = 0; k < 3; ++k) cen[k] = 0.5f * tn.box.lo[k] + 0.5f * tn.box.hi[k];
            bool slot_used[8] = { false, false, false, false, false, false, false, false };
            bool child_done[8] = { false, false, false, false, false, false, false, false };
            float cost[8][8];
            for (uint32_t c = 0; c < tn.nchild; ++c) {
                const Box& cb = B.nodes[tn.child[c]].box;
                float v[3];
                for (int k = 0; k < 3; ++k) v[k] = (0.5f * cb.lo[k] + 0.5f * cb.hi[k]) - cen[k];
                for (int sl = 0; sl < 8; ++sl)
                    cost[c][sl] = ((sl & 1) ? v[0] : -v[0]) + ((sl & 2) ? v[1] : -v[1]) + ((sl & 4) ? v[2] : -v[2]);
            }
            for (uint32_t r = 0; r < tn.nchild; ++r) {
                int bc = -1, bs = -1; float bv = -INFINITY;
                for (uint32_t c = 0; c < tn.nchild; ++c) {
                    if (child_done[c]) continue;
                    for (int sl = 0; sl < 8; ++sl)
                        if (!slot_used[sl] && (bc < 0 || cost[c][sl] > bv)) { bc = (int)c; bs = sl; bv = cost[c][sl]; }
                }
                child_done[bc] = true; slot_used[bs] = true;
                child_in_all[(size_t)ti * 8 + (size_t)bs] = (uint8_t)bc;
                const TNode& cn = B.nodes[tn.child[bc]];
                if (cn.nchild == 0 && (cn.count < 1 || cn.count > kLeafMax)) err.store(-2);
                kind_all[(size_t)ti * 8 + (size_t)bs] = cn.nchild != 0 ? (uint8_t)0xFF : (uint8_t)cn.count;
            }
        }
    });
    if (err.load() != 0) return err.load();
    if (trace) { fprintf(stderr, "[bvh] slot assignment %.1f ms\n", ms_since(t_phase)); t_phase = now(); }

    std::vector<uint32_t> o_tnode, o_child_base, o_tri_base;             // per OUTPUT node (= position in the breadth-first queue)
    o_tnode.reserve(tn_count / 4 + 16); o_child_base.reserve(tn_count / 4 + 16); o_tri_base.reserve(tn_count / 4 + 16);
    o_tnode.push_back(0u);
    uint32_t next_node = 1, next_tri = 0;
    for (size_t qi = 0; qi < o_tnode.size(); ++qi) {
        const uint32_t ti = o_tnode[qi];
        const uint8_t* child_in = &child_in_all[(size_t)ti * 8];
        const uint8_t* kind = &kind_all[(size_t)ti * 8];
        o_child_base.push_back(next_node);
        o_tri_base.push_back(next_tri);
        uint32_t n_tris = 0;
        for (int sl = 0; sl < 8; ++sl) {
            if (kind[sl] == 0xFF) { o_tnode.push_back((uint32_t)B.nodes[ti].child[child_in[sl]]); next_node += 1; }
            else n_tris += kind[sl];
        }
        if (n_tris > 24) return -2;
        next_tri += n_tris;
    }
    if (next_tri != n) return -3;
    out.order.resize(n);
    if (trace) { fprintf(stderr, "[bvh] breadth-first numbering %.1f ms\n", ms_since(t_phase)); t_phase = now(); }

    out.nodes.assign(24 * (size_t)next_node, 0u);
    parallel_chunks(next_node, threads, [&](int, uint32_t a0, uint32_t b0) {
        for (uint32_t oi = a0; oi < b0; ++oi) {
            const TNode& tn = B.nodes[o_tnode[oi]];
            const uint8_t* child_in = &child_in_all[(size_t)o_tnode[oi] * 8];
            // quantisation frame
            uint32_t w[20];
            memset(w, 0, sizeof w);
            float p[3]; uint32_t ebits[3]; double inv_scale[3];
            for (int k = 0; k < 3; ++k) {
                p[k] = tn.box.lo[k];
                const double ext = (double)tn.box.hi[k] - (double)p[k];
                int e = -126;
                if (ext > 0.0) { e = (int)std::ceil(std::log2(ext / 255.0)); if (e < -126) e = -126; }
                while (std::ldexp(255.0, e) < ext) e += 1;
                if (e > 127) { err.store(-1); e = 127; }
                ebits[k] = (uint32_t)(e + 127);
                inv_scale[k] = std::ldexp(1.0, -e);
            }
            memcpy(&w[0], p, 12);
            uint32_t imask = 0, n_tris = 0;
            uint8_t meta[8] = { 0 }, q[6][8];
            memset(q, 0, sizeof q);
            for (int sl = 0; sl < 8; ++sl) {
                if (child_in[sl] == 0xFF) continue;
                const TNode& cn = B.nodes[(uint32_t)tn.child[child_in[sl]]];
                for (int k = 0; k < 3; ++k) {
                    double lo = std::floor(((double)cn.box.lo[k] - (double)p[k]) * inv_scale[k]);
                    double hi = std::ceil(((double)cn.box.hi[k] - (double)p[k]) * inv_scale[k]);
                    lo = lo < 0.0 ? 0.0 : (lo > 255.0 ? 255.0 : lo);
                    hi = hi < 0.0 ? 0.0 : (hi > 255.0 ? 255.0 : hi);
                    q[k][sl] = (uint8_t)lo; q[3 + k][sl] = (uint8_t)hi;
                }
                if (cn.nchild != 0) {
                    imask |= 1u << sl;
                    meta[sl] = (uint8_t)(0x20u | (24u + (uint32_t)sl));
                } else {
                    meta[sl] = (uint8_t)((((1u << cn.count) - 1u) << 5) | n_tris);
                    uint32_t ids[kLeafMax];
                    for (uint32_t i = 0; i < cn.count; ++i) ids[i] = B.idx[cn.first + i];
                    std::sort(ids, ids + cn.count);
                    for (uint32_t i = 0; i < cn.count; ++i) out.order[(size_t)o_tri_base[oi] + n_tris + i] = ids[i];
                    n_tris += cn.count;
                }
            }
            w[3] = ebits[0] | (ebits[1] << 8) | (ebits[2] << 16) | (imask << 24);
            w[4] = o_child_base[oi]; w[5] = o_tri_base[oi];
            memcpy(&w[6], meta, 8);
            memcpy(&w[8], q[0], 8); memcpy(&w[10], q[1], 8); memcpy(&w[12], q[2], 8);
            memcpy(&w[14], q[3], 8); memcpy(&w[16], q[4], 8); memcpy(&w[18], q[5], 8);
            memcpy(&out.nodes[24 * (size_t)oi], w, sizeof w);             // 80 bytes of content at a 96-byte stride
        }
    });
    if (err.load() != 0) return err.load();
    out.num_nodes = next_node;
    if (trace) fprintf(stderr, "[bvh] quantisation %.1f ms (%u nodes)\n", ms_since(t_phase), next_node);
    if (out.order.size() != n) return -3;
    return 0;
}
