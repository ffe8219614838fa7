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
//      built on all host threads. From depth 28 on splits are medians, which bounds the depth under the traversal
//      stack (kBvhStack);
//   3. breadth-first flattening: the inner children of a node get consecutive node indices, the triangles of its
//      leaf children consecutive slots of the BVH-ordered triangle list; children are assigned to the eight slots
//      greedily by (child centre - node centre) . (+-1, +-1, +-1) so that slot ^ ray octant orders them front to
//      back; child boxes are quantised to 8 bits per bound in the node's frame p + q * 2^e, rounded outwards.
#include "rtc_bvh_host.h"

#include <algorithm>
#include <atomic>
#include <cmath>
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

struct Builder {
    const Box* tbox = nullptr;
    const float (*tcen)[3] = nullptr;
    std::vector<uint32_t> idx;
    std::unique_ptr<TNode[]> nodes;
    std::atomic<uint32_t> num_nodes{ 0 };
    std::atomic<int> spare_threads{ 0 };
    std::atomic<uint32_t> max_depth{ 0 };

    Box range_box(uint32_t first, uint32_t count) const
    {
        Box b; b.reset();
        for (uint32_t i = first; i < first + count; ++i) b.grow(tbox[idx[i]]);
        return b;
    }

    // Splits idx[first, first+count) in place; returns the size of the left part (0 < left < count).
    uint32_t split(uint32_t first, uint32_t count, bool median_only)
    {
        Box cb; cb.reset();
        for (uint32_t i = first; i < first + count; ++i) cb.grow(tcen[idx[i]]);
        int best_axis = -1, best_plane = 0;
        float best_cost = INFINITY;
        float scale[3] = { 0, 0, 0 };
        if (!median_only) {
            for (int a = 0; a < 3; ++a) {
                const float ext = cb.hi[a] - cb.lo[a];
                if (!(ext > 0.0f) || !std::isfinite(ext)) continue;
                scale[a] = (float)kBins / ext;
                Box bb[kBins]; uint32_t bn[kBins];
                for (int k = 0; k < kBins; ++k) { bb[k].reset(); bn[k] = 0; }
                for (uint32_t i = first; i < first + count; ++i) {
                    const uint32_t t = idx[i];
                    int k = (int)((tcen[t][a] - cb.lo[a]) * scale[a]);
                    k = k < 0 ? 0 : (k >= kBins ? kBins - 1 : k);
                    bb[k].grow(tbox[t]); bn[k] += 1;
                }
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
            uint32_t* b = idx.data() + first;
            uint32_t* m = std::partition(b, b + count, [&](uint32_t t) {
                int k = (int)((tcen[t][a] - cb.lo[a]) * scale[a]);
                k = k < 0 ? 0 : (k >= kBins ? kBins - 1 : k);
                return k <= best_plane;
            });
            const uint32_t left = (uint32_t)(m - b);
            if (left > 0 && left < count) return left;
        }
        // median along the widest centroid axis (ties / identical centroids: by index)
        int a = 0;
        for (int k = 1; k < 3; ++k) if (cb.hi[k] - cb.lo[k] > cb.hi[a] - cb.lo[a]) a = k;
        uint32_t* b = idx.data() + first;
        std::nth_element(b, b + count / 2, b + count, [&](uint32_t x, uint32_t y) {
            return tcen[x][a] < tcen[y][a] || (tcen[x][a] == tcen[y][a] && x < y);
        });
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
            int pick = -1; float key = -1.0f; bool pick_big = false;
            static const int min_split = getenv("RTC_BVH_MINSPLIT") ? atoi(getenv("RTC_BVH_MINSPLIT")) : 2;
            static const float small_w = getenv("RTC_BVH_SMALLW") ? (float)atof(getenv("RTC_BVH_SMALLW")) : 0.5f;
            for (uint32_t k = 0; k < np; ++k) {
                if ((int)parts[k].count < min_split) continue;
                const bool big = parts[k].count > kLeafMax;
                const float v = parts[k].box.area() * (big ? 1.0f : small_w);
                if ((big && !pick_big) || (big == pick_big && v > key)) { pick = (int)k; key = v; pick_big = big; }
            }
            if (pick < 0) break;
            const Part p = parts[pick];
            const uint32_t left = split(p.first, p.count, median_only);
            parts[pick] = { p.first, left, range_box(p.first, left) };
            parts[np++] = { p.first + left, p.count - left, range_box(p.first + left, p.count - left) };
        }
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

    // 1. boxes and centroids
    std::vector<Box> tbox(n);
    std::vector<float> tcen((size_t)n * 3);
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
        std::vector<std::thread> th;
        const uint32_t per = (n + threads - 1) / threads;
        for (int k = 0; k < threads; ++k) {
            const uint32_t a = std::min<uint64_t>((uint64_t)k * per, n), b = std::min<uint64_t>((uint64_t)a + per, n);
            if (a < b) th.emplace_back(work, a, b);
        }
        for (auto& t : th) t.join();
        for (uint32_t i = 0; i < n; ++i) scene.grow(tbox[i]);
        float size = 0.0f;
        for (int k = 0; k < 3; ++k) size = std::max(size, std::max(std::fabs(scene.lo[k]), std::fabs(scene.hi[k])));
        const float pad = size * 9.5367431640625e-07f;     // 2^-20 of the largest coordinate: ~8 fp32 ulps there
        for (uint32_t i = 0; i < n; ++i)
            for (int k = 0; k < 3; ++k) {
                tbox[i].lo[k] -= pad; tbox[i].hi[k] += pad;
                tcen[3 * (size_t)i + k] = 0.5f * tbox[i].lo[k] + 0.5f * tbox[i].hi[k];
            }
        for (int k = 0; k < 3; ++k) { scene.lo[k] -= pad; scene.hi[k] += pad; }
    }

    // 2. the tree
    Builder B;
    B.tbox = tbox.data();
    B.tcen = reinterpret_cast<const float (*)[3]>(tcen.data());
    B.idx.resize(n);
    for (uint32_t i = 0; i < n; ++i) B.idx[i] = i;
    B.nodes.reset(new TNode[(size_t)2 * n + 16]);
    B.num_nodes = 1;
    B.spare_threads = threads - 1;
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

    // 3. flatten
    struct Item { uint32_t tnode, onode; };
    std::vector<Item> queue;
    queue.push_back({ 0u, 0u });
    uint32_t next_node = 1;
    out.nodes.assign(24, 0u);
    out.order.reserve(n);
    for (size_t qi = 0; qi < queue.size(); ++qi) {
        const Item it = queue[qi];
        const TNode& tn = B.nodes[it.tnode];
        // slot assignment
        float cen[3];
        for (int k = 0; k < 3; ++k) cen[k] = 0.5f * tn.box.lo[k] + 0.5f * tn.box.hi[k];
        int slot_of[8]; bool slot_used[8] = { false, false, false, false, false, false, false, false };
        bool child_done[8] = { false, false, false, false, false, false, false, false };
        float cost[8][8];
        for (uint32_t c = 0; c < tn.nchild; ++c) {
            const Box& cb = B.nodes[tn.child[c]].box;
            float v[3];
            for (int k = 0; k < 3; ++k) v[k] = (0.5f * cb.lo[k] + 0.5f * cb.hi[k]) - cen[k];
            for (int s = 0; s < 8; ++s)
                cost[c][s] = ((s & 1) ? v[0] : -v[0]) + ((s & 2) ? v[1] : -v[1]) + ((s & 4) ? v[2] : -v[2]);
        }
        for (uint32_t r = 0; r < tn.nchild; ++r) {
            int bc = -1, bs = -1; float bv = -INFINITY;
            for (uint32_t c = 0; c < tn.nchild; ++c) {
                if (child_done[c]) continue;
                for (int s = 0; s < 8; ++s)
                    if (!slot_used[s] && (bc < 0 || cost[c][s] > bv)) { bc = (int)c; bs = s; bv = cost[c][s]; }
            }
            child_done[bc] = true; slot_used[bs] = true; slot_of[bc] = bs;
        }
        int child_in[8];
        for (int s = 0; s < 8; ++s) child_in[s] = -1;
        for (uint32_t c = 0; c < tn.nchild; ++c) child_in[slot_of[c]] = (int)c;

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
            if (e > 127) return -1;
            ebits[k] = (uint32_t)(e + 127);
            inv_scale[k] = std::ldexp(1.0, -e);
        }
        memcpy(&w[0], p, 12);
        uint32_t imask = 0, n_inner = 0, n_tris = 0;
        uint8_t meta[8] = { 0 }, q[6][8];
        memset(q, 0, sizeof q);
        const uint32_t child_base = next_node, tri_base = (uint32_t)out.order.size();
        for (int s = 0; s < 8; ++s) {
            if (child_in[s] < 0) continue;
            const uint32_t ci = (uint32_t)tn.child[child_in[s]];
            const TNode& cn = B.nodes[ci];
            for (int k = 0; k < 3; ++k) {
                double lo = std::floor(((double)cn.box.lo[k] - (double)p[k]) * inv_scale[k]);
                double hi = std::ceil(((double)cn.box.hi[k] - (double)p[k]) * inv_scale[k]);
                lo = lo < 0.0 ? 0.0 : (lo > 255.0 ? 255.0 : lo);
                hi = hi < 0.0 ? 0.0 : (hi > 255.0 ? 255.0 : hi);
                q[k][s] = (uint8_t)lo; q[3 + k][s] = (uint8_t)hi;
            }
            if (cn.nchild != 0) {
                imask |= 1u << s;
                meta[s] = (uint8_t)(0x20u | (24u + (uint32_t)s));
                queue.push_back({ ci, child_base + n_inner });
                n_inner += 1;
            } else {
                if (cn.count < 1 || cn.count > kLeafMax || n_tris + cn.count > 24) return -2;
                meta[s] = (uint8_t)((((1u << cn.count) - 1u) << 5) | n_tris);
                for (uint32_t i = 0; i < cn.count; ++i) out.order.push_back(B.idx[cn.first + i]);
                n_tris += cn.count;
            }
        }
        next_node += n_inner;
        w[3] = ebits[0] | (ebits[1] << 8) | (ebits[2] << 16) | (imask << 24);
        w[4] = child_base; w[5] = tri_base;
        memcpy(&w[6], meta, 8);
        memcpy(&w[8], q[0], 8); memcpy(&w[10], q[1], 8); memcpy(&w[12], q[2], 8);
        memcpy(&w[14], q[3], 8); memcpy(&w[16], q[4], 8); memcpy(&w[18], q[5], 8);
        if (out.nodes.size() < 24 * (size_t)next_node) out.nodes.resize(24 * (size_t)next_node, 0u);
        memcpy(&out.nodes[24 * (size_t)it.onode], w, sizeof w);       // 80 bytes of content at a 96-byte stride
    }
    out.num_nodes = next_node;
    if (out.order.size() != n) return -3;
    return 0;
}
