// TEST INFRASTRUCTURE (host only, built by tests/test_bvh_host.py with nvcc): builds the product's 8-wide BVH
// (csrc/rtc_bvh_build.cpp) over a triangle list and walks rays through it with the product's own traversal functions
// (csrc/rtc_bvh.cuh compiled for the host), next to a brute-force loop over ALL triangles with the same triangle test
// and the same acceptance rule (smallest t > 0, exact ties to the smaller source id). Any difference is a box that
// failed to bound its triangle or a traversal bug.
#include "../../zig_raytracing_contest_b200/csrc/rtc_bvh.cuh"
#include "../../zig_raytracing_contest_b200/csrc/rtc_bvh_host.h"

#include <cstdio>
#include <vector>

using namespace rtc;

extern "C" {

// tris: n x 12 floats {v0 _ e1 _ e2 _}; rays: m x 6 floats {o, d}. out_tri[m] = input triangle index or -1, out_t[m].
// Returns the number of rays whose BVH answer differs from the brute-force answer; stats = {nodes, depth, node
// visits, triangle tests, max stack}.
int bvh_check(const float* tris, uint32_t n, const float* rays, uint32_t m, int threads, int32_t* out_tri, float* out_t,
              int32_t* brute_tri, float* brute_t, uint64_t* stats)
{
    RtcBvhHost h;
    const int rc = rtc_bvh_build_host(tris, n, threads, h);
    if (rc != 0) return -1000 + rc;
    // BVH-ordered triangle records with the input index as the source id
    std::vector<float4> bt(kBvhTriQuads * (size_t)n + kBvhTriQuads);
    for (uint32_t k = 0; k < n; ++k) {
        const uint32_t i = h.order[k];
        const float* t = tris + 12 * (size_t)i;
        float w; memcpy(&w, &i, 4);
        bt[kBvhTriQuads * (size_t)k] = make_float4(t[0], t[1], t[2], w);
        bt[kBvhTriQuads * (size_t)k + 1] = make_float4(t[4], t[5], t[6], 0.0f);
        bt[kBvhTriQuads * (size_t)k + 2] = make_float4(t[8], t[9], t[10], 0.0f);
    }
    // every input triangle exactly once
    {
        std::vector<uint8_t> seen(n, 0);
        for (uint32_t k = 0; k < n; ++k) { if (h.order[k] >= n || seen[h.order[k]]) return -2000; seen[h.order[k]] = 1; }
    }
    BvhView view; view.num_nodes = h.num_nodes; view.nodes = reinterpret_cast<const float4*>(h.nodes.data()); view.tris = bt.data();
    uint64_t visits = 0, tests = 0, max_sp = 0;
    int bad = 0;
    for (uint32_t r = 0; r < m; ++r) {
        Ray ray; ray.o = mk(rays[6 * r], rays[6 * r + 1], rays[6 * r + 2]); ray.d = mk(rays[6 * r + 3], rays[6 * r + 4], rays[6 * r + 5]);
        BvhRay br; bvh_ray_init(br, ray.o, ray.d);
        BvhTrav s; bvh_trav_begin(s);
        uint2 stack[kBvhStack];
        uint32_t nv = 0, nt = 0;
        for (;;) {
            const bool done = bvh_step<true>(view, br, ray, s, stack, 1u, nv, nt);
            if (s.sp > max_sp) max_sp = s.sp;
            if (s.sp >= kBvhStack) return -3000;
            if (done) break;
        }
        visits += nv; tests += nt;
        out_t[r] = s.best_t;
        out_tri[r] = s.best_tri == 0xFFFFFFFFu ? -1 : (int32_t)h.order[s.best_tri];
        // brute force over the input order (ascending source id: a strict compare keeps the smaller id on ties)
        float bt_best = f_inf(); int32_t bi = -1;
        for (uint32_t i = 0; i < n; ++i) {
            const float* t = tris + 12 * (size_t)i;
            float tt, u, v;
            if (tri_ray(mk(t[0], t[1], t[2]), mk(t[4], t[5], t[6]), mk(t[8], t[9], t[10]), ray, tt, u, v) && tt > 0.0f && bt_best > tt) {
                bt_best = tt; bi = (int32_t)i;
            }
        }
        brute_t[r] = bt_best; brute_tri[r] = bi;
        if (bi != out_tri[r] || f_bits(bt_best) != f_bits(out_t[r])) bad += 1;
    }
    stats[0] = h.num_nodes; stats[1] = h.depth; stats[2] = visits; stats[3] = tests; stats[4] = max_sp;
    return bad;
}

}  // extern "C"
