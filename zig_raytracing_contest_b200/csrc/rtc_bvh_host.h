// rtc_bvh_host.h — interface of the host BVH builder (rtc_bvh_build.cpp) used by rtc_api.cu and by the host-side
// checker tests/native/bvh_check.cu.
#pragma once
#include <cstdint>
#include <vector>

struct RtcBvhHost {
    std::vector<uint32_t> nodes;     // 24 words per node (80 bytes of content at a 96-byte stride), layout in rtc_bvh.cuh; node 0 = root
    std::vector<uint32_t> order;     // BVH triangle slot -> index of the input triangle
    uint32_t num_nodes = 0;
    uint32_t depth = 0;              // levels of inner nodes (bounds the traversal stack)
};

// tris: n records of 12 floats {v0.xyz, _, e1.xyz, _, e2.xyz, _} (rtc_tri_pos). threads < 1 = all host threads.
// Returns 0, or a negative code on inputs the layout cannot hold (coordinates beyond 2^127).
int rtc_bvh_build_host(const float* tris, uint32_t n, int threads, RtcBvhHost& out);
