// scratch: times rtc_bvh_build_host on a synthetic terrain-like soup (2 * side^2 triangles), host only:
//   g++ -O3 -std=c++17 -o /tmp/bvhtime tools/scratch/bvhtime.cpp zig_raytracing_contest_b200/csrc/rtc_bvh_build.cpp -lpthread
//   RTC_TRACE_BVH=1 /tmp/bvhtime 1581 16        (side, threads; profiles/r2zz_bvhtime_*.txt)
#include "../../zig_raytracing_contest_b200/csrc/rtc_bvh_host.h"
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
int main(int argc, char** argv)
{
    const uint32_t side = argc > 1 ? atoi(argv[1]) : 1581;      // 2 * side^2 triangles
    const int threads = argc > 2 ? atoi(argv[2]) : 8;
    std::vector<float> t; t.reserve((size_t)side * side * 24);
    auto h = [&](float x, float z) { return 3.0f * sinf(x * 0.05f) * cosf(z * 0.07f) + 0.7f * sinf(x * 0.31f + z * 0.17f); };
    for (uint32_t i = 0; i < side; ++i) for (uint32_t j = 0; j < side; ++j) {
        const float x0 = i * 0.1f, z0 = j * 0.1f, x1 = x0 + 0.1f, z1 = z0 + 0.1f;
        const float p[4][3] = { { x0, h(x0, z0), z0 }, { x1, h(x1, z0), z0 }, { x1, h(x1, z1), z1 }, { x0, h(x0, z1), z1 } };
        const int tri[2][3] = { { 0, 1, 2 }, { 0, 2, 3 } };
        for (auto& q : tri) {
            for (int k = 0; k < 3; ++k) t.push_back(p[q[0]][k]); t.push_back(0);
            for (int k = 0; k < 3; ++k) t.push_back(p[q[1]][k] - p[q[0]][k]); t.push_back(0);
            for (int k = 0; k < 3; ++k) t.push_back(p[q[2]][k] - p[q[0]][k]); t.push_back(0);
        }
    }
    const uint32_t n = (uint32_t)(t.size() / 12);
    RtcBvhHost out;
    const auto t0 = std::chrono::steady_clock::now();
    const int rc = rtc_bvh_build_host(t.data(), n, threads, out);
    const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    unsigned long long hsh = 1469598103934665603ull;
    for (uint32_t w : out.nodes) { hsh ^= w; hsh *= 1099511628211ull; }
    for (uint32_t w : out.order) { hsh ^= w; hsh *= 1099511628211ull; }
    printf("rc %d tris %u threads %d secs %.3f nodes %u depth %u hash %016llx\n", rc, n, threads, s, out.num_nodes, out.depth, hsh);
    return 0;
}
