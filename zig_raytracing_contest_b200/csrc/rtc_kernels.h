// rtc_kernels.h — launch entry points of the render kernels. rtc_wavefront.cu is compiled twice,
// once per arithmetic contract (namespace exact: -fmad=false; namespace fast: -fmad=true).
#pragma once
#include "rtc_device.cuh"

namespace rtc {

// CUDA-event brackets around every k_trace launch (the dominant kernel): the roofline's denominator is
// measured live, on the launching stream, per launch. Pairs are read and recycled at rtc_scene_stats().
struct TraceTimer {
    cudaEvent_t* ev;       // 2 * capacity events
    uint32_t capacity;
    uint32_t used;         // pairs recorded since the last read
    uint32_t dropped;      // launches not bracketed because the pool was full
};

// Launch state that belongs to ONE device (occupancy of the traversal kernels there, the shared-memory carveout set
// there): lives in the rtc_scene, which is bound to a device.
struct LaunchCache {
    int trace_per_sm[2] = { 0, 0 };
    size_t trace_smem[2] = { ~(size_t)0, ~(size_t)0 };
    int bvh_per_sm[2] = { 0, 0 };
};

struct LaunchCfg {
    int sm_count;          // SMs of the device the launch goes to
    cudaStream_t stream;
    bool stats;            // instrumented variant (cells/tris counters)
    TraceTimer* timer;     // may be null
    LaunchCache* cache;    // never null
};

struct BvhView;

// the accumulators of the devices of an in-process multi-GPU render, as seen from one of them (peer pointers)
struct PeerAccums {
    const float4* p[16];
    uint32_t n;
};

struct PrimaryOut {
    uint32_t* hit_id;      // num_pixels, source-triangle id or 0xFFFFFFFF (may be null)
    float* hit_tuv;        // 3*num_pixels (may be null)
};

#define RTC_DECLARE_LAUNCHERS(NS)                                                                     \
    namespace NS {                                                                                    \
    /* renderWorker + traceRayRecursive (src/stage3.zig:188-245) as a wavefront: adds                */ \
    /* prm.num_samples samples of every pixel in [first_pixel, first_pixel+num_pixels) into          */ \
    /* accum[pixel] (rgb sum, count). primary: the deterministic slice (centre ray, depth 1:         */ \
    /* hit -> emissive, miss -> sky) + hit buffers. *launches += kernels launched.                   */ \
    cudaError_t launch_render(const DevScene& sc, const DevCamera& cam, const DevParams& prm,         \
                              bool primary, float4* accum, PrimaryOut out,                            \
                              unsigned long long* counters, const WaveScratch& ws,                    \
                              const LaunchCfg& cfg, uint32_t* launches,                               \
                              const BvhView* bvh /* null = RTC_ACCEL_GRID */);                        \
    /* pixel.mul(inv_num_samples).toRGB() (src/stage3.zig:242).                                      */ \
    cudaError_t launch_resolve(const float4* accum, uint32_t num_pixels, uint32_t total_samples,      \
                               uint8_t* rgb, float* radiance, cudaStream_t stream);                   \
    /* sample sharding: sum of the devices' accumulators over peer memory + resolve, pixels           */ \
    /* [first, first+count) (in-process); rgb-only packing and the resolve from it (rtc_comm)         */ \
    cudaError_t launch_combine_resolve(const PeerAccums& pa, uint32_t first, uint32_t count,          \
                                       uint32_t total_samples, uint8_t* rgb, float* radiance,         \
                                       cudaStream_t stream);                                          \
    cudaError_t launch_pack_rgb(const float4* accum, uint32_t num_pixels, float* rgb3,                \
                                cudaStream_t stream);                                                 \
    cudaError_t launch_resolve3(const float* sum3, uint32_t num_pixels, uint32_t total_samples,       \
                                uint8_t* rgb, cudaStream_t stream);                                   \
    }

RTC_DECLARE_LAUNCHERS(exact)
RTC_DECLARE_LAUNCHERS(fast)

}  // namespace rtc
