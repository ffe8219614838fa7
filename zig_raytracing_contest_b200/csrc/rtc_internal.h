// rtc_internal.h — what rtc_api.cu and rtc_build.cu share: the device-resident scene object behind the
// opaque rtc_scene of include/rtc.h, error plumbing, and the scene set-up steps common to both ways a scene
// comes to exist (uploaded baked arrays / built on the device from a triangle soup).
#pragma once
#include "../../include/rtc.h"
#include "rtc_kernels.h"
#include "rtc_bvh.cuh"
#include "rtc_geometry_owner.h"

#include <cuda_runtime.h>

#include <string>
#include <vector>

int rtc_fail_(int code, const char* fmt, ...);     // sets rtc_last_error(), returns code (rtc_api.cu)

#define RTC_CU(call)                                                                              \
    do {                                                                                          \
        cudaError_t e_ = (call);                                                                  \
        if (e_ != cudaSuccess)                                                                    \
            return rtc_fail_(e_ == cudaErrorMemoryAllocation ? RTC_ERR_OUT_OF_MEMORY : RTC_ERR_CUDA, \
                             "%s: %s", #call, cudaGetErrorString(e_));                            \
    } while (0)

struct RtcDeviceGuard {
    int prev = -1;
    explicit RtcDeviceGuard(int dev) { cudaGetDevice(&prev); if (dev != prev) cudaSetDevice(dev); }
    ~RtcDeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// Device-resident scene + per-scene scratch reused across rtc_render calls.
struct rtc_scene {
    int device = 0;
    int sm_count = 0;
    rtc::DevScene dev{};
    std::vector<void*> allocs;          // everything cudaMalloc'ed for the scene
    unsigned long long* counters = nullptr;
    uint32_t* work = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // frame scratch (grown on demand)
    uint32_t cap_pixels = 0;
    float4* accum = nullptr;
    uint8_t* d_rgb = nullptr;
    float* d_radiance = nullptr;
    uint32_t* d_hit_id = nullptr;
    float* d_hit_tuv = nullptr;
    uint8_t* h_pinned = nullptr;        // staging for the D2H copies (rgb | radiance | hit_id | tuv)
    size_t h_pinned_bytes = 0;
    uint64_t num_tris = 0;
    uint32_t num_src_tris = 0;
    // wavefront scratch (grown on demand) + kernels launched since the counters were last read
    rtc::WaveScratch wave{};
    std::vector<void*> wave_allocs;
    bool wave_budget_limited = false;
    uint64_t launches = 0;
    std::vector<cudaEvent_t> timer_events;
    rtc::TraceTimer timer{};
    rtc::LaunchCache cache{};           // per-device launch state of the traversal kernels
    // RTC_ACCEL_BVH (rtc_scene_build_bvh): nodes + BVH-ordered {v0,e1,e2} records, in `allocs`
    bool has_bvh = false;
    rtc::BvhView bvh{};
    rtc_bvh_info bvh_info{};

    template <typename T>
    int alloc(T** out, size_t count)
    {
        void* p = nullptr;
        cudaError_t e = cudaMalloc(&p, count ? count * sizeof(T) : sizeof(T));
        if (e != cudaSuccess)
            return rtc_fail_(RTC_ERR_OUT_OF_MEMORY, "cudaMalloc(%zu bytes): %s", count * sizeof(T), cudaGetErrorString(e));
        allocs.push_back(p);
        *out = static_cast<T*>(p);
        return RTC_OK;
    }
};

// rtc_api.cu
int  rtc_scene_new_(int device, rtc_scene** out);                  // device checks + empty scene object
void rtc_scene_free_(rtc_scene* s);
void rtc_keep_error_free_(rtc_scene* s);                            // frees the scene, keeps rtc_last_error()
// preset: NULL, or 3 entries per material (base colour, emissive, transparency): non-null = that texture's texels are
// already on the device in the device layout (float4 rgb0 / float) and are used as they are.
int  rtc_scene_upload_materials_(rtc_scene* s, const rtc_material* materials, uint32_t num_materials, const void* const* preset = nullptr);
int  rtc_scene_finish_(rtc_scene* s);                              // counters, stream, events
// packed-cell geometry of a grid resolution (rtc_device.cuh): fills dev.mask/shift, returns the padded block count
size_t rtc_scene_set_packing_(rtc::DevScene& d, const uint32_t res[3]);
