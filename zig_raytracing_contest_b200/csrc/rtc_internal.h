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
    struct Alloc { void* p; size_t bytes; };
    std::vector<Alloc> allocs;          // everything cudaMalloc'ed for the scene (sizes kept: replicas are made by copying these)
    unsigned long long* counters = nullptr;
    uint32_t* work = nullptr;
    cudaStream_t stream = nullptr;      // the stream every launch goes to: own_stream, or the host's (rtc_scene_set_stream)
    cudaStream_t own_stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_frame = nullptr;   // start of the render kernels / end of them (+ local resolve) / end of combine + resolve
    uint8_t* d_stage = nullptr;         // rtc_render_sharded: packed strips in flight (3 B / pixel)
    size_t d_stage_bytes = 0;
    // frame scratch (grown on demand)
    uint32_t cap_pixels = 0;
    float4* accum = nullptr;
    uint8_t* d_rgb = nullptr;
    float* d_radiance = nullptr;
    uint32_t* d_hit_id = nullptr;
    float* d_hit_tuv = nullptr;
    uint8_t* h_pinned = nullptr;        // full-frame host staging (rgb | radiance | hit_id | tuv), pinned + portable: every
    size_t h_pinned_bytes = 0;          // device of a multi-GPU render copies its pixels straight to their final place in it
    uint32_t h_pinned_pixels = 0;
    // multi-GPU (rtc_params.num_gpus > 1): replicas of this scene on the other devices, made on first use
    std::vector<rtc_scene*> replicas;   // replicas[k] renders on its own device; owned by this scene
    std::vector<int> multi_devices;     // the device list of the last num_gpus > 1 render (this scene's device first)
    bool peers_enabled = false;
    cudaEvent_t ev_part = nullptr;      // "this device's part of the frame is rendered" (sample sharding: peers wait on it)
    bool is_replica = false;
    uint64_t num_tris = 0;
    uint32_t num_src_tris = 0;
    // wavefront scratch (grown on demand) + kernels launched since the counters were last read
    rtc::WaveScratch wave{};
    std::vector<void*> wave_allocs;
    bool wave_budget_limited = false;
    int wave_stage = 0;                 // budget stage the scratch was sized with (0: first frame, 1: render loop)
    uint64_t render_calls = 0;          // frames (parts) launched on this scene so far
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
        allocs.push_back(Alloc{ p, count ? count * sizeof(T) : sizeof(T) });
        *out = static_cast<T*>(p);
        return RTC_OK;
    }
};

// rtc_api.cu
// Where the parts of a frame land on the host: pointers into ONE pinned, portable full-frame buffer (any may be null).
struct RtcHostFrame {
    uint8_t* rgb = nullptr; float* radiance = nullptr; uint32_t* hit_id = nullptr; float* hit_tuv = nullptr;
};
int  rtc_scene_new_(int device, rtc_scene** out);                  // device checks + empty scene object
void rtc_scene_free_(rtc_scene* s);
void rtc_keep_error_free_(rtc_scene* s);                            // frees the scene, keeps rtc_last_error()
// preset: NULL, or 3 entries per material (base colour, emissive, transparency): non-null = that texture's texels are
// already on the device in the device layout (float4 rgb0 / float) and are used as they are.
int  rtc_scene_upload_materials_(rtc_scene* s, const rtc_material* materials, uint32_t num_materials, const void* const* preset = nullptr);
int  rtc_scene_finish_(rtc_scene* s);                              // counters, stream, events
// packed-cell geometry of a grid resolution (rtc_device.cuh): fills dev.mask/shift, returns the padded block count
size_t rtc_scene_set_packing_(rtc::DevScene& d, const uint32_t res[3]);
// rtc_multi.cu / rtc_api.cu: pieces of rtc_render shared by the one-device, the in-process multi-device and the
// one-process-per-device (rtc_comm) paths
int  rtc_check_render_args_(const rtc_scene* s, const rtc_camera* cam, const rtc_params* p);
int  rtc_ensure_frame_(rtc_scene* s, uint32_t pixels);                       // device frame buffers
int  rtc_ensure_host_frame_(rtc_scene* s, uint32_t pixels, RtcHostFrame* hf); // the pinned host frame (primary scene)
// Renders this call's pixels / samples on s's device and queues the D2H copies of its pixels to their final place in
// `hf` (nothing is copied for null members). Asynchronous on s->stream; the caller synchronises.
int  rtc_render_part_(rtc_scene* s, const rtc_camera* cam, const rtc_params* p, uint32_t total_samples, const RtcHostFrame& hf,
                      bool want_hits, bool resolve_and_fetch, bool clear_counters);
// pinned frame -> caller's pageable buffers for the pixels params select (all of them when tile_world <= 1)
void rtc_copy_out_(const rtc_camera* cam, const rtc_params* p, const RtcHostFrame& hf, uint8_t* rgb_out, rtc_aux* aux, bool want_hits);
int  rtc_read_counters_(rtc_scene* s, rtc_stats* out);
int  rtc_scene_clone_(const rtc_scene* src, int device, rtc_scene** out);     // a replica on another device (peer copies)
int  rtc_render_multi_(rtc_scene* s, const rtc_camera* cam, const rtc_params* p, uint8_t* rgb_out, rtc_aux* aux);
