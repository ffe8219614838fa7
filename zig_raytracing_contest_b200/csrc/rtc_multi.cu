// rtc_multi.cu — one frame on several GPUs behind the C ABI (SURVEY §8(b) num_gpus / shard, §8(e)).
//
// The reference's caller is ONE blocking call, `try scene.render(threads, camera, img)` (src/main.zig:126), and its
// own parallel split is contiguous pixel ranges per worker with no communication (src/stage3.zig:228-233,247-255).
// Two hosts are served:
//   * a single process (the Zig program, bin/rtc, any C host): rtc_render with params.num_gpus = N. The scene is
//     replicated onto the other devices by peer copies on first use; one host thread per device issues that device's
//     launches. STRIPS: every device renders, resolves and copies ITS strips straight to their final place in one
//     pinned, portable host frame — no inter-GPU traffic at all. SAMPLES: every device accumulates its sample range,
//     then device g sums slice g of all accumulators through NVLink peer loads inside the resolve kernel
//     (k_combine_resolve: reduce-scatter + stage3.zig:242 in one kernel) and copies its slice to the host frame.
//   * one process per GPU (MPI / torchrun style): rtc_comm_* + rtc_render_sharded, the frame combined with NCCL
//     (ncclSend/ncclRecv of the RGB8 strips, or ncclReduce(sum) of the rgb accumulators), loaded at run time so that
//     librtc.so carries no link-time dependency on it.
#include "rtc_internal.h"

#include <dlfcn.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#define CU(call) RTC_CU(call)

namespace {

uint32_t frame_total(const rtc_params* p)
{
    return p->mode == RTC_MODE_PRIMARY ? 1u : p->num_samples;
}

void enable_peer(int from, int to)
{
    if (from == to) return;
    RtcDeviceGuard g(from);
    int can = 0;
    if (cudaDeviceCanAccessPeer(&can, from, to) == cudaSuccess && can) cudaDeviceEnablePeerAccess(to, 0);
    cudaGetLastError();          // already enabled (by us or by the host framework) is fine
}

}  // namespace

// A replica of `src` on `device`: every allocation of the scene copied peer to peer, every device pointer inside the
// scene (DevScene, the material table's texel pointers, the BVH view) re-based onto the copies.
int rtc_scene_clone_(const rtc_scene* src, int device, rtc_scene** out)
{
    rtc_scene* r = nullptr;
    if (int rc = rtc_scene_new_(device, &r)) return rc;
    RtcDeviceGuard guard(device);
    enable_peer(device, src->device);
    enable_peer(src->device, device);
    struct Map { const char* from; size_t bytes; char* to; };
    std::vector<Map> map;
    auto bail = [&](int rc) { rtc_keep_error_free_(r); return rc; };
    for (const rtc_scene::Alloc& a : src->allocs) {
        void* p = nullptr;
        cudaError_t e = cudaMalloc(&p, a.bytes);
        if (e != cudaSuccess) return bail(rtc_fail_(RTC_ERR_OUT_OF_MEMORY, "replica on device %d: cudaMalloc(%zu): %s", device, a.bytes, cudaGetErrorString(e)));
        r->allocs.push_back(rtc_scene::Alloc{ p, a.bytes });
        e = cudaMemcpyPeer(p, device, a.p, src->device, a.bytes);
        if (e != cudaSuccess) return bail(rtc_fail_(RTC_ERR_CUDA, "replica on device %d: cudaMemcpyPeer: %s", device, cudaGetErrorString(e)));
        map.push_back(Map{ (const char*)a.p, a.bytes, (char*)p });
    }
    bool lost = false;
    auto tr = [&](const void* p) -> void* {
        if (!p) return nullptr;
        for (const Map& m : map)
            if ((const char*)p >= m.from && (const char*)p < m.from + m.bytes) return m.to + ((const char*)p - m.from);
        lost = true;
        return nullptr;
    };
    r->dev = src->dev;
    rtc::DevScene& d = r->dev;
    d.cells = (const uint2*)tr(d.cells);
    d.tri_v0 = (const float4*)tr(d.tri_v0); d.tri_e1 = (const float4*)tr(d.tri_e1); d.tri_e2 = (const float4*)tr(d.tri_e2);
    d.tri_data = (const float4*)tr(d.tri_data);
    d.fine = (const unsigned long long*)tr(d.fine);
    d.coarse = (const uint32_t*)tr(d.coarse);
    d.materials = (const rtc::DevMaterial*)tr(d.materials);
    if (src->dev.num_materials) {
        std::vector<rtc::DevMaterial> mats(src->dev.num_materials);
        cudaError_t e = cudaMemcpy(mats.data(), src->dev.materials, mats.size() * sizeof(rtc::DevMaterial), cudaMemcpyDefault);
        if (e != cudaSuccess) return bail(rtc_fail_(RTC_ERR_CUDA, "replica: reading the material table: %s", cudaGetErrorString(e)));
        for (rtc::DevMaterial& m : mats) {
            m.base.texels = tr(m.base.texels); m.emissive.texels = tr(m.emissive.texels); m.alpha.texels = tr(m.alpha.texels);
        }
        if (d.materials) {
            e = cudaMemcpy(const_cast<rtc::DevMaterial*>(d.materials), mats.data(), mats.size() * sizeof(rtc::DevMaterial), cudaMemcpyDefault);
            if (e != cudaSuccess) return bail(rtc_fail_(RTC_ERR_CUDA, "replica: writing the material table: %s", cudaGetErrorString(e)));
        }
    }
    r->has_bvh = src->has_bvh;
    r->bvh_info = src->bvh_info;
    if (src->has_bvh) { r->bvh.nodes = (const float4*)tr(src->bvh.nodes); r->bvh.tris = (const float4*)tr(src->bvh.tris); r->bvh.num_nodes = src->bvh.num_nodes; }
    if (lost) return bail(rtc_fail_(RTC_ERR_CUDA, "replica: a scene pointer lies outside the scene's allocations"));
    r->num_tris = src->num_tris; r->num_src_tris = src->num_src_tris;
    r->is_replica = true;
    if (int rc = rtc_scene_finish_(r)) return bail(rc);
    *out = r;
    return RTC_OK;
}

namespace {

// The devices of an in-process multi-GPU render: the scene's own first, then the other sm_100 devices by index.
int pick_devices(rtc_scene* s, int n, std::vector<int>& out)
{
    // cudaGetDeviceProperties costs milliseconds per device: the list is made once per scene and device count
    if ((int)s->multi_devices.size() == n) { out = s->multi_devices; return RTC_OK; }
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess) { cudaGetLastError(); count = 0; }
    out.assign(1, s->device);
    for (int d = 0; d < count && (int)out.size() < n; ++d) {
        if (d == s->device) continue;
        cudaDeviceProp p;
        if (cudaGetDeviceProperties(&p, d) == cudaSuccess && p.major == 10) out.push_back(d);
    }
    if ((int)out.size() < n) return rtc_fail_(RTC_ERR_NO_DEVICE, "num_gpus = %d but only %zu sm_100 device(s) are visible", n, out.size());
    s->multi_devices = out;
    s->peers_enabled = false;
    return RTC_OK;
}

int ensure_replicas(rtc_scene* s, const std::vector<int>& devices)
{
    for (size_t g = 1; g < devices.size(); ++g) {
        if (g - 1 < s->replicas.size()) {
            rtc_scene* r = s->replicas[g - 1];
            if (r->device != devices[g]) return rtc_fail_(RTC_ERR_INVALID_ARGUMENT, "replica %zu lives on device %d, expected %d", g, r->device, devices[g]);
            // the BVH may have been built on the primary after the replica was made: replicas are re-made then
            if (r->has_bvh == s->has_bvh) continue;
            rtc_scene_free_(r);
            s->replicas[g - 1] = nullptr;
            rtc_scene* fresh = nullptr;
            if (int rc = rtc_scene_clone_(s, devices[g], &fresh)) { s->replicas.erase(s->replicas.begin() + (long)(g - 1)); return rc; }
            s->replicas[g - 1] = fresh;
            continue;
        }
        rtc_scene* r = nullptr;
        if (int rc = rtc_scene_clone_(s, devices[g], &r)) return rc;
        s->replicas.push_back(r);
    }
    // all pairs see each other (sample sharding reads every accumulator from every device); once per device list
    if (!s->peers_enabled) {
        for (int a : devices) for (int b : devices) enable_peer(a, b);
        s->peers_enabled = true;
    }
    return RTC_OK;
}

struct PartResult { int rc = RTC_OK; std::string msg; };

}  // namespace

int rtc_render_multi_(rtc_scene* s, const rtc_camera* cam, const rtc_params* p, uint8_t* rgb_out, rtc_aux* aux)
{
    const int N = p->num_gpus;
    const uint32_t total = frame_total(p);
    if (total == 0) return rtc_fail_(RTC_ERR_INVALID_ARGUMENT, "num_samples is zero");
    if (s->is_replica) return rtc_fail_(RTC_ERR_INVALID_ARGUMENT, "num_gpus > 1 on a replica");
    // RTC_TRACE_MULTI=1: host-side phase times of this call on stderr
    static const bool trace = getenv("RTC_TRACE_MULTI") && getenv("RTC_TRACE_MULTI")[0] == '1';
    const auto t_start = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (trace) fprintf(stderr, "rtc multi: %-22s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_start).count());
    };
    std::vector<int> devices;
    if (int rc = pick_devices(s, N, devices)) return rc;
    if (p->accel == RTC_ACCEL_BVH && !s->has_bvh) { if (int rc = rtc_scene_build_bvh(s, 0, nullptr)) return rc; }
    if (int rc = ensure_replicas(s, devices)) return rc;
    lap("devices + replicas");
    std::vector<rtc_scene*> sc(N);
    sc[0] = s;
    for (int g = 1; g < N; ++g) sc[g] = s->replicas[g - 1];

    const uint32_t pixels = cam->w * cam->h;
    const bool want_hits = p->mode == RTC_MODE_PRIMARY && aux && (aux->hit_id || aux->hit_tuv);
    const bool want_stats = aux && aux->stats;
    RtcHostFrame hf;
    {
        RtcDeviceGuard guard(s->device);
        if (int rc = rtc_ensure_host_frame_(s, pixels, &hf)) return rc;
    }
    if (!(aux && aux->radiance)) hf.radiance = nullptr;
    if (!(want_hits && aux->hit_id)) hf.hit_id = nullptr;
    if (!(want_hits && aux->hit_tuv)) hf.hit_tuv = nullptr;
    // the deterministic slice has one sample: it is always dealt by pixels
    const bool by_samples = p->shard == RTC_SHARD_SAMPLES && p->mode == RTC_MODE_PATH;

    std::vector<PartResult> res(N);
    auto run_parts = [&](auto&& body) {
        std::vector<std::thread> th;
        for (int g = 1; g < N; ++g)
            th.emplace_back([&, g] {
                RtcDeviceGuard guard(sc[g]->device);
                res[g].rc = body(g);
                if (res[g].rc != RTC_OK) res[g].msg = rtc_last_error();
            });
        {
            RtcDeviceGuard guard(sc[0]->device);
            res[0].rc = body(0);
            if (res[0].rc != RTC_OK) res[0].msg = rtc_last_error();
        }
        for (auto& t : th) t.join();
        for (int g = 0; g < N; ++g)
            if (res[g].rc != RTC_OK) return rtc_fail_(res[g].rc, "device %d: %s", sc[g]->device, res[g].msg.c_str());
        return (int)RTC_OK;
    };

    if (!by_samples) {
        const int rc = run_parts([&](int g) -> int {
            rtc_params pg = *p;
            pg.num_gpus = 0; pg.tile_rank = (uint32_t)g; pg.tile_world = (uint32_t)N; pg.tile_rows = p->tile_rows ? p->tile_rows : 8u;
            pg.total_samples = total;
            if (int rc = rtc_render_part_(sc[g], cam, &pg, total, hf, want_hits, true, want_stats)) return rc;
            CU(cudaEventRecord(sc[g]->ev_frame, sc[g]->stream));
            CU(cudaStreamSynchronize(sc[g]->stream));
            return RTC_OK;
        });
        if (rc != RTC_OK) return rc;
    } else {
        // phase 1: every device accumulates its own sample range of every pixel
        int rc = run_parts([&](int g) -> int {
            rtc_params pg = *p;
            const uint32_t base = total / (uint32_t)N, extra = total % (uint32_t)N;
            const uint32_t cnt = base + ((uint32_t)g < extra ? 1u : 0u);
            const uint32_t off = (uint32_t)g * base + ((uint32_t)g < extra ? (uint32_t)g : extra);
            pg.num_gpus = 0; pg.num_samples = (uint16_t)cnt; pg.sample_offset = off; pg.total_samples = total;
            if (int rc = rtc_render_part_(sc[g], cam, &pg, total, RtcHostFrame{}, false, false, want_stats)) return rc;
            CU(cudaEventRecord(sc[g]->ev_part, sc[g]->stream));
            return RTC_OK;
        });
        if (rc != RTC_OK) return rc;
        // phase 2 (every ev_part is recorded by now): device g waits for all parts, sums and resolves slice g of the
        // pixels over peer memory and copies the slice to the host frame
        rtc::PeerAccums pa;
        memset(&pa, 0, sizeof pa);
        pa.n = (uint32_t)N;
        for (int g = 0; g < N; ++g) pa.p[g] = sc[g]->accum;
        for (int g = 0; g < N; ++g) {
            RtcDeviceGuard guard(sc[g]->device);
            cudaStream_t st = sc[g]->stream;
            for (int j = 0; j < N; ++j) if (j != g) CU(cudaStreamWaitEvent(st, sc[j]->ev_part, 0));
            const uint32_t first = (uint32_t)(((uint64_t)pixels * (uint32_t)g) / (uint32_t)N);
            const uint32_t last = (uint32_t)(((uint64_t)pixels * (uint32_t)(g + 1)) / (uint32_t)N);
            const cudaError_t e = p->math == RTC_MATH_FAST
                ? rtc::fast::launch_combine_resolve(pa, first, last - first, total, sc[g]->d_rgb, hf.radiance ? sc[g]->d_radiance : nullptr, st)
                : rtc::exact::launch_combine_resolve(pa, first, last - first, total, sc[g]->d_rgb, hf.radiance ? sc[g]->d_radiance : nullptr, st);
            if (e != cudaSuccess) return rtc_fail_(RTC_ERR_CUDA, "combine launch on device %d: %s", sc[g]->device, cudaGetErrorString(e));
            sc[g]->launches += 1;
            CU(cudaEventRecord(sc[g]->ev_frame, st));
            if (last > first) {
                CU(cudaMemcpyAsync(hf.rgb + 3 * (size_t)first, sc[g]->d_rgb + 3 * (size_t)first, 3 * (size_t)(last - first), cudaMemcpyDeviceToHost, st));
                if (hf.radiance)
                    CU(cudaMemcpyAsync(hf.radiance + 3 * (size_t)first, sc[g]->d_radiance + 3 * (size_t)first, 12 * (size_t)(last - first), cudaMemcpyDeviceToHost, st));
            }
        }
        for (int g = 0; g < N; ++g) {
            RtcDeviceGuard guard(sc[g]->device);
            CU(cudaStreamSynchronize(sc[g]->stream));
        }
    }
    lap("parts rendered + fetched");
    rtc_params all = *p;
    all.tile_world = 0; all.tile_rank = 0;
    rtc_copy_out_(cam, &all, hf, rgb_out, aux, want_hits);
    lap("copied to the caller");
    if (want_stats) {
        rtc_stats sum;
        memset(&sum, 0, sizeof sum);
        for (int g = 0; g < N; ++g) {
            RtcDeviceGuard guard(sc[g]->device);
            rtc_stats st;
            memset(&st, 0, sizeof st);
            if (int rc = rtc_read_counters_(sc[g], &st)) return rc;
            float k_ms = 0.0f, f_ms = 0.0f;
            CU(cudaEventElapsedTime(&k_ms, sc[g]->ev0, sc[g]->ev1));
            CU(cudaEventElapsedTime(&f_ms, sc[g]->ev0, sc[g]->ev_frame));
            sum.rays += st.rays; sum.cells_visited += st.cells_visited; sum.tris_tested += st.tris_tested;
            sum.hits_shaded += st.hits_shaded; sum.paths += st.paths; sum.launches += st.launches;
            sum.trace_ms += st.trace_ms; sum.trace_launches += st.trace_launches; sum.trace_launches_untimed += st.trace_launches_untimed;
            if (k_ms > sum.kernel_ms) sum.kernel_ms = k_ms;          // the devices run side by side: the slowest one
            if (f_ms > sum.frame_ms) sum.frame_ms = f_ms;
        }
        *aux->stats = sum;
    }
    return RTC_OK;
}

// ---- NCCL, loaded at run time -------------------------------------------------------------------------------
namespace {

// the few declarations of nccl.h this file needs (stable since NCCL 2.7; the library is the one the process already
// has loaded — a host framework's — or the system's libnccl.so.2)
struct NcclUniqueId { char internal[128]; };
typedef struct ncclComm* NcclComm;
enum { kNcclUint8 = 1, kNcclFloat32 = 7, kNcclSum = 0 };
static_assert(sizeof(NcclUniqueId) == RTC_COMM_ID_BYTES, "ncclUniqueId is 128 bytes");

struct Nccl {
    void* h = nullptr;
    int (*GetUniqueId)(NcclUniqueId*) = nullptr;
    int (*CommInitRank)(NcclComm*, int, NcclUniqueId, int) = nullptr;
    int (*CommDestroy)(NcclComm) = nullptr;
    int (*Reduce)(const void*, void*, size_t, int, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*Send)(const void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    bool ok = false;
};

Nccl* nccl()
{
    static Nccl n;
    static std::once_flag once;
    std::call_once(once, [] {
        const char* names[] = { "libnccl.so.2", "libnccl.so" };
        for (const char* nm : names) { n.h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL); if (n.h) break; }
        if (!n.h) return;
        auto sym = [&](const char* s) { return dlsym(n.h, s); };
        n.GetUniqueId = (decltype(n.GetUniqueId))sym("ncclGetUniqueId");
        n.CommInitRank = (decltype(n.CommInitRank))sym("ncclCommInitRank");
        n.CommDestroy = (decltype(n.CommDestroy))sym("ncclCommDestroy");
        n.Reduce = (decltype(n.Reduce))sym("ncclReduce");
        n.Send = (decltype(n.Send))sym("ncclSend");
        n.Recv = (decltype(n.Recv))sym("ncclRecv");
        n.GroupStart = (decltype(n.GroupStart))sym("ncclGroupStart");
        n.GroupEnd = (decltype(n.GroupEnd))sym("ncclGroupEnd");
        n.GetErrorString = (decltype(n.GetErrorString))sym("ncclGetErrorString");
        n.ok = n.GetUniqueId && n.CommInitRank && n.CommDestroy && n.Reduce && n.Send && n.Recv && n.GroupStart && n.GroupEnd && n.GetErrorString;
    });
    return n.ok ? &n : nullptr;
}

int need_nccl(Nccl** out)
{
    *out = nccl();
    if (!*out) return rtc_fail_(RTC_ERR_UNSUPPORTED, "NCCL (libnccl.so.2) cannot be loaded: %s", dlerror() ? dlerror() : "symbols missing");
    return RTC_OK;
}

#define NC(call)                                                                                          \
    do {                                                                                                  \
        const int r_ = (call);                                                                            \
        if (r_ != 0) return rtc_fail_(RTC_ERR_CUDA, "%s: %s", #call, nc->GetErrorString(r_));             \
    } while (0)

}  // namespace

struct rtc_comm {
    NcclComm comm = nullptr;
    int rank = 0, world = 1, device = 0;
};

extern "C" {

int rtc_comm_unique_id(uint8_t id_out[RTC_COMM_ID_BYTES])
{
    if (!id_out) return rtc_fail_(RTC_ERR_INVALID_ARGUMENT, "null id");
    Nccl* nc;
    if (int rc = need_nccl(&nc)) return rc;
    NcclUniqueId id;
    NC(nc->GetUniqueId(&id));
    memcpy(id_out, &id, sizeof id);
    return RTC_OK;
}

int rtc_comm_create(const uint8_t id_in[RTC_COMM_ID_BYTES], int rank, int world, int device, rtc_comm** out)
{
    if (!id_in || !out || world < 1 || rank < 0 || rank >= world) return rtc_fail_(RTC_ERR_INVALID_ARGUMENT, "rtc_comm_create: rank %d of %d", rank, world);
    *out = nullptr;
    Nccl* nc;
    if (int rc = need_nccl(&nc)) return rc;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) { cudaGetLastError(); return rtc_fail_(RTC_ERR_NO_DEVICE, "no CUDA device"); }
    if (device < 0) CU(cudaGetDevice(&device));
    if (device >= count) return rtc_fail_(RTC_ERR_INVALID_ARGUMENT, "device %d of %d", device, count);
    RtcDeviceGuard guard(device);
    NcclUniqueId id;
    memcpy(&id, id_in, sizeof id);
    rtc_comm* c = new rtc_comm();
    c->rank = rank; c->world = world; c->device = device;
    const int r = nc->CommInitRank(&c->comm, world, id, rank);
    if (r != 0) { delete c; return rtc_fail_(RTC_ERR_CUDA, "ncclCommInitRank: %s", nc->GetErrorString(r)); }
    *out = c;
    return RTC_OK;
}

void rtc_comm_destroy(rtc_comm* c)
{
    if (!c) return;
    if (Nccl* nc = nccl()) { RtcDeviceGuard guard(c->device); if (c->comm) nc->CommDestroy(c->comm); }
    delete c;
}

int rtc_scene_prepare_gpus(rtc_scene* s, int num_gpus, uint32_t accel)
{
    if (!s) return rtc_fail_(RTC_ERR_INVALID_ARGUMENT, "null scene");
    if (num_gpus < 0 || num_gpus > RTC_MAX_GPUS) return rtc_fail_(RTC_ERR_INVALID_ARGUMENT, "num_gpus %d outside 0..%d", num_gpus, RTC_MAX_GPUS);
    if (s->is_replica) return rtc_fail_(RTC_ERR_INVALID_ARGUMENT, "rtc_scene_prepare_gpus on a replica");
    if (accel == RTC_ACCEL_BVH && !s->has_bvh) { if (int rc = rtc_scene_build_bvh(s, 0, nullptr)) return rc; }
    if (num_gpus <= 1) return RTC_OK;
    std::vector<int> devices;
    if (int rc = pick_devices(s, num_gpus, devices)) return rc;
    return ensure_replicas(s, devices);
}

int rtc_scene_set_stream(rtc_scene* s, void* stream)
{
    if (!s) return rtc_fail_(RTC_ERR_INVALID_ARGUMENT, "null scene");
    RtcDeviceGuard guard(s->device);
    CU(cudaStreamSynchronize(s->stream));             // nothing of this scene is in flight on the old stream any more
    s->stream = stream ? (cudaStream_t)stream : s->own_stream;
    return RTC_OK;
}

int rtc_render_sharded(rtc_scene* s, rtc_comm* c, const rtc_camera* cam, const rtc_params* p, uint8_t* rgb_out, rtc_aux* aux)
{
    if (int rc = rtc_check_render_args_(s, cam, p)) return rc;
    const int world = c ? c->world : 1, rank = c ? c->rank : 0;
    if (c && c->device != s->device) return rtc_fail_(RTC_ERR_INVALID_ARGUMENT, "communicator on device %d, scene on device %d", c->device, s->device);
    Nccl* nc = nullptr;
    if (world > 1) { if (int rc = need_nccl(&nc)) return rc; }
    const uint32_t total = frame_total(p);
    if (total == 0) return rtc_fail_(RTC_ERR_INVALID_ARGUMENT, "num_samples is zero");
    RtcDeviceGuard guard(s->device);
    const uint32_t pixels = cam->w * cam->h;
    const bool want_stats = aux && aux->stats;
    const bool by_samples = p->shard == RTC_SHARD_SAMPLES && p->mode == RTC_MODE_PATH && world > 1;
    cudaStream_t st = s->stream;
    rtc_params pg = *p;
    pg.num_gpus = 0; pg.sample_offset = 0; pg.total_samples = total; pg.tile_rank = 0; pg.tile_world = 0;
    if (!by_samples) {
        // ---- strips: render + resolve own strips; rank 0 receives everybody's, packed, and puts them in place ----
        const uint32_t rows = p->tile_rows ? p->tile_rows : 8u;
        if (world > 1) { pg.tile_rank = (uint32_t)rank; pg.tile_world = (uint32_t)world; pg.tile_rows = rows; }
        if (int rc = rtc_render_part_(s, cam, &pg, total, RtcHostFrame{}, false, true, want_stats)) return rc;
        if (world > 1) {
            const size_t need = 3 * (size_t)pixels;
            if (s->d_stage_bytes < need) {
                cudaFree(s->d_stage); s->d_stage = nullptr; s->d_stage_bytes = 0;
                CU(cudaMalloc(&s->d_stage, need));
                s->d_stage_bytes = need;
            }
            const uint32_t strip_rows = rows < cam->h ? rows : cam->h;
            const uint32_t nstrips = (cam->h + strip_rows - 1) / strip_rows;
            const size_t strip_b = 3 * (size_t)strip_rows * cam->w, frame_b = 3 * (size_t)pixels;
            // strips of rank r: r, r + world, ...; all full except possibly the frame's last one
            auto strips_of = [&](int r, uint32_t* full, size_t* tail_b) {
                *full = 0; *tail_b = 0;
                if ((uint32_t)r >= nstrips) return;
                const uint32_t mine = (nstrips - 1 - (uint32_t)r) / (uint32_t)world + 1;
                const uint32_t last = (uint32_t)r + (mine - 1) * (uint32_t)world;
                const size_t last_b = last == nstrips - 1 ? frame_b - (size_t)last * strip_b : strip_b;
                *full = last_b == strip_b ? mine : mine - 1;
                *tail_b = last_b == strip_b ? 0 : last_b;
            };
            auto packed_bytes = [&](int r) { uint32_t f; size_t t; strips_of(r, &f, &t); return (size_t)f * strip_b + t; };
            if (rank != 0) {
                uint32_t full; size_t tail_b;
                strips_of(rank, &full, &tail_b);
                const uint8_t* src = s->d_rgb + (size_t)rank * strip_b;
                if (full) CU(cudaMemcpy2DAsync(s->d_stage, strip_b, src, strip_b * world, strip_b, full, cudaMemcpyDeviceToDevice, st));
                if (tail_b) CU(cudaMemcpyAsync(s->d_stage + (size_t)full * strip_b, src + (size_t)full * strip_b * world, tail_b, cudaMemcpyDeviceToDevice, st));
                const size_t nb = packed_bytes(rank);
                if (nb) NC(nc->Send(s->d_stage, nb, kNcclUint8, 0, c->comm, st));
            } else {
                size_t off = 0;
                NC(nc->GroupStart());
                std::vector<size_t> offs(world, 0);
                for (int r = 1; r < world; ++r) {
                    const size_t nb = packed_bytes(r);
                    offs[r] = off;
                    if (nb) NC(nc->Recv(s->d_stage + off, nb, kNcclUint8, r, c->comm, st));
                    off += nb;
                }
                NC(nc->GroupEnd());
                for (int r = 1; r < world; ++r) {
                    uint32_t full; size_t tail_b;
                    strips_of(r, &full, &tail_b);
                    uint8_t* dst = s->d_rgb + (size_t)r * strip_b;
                    const uint8_t* src = s->d_stage + offs[r];
                    if (full) CU(cudaMemcpy2DAsync(dst, strip_b * world, src, strip_b, strip_b, full, cudaMemcpyDeviceToDevice, st));
                    if (tail_b) CU(cudaMemcpyAsync(dst + (size_t)full * strip_b * world, src + (size_t)full * strip_b, tail_b, cudaMemcpyDeviceToDevice, st));
                }
            }
        }
    } else {
        // ---- sample ranges: accumulate own range, rgb-only ncclReduce(sum) to rank 0, resolve there ----
        const uint32_t base = total / (uint32_t)world, extra = total % (uint32_t)world;
        pg.num_samples = (uint16_t)(base + ((uint32_t)rank < extra ? 1u : 0u));
        pg.sample_offset = (uint32_t)rank * base + ((uint32_t)rank < extra ? (uint32_t)rank : extra);
        if (int rc = rtc_render_part_(s, cam, &pg, total, RtcHostFrame{}, false, false, want_stats)) return rc;
        float* sum3 = s->d_radiance;                    // 3 floats per pixel: the frame's radiance buffer doubles as the wire format
        cudaError_t e = rtc::exact::launch_pack_rgb(s->accum, pixels, sum3, st);
        if (e != cudaSuccess) return rtc_fail_(RTC_ERR_CUDA, "pack launch: %s", cudaGetErrorString(e));
        NC(nc->Reduce(sum3, sum3, 3 * (size_t)pixels, kNcclFloat32, kNcclSum, 0, c->comm, st));
        s->launches += 1;
        if (rank == 0) {
            e = p->math == RTC_MATH_FAST ? rtc::fast::launch_resolve3(sum3, pixels, total, s->d_rgb, st)
                                         : rtc::exact::launch_resolve3(sum3, pixels, total, s->d_rgb, st);
            if (e != cudaSuccess) return rtc_fail_(RTC_ERR_CUDA, "resolve launch: %s", cudaGetErrorString(e));
            s->launches += 1;
        }
    }
    CU(cudaEventRecord(s->ev_frame, st));
    if (rank == 0 && rgb_out) {
        RtcHostFrame hf;
        if (int rc = rtc_ensure_host_frame_(s, pixels, &hf)) return rc;
        CU(cudaMemcpyAsync(hf.rgb, s->d_rgb, 3 * (size_t)pixels, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        memcpy(rgb_out, hf.rgb, 3 * (size_t)pixels);
    } else {
        CU(cudaStreamSynchronize(st));
    }
    if (want_stats) {
        if (int rc = rtc_read_counters_(s, aux->stats)) return rc;
        float k_ms = 0.0f, f_ms = 0.0f;
        CU(cudaEventElapsedTime(&k_ms, s->ev0, s->ev1));
        CU(cudaEventElapsedTime(&f_ms, s->ev0, s->ev_frame));
        aux->stats->kernel_ms = k_ms;
        aux->stats->frame_ms = f_ms;
    }
    return RTC_OK;
}

}  // extern "C"
