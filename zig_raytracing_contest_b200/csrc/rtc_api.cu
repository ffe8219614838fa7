// rtc_api.cu — the C ABI of include/rtc.h over the sm_100a kernels in rtc_kernels.cu.
// Replaces `scene.render(threads, camera, img)` (src/main.zig:126, src/stage3.zig:247).
// No CPU fallback: every compute entry point fails with RTC_ERR_NO_DEVICE without an sm_100 GPU.
#include "rtc_internal.h"
#include "rtc_bvh_host.h"

#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

namespace {
thread_local std::string g_error;
}  // namespace

int rtc_fail_(int code, const char* fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_error = buf;
    return code;
}
#define fail rtc_fail_

// used by rtc_geometry.cpp
int rtc_set_error_(int code, const char* msg) { g_error = msg; return code; }

namespace {

#define CU(call) RTC_CU(call)

float __uint_as_float_host(uint32_t u) { float f; memcpy(&f, &u, sizeof f); return f; }

bool device_is_sm100(int dev)
{
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, dev) != cudaSuccess) return false;
    return p.major == 10;
}

using DeviceGuard = RtcDeviceGuard;

}  // namespace

extern "C" {

const char* rtc_last_error(void) { return g_error.c_str(); }
uint32_t rtc_abi_version(void) { return RTC_ABI_VERSION; }

int rtc_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    int ok = 0;
    for (int i = 0; i < n; ++i) ok += device_is_sm100(i) ? 1 : 0;
    return ok;
}

}  // extern "C"

namespace {

int upload_texture(rtc_scene* s, const rtc_texture& t, int want_channels, rtc::DevTexture* out,
                   std::vector<std::pair<const float*, void*>>& seen, const void* preset)
{
    if ((!t.texels && !preset) || t.w <= 0 || t.h <= 0) return fail(RTC_ERR_INVALID_ARGUMENT, "texture without texels or with non-positive size");
    if (t.channels != want_channels) return fail(RTC_ERR_INVALID_ARGUMENT, "texture has %d channels, expected %d", t.channels, want_channels);
    if ((int64_t)t.w * t.h > (int64_t)1 << 30) return fail(RTC_ERR_UNSUPPORTED, "texture larger than 2^30 texels");
    out->w = t.w; out->h = t.h;
    out->u_min = t.u_min; out->u_max = t.u_max; out->v_min = t.v_min; out->v_max = t.v_max;
    out->wf = (float)t.w; out->hf = (float)t.h;            // Texture.w/h: f32 (src/stage3.zig:85-86)
    if (preset) { out->texels = preset; return RTC_OK; }   // already on the device, in the device layout
    for (auto& kv : seen)
        if (kv.first == t.texels) { out->texels = kv.second; return RTC_OK; }
    const size_t n = (size_t)t.w * (size_t)t.h;
    void* d = nullptr;
    if (want_channels == 3) {
        std::vector<float4> staged(n);
        for (size_t i = 0; i < n; ++i) staged[i] = make_float4(t.texels[3 * i], t.texels[3 * i + 1], t.texels[3 * i + 2], 0.0f);
        float4* dp = nullptr;
        if (int rc = s->alloc(&dp, n)) return rc;
        CU(cudaMemcpy(dp, staged.data(), n * sizeof(float4), cudaMemcpyHostToDevice));
        d = dp;
    } else {
        float* dp = nullptr;
        if (int rc = s->alloc(&dp, n)) return rc;
        CU(cudaMemcpy(dp, t.texels, n * sizeof(float), cudaMemcpyHostToDevice));
        d = dp;
    }
    seen.emplace_back(t.texels, d);
    out->texels = d;
    return RTC_OK;
}

int scene_upload(rtc_scene* s, const rtc_scene_desc* desc)
{
    const rtc_grid& g = desc->grid;
    const uint64_t num_cells = (uint64_t)g.resolution[0] * g.resolution[1] * g.resolution[2];
    if (g.resolution[0] == 0 || g.resolution[1] == 0 || g.resolution[2] == 0 || num_cells != desc->num_cells)
        return fail(RTC_ERR_INVALID_ARGUMENT, "num_cells (%llu) does not match the grid resolution", (unsigned long long)desc->num_cells);
    if (num_cells >= ((uint64_t)1 << 32) || desc->num_tris >= ((uint64_t)1 << 32))
        return fail(RTC_ERR_UNSUPPORTED, "more than 2^32 cells or baked triangles");
    if (!desc->cells || (desc->num_tris && (!desc->tri_pos || !desc->tri_data)) || !desc->materials || desc->num_materials == 0)
        return fail(RTC_ERR_INVALID_ARGUMENT, "scene description has null arrays");

    for (int a = 0; a < 3; ++a)
        if (g.resolution[a] > rtc::kMaxGridRes)
            return fail(RTC_ERR_UNSUPPORTED, "grid resolution %u on axis %d exceeds %u (cell coordinates are packed 10 bits per axis)",
                        g.resolution[a], a, rtc::kMaxGridRes);
    // validate cells; re-lay them out by packed cell (rtc_device.cuh) and derive the two occupancy levels
    rtc::DevScene& d = s->dev;
    const size_t nb = rtc_scene_set_packing_(d, g.resolution);
    const uint32_t* shift = d.shift;
    std::vector<rtc_cell> cells_b(nb * 64, rtc_cell{ 0u, 0u });
    std::vector<unsigned long long> fine(nb, 0ull);
    std::vector<uint32_t> coarse((nb + 31) / 32, 0u);
    {
        uint64_t c = 0;
        for (uint32_t z = 0; z < g.resolution[2]; ++z)
            for (uint32_t y = 0; y < g.resolution[1]; ++y)
                for (uint32_t x = 0; x < g.resolution[0]; ++x, ++c) {
                    const rtc_cell& cell = desc->cells[c];
                    if (cell.begin > cell.end || cell.end > desc->num_tris)
                        return fail(RTC_ERR_INVALID_ARGUMENT, "cell %llu has range [%u,%u) outside the %llu baked triangles",
                                    (unsigned long long)c, cell.begin, cell.end, (unsigned long long)desc->num_tris);
                    const uint32_t pc = rtc::pack_cell(shift, x, y, z);
                    const size_t bi = pc >> 6;
                    cells_b[pc] = cell;
                    if (cell.end > cell.begin) { fine[bi] |= 1ull << (pc & 63u); coarse[bi >> 5] |= 1u << (bi & 31); }
                }
    }
    uint32_t num_src = desc->num_src_tris;
    uint32_t max_src = 0;
    for (uint64_t i = 0; i < desc->num_tris; ++i) {
        if (desc->tri_data[i].material >= desc->num_materials)
            return fail(RTC_ERR_INVALID_ARGUMENT, "baked triangle %llu references material %u of %u",
                        (unsigned long long)i, desc->tri_data[i].material, desc->num_materials);
        if (desc->tri_data[i].src_tri > max_src) max_src = desc->tri_data[i].src_tri;
    }
    if (num_src == 0) num_src = desc->num_tris ? max_src + 1 : 0;
    if (desc->num_tris && max_src >= num_src) return fail(RTC_ERR_INVALID_ARGUMENT, "src_tri %u >= num_src_tris %u", max_src, num_src);

    for (int i = 0; i < 3; ++i) {
        d.bmin[i] = g.bbox_min[i]; d.bmax[i] = g.bbox_max[i];
        d.res[i] = g.resolution[i]; d.cell[i] = g.cell_size[i];
    }
    uint2* d_cells = nullptr; unsigned long long* d_fine = nullptr; uint32_t* d_coarse = nullptr;
    float4* d_pos = nullptr; float4* d_data = nullptr;
    const size_t M = (size_t)desc->num_tris;
    if (int rc = s->alloc(&d_cells, cells_b.size())) return rc;
    if (int rc = s->alloc(&d_fine, fine.size())) return rc;
    if (int rc = s->alloc(&d_coarse, coarse.size())) return rc;
    if (int rc = s->alloc(&d_pos, 3 * M)) return rc;
    if (int rc = s->alloc(&d_data, 4 * (size_t)num_src)) return rc;
    static_assert(sizeof(rtc_cell) == sizeof(uint2), "rtc_cell layout");
    CU(cudaMemcpy(d_cells, cells_b.data(), cells_b.size() * sizeof(uint2), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_fine, fine.data(), fine.size() * sizeof(unsigned long long), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_coarse, coarse.data(), coarse.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    {
        // SoA {v0}{e1}{e2} float4 arrays; v0.w carries the source-triangle id (the reference's rows have
        // an unused 4th lane: @Vector(3,f32) is 16 bytes), so a hit needs no extra index fetch.
        std::vector<float4> pos(3 * M);
        std::vector<float4> data(4 * (size_t)num_src, make_float4(0, 0, 0, 0));
        for (size_t i = 0; i < M; ++i) {
            const rtc_tri_pos& p = desc->tri_pos[i];
            const rtc_tri_data& t = desc->tri_data[i];
            pos[i] = make_float4(p.v0[0], p.v0[1], p.v0[2], __uint_as_float_host(t.src_tri));
            pos[M + i] = make_float4(p.e1[0], p.e1[1], p.e1[2], 0.0f);
            pos[2 * M + i] = make_float4(p.e2[0], p.e2[1], p.e2[2], 0.0f);
            float w[16] = { 0 };
            memcpy(w, t.normal, 9 * sizeof(float));
            memcpy(w + 9, t.texcoord, 6 * sizeof(float));
            memcpy(w + 15, &t.material, sizeof(uint32_t));
            memcpy(&data[4 * (size_t)t.src_tri], w, sizeof w);
        }
        CU(cudaMemcpy(d_pos, pos.data(), pos.size() * sizeof(float4), cudaMemcpyHostToDevice));
        CU(cudaMemcpy(d_data, data.data(), data.size() * sizeof(float4), cudaMemcpyHostToDevice));
    }
    if (int rc = rtc_scene_upload_materials_(s, desc->materials, desc->num_materials)) return rc;
    d.cells = d_cells; d.fine = d_fine; d.coarse = d_coarse; d.coarse_words = (uint32_t)coarse.size();
    d.tri_v0 = d_pos; d.tri_e1 = d_pos + M; d.tri_e2 = d_pos + 2 * M; d.tri_data = d_data;
    s->num_tris = desc->num_tris;
    s->num_src_tris = num_src;

    return rtc_scene_finish_(s);
}

}  // namespace

size_t rtc_scene_set_packing_(rtc::DevScene& d, const uint32_t res[3])
{
    const uint32_t nbx = (res[0] + 3) / 4, nby = (res[1] + 3) / 4, nbz = (res[2] + 3) / 4;
    auto clog2 = [](uint32_t n) { uint32_t l = 0; while ((1u << l) < n) ++l; return l; };
    const uint32_t lx = clog2(nbx), ly = clog2(nby), lz = clog2(nbz);
    d.shift[0] = 6u; d.shift[1] = 6u + lx; d.shift[2] = 6u + lx + ly;
    d.mask[0] = 0x03u | (((1u << lx) - 1u) << d.shift[0]);
    d.mask[1] = 0x0Cu | (((1u << ly) - 1u) << d.shift[1]);
    d.mask[2] = 0x30u | (((1u << lz) - 1u) << d.shift[2]);
    const size_t nb = (size_t)1 << (lx + ly + lz);
    d.coarse_words = (uint32_t)((nb + 31) / 32);
    return nb;
}

int rtc_scene_upload_materials_(rtc_scene* s, const rtc_material* materials, uint32_t num_materials, const void* const* preset)
{
    if (!materials || num_materials == 0) return fail(RTC_ERR_INVALID_ARGUMENT, "scene has no materials");
    std::vector<rtc::DevMaterial> mats(num_materials);
    std::vector<std::pair<const float*, void*>> seen;
    for (uint32_t m = 0; m < num_materials; ++m) {
        const void* const* pm = preset ? preset + 3 * (size_t)m : nullptr;
        if (int rc = upload_texture(s, materials[m].base_color, 3, &mats[m].base, seen, pm ? pm[0] : nullptr)) return rc;
        if (int rc = upload_texture(s, materials[m].emissive, 3, &mats[m].emissive, seen, pm ? pm[1] : nullptr)) return rc;
        if (int rc = upload_texture(s, materials[m].transparency, 1, &mats[m].alpha, seen, pm ? pm[2] : nullptr)) return rc;
    }
    rtc::DevMaterial* d_mats = nullptr;
    if (int rc = s->alloc(&d_mats, mats.size())) return rc;
    CU(cudaMemcpy(d_mats, mats.data(), mats.size() * sizeof(rtc::DevMaterial), cudaMemcpyHostToDevice));
    s->dev.materials = d_mats;
    s->dev.num_materials = num_materials;
    return RTC_OK;
}

int rtc_scene_finish_(rtc_scene* s)
{
    if (int rc = s->alloc(&s->counters, (size_t)rtc::CNT_N)) return rc;
    if (int rc = s->alloc(&s->work, (size_t)4)) return rc;
    CU(cudaMemset(s->counters, 0, rtc::CNT_N * sizeof(unsigned long long)));
    CU(cudaStreamCreateWithFlags(&s->own_stream, cudaStreamNonBlocking));
    s->stream = s->own_stream;
    CU(cudaEventCreate(&s->ev0));
    CU(cudaEventCreate(&s->ev1));
    CU(cudaEventCreate(&s->ev_frame));
    CU(cudaEventCreateWithFlags(&s->ev_part, cudaEventDisableTiming));
    s->timer_events.resize(2 * 4096);
    for (cudaEvent_t& e : s->timer_events) CU(cudaEventCreate(&e));
    s->timer.ev = s->timer_events.data(); s->timer.capacity = 4096; s->timer.used = 0; s->timer.dropped = 0;
    return RTC_OK;
}

void rtc_scene_free_(rtc_scene* s)
{
    if (!s) return;
    DeviceGuard guard(s->device);
    cudaDeviceSynchronize();
    for (rtc_scene* r : s->replicas) rtc_scene_free_(r);
    s->replicas.clear();
    for (const rtc_scene::Alloc& a : s->allocs) cudaFree(a.p);
    for (void* p : s->wave_allocs) cudaFree(p);
    cudaFree(s->accum); cudaFree(s->d_rgb); cudaFree(s->d_radiance); cudaFree(s->d_hit_id); cudaFree(s->d_hit_tuv);
    if (s->h_pinned) cudaFreeHost(s->h_pinned);
    for (cudaEvent_t e : s->timer_events) cudaEventDestroy(e);
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    if (s->ev_part) cudaEventDestroy(s->ev_part);
    if (s->ev_frame) cudaEventDestroy(s->ev_frame);
    if (s->own_stream) cudaStreamDestroy(s->own_stream);
    cudaFree(s->d_stage);
    delete s;
}

int rtc_scene_new_(int device, rtc_scene** out)
{
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) { cudaGetLastError(); return fail(RTC_ERR_NO_DEVICE, "no CUDA device (this library has no CPU fallback)"); }
    if (device < 0) CU(cudaGetDevice(&device));
    if (device >= n) return fail(RTC_ERR_INVALID_ARGUMENT, "device %d of %d", device, n);
    if (!device_is_sm100(device)) return fail(RTC_ERR_NO_DEVICE, "device %d is not sm_100 (kernels are built for sm_100a only)", device);
    rtc_scene* s = new (std::nothrow) rtc_scene();
    if (!s) return fail(RTC_ERR_OUT_OF_MEMORY, "host allocation failed");
    s->device = device;
    RtcDeviceGuard guard(device);
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, device);
    s->sm_count = prop.multiProcessorCount;
    *out = s;
    return RTC_OK;
}

void rtc_keep_error_free_(rtc_scene* s)
{
    const std::string keep = g_error;
    rtc_scene_free_(s);
    g_error = keep;
}

extern "C" {

int rtc_scene_create(const rtc_scene_desc* desc, int device, rtc_scene** out)
{
    if (!desc || !out) return fail(RTC_ERR_INVALID_ARGUMENT, "rtc_scene_create: null argument");
    rtc_scene* s = nullptr;
    if (int rc = rtc_scene_new_(device, &s)) return rc;
    DeviceGuard guard(s->device);
    const int rc = scene_upload(s, desc);
    if (rc != RTC_OK) { rtc_keep_error_free_(s); return rc; }
    *out = s;
    return RTC_OK;
}

void rtc_scene_destroy(rtc_scene* scene) { rtc_scene_free_(scene); }

}  // extern "C"

namespace {

rtc::DevCamera to_dev(const rtc_camera& c)
{
    rtc::DevCamera d;
    d.w = c.w; d.h = c.h;
    for (int i = 0; i < 3; ++i) { d.origin[i] = c.origin[i]; d.llc[i] = c.lower_left_corner[i]; d.right[i] = c.right[i]; d.up[i] = c.up[i]; }
    return d;
}

}  // namespace

int rtc_check_render_args_(const rtc_scene* s, const rtc_camera* cam, const rtc_params* p)
{
    if (!s || !cam || !p) return fail(RTC_ERR_INVALID_ARGUMENT, "null scene, camera or params");
    if (cam->w == 0 || cam->h == 0) return fail(RTC_ERR_INVALID_ARGUMENT, "camera has zero size");
    if ((uint64_t)cam->w * cam->h >= ((uint64_t)1 << 31)) return fail(RTC_ERR_UNSUPPORTED, "more than 2^31 pixels");
    if (p->mode != RTC_MODE_PATH && p->mode != RTC_MODE_PRIMARY) return fail(RTC_ERR_INVALID_ARGUMENT, "unknown mode %u", p->mode);
    if (p->accel != RTC_ACCEL_GRID && p->accel != RTC_ACCEL_BVH) return fail(RTC_ERR_UNSUPPORTED, "unknown accel %u", p->accel);
    if (p->math != RTC_MATH_EXACT && p->math != RTC_MATH_FAST) return fail(RTC_ERR_INVALID_ARGUMENT, "unknown math contract %u", p->math);
    if (p->mode == RTC_MODE_PATH && p->max_bounce > rtc::kMaxBounce)
        return fail(RTC_ERR_UNSUPPORTED, "max_bounce %u > %u (per-bounce queue bookkeeping of the kernels)", (unsigned)p->max_bounce, rtc::kMaxBounce);
    if (p->tile_world > 1 && p->tile_rank >= p->tile_world) return fail(RTC_ERR_INVALID_ARGUMENT, "tile_rank %u >= tile_world %u", p->tile_rank, p->tile_world);
    if (p->num_gpus < 0 || p->num_gpus > RTC_MAX_GPUS) return fail(RTC_ERR_INVALID_ARGUMENT, "num_gpus %d outside 0..%d", p->num_gpus, RTC_MAX_GPUS);
    if (p->shard != RTC_SHARD_STRIPS && p->shard != RTC_SHARD_SAMPLES) return fail(RTC_ERR_INVALID_ARGUMENT, "unknown shard %u", p->shard);
    if (p->num_gpus > 1 && (p->tile_world > 1 || p->sample_offset != 0 || (p->total_samples != 0 && p->total_samples != p->num_samples)))
        return fail(RTC_ERR_INVALID_ARGUMENT, "num_gpus > 1 shards the frame itself: tile_rank/tile_world/sample_offset/total_samples must be unset");
    return RTC_OK;
}

namespace {
#define check_render_args rtc_check_render_args_

// The pixels of the frame that this call renders (partitioning A). tile_world <= 1: all.
// tile_rows == 0: one contiguous band of rows per rank, rank r of n owns rows [r*h/n, (r+1)*h/n) (the reference's
// own split is contiguous pixel ranges per worker, src/stage3.zig:228-233). tile_rows > 0: strips of tile_rows
// full rows dealt round-robin, strip k belongs to rank k % n — sky rows and geometry rows spread over every rank.
struct Shard {
    uint32_t first = 0, count = 0;          // first pixel; number of pixels this call renders
    uint32_t strip_px = 0, stride_px = 0;   // stride_px == 0: contiguous
    uint32_t full_strips = 0, tail_px = 0;  // interleaved: count = full_strips * strip_px + tail_px
};
Shard shard_pixels(const rtc_camera* cam, const rtc_params* p)
{
    Shard sh;
    const uint32_t P = cam->w * cam->h;
    if (p->tile_world <= 1) { sh.count = P; return sh; }
    if (p->tile_rows == 0) {
        const uint32_t y0 = (uint32_t)(((uint64_t)cam->h * p->tile_rank) / p->tile_world);
        const uint32_t y1 = (uint32_t)(((uint64_t)cam->h * (p->tile_rank + 1)) / p->tile_world);
        sh.first = y0 * cam->w; sh.count = (y1 - y0) * cam->w;
        return sh;
    }
    const uint32_t rows = p->tile_rows < cam->h ? p->tile_rows : cam->h;
    const uint32_t nstrips = (cam->h + rows - 1) / rows;
    sh.strip_px = rows * cam->w;
    sh.stride_px = sh.strip_px * p->tile_world;
    sh.first = sh.strip_px * p->tile_rank;
    if (p->tile_rank < nstrips) {
        const uint32_t mine = (nstrips - 1 - p->tile_rank) / p->tile_world + 1;     // strips rank, rank+n, ...
        const uint32_t last = p->tile_rank + (mine - 1) * p->tile_world;
        const uint32_t last_px = last == nstrips - 1 ? P - last * sh.strip_px : sh.strip_px;
        sh.full_strips = last_px == sh.strip_px ? mine : mine - 1;
        sh.tail_px = last_px == sh.strip_px ? 0 : last_px;
        sh.count = sh.full_strips * sh.strip_px + sh.tail_px;
    }
    return sh;
}

// Wavefront scratch for `slots` path slots and `levels` scattering levels (rtc_device.cuh WaveScratch).
// The slot count is capped by a byte budget; frames with more paths than slots are rendered in batches. The budget has
// two stages: 6 GiB for a scene's FIRST frame (a one-shot program does not pay a 30 GB allocation inside its render
// bracket), 48 GiB from the second frame on when the first one had to be batched (a render loop amortises the
// allocation: cfg3 in one batch and 11 launches instead of 6.4 batches and 61 launches, +1.8 %, profiles/r2p_*).
// Always at most half of the free HBM; RTC_WAVE_BYTES fixes it for both stages. Batching never changes a result.
int ensure_wave(rtc_scene* s, uint64_t pixels, uint32_t samples, uint32_t levels)
{
    if (levels < 1) levels = 1;
    const char* env = getenv("RTC_WAVE_BYTES");
    const long long env_bytes = env ? atoll(env) : 0;
    const int stage = s->render_calls >= 1 ? 1 : 0;
    const bool grow = env_bytes <= 0 && s->wave_budget_limited && stage > s->wave_stage;
    // fast path: the scratch already holds this frame, or is already as large as the budget allowed
    if (!grow && s->wave.cap && s->wave.max_levels >= levels &&
        (s->wave.cap >= pixels * (uint64_t)samples || (s->wave_budget_limited && s->wave.max_levels == levels)))
        return RTC_OK;
    const size_t per_slot = 2 * 64 + sizeof(uint2) + sizeof(float4) + (size_t)levels * (sizeof(float4) + sizeof(float2));
    size_t budget = stage ? (size_t)48 << 30 : (size_t)6 << 30;
    if (env_bytes > 0) budget = (size_t)env_bytes;
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) {
        size_t held = 0;
        if (s->wave.cap) held = (size_t)s->wave.cap * (2 * 64 + sizeof(uint2) + sizeof(float4) + (size_t)s->wave.max_levels * 24);
        if (budget > (free_b + held) / 2) budget = (free_b + held) / 2;
    }
    uint64_t want = pixels * (uint64_t)samples;
    const uint64_t fit = budget / per_slot;
    const bool limited = want > fit;
    if (want > fit) want = fit;
    if (want > 0x7FFFFFFFull) want = 0x7FFFFFFFull;
    if (want < 1024) want = 1024;
    if (s->wave.cap >= want && s->wave.max_levels >= levels) { s->wave_stage = stage; return RTC_OK; }
    for (void* p : s->wave_allocs) cudaFree(p);
    s->wave_allocs.clear();
    s->wave = rtc::WaveScratch{};
    const size_t cap = (size_t)want;
    auto grab = [&](void** out, size_t bytes) -> int {
        cudaError_t e = cudaMalloc(out, bytes);
        if (e != cudaSuccess) return fail(RTC_ERR_OUT_OF_MEMORY, "cudaMalloc(%zu bytes of wavefront scratch): %s", bytes, cudaGetErrorString(e));
        s->wave_allocs.push_back(*out);
        return RTC_OK;
    };
    rtc::WaveScratch w{};
    for (int k = 0; k < 2; ++k) {
        float4** arr[4] = { &w.q[k].a0, &w.q[k].a1, &w.q[k].a2, &w.q[k].a3 };
        for (int j = 0; j < 4; ++j)
            if (int rc = grab((void**)arr[j], cap * sizeof(float4))) return rc;
    }
    if (int rc = grab((void**)&w.hits, cap * sizeof(uint2))) return rc;
    if (int rc = grab((void**)&w.stack_e, cap * levels * sizeof(float4))) return rc;
    if (int rc = grab((void**)&w.stack_a, cap * levels * sizeof(float2))) return rc;
    if (int rc = grab((void**)&w.leaf, cap * sizeof(float4))) return rc;
    if (int rc = grab((void**)&w.counts, 2 * (rtc::kMaxBounce + 1) * sizeof(uint32_t))) return rc;
    w.cap = (uint32_t)cap; w.max_levels = levels;
    s->wave = w;
    s->wave_budget_limited = limited;
    s->wave_stage = stage;
    return RTC_OK;
}

int launch_render(rtc_scene* s, const rtc_camera* cam, const rtc_params* p, float4* accum,
                  rtc::PrimaryOut pout, cudaStream_t stream)
{
    const Shard sh = shard_pixels(cam, p);
    rtc::DevParams dp;
    dp.num_samples = p->num_samples; dp.max_bounce = p->max_bounce; dp.sample_offset = p->sample_offset;
    dp.seed_lo = (uint32_t)p->seed; dp.seed_hi = (uint32_t)(p->seed >> 32);
    dp.first_pixel = sh.first; dp.num_pixels = sh.count; dp.strip_px = sh.strip_px; dp.stride_px = sh.stride_px;
    const bool primary = p->mode == RTC_MODE_PRIMARY;
    if (int rc = ensure_wave(s, dp.num_pixels, primary ? 1u : p->num_samples, primary ? 1u : p->max_bounce)) return rc;
    const rtc::DevCamera dc = to_dev(*cam);
    rtc::LaunchCfg cfg; cfg.sm_count = s->sm_count; cfg.stream = stream; cfg.stats = p->collect_stats != 0;
    cfg.timer = &s->timer;
    cfg.cache = &s->cache;
    uint32_t launches = 0;
    const rtc::BvhView* bvh = nullptr;
    const rtc::DevScene& dev = s->dev;
    if (p->accel == RTC_ACCEL_BVH) {
        if (!s->has_bvh) { if (int rc = rtc_scene_build_bvh(s, 0, nullptr)) return rc; }
        bvh = &s->bvh;
    }
    const cudaError_t e = p->math == RTC_MATH_FAST
        ? rtc::fast::launch_render(dev, dc, dp, primary, accum, pout, s->counters, s->wave, cfg, &launches, bvh)
        : rtc::exact::launch_render(dev, dc, dp, primary, accum, pout, s->counters, s->wave, cfg, &launches, bvh);
    s->launches += launches;
    if (e != cudaSuccess) return fail(RTC_ERR_CUDA, "kernel launch: %s", cudaGetErrorString(e));
    s->render_calls += 1;
    return RTC_OK;
}

}  // namespace

int rtc_ensure_frame_(rtc_scene* s, uint32_t pixels)
{
    if (pixels <= s->cap_pixels) return RTC_OK;
    cudaFree(s->accum); cudaFree(s->d_rgb); cudaFree(s->d_radiance); cudaFree(s->d_hit_id); cudaFree(s->d_hit_tuv);
    s->accum = nullptr; s->d_rgb = nullptr; s->d_radiance = nullptr; s->d_hit_id = nullptr; s->d_hit_tuv = nullptr;
    s->cap_pixels = 0;
    CU(cudaMalloc(&s->accum, (size_t)pixels * sizeof(float4)));
    CU(cudaMalloc(&s->d_rgb, (size_t)pixels * 3));
    CU(cudaMalloc(&s->d_radiance, (size_t)pixels * 3 * sizeof(float)));
    CU(cudaMalloc(&s->d_hit_id, (size_t)pixels * sizeof(uint32_t)));
    CU(cudaMalloc(&s->d_hit_tuv, (size_t)pixels * 3 * sizeof(float)));
    s->cap_pixels = pixels;
    return RTC_OK;
}

// The pinned host frame: radiance | hit_tuv | hit_id | rgb (every section 4-byte aligned). Portable, so that every
// device of a multi-GPU render can copy into it.
int rtc_ensure_host_frame_(rtc_scene* s, uint32_t pixels, RtcHostFrame* hf)
{
    if (pixels > s->h_pinned_pixels) {
        if (s->h_pinned) cudaFreeHost(s->h_pinned);
        s->h_pinned = nullptr; s->h_pinned_pixels = 0;
        s->h_pinned_bytes = (size_t)pixels * (12 + 12 + 4 + 3);
        CU(cudaHostAlloc(&s->h_pinned, s->h_pinned_bytes, cudaHostAllocPortable));
        s->h_pinned_pixels = pixels;
    }
    const size_t P = s->h_pinned_pixels;
    hf->radiance = reinterpret_cast<float*>(s->h_pinned);
    hf->hit_tuv = reinterpret_cast<float*>(s->h_pinned + 12 * P);
    hf->hit_id = reinterpret_cast<uint32_t*>(s->h_pinned + 24 * P);
    hf->rgb = s->h_pinned + 28 * P;
    return RTC_OK;
}

namespace {
int read_counters(rtc_scene* s, rtc_stats* out)
{
    unsigned long long c[rtc::CNT_N];
    CU(cudaMemcpy(c, s->counters, sizeof c, cudaMemcpyDeviceToHost));
    CU(cudaMemset(s->counters, 0, sizeof c));
    out->rays = c[rtc::CNT_RAYS]; out->cells_visited = c[rtc::CNT_CELLS]; out->tris_tested = c[rtc::CNT_TRIS];
    out->hits_shaded = c[rtc::CNT_HITS]; out->paths = c[rtc::CNT_PATHS];
    out->launches = s->launches;
    s->launches = 0;
    // k_trace brackets (the caller has synchronised the stream the launches went to)
    double ms = 0.0;
    uint32_t timed = 0;
    for (uint32_t i = 0; i < s->timer.used; ++i) {
        float t = 0.0f;
        if (cudaEventElapsedTime(&t, s->timer.ev[2 * i], s->timer.ev[2 * i + 1]) == cudaSuccess) { ms += t; timed += 1; }
        else cudaGetLastError();
    }
    out->trace_ms = ms;
    out->trace_launches = timed;
    out->trace_launches_untimed = s->timer.dropped;
    s->timer.used = 0; s->timer.dropped = 0;
    return RTC_OK;
}

}  // namespace

int rtc_read_counters_(rtc_scene* s, rtc_stats* out) { return read_counters(s, out); }

int rtc_render_part_(rtc_scene* s, const rtc_camera* cam, const rtc_params* p, uint32_t total, const RtcHostFrame& hf,
                     bool want_hits, bool resolve_and_fetch, bool clear_counters)
{
    const uint32_t pixels = cam->w * cam->h;
    if (int rc = rtc_ensure_frame_(s, pixels)) return rc;
    const Shard sh = shard_pixels(cam, p);
    // the span of the frame buffers this call touches: its band, or (interleaved strips) the whole frame
    const bool strips = sh.stride_px != 0;
    const size_t first = strips ? 0 : sh.first, count = strips ? pixels : sh.count;
    cudaStream_t st = s->stream;
    if (clear_counters) CU(cudaMemsetAsync(s->counters, 0, rtc::CNT_N * sizeof(unsigned long long), st));
    if (count) CU(cudaMemsetAsync(s->accum + first, 0, count * sizeof(float4), st));
    CU(cudaEventRecord(s->ev0, st));
    rtc::PrimaryOut pout; pout.hit_id = want_hits ? s->d_hit_id : nullptr; pout.hit_tuv = want_hits ? s->d_hit_tuv : nullptr;
    if (int rc = launch_render(s, cam, p, s->accum, pout, st)) return rc;
    if (!resolve_and_fetch) { CU(cudaEventRecord(s->ev1, st)); return RTC_OK; }
    const bool want_rad = hf.radiance != nullptr;
    cudaError_t e = p->math == RTC_MATH_FAST
        ? rtc::fast::launch_resolve(s->accum + first, (uint32_t)count, total, s->d_rgb + 3 * first, want_rad ? s->d_radiance + 3 * first : nullptr, st)
        : rtc::exact::launch_resolve(s->accum + first, (uint32_t)count, total, s->d_rgb + 3 * first, want_rad ? s->d_radiance + 3 * first : nullptr, st);
    if (e != cudaSuccess) return fail(RTC_ERR_CUDA, "resolve launch: %s", cudaGetErrorString(e));
    s->launches += 1;
    CU(cudaEventRecord(s->ev1, st));
    // D2H of this call's pixels to their FINAL place in the pinned frame: a band is one copy, interleaved strips are one
    // 2-D copy strided on both sides (+ the short tail)
    auto fetch = [&](const void* d_base, size_t elem, void* h_base) -> cudaError_t {
        const uint8_t* src = (const uint8_t*)d_base + elem * sh.first;
        uint8_t* dst = (uint8_t*)h_base + elem * sh.first;
        if (!strips) return cudaMemcpyAsync(dst, src, elem * sh.count, cudaMemcpyDeviceToHost, st);
        cudaError_t r = cudaSuccess;
        if (sh.full_strips)
            r = cudaMemcpy2DAsync(dst, elem * sh.stride_px, src, elem * sh.stride_px, elem * sh.strip_px, sh.full_strips, cudaMemcpyDeviceToHost, st);
        if (r == cudaSuccess && sh.tail_px)
            r = cudaMemcpyAsync(dst + elem * (size_t)sh.stride_px * sh.full_strips, src + elem * (size_t)sh.stride_px * sh.full_strips,
                                elem * sh.tail_px, cudaMemcpyDeviceToHost, st);
        return r;
    };
    if (sh.count) {
        if (hf.rgb) CU(fetch(s->d_rgb, 3, hf.rgb));
        if (want_rad) CU(fetch(s->d_radiance, 12, hf.radiance));
        if (want_hits && hf.hit_id) CU(fetch(s->d_hit_id, 4, hf.hit_id));
        if (want_hits && hf.hit_tuv) CU(fetch(s->d_hit_tuv, 12, hf.hit_tuv));
    }
    return RTC_OK;
}

void rtc_copy_out_(const rtc_camera* cam, const rtc_params* p, const RtcHostFrame& hf, uint8_t* rgb_out, rtc_aux* aux, bool want_hits)
{
    const Shard sh = shard_pixels(cam, p);
    if (!sh.count) return;
    const bool strips = sh.stride_px != 0;
    auto copy = [&](void* user, size_t elem, const void* pinned) {
        uint8_t* dst = (uint8_t*)user + elem * sh.first;
        const uint8_t* src = (const uint8_t*)pinned + elem * sh.first;
        if (!strips) { memcpy(dst, src, elem * sh.count); return; }
        for (uint32_t k = 0; k < sh.full_strips; ++k) memcpy(dst + elem * (size_t)sh.stride_px * k, src + elem * (size_t)sh.stride_px * k, elem * sh.strip_px);
        if (sh.tail_px) memcpy(dst + elem * (size_t)sh.stride_px * sh.full_strips, src + elem * (size_t)sh.stride_px * sh.full_strips, elem * sh.tail_px);
    };
    if (rgb_out && hf.rgb) copy(rgb_out, 3, hf.rgb);
    if (aux && aux->radiance && hf.radiance) copy(aux->radiance, 12, hf.radiance);
    if (want_hits && aux && aux->hit_id && hf.hit_id) copy(aux->hit_id, 4, hf.hit_id);
    if (want_hits && aux && aux->hit_tuv && hf.hit_tuv) copy(aux->hit_tuv, 12, hf.hit_tuv);
}

extern "C" {

int rtc_render(rtc_scene* s, const rtc_camera* cam, const rtc_params* p, uint8_t* rgb_out, rtc_aux* aux)
{
    if (int rc = check_render_args(s, cam, p)) return rc;
    if (!rgb_out) return fail(RTC_ERR_INVALID_ARGUMENT, "rgb_out is null");
    if (p->num_gpus > 1) return rtc_render_multi_(s, cam, p, rgb_out, aux);
    DeviceGuard guard(s->device);
    const uint32_t pixels = cam->w * cam->h;
    const uint32_t total = p->mode == RTC_MODE_PRIMARY ? 1u : (p->total_samples ? p->total_samples : p->num_samples);
    if (total == 0) return fail(RTC_ERR_INVALID_ARGUMENT, "num_samples is zero");
    const bool want_hits = p->mode == RTC_MODE_PRIMARY && aux && (aux->hit_id || aux->hit_tuv);
    RtcHostFrame hf;
    if (int rc = rtc_ensure_host_frame_(s, pixels, &hf)) return rc;
    if (!(aux && aux->radiance)) hf.radiance = nullptr;
    if (!(want_hits && aux->hit_id)) hf.hit_id = nullptr;
    if (!(want_hits && aux->hit_tuv)) hf.hit_tuv = nullptr;
    if (int rc = rtc_render_part_(s, cam, p, total, hf, want_hits, true, aux && aux->stats)) return rc;
    CU(cudaStreamSynchronize(s->stream));
    rtc_copy_out_(cam, p, hf, rgb_out, aux, want_hits);
    if (aux && aux->stats) {
        if (int rc = read_counters(s, aux->stats)) return rc;
        float ms = 0.0f;
        CU(cudaEventElapsedTime(&ms, s->ev0, s->ev1));
        aux->stats->kernel_ms = ms;
        aux->stats->frame_ms = ms;
    }
    return RTC_OK;
}

int rtc_render_accumulate(rtc_scene* s, const rtc_camera* cam, const rtc_params* p, void* d_accum, void* stream)
{
    if (int rc = check_render_args(s, cam, p)) return rc;
    if (!d_accum) return fail(RTC_ERR_INVALID_ARGUMENT, "d_accum is null");
    DeviceGuard guard(s->device);
    rtc::PrimaryOut pout; pout.hit_id = nullptr; pout.hit_tuv = nullptr;
    return launch_render(s, cam, p, (float4*)d_accum, pout, (cudaStream_t)stream);
}

int rtc_resolve(const void* d_accum, uint32_t num_pixels, uint32_t total_samples, void* d_rgb, void* d_radiance,
                int device, void* stream)
{
    if (!d_accum || total_samples == 0) return fail(RTC_ERR_INVALID_ARGUMENT, "rtc_resolve: null accumulator or zero samples");
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) { cudaGetLastError(); return fail(RTC_ERR_NO_DEVICE, "no CUDA device"); }
    if (device < 0) CU(cudaGetDevice(&device));
    DeviceGuard guard(device);
    cudaError_t e = rtc::exact::launch_resolve((const float4*)d_accum, num_pixels, total_samples, (uint8_t*)d_rgb,
                                               (float*)d_radiance, (cudaStream_t)stream);
    if (e != cudaSuccess) return fail(RTC_ERR_CUDA, "resolve launch: %s", cudaGetErrorString(e));
    return RTC_OK;
}

int rtc_scene_build_bvh(rtc_scene* s, int num_threads, rtc_bvh_info* info)
{
    if (!s) return fail(RTC_ERR_INVALID_ARGUMENT, "null scene");
    if (s->has_bvh) { if (info) *info = s->bvh_info; return RTC_OK; }
    DeviceGuard guard(s->device);
    const auto t0 = std::chrono::steady_clock::now();
    // the scene's own records, whichever way it was created: baked (duplicated, cell-ordered) {v0 | src} {e1} {e2};
    // every source triangle is taken once, from its first baked copy
    const size_t M = (size_t)s->num_tris;
    std::vector<float4> v0(M), e1(M), e2(M);
    if (M) {
        CU(cudaMemcpy(v0.data(), s->dev.tri_v0, M * sizeof(float4), cudaMemcpyDeviceToHost));
        CU(cudaMemcpy(e1.data(), s->dev.tri_e1, M * sizeof(float4), cudaMemcpyDeviceToHost));
        CU(cudaMemcpy(e2.data(), s->dev.tri_e2, M * sizeof(float4), cudaMemcpyDeviceToHost));
    }
    std::vector<uint32_t> first(s->num_src_tris, 0xFFFFFFFFu);
    std::vector<uint32_t> uniq;
    uniq.reserve(s->num_src_tris);
    for (size_t i = 0; i < M; ++i) {
        uint32_t src; memcpy(&src, &v0[i].w, sizeof src);
        if (src >= s->num_src_tris) return fail(RTC_ERR_INVALID_ARGUMENT, "baked triangle %zu has source id %u of %u", i, src, s->num_src_tris);
        if (first[src] == 0xFFFFFFFFu) { first[src] = (uint32_t)i; uniq.push_back((uint32_t)i); }
    }
    const uint32_t N = (uint32_t)uniq.size();
    std::vector<float> rec((size_t)N * 12);
    for (uint32_t k = 0; k < N; ++k) {
        const uint32_t i = uniq[k];
        float* r = &rec[12 * (size_t)k];
        r[0] = v0[i].x; r[1] = v0[i].y; r[2] = v0[i].z; r[3] = v0[i].w;
        r[4] = e1[i].x; r[5] = e1[i].y; r[6] = e1[i].z; r[7] = 0.0f;
        r[8] = e2[i].x; r[9] = e2[i].y; r[10] = e2[i].z; r[11] = 0.0f;
    }
    RtcBvhHost h;
    const int brc = rtc_bvh_build_host(rec.data(), N, num_threads, h);
    if (brc != 0) return fail(RTC_ERR_UNSUPPORTED, "BVH build failed (code %d): coordinates outside the quantisation range or a leaf overflow", brc);
    if (h.depth + 1 >= rtc::kBvhStack) return fail(RTC_ERR_UNSUPPORTED, "BVH depth %u exceeds the traversal stack (%u)", h.depth, rtc::kBvhStack);
    const size_t TQ = rtc::kBvhTriQuads, NQ = rtc::kBvhNodeQuads;
    std::vector<float4> bt(TQ * (size_t)N + TQ, make_float4(0, 0, 0, 0));
    for (uint32_t k = 0; k < N; ++k) {
        const float* r = &rec[12 * (size_t)h.order[k]];
        bt[TQ * (size_t)k] = make_float4(r[0], r[1], r[2], r[3]);
        bt[TQ * (size_t)k + 1] = make_float4(r[4], r[5], r[6], 0.0f);
        bt[TQ * (size_t)k + 2] = make_float4(r[8], r[9], r[10], 0.0f);
    }
    float4* d_nodes = nullptr; float4* d_tris = nullptr;
    if (h.nodes.size() != 4 * NQ * (size_t)h.num_nodes) return fail(RTC_ERR_CUDA, "BVH builder returned %zu words for %u nodes", h.nodes.size(), h.num_nodes);
    if (int rc = s->alloc(&d_nodes, NQ * (size_t)h.num_nodes)) return rc;
    if (int rc = s->alloc(&d_tris, bt.size())) return rc;
    CU(cudaMemcpy(d_nodes, h.nodes.data(), 16 * NQ * (size_t)h.num_nodes, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_tris, bt.data(), bt.size() * sizeof(float4), cudaMemcpyHostToDevice));
    s->bvh.nodes = d_nodes; s->bvh.tris = d_tris; s->bvh.num_nodes = h.num_nodes;
    s->bvh_info.num_nodes = h.num_nodes; s->bvh_info.num_tris = N; s->bvh_info.depth = h.depth; s->bvh_info.reserved = 0;
    s->bvh_info.bytes = 16ull * NQ * h.num_nodes + 16ull * TQ * N;
    s->bvh_info.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    s->has_bvh = true;
    if (info) *info = s->bvh_info;
    return RTC_OK;
}

int rtc_scene_stats(rtc_scene* s, rtc_stats* out)
{
    if (!s || !out) return fail(RTC_ERR_INVALID_ARGUMENT, "null argument");
    memset(out, 0, sizeof *out);
    {
        DeviceGuard guard(s->device);
        CU(cudaDeviceSynchronize());
        if (int rc = read_counters(s, out)) return rc;
    }
    // a scene that has rendered with num_gpus > 1: its replicas' counters belong to it (summed, cleared)
    for (rtc_scene* r : s->replicas) {
        DeviceGuard guard(r->device);
        CU(cudaDeviceSynchronize());
        rtc_stats st;
        memset(&st, 0, sizeof st);
        if (int rc = read_counters(r, &st)) return rc;
        out->rays += st.rays; out->cells_visited += st.cells_visited; out->tris_tested += st.tris_tested;
        out->hits_shaded += st.hits_shaded; out->paths += st.paths; out->launches += st.launches;
        out->trace_ms += st.trace_ms; out->trace_launches += st.trace_launches; out->trace_launches_untimed += st.trace_launches_untimed;
    }
    return RTC_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// rtc_render_accumulate_static: rtc_render_accumulate with the wavefront scratch (ray queues, hit records, throughput
// stacks, queue counters) in a file-scope device array instead of cudaMalloc'd memory — for callers that cannot allocate
// (CUDA graph capture, serving loops, harnesses that forbid it). No counterpart in the reference.
// Nothing on this path calls cudaMalloc / cudaFree / an allocating library routine.
//
// g_static_scratch below is, in HOST code, a stand-in for the symbol and not a device address: the device address is
// taken once per device with cudaGetSymbolAddress (that device current) and only that pointer goes to the kernels.
// Every GPU that loads the library holds the array until exit. All calls share it, and it keeps the last call's
// contents: the only part that must start at zero (the queue counters) is cleared on the caller's stream by the
// launcher before every batch; the rest is written before it is read. Calls that overlap on a GPU race.
// ---------------------------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_static_scratch[RTC_STATIC_SCRATCH_BYTES];

namespace {

struct StaticScratchAddr { bool known = false; unsigned char* ptr = nullptr; };
StaticScratchAddr g_static_addr[RTC_MAX_GPUS];
std::mutex g_static_mu;

constexpr size_t kStaticAlign = 256;
constexpr size_t kStaticCountsBytes = 2 * (rtc::kMaxBounce + 1) * sizeof(uint32_t);
constexpr size_t kStaticSections = 13;      // 8 queue planes, hits, stack_e, stack_a, leaf, counts

size_t static_slot_bytes(uint32_t levels)
{
    return 2 * 64 + sizeof(uint2) + sizeof(float4) + (size_t)levels * (sizeof(float4) + sizeof(float2));
}

// device address of g_static_scratch on `device` (which is current): asked for once per device
int static_scratch_base(int device, unsigned char** out)
{
    if (device < 0 || device >= RTC_MAX_GPUS) return fail(RTC_ERR_INVALID_ARGUMENT, "device %d outside [0, %d)", device, RTC_MAX_GPUS);
    std::lock_guard<std::mutex> lock(g_static_mu);
    StaticScratchAddr& a = g_static_addr[device];
    if (!a.known) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return fail(RTC_ERR_CUDA, "cudaGetSymbolAddress(static scratch) on device %d: %s", device, cudaGetErrorString(e));
        }
        a.ptr = (unsigned char*)p;
        a.known = true;
    }
    *out = a.ptr;
    return RTC_OK;
}

}  // namespace

extern "C" {

int rtc_render_accumulate_static(rtc_scene* s, const rtc_camera* cam, const rtc_params* p, void* d_accum, void* stream)
{
    if (int rc = check_render_args(s, cam, p)) return rc;
    if (!d_accum) return fail(RTC_ERR_INVALID_ARGUMENT, "d_accum is null");
    if (p->num_gpus > 1) return fail(RTC_ERR_INVALID_ARGUMENT, "rtc_render_accumulate_static renders on the scene's device only (num_gpus %d)", (int)p->num_gpus);
    const bool primary = p->mode == RTC_MODE_PRIMARY;
    if (p->accel == RTC_ACCEL_BVH && !s->has_bvh)
        return fail(RTC_ERR_INVALID_ARGUMENT, "rtc_render_accumulate_static allocates nothing: call rtc_scene_build_bvh before asking for RTC_ACCEL_BVH");
    const Shard sh = shard_pixels(cam, p);
    const uint32_t samples = primary ? 1u : p->num_samples;
    const uint32_t levels = primary ? 1u : (p->max_bounce < 1 ? 1u : p->max_bounce);
    const uint64_t paths = (uint64_t)sh.count * samples;
    const size_t usable = (size_t)RTC_STATIC_SCRATCH_BYTES - kStaticCountsBytes - kStaticSections * kStaticAlign;
    const uint64_t cap = usable / static_slot_bytes(levels);
    if (paths > cap)
        return fail(RTC_ERR_UNSUPPORTED, "rtc_render_accumulate_static: %llu paths (%u pixels x %u samples) exceed the %llu that the %u MiB static scratch "
                    "holds at max_bounce %u (%zu bytes per path); render fewer pixels or samples per call, or use rtc_render_accumulate",
                    (unsigned long long)paths, sh.count, samples, (unsigned long long)cap, (unsigned)(RTC_STATIC_SCRATCH_BYTES >> 20), levels,
                    static_slot_bytes(levels));
    if (paths == 0) return RTC_OK;
    DeviceGuard guard(s->device);
    unsigned char* base = nullptr;
    if (int rc = static_scratch_base(s->device, &base)) return rc;
    // carve the sections (rtc_device.cuh WaveScratch), each 256-byte aligned
    size_t off = 0;
    auto take = [&](size_t bytes) { unsigned char* q = base + off; off += (bytes + kStaticAlign - 1) / kStaticAlign * kStaticAlign; return q; };
    const size_t n = (size_t)cap;
    rtc::WaveScratch w{};
    for (int k = 0; k < 2; ++k) {
        w.q[k].a0 = (float4*)take(n * sizeof(float4)); w.q[k].a1 = (float4*)take(n * sizeof(float4));
        w.q[k].a2 = (float4*)take(n * sizeof(float4)); w.q[k].a3 = (float4*)take(n * sizeof(float4));
    }
    w.hits = (uint2*)take(n * sizeof(uint2));
    w.stack_e = (float4*)take(n * levels * sizeof(float4));
    w.stack_a = (float2*)take(n * levels * sizeof(float2));
    w.leaf = (float4*)take(n * sizeof(float4));
    w.counts = (uint32_t*)take(kStaticCountsBytes);
    if (off > (size_t)RTC_STATIC_SCRATCH_BYTES) return fail(RTC_ERR_CUDA, "static scratch layout: %zu bytes > %zu", off, (size_t)RTC_STATIC_SCRATCH_BYTES);
    w.cap = (uint32_t)cap; w.max_levels = levels;

    rtc::DevParams dp;
    dp.num_samples = p->num_samples; dp.max_bounce = p->max_bounce; dp.sample_offset = p->sample_offset;
    dp.seed_lo = (uint32_t)p->seed; dp.seed_hi = (uint32_t)(p->seed >> 32);
    dp.first_pixel = sh.first; dp.num_pixels = sh.count; dp.strip_px = sh.strip_px; dp.stride_px = sh.stride_px;
    const rtc::DevCamera dc = to_dev(*cam);
    rtc::LaunchCfg cfg; cfg.sm_count = s->sm_count; cfg.stream = (cudaStream_t)stream; cfg.stats = p->collect_stats != 0;
    cfg.timer = nullptr;                       // no event brackets on this path
    cfg.cache = &s->cache;
    rtc::PrimaryOut pout; pout.hit_id = nullptr; pout.hit_tuv = nullptr;
    uint32_t launches = 0;
    const rtc::BvhView* bvh = p->accel == RTC_ACCEL_BVH ? &s->bvh : nullptr;
    const cudaError_t e = p->math == RTC_MATH_FAST
        ? rtc::fast::launch_render(s->dev, dc, dp, primary, (float4*)d_accum, pout, s->counters, w, cfg, &launches, bvh)
        : rtc::exact::launch_render(s->dev, dc, dp, primary, (float4*)d_accum, pout, s->counters, w, cfg, &launches, bvh);
    s->launches += launches;
    if (e != cudaSuccess) return fail(RTC_ERR_CUDA, "kernel launch: %s", cudaGetErrorString(e));
    return RTC_OK;
}

}  // extern "C"
