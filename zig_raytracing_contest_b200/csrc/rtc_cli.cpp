// rtc_cli.cpp — `rtc`: the reference's program (src/main.zig:73-143) with scene.render on the B200.
//
// Same command line (zig-args over CmdlineArgs, src/main.zig:33-39: --in --out --camera --width --height,
// `--name value` and `--name=value`, unknown option -> message on stderr + failure), same config.json read from
// the current directory (src/main.zig:87), glTF in, PNG out, and the same phase log on stderr
// (Loaded / Preprocessed / Compiled / Rendered / Saved / Done, src/main.zig:106-142). Everything it calls is
// the C ABI of include/rtc.h; there is no CPU render fallback.
//
// Options beyond the reference's, all optional: --device N, --seed N, --mode path|primary, --math exact|fast,
// --accel grid|bvh, --gpus N (one process, N devices, one rtc_render call), --shard strips|samples.
#include "../../include/rtc.h"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

namespace {

using Clock = std::chrono::steady_clock;

// std.fmt.fmtDuration-style: the largest unit that keeps the value >= 1, three decimals.
std::string fmt_ns(double ns)
{
    char buf[64];
    if (ns >= 60e9) snprintf(buf, sizeof buf, "%dm%.3fs", (int)(ns / 60e9), (ns - 60e9 * (int)(ns / 60e9)) / 1e9);
    else if (ns >= 1e9) snprintf(buf, sizeof buf, "%.3fs", ns / 1e9);
    else if (ns >= 1e6) snprintf(buf, sizeof buf, "%.3fms", ns / 1e6);
    else if (ns >= 1e3) snprintf(buf, sizeof buf, "%.3fus", ns / 1e3);
    else snprintf(buf, sizeof buf, "%.0fns", ns);
    return buf;
}
std::string since(Clock::time_point t0)
{
    return fmt_ns((double)std::chrono::duration_cast<std::chrono::nanoseconds>(Clock::now() - t0).count());
}

struct Args {
    std::string in = "input.gltf", out = "output.png", camera;
    bool has_camera = false;
    uint32_t width = 0, height = 0;            // 0 = not given (?u16 = null)
    int device = -1;
    uint64_t seed = 0;
    uint32_t mode = RTC_MODE_PATH, math = RTC_MATH_EXACT, accel = RTC_ACCEL_GRID, shard = RTC_SHARD_STRIPS;
    int gpus = 1;
};

bool parse_u(const std::string& v, uint64_t max, uint64_t* out)
{
    if (v.empty()) return false;
    char* e = nullptr;
    const unsigned long long x = strtoull(v.c_str(), &e, 0);
    if (!e || *e || x > max) return false;
    *out = x;
    return true;
}

// zig-args semantics (libs/zig-args/args.zig:102-168): "--" ends the options; "--name=value" or "--name value";
// anything not starting with "--" is a positional (the reference ignores them).
bool parse_args(int argc, char** argv, Args* a)
{
    bool ok = true;
    for (int i = 1; i < argc; ++i) {
        std::string arg = argv[i];
        if (arg == "--") break;
        if (arg.compare(0, 2, "--") != 0) continue;
        std::string name = arg.substr(2), value;
        const size_t eq = name.find('=');
        bool has_value = eq != std::string::npos;
        if (has_value) { value = name.substr(eq + 1); name = name.substr(0, eq); }
        static const char* known[] = { "in", "out", "camera", "width", "height", "device", "seed", "mode", "math", "accel", "gpus", "shard" };
        bool is_known = false;
        for (const char* k : known) is_known = is_known || name == k;
        if (!is_known) { fprintf(stderr, "Unknown command line option: %s\n", name.c_str()); ok = false; continue; }
        if (!has_value) {
            if (i + 1 >= argc) { fprintf(stderr, "Missing argument for option %s\n", name.c_str()); ok = false; continue; }
            value = argv[++i];
        }
        uint64_t u = 0;
        if (name == "in") a->in = value;
        else if (name == "out") a->out = value;
        else if (name == "camera") { a->camera = value; a->has_camera = true; }
        else if (name == "width" || name == "height") {
            if (!parse_u(value, 65535, &u)) { fprintf(stderr, "Invalid value '%s' for option %s\n", value.c_str(), name.c_str()); ok = false; continue; }
            (name == "width" ? a->width : a->height) = (uint32_t)u;
        } else if (name == "device") { if (!parse_u(value, 1023, &u)) { ok = false; continue; } a->device = (int)u; }
        else if (name == "seed") { if (!parse_u(value, ~0ull, &u)) { ok = false; continue; } a->seed = u; }
        else if (name == "mode") {
            if (value == "path") a->mode = RTC_MODE_PATH; else if (value == "primary") a->mode = RTC_MODE_PRIMARY;
            else { fprintf(stderr, "Invalid value '%s' for option mode\n", value.c_str()); ok = false; }
        } else if (name == "gpus") { if (!parse_u(value, RTC_MAX_GPUS, &u) || u == 0) { fprintf(stderr, "Invalid value '%s' for option gpus\n", value.c_str()); ok = false; continue; } a->gpus = (int)u; }
        else if (name == "shard") {
            if (value == "strips") a->shard = RTC_SHARD_STRIPS; else if (value == "samples") a->shard = RTC_SHARD_SAMPLES;
            else { fprintf(stderr, "Invalid value '%s' for option shard\n", value.c_str()); ok = false; }
        } else if (name == "accel") {
            if (value == "grid") a->accel = RTC_ACCEL_GRID; else if (value == "bvh") a->accel = RTC_ACCEL_BVH;
            else { fprintf(stderr, "Invalid value '%s' for option accel\n", value.c_str()); ok = false; }
        } else if (name == "math") {
            if (value == "exact") a->math = RTC_MATH_EXACT; else if (value == "fast") a->math = RTC_MATH_FAST;
            else { fprintf(stderr, "Invalid value '%s' for option math\n", value.c_str()); ok = false; }
        }
    }
    return ok;
}

int die(const char* what)
{
    fprintf(stderr, "error: %s: %s\n", what, rtc_last_error());
    return 1;
}

}  // namespace

int main(int argc, char** argv)
{
    const Clock::time_point start = Clock::now();
    Args args;
    if (!parse_args(argc, argv, &args)) { fprintf(stderr, "error: InvalidArguments\n"); return 1; }

    rtc_config config;
    if (rtc_config_load("config.json", &config) != RTC_OK) return die("config.json");
    fprintf(stderr, "info: Num samples: %u, max bounce %u\n", (unsigned)config.num_samples, (unsigned)config.max_bounce);
    int num_threads = (int)config.num_threads;
    if (num_threads == 0) num_threads = (int)std::thread::hardware_concurrency();
    if (num_threads < 1) num_threads = 1;
    fprintf(stderr, "info: Num threads: %d\n", num_threads);

    fprintf(stderr, "info: Grid resolution: { %u, %u, %u }\n", config.grid_resolution[0], config.grid_resolution[1], config.grid_resolution[2]);
    const char* camera_name = args.has_camera ? args.camera.c_str() : nullptr;
    const char* env_load = getenv("RTC_HOST_LOAD");
    const char* env_build = getenv("RTC_HOST_BUILD");
    const bool host_build = env_build && env_build[0] == '1';
    const bool host_load = host_build || (env_load && env_load[0] == '1');
    Clock::time_point t = Clock::now();
    rtc_scene* scene = nullptr;
    rtc_camera camera;
    rtc_build_info bi;
    memset(&bi, 0, sizeof bi);
    memset(&camera, 0, sizeof camera);
    if (!host_load) {
        // Default: files + image decode on host threads, then triangle assembly (stage1), grid build and bake
        // (stage2) on the device in one call; the soup and the baked arrays never exist on the host.
        if (rtc_scene_create_from_gltf(args.in.c_str(), camera_name, args.width, args.height, num_threads, config.grid_resolution,
                                       args.device, &scene, &camera, &bi) != RTC_OK)
            return die(args.in.c_str());
        fprintf(stderr, "info: Pixels count: %u\n", camera.w * camera.h);
        // The reference's three phase lines (src/main.zig:106,113,119), in its order. The call above fuses the phases, so
        // the brackets are its own clocks: everything that is not a device kernel (files, JSON, image decode, uploads)
        // is the load; triangle assembly + texel linearisation on the device is the preprocessing; the grid build +
        // bake kernels are the compile.
        const double fused_ms = std::chrono::duration<double, std::milli>(Clock::now() - t).count();
        const double pre_ms = (double)bi.assemble_ms, comp_ms = (double)bi.device_ms;
        const double load_ms = fused_ms > pre_ms + comp_ms ? fused_ms - pre_ms - comp_ms : 0.0;
        fprintf(stderr, "info: Loaded in %s\n", fmt_ns(load_ms * 1e6).c_str());
        fprintf(stderr, "info: Preprocessed in %s\n", fmt_ns(pre_ms * 1e6).c_str());
        fprintf(stderr, "info: Compiled in %s\n", fmt_ns(comp_ms * 1e6).c_str());
    } else {
        // RTC_HOST_LOAD=1: stage1 on host threads, the grid built and baked on the device from the uploaded soup.
        // RTC_HOST_BUILD=1: stage2 on a host thread too, the baked arrays uploaded. Same scene bit for bit.
        rtc_host_scene host;
        if (rtc_gltf_load(args.in.c_str(), camera_name, args.width, args.height, num_threads, &host) != RTC_OK) return die(args.in.c_str());
        camera = host.camera;
        fprintf(stderr, "info: Pixels count: %u\n", camera.w * camera.h);
        fprintf(stderr, "info: Materials count: %u\n", host.num_materials);
        fprintf(stderr, "info: Loaded in %s\n", since(t).c_str());
        fprintf(stderr, "info: Preprocessed in 0ns (camera, materials and triangles are part of the load call above)\n");
        t = Clock::now();
        if (host_build) {
            rtc_geometry geo;
            if (rtc_geometry_build(host.positions, host.tri_data, host.num_tris, config.grid_resolution, &geo) != RTC_OK) return die("grid build");
            bi.grid = geo.grid; bi.num_cells = geo.num_cells; bi.num_tris = geo.num_tris; bi.num_src_tris = geo.num_src_tris;
            bi.empty_cells = geo.empty_cells; bi.min_tris = geo.min_tris; bi.max_tris = geo.max_tris;
            rtc_scene_desc desc;
            memset(&desc, 0, sizeof desc);
            desc.grid = geo.grid; desc.cells = geo.cells; desc.num_cells = geo.num_cells;
            desc.tri_pos = geo.tri_pos; desc.tri_data = geo.tri_data; desc.num_tris = geo.num_tris;
            desc.materials = host.materials; desc.num_materials = host.num_materials; desc.num_src_tris = geo.num_src_tris;
            if (rtc_scene_create(&desc, args.device, &scene) != RTC_OK) return die("scene upload");
            rtc_geometry_free(&geo);
        } else {
            if (rtc_scene_create_from_soup(host.positions, host.tri_data, host.num_tris, config.grid_resolution, host.materials,
                                           host.num_materials, args.device, &scene, &bi) != RTC_OK)
                return die("grid build");
        }
        rtc_host_scene_free(&host);
        fprintf(stderr, "info: Compiled in %s\n", since(t).c_str());
    }
    fprintf(stderr, "info: Empty cells: %u/%llu (%.2f%%) min triangles: %u max triangles: %u mean_triangles: %llu\n", bi.empty_cells,
            (unsigned long long)bi.num_cells, bi.num_cells ? 100.0 * bi.empty_cells / (double)bi.num_cells : 0.0, bi.min_tris, bi.max_tris,
            (unsigned long long)(bi.num_cells > bi.empty_cells ? bi.num_tris / (bi.num_cells - bi.empty_cells) : 0));
    fprintf(stderr, "info: Unique triangle count: %u/%llu (%.2f%%)\n", bi.num_src_tris, (unsigned long long)bi.num_tris,
            bi.num_tris ? 100.0 * bi.num_src_tris / (double)bi.num_tris : 0.0);

    if (args.gpus > 1 || args.accel == RTC_ACCEL_BVH) {
        // set-up that belongs to "Compiled", not to "Rendered": contexts + replicas on the other devices, the BVH
        t = Clock::now();
        if (rtc_scene_prepare_gpus(scene, args.gpus, args.accel) != RTC_OK) return die("prepare");
        fprintf(stderr, "info: Prepared %d device(s)%s in %s\n", args.gpus, args.accel == RTC_ACCEL_BVH ? " + BVH" : "", since(t).c_str());
    }
    std::vector<uint8_t> img((size_t)camera.w * camera.h * 3);
    t = Clock::now();
    rtc_params prm;
    memset(&prm, 0, sizeof prm);
    prm.num_samples = config.num_samples; prm.max_bounce = config.max_bounce;
    prm.mode = args.mode; prm.seed = args.seed; prm.accel = args.accel; prm.math = args.math;
    prm.num_gpus = args.gpus > 1 ? args.gpus : 0; prm.shard = args.shard;      // still ONE blocking call (src/main.zig:126)
    rtc_stats stats;
    memset(&stats, 0, sizeof stats);
    rtc_aux aux;
    memset(&aux, 0, sizeof aux);
    aux.stats = &stats;
    if (rtc_render(scene, &camera, &prm, img.data(), &aux) != RTC_OK) return die("render");
    const double render_s = std::chrono::duration<double>(Clock::now() - t).count();
    fprintf(stderr, "info: Rendered in %s\n", since(t).c_str());
    fprintf(stderr, "info: Rays: %llu (%.1f Mrays/s), paths: %llu, kernels launched: %llu\n", (unsigned long long)stats.rays,
            render_s > 0 ? (double)stats.rays / render_s / 1e6 : 0.0, (unsigned long long)stats.paths, (unsigned long long)stats.launches);

    t = Clock::now();
    if (rtc_png_write(args.out.c_str(), camera.w, camera.h, img.data()) != RTC_OK) return die("WritePngFail");
    fprintf(stderr, "info: Saved in %s\n", since(t).c_str());
    rtc_scene_destroy(scene);
    fprintf(stderr, "info: Done in %s\n", since(start).c_str());
    return 0;
}
