// rtc_gltf.cpp — the reference's input/output file contracts for hosts that are not the Zig program:
//   config.json                      src/main.zig:56-69
//   glTF 2.0 (.gltf / .glb) -> camera, materials, world-space triangle soup
//                                    src/stage1.zig:75-107 (files), :217-259 (triangles), :309-371 (camera),
//                                    :381-496 (materials); libs/zgltf/src/main.zig:248-280 + helpers.zig:17-99
//                                    (node transforms: column-major, parent * child)
//   PNG out                          src/main.zig:130-135 (stbi_write_png, 3 channels, stride 3*w)
// Host code only; compiled with -ffp-contract=off because the transform arithmetic decides the bits of every
// vertex the render path sees, and must evaluate like the reference's fp32 code.
#include "../../include/rtc.h"
#include "rtc_gltf_plan.h"
#include "rtc_jpeg.h"
#include "rtc_json.h"
#include "rtc_png.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <limits>
#include <string>
#include <thread>
#include <vector>

int rtc_set_error_(int code, const char* msg);   // rtc_api.cu

namespace {

using rtc_host::Json;

struct HostError {
    int code;
    std::string msg;
};
[[noreturn]] void fail(int code, const std::string& msg) { throw HostError{ code, msg }; }

// The file as a read-only mapping (the 100 MB-class .bin / .glb files the loader exists for); anything that
// cannot be mapped (pipes, empty files) is read.
RtcBytes read_file(const std::string& path)
{
    const int fd = open(path.c_str(), O_RDONLY | O_CLOEXEC);
    if (fd < 0) fail(RTC_ERR_IO, "cannot open '" + path + "'");
    struct stat st;
    if (fstat(fd, &st) == 0 && S_ISREG(st.st_mode) && st.st_size > 0) {
        const size_t n = (size_t)st.st_size;
        void* m = mmap(nullptr, n, PROT_READ, MAP_PRIVATE | MAP_POPULATE, fd, 0);
        if (m != MAP_FAILED) {
            close(fd);
            RtcBytes r;
            r.map = std::shared_ptr<const void>(m, [n](const void* q) { munmap(const_cast<void*>(q), n); });
            r.p = static_cast<const uint8_t*>(m); r.n = n;
            return r;
        }
    }
    close(fd);
    std::ifstream f(path, std::ios::binary);
    if (!f) fail(RTC_ERR_IO, "cannot open '" + path + "'");
    f.seekg(0, std::ios::end);
    const std::streamoff n = f.tellg();
    f.seekg(0, std::ios::beg);
    std::vector<uint8_t> data((size_t)n);
    if (n && !f.read((char*)data.data(), n)) fail(RTC_ERR_IO, "cannot read '" + path + "'");
    return RtcBytes(std::move(data));
}

std::string dir_of(const std::string& path)
{
    const size_t k = path.find_last_of('/');
    return k == std::string::npos ? std::string(".") : path.substr(0, k);
}

std::vector<uint8_t> base64_decode(const std::string& s, size_t from)
{
    std::vector<uint8_t> out;
    uint32_t acc = 0;
    int bits = 0;
    for (size_t i = from; i < s.size(); ++i) {
        const char c = s[i];
        int v;
        if (c >= 'A' && c <= 'Z') v = c - 'A';
        else if (c >= 'a' && c <= 'z') v = c - 'a' + 26;
        else if (c >= '0' && c <= '9') v = c - '0' + 52;
        else if (c == '+' || c == '-') v = 62;
        else if (c == '/' || c == '_') v = 63;
        else continue;                                    // '=', whitespace
        acc = (acc << 6) | (uint32_t)v;
        bits += 6;
        if (bits >= 8) { bits -= 8; out.push_back((uint8_t)(acc >> bits)); }
    }
    return out;
}

// A uri of a buffer or image: a data: URI (base64) or a path relative to the asset (src/stage1.zig:47,92).
RtcBytes load_uri(const std::string& dir, const std::string& uri)
{
    if (uri.compare(0, 5, "data:") == 0) {
        const size_t k = uri.find("base64,");
        if (k == std::string::npos) fail(RTC_ERR_FORMAT, "data: URI without base64 payload");
        return RtcBytes(base64_decode(uri, k + 7));
    }
    std::string rel;
    for (size_t i = 0; i < uri.size(); ++i) {             // percent-decoding of the relative path
        if (uri[i] == '%' && i + 2 < uri.size()) { rel += (char)strtol(uri.substr(i + 1, 2).c_str(), nullptr, 16); i += 2; }
        else rel += uri[i];
    }
    return read_file(dir + "/" + rel);
}

// ---- fp32 math of the loader ---------------------------------------------------------------------------
struct V3 { float x, y, z; };
inline V3 v3(float x, float y, float z) { return V3{ x, y, z }; }
inline V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 scale(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
inline float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }                  // linalg.zig:189-191
inline V3 cross(V3 a, V3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline V3 normalize(V3 a) { return scale(a, 1.0f / sqrtf(dot(a, a))); }                      // linalg.zig:119-125

struct Mat4 { float m[4][4]; };      // m[column][row], as zgltf's Mat4 = [4][4]f32 (helpers.zig:4)

// helpers.zig:81-99: result[column][row] = sum over k (in order 0..3, starting from 0) of left[k][row] * right[column][k]
Mat4 mul(const Mat4& left, const Mat4& right)
{
    Mat4 r;
    for (int c = 0; c < 4; ++c)
        for (int row = 0; row < 4; ++row) {
            float sum = 0.0f;
            for (int k = 0; k < 4; ++k) sum += left.m[k][row] * right.m[c][k];
            r.m[c][row] = sum;
        }
    return r;
}
Mat4 identity()
{
    Mat4 r;
    for (int c = 0; c < 4; ++c)
        for (int row = 0; row < 4; ++row) r.m[c][row] = c == row ? 1.0f : 0.0f;
    return r;
}
// helpers.zig:17-78: T * (R * S)
Mat4 recompose(const float t[3], const float q[4], const float s[3])
{
    Mat4 T = identity(), R = identity(), S = identity();
    T.m[3][0] = t[0]; T.m[3][1] = t[1]; T.m[3][2] = t[2];
    const float x = q[0], y = q[1], z = q[2], w = q[3];
    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, xz = x * z, yz = y * z, wx = w * x, wy = w * y, wz = w * z;
    R.m[0][0] = 1.0f - 2.0f * (yy + zz); R.m[0][1] = 2.0f * (xy + wz);        R.m[0][2] = 2.0f * (xz - wy);
    R.m[1][0] = 2.0f * (xy - wz);        R.m[1][1] = 1.0f - 2.0f * (xx + zz); R.m[1][2] = 2.0f * (yz + wx);
    R.m[2][0] = 2.0f * (xz + wy);        R.m[2][1] = 2.0f * (yz - wx);        R.m[2][2] = 1.0f - 2.0f * (xx + yy);
    S.m[0][0] = s[0]; S.m[1][1] = s[1]; S.m[2][2] = s[2];
    return mul(T, mul(R, S));
}
inline V3 col3(const Mat4& a, int c) { return v3(a.m[c][0], a.m[c][1], a.m[c][2]); }
// linalg.zig:263-278: the @Vector adds associate left to right
inline V3 transform_position(const Mat4& a, V3 v) { return ((scale(col3(a, 0), v.x) + scale(col3(a, 1), v.y)) + scale(col3(a, 2), v.z)) + col3(a, 3); }
inline V3 transform_direction(const Mat4& a, V3 v) { return (scale(col3(a, 0), v.x) + scale(col3(a, 1), v.y)) + scale(col3(a, 2), v.z); }

}  // namespace

namespace {

// ---- the asset ---------------------------------------------------------------------------------------------
struct Image {
    uint32_t w = 0, h = 0;
    int actual_c = 0;
    std::vector<float> data;           // w*h*4 linear floats, what stbi_loadf_from_memory(..., 4) returns
    std::vector<uint8_t> rgba8;        // deferred mode: the decoded 8-bit image, linearised on the device instead
};

struct Asset {
    Json doc;
    std::string dir;
    std::vector<RtcBytes> buffers;
    std::vector<Image> images;
    bool defer_texels = false;         // images stay 8-bit, textures become recipes (rtc_gltf_plan.h)
    std::vector<int> parent;           // per node, -1 = root (zgltf main.zig:1246-1295)
};

const Json& arr_of(const Json& doc, const char* key)
{
    static const Json empty = [] { Json j; j.kind = Json::Array; return j; }();
    const Json* j = doc.find(key);
    return j && j->kind == Json::Array ? *j : empty;
}
const Json& item(const Json& doc, const char* key, long long idx)
{
    const Json& a = arr_of(doc, key);
    if (idx < 0 || (size_t)idx >= a.arr.size()) fail(RTC_ERR_FORMAT, std::string("glTF: index out of range in '") + key + "'");
    return a.arr[(size_t)idx];
}

void fill_parents(const Json& nodes, std::vector<int>& parent, size_t node, int depth)
{
    if (depth > 4096) fail(RTC_ERR_FORMAT, "glTF: node hierarchy too deep (cycle?)");
    const Json* ch = nodes.arr[node].find("children");
    if (!ch || ch->kind != Json::Array) return;
    for (const Json& c : ch->arr) {
        const size_t k = (size_t)c.num;
        if (k >= nodes.arr.size()) fail(RTC_ERR_FORMAT, "glTF: child index out of range");
        parent[k] = (int)node;
        fill_parents(nodes, parent, k, depth + 1);
    }
}

// stb_image.h:1856-1877 (stbi__ldr_to_hdr): colour channels pow(c/255.0f, 2.2f) * 1.0f evaluated in double and
// rounded to float; the alpha channel is c/255 in float.
void linear_lut(float lut[256])
{
    // stb_image.h:1571,1867: the gamma and scale are FLOAT statics (2.2f, 1.0f) promoted to double inside pow —
    // pow(c/255.0f, (double)2.2f), not pow(., 2.2); pinned against the reference's own unit in tests/test_ref_stb.py
    const float gamma = 2.2f, scale = 1.0f;
    for (int i = 0; i < 256; ++i) lut[i] = (float)(pow((double)((float)i / 255.0f), (double)gamma) * (double)scale);
}
void linearise(const rtc_host::DecodedImage& src, Image& dst)
{
    static float lut[256];
    static std::atomic<bool> ready{ false };
    if (!ready.load()) {
        linear_lut(lut);
        ready.store(true);
    }
    dst.w = src.w; dst.h = src.h; dst.actual_c = src.actual_c;
    const size_t n = (size_t)src.w * src.h;
    dst.data.resize(n * 4);
    for (size_t i = 0; i < n; ++i) {
        dst.data[4 * i + 0] = lut[src.rgba[4 * i + 0]];
        dst.data[4 * i + 1] = lut[src.rgba[4 * i + 1]];
        dst.data[4 * i + 2] = lut[src.rgba[4 * i + 2]];
        dst.data[4 * i + 3] = (float)src.rgba[4 * i + 3] / 255.0f;
    }
}

struct BufferSpan { const uint8_t* p; size_t n; };
BufferSpan view_bytes(const Asset& a, long long view_idx)
{
    const Json& v = item(a.doc, "bufferViews", view_idx);
    const size_t b = (size_t)v.integer("buffer", -1), off = (size_t)v.integer("byteOffset", 0), len = (size_t)v.integer("byteLength", 0);
    if (b >= a.buffers.size() || off + len > a.buffers[b].size()) fail(RTC_ERR_FORMAT, "glTF: bufferView outside its buffer");
    return BufferSpan{ a.buffers[b].data() + off, len };
}

// loadGltfFile (src/stage1.zig:75-107): parse, load the buffers, decode the images on `threads` threads
// (image i is decoded by thread i % threads, src/stage1.zig:36-37).
void load_asset(const std::string& path, int threads, Asset& a)
{
    RtcBytes file = read_file(path);
    a.dir = dir_of(path);
    RtcBytes glb_bin;
    bool has_glb_bin = false;
    if (file.size() >= 12 && memcmp(file.data(), "glTF", 4) == 0) {                     // GLB container
        size_t p = 12;
        RtcBytes json;
        while (p + 8 <= file.size()) {
            uint32_t len, type;
            memcpy(&len, file.data() + p, 4); memcpy(&type, file.data() + p + 4, 4);
            if (p + 8 + (size_t)len > file.size()) fail(RTC_ERR_FORMAT, "GLB: truncated chunk");
            if (type == 0x4E4F534Au) json = file.window(p + 8, len);
            else if (type == 0x004E4942u && !has_glb_bin) { glb_bin = file.window(p + 8, len); has_glb_bin = true; }
            p += 8 + (size_t)len;
        }
        if (json.empty()) fail(RTC_ERR_FORMAT, "GLB: no JSON chunk");
        file = std::move(json);
    }
    try {
        a.doc = rtc_host::JsonParser((const char*)file.data(), file.size()).parse();
    } catch (const std::exception& e) {
        fail(RTC_ERR_FORMAT, std::string("'") + path + "': " + e.what());
    }
    if (a.doc.kind != Json::Object) fail(RTC_ERR_FORMAT, "glTF: top level is not an object");
    const Json& bufs = arr_of(a.doc, "buffers");
    for (size_t i = 0; i < bufs.arr.size(); ++i) {
        if (i == 0 && !bufs.arr[i].has("uri")) {                                        // src/stage1.zig:87-90
            if (!has_glb_bin) fail(RTC_ERR_FORMAT, "glTF: buffer 0 has no uri and the file has no GLB binary chunk");
            a.buffers.push_back(std::move(glb_bin));
        } else {
            if (!bufs.arr[i].has("uri")) fail(RTC_ERR_FORMAT, "glTF: buffer without uri");
            a.buffers.push_back(load_uri(a.dir, bufs.arr[i].string("uri", "")));
        }
    }
    const Json& nodes = arr_of(a.doc, "nodes");
    a.parent.assign(nodes.arr.size(), -1);
    for (const Json& scene : arr_of(a.doc, "scenes").arr)
        if (const Json* roots = scene.find("nodes"))
            for (const Json& r : roots->arr) {
                if ((size_t)r.num >= nodes.arr.size()) fail(RTC_ERR_FORMAT, "glTF: scene node index out of range");
                fill_parents(nodes, a.parent, (size_t)r.num, 0);
            }
    const Json& imgs = arr_of(a.doc, "images");
    a.images.resize(imgs.arr.size());
    const int T = threads < 1 ? 1 : threads;
    std::vector<std::thread> pool;
    std::vector<HostError> errors((size_t)T, HostError{ 0, "" });
    for (int t = 0; t < T; ++t)
        pool.emplace_back([&, t] {
            try {
                for (size_t i = (size_t)t; i < imgs.arr.size(); i += (size_t)T) {
                    const Json& im = imgs.arr[i];
                    RtcBytes owned;
                    BufferSpan bytes{ nullptr, 0 };
                    if (im.has("bufferView")) bytes = view_bytes(a, im.integer("bufferView", -1));
                    else if (im.has("uri")) { owned = load_uri(a.dir, im.string("uri", "")); bytes = BufferSpan{ owned.data(), owned.size() }; }
                    else fail(RTC_ERR_FORMAT, "glTF: image without bufferView or uri");
                    try {
                        const bool jpg = rtc_host::is_jpeg(bytes.p, bytes.n);
                        if (a.defer_texels) {
                            rtc_host::DecodedImage dec = jpg ? rtc_host::jpeg_decode(bytes.p, bytes.n) : rtc_host::png_decode(bytes.p, bytes.n);
                            Image& im8 = a.images[i];
                            im8.w = dec.w; im8.h = dec.h; im8.actual_c = dec.actual_c; im8.rgba8 = std::move(dec.rgba);
                        } else
                            linearise(jpg ? rtc_host::jpeg_decode(bytes.p, bytes.n) : rtc_host::png_decode(bytes.p, bytes.n), a.images[i]);
                    } catch (const std::runtime_error& e) {
                        fail(RTC_ERR_UNSUPPORTED, "glTF: image " + std::to_string(i) + ": " + e.what());
                    }
                }
            } catch (const HostError& e) { errors[(size_t)t] = e; }
        });
    for (auto& th : pool) th.join();
    for (const HostError& e : errors)
        if (e.code) throw e;
}

// getLocalTransform / getGlobalTransform (zgltf main.zig:248-280): parent * child up to the root.
Mat4 local_transform(const Json& node)
{
    if (const Json* m = node.find("matrix")) {
        if (m->arr.size() != 16) fail(RTC_ERR_FORMAT, "glTF: node.matrix needs 16 numbers");
        Mat4 r;
        for (int i = 0; i < 16; ++i) r.m[i / 4][i % 4] = (float)m->arr[(size_t)i].num;
        return r;
    }
    float t[3] = { 0, 0, 0 }, q[4] = { 0, 0, 0, 1 }, s[3] = { 1, 1, 1 };
    if (const Json* j = node.find("translation")) for (size_t i = 0; i < 3 && i < j->arr.size(); ++i) t[i] = (float)j->arr[i].num;
    if (const Json* j = node.find("rotation")) for (size_t i = 0; i < 4 && i < j->arr.size(); ++i) q[i] = (float)j->arr[i].num;
    if (const Json* j = node.find("scale")) for (size_t i = 0; i < 3 && i < j->arr.size(); ++i) s[i] = (float)j->arr[i].num;
    return recompose(t, q, s);
}
Mat4 global_transform(const Asset& a, size_t node)
{
    const Json& nodes = arr_of(a.doc, "nodes");
    Mat4 m = local_transform(nodes.arr[node]);
    for (int p = a.parent[node]; p >= 0; p = a.parent[(size_t)p]) m = mul(local_transform(nodes.arr[(size_t)p]), m);
    return m;
}

}  // namespace

namespace {

// ---- accessors (src/stage1.zig:130-201) -------------------------------------------------------------------
// The reference reads f32 vec3 / vec2 and u16 scalars at the accessor's tight element stride. This reader also
// honours bufferView.byteStride and u8 / u32 indices; on the assets the reference renders correctly the
// results are the same.
struct Accessor {
    const uint8_t* base = nullptr;
    size_t count = 0, stride = 0;
    int component = 0;
    size_t buffer = 0, offset = 0;      // where base lies: buffers[buffer] + offset
};
Accessor accessor(const Asset& a, long long idx, const char* want_type, bool want_float)
{
    const Json& acc = item(a.doc, "accessors", idx);
    if (acc.string("type", "") != want_type) fail(RTC_ERR_FORMAT, std::string("glTF: accessor type is not ") + want_type);
    Accessor r;
    r.component = (int)acc.integer("componentType", 0);
    size_t comp_size;
    switch (r.component) {
        case 5121: comp_size = 1; break;
        case 5123: comp_size = 2; break;
        case 5125: comp_size = 4; break;
        case 5126: comp_size = 4; break;
        default: fail(RTC_ERR_UNSUPPORTED, "glTF: unsupported accessor componentType");
    }
    if (want_float != (r.component == 5126)) fail(RTC_ERR_FORMAT, "glTF: accessor componentType does not fit its use");
    const size_t ncomp = !strcmp(want_type, "VEC3") ? 3 : (!strcmp(want_type, "VEC2") ? 2 : 1);
    if (!acc.has("bufferView")) fail(RTC_ERR_UNSUPPORTED, "glTF: accessor without bufferView");
    const Json& view = item(a.doc, "bufferViews", acc.integer("bufferView", -1));
    const BufferSpan bytes = view_bytes(a, acc.integer("bufferView", -1));
    const size_t off = (size_t)acc.integer("byteOffset", 0);
    r.count = (size_t)acc.integer("count", 0);
    r.stride = (size_t)view.integer("byteStride", 0);
    if (r.stride == 0) r.stride = comp_size * ncomp;
    if (r.count && off + (r.count - 1) * r.stride + comp_size * ncomp > bytes.n) fail(RTC_ERR_FORMAT, "glTF: accessor outside its bufferView");
    r.base = bytes.p + off;
    r.buffer = (size_t)view.integer("buffer", -1);
    r.offset = (size_t)view.integer("byteOffset", 0) + off;
    return r;
}
inline V3 read_v3(const Accessor& a, size_t i) { float f[3]; memcpy(f, a.base + i * a.stride, 12); return v3(f[0], f[1], f[2]); }
inline uint32_t read_index(const Accessor& a, size_t i)
{
    const uint8_t* p = a.base + i * a.stride;
    if (a.component == 5123) { uint16_t v; memcpy(&v, p, 2); return v; }
    if (a.component == 5125) { uint32_t v; memcpy(&v, p, 4); return v; }
    return *p;
}

using HostScene = RtcHostSceneOwner;      // rtc_gltf_plan.h

// loadTriangles (src/stage1.zig:217-259): for every node with a mesh, for every primitive, in file order.
// f(matrix, material, pos, nrm, uv, has_uv, ind) sees every primitive once, validated.
template <typename F>
void for_each_primitive(const Asset& a, F&& f)
{
    const Json& nodes = arr_of(a.doc, "nodes");
    const size_t num_materials = arr_of(a.doc, "materials").arr.size();
    for (size_t n = 0; n < nodes.arr.size(); ++n) {
        if (!nodes.arr[n].has("mesh")) continue;
        const Json& mesh = item(a.doc, "meshes", nodes.arr[n].integer("mesh", -1));
        const Mat4 matrix = global_transform(a, n);
        for (const Json& prim : arr_of(mesh, "primitives").arr) {
            if (prim.integer("mode", 4) != 4) fail(RTC_ERR_UNSUPPORTED, "glTF: primitive.mode is not TRIANGLES (src/stage1.zig:226)");
            if (!prim.has("material")) fail(RTC_ERR_FORMAT, "glTF: primitive without material (src/stage1.zig:239)");
            const long long material = prim.integer("material", -1);
            if (material < 0 || (size_t)material >= num_materials) fail(RTC_ERR_FORMAT, "glTF: primitive.material out of range");
            if (!prim.has("indices")) fail(RTC_ERR_FORMAT, "glTF: primitive without indices (src/stage1.zig:209)");
            const Json& attrs = prim.at("attributes");
            if (!attrs.has("POSITION") || !attrs.has("NORMAL")) fail(RTC_ERR_FORMAT, "glTF: primitive needs POSITION and NORMAL");
            const Accessor pos = accessor(a, attrs.integer("POSITION", -1), "VEC3", true);
            const Accessor nrm = accessor(a, attrs.integer("NORMAL", -1), "VEC3", true);
            Accessor uv;                                              // absent: texcoords read as zeroes
            const bool has_uv = attrs.has("TEXCOORD_0");
            if (has_uv) uv = accessor(a, attrs.integer("TEXCOORD_0", -1), "VEC2", true);
            const Accessor ind = accessor(a, prim.integer("indices", -1), "SCALAR", false);
            f(matrix, (uint32_t)material, pos, nrm, uv, has_uv, ind);
        }
    }
}

// The reference fills the soup serially on its main thread; here the primitives are collected first and the
// triangles are filled by `threads` workers over equal slices of the soup (same values, same order).
void load_geometry(const Asset& a, HostScene& out, int threads)
{
    struct Prim { Mat4 matrix; uint32_t material; Accessor pos, nrm, uv, ind; bool has_uv; size_t first, tris; };
    std::vector<Prim> prims;
    size_t total = 0;
    for_each_primitive(a, [&](const Mat4& matrix, uint32_t material, const Accessor& pos, const Accessor& nrm,
                              const Accessor& uv, bool has_uv, const Accessor& ind) {
        const size_t tris = ind.count / 3;
        if (!tris) return;
        prims.push_back(Prim{ matrix, material, pos, nrm, uv, ind, has_uv, total, tris });
        total += tris;
    });
    if (total > 0xFFFFFFFFull) fail(RTC_ERR_UNSUPPORTED, "glTF: more than 2^32 - 1 triangles");
    const auto tr0 = std::chrono::steady_clock::now();
    out.positions.resize(total * 9);
    out.tri_data.resize(total);
    if (getenv("RTC_TRACE_LOAD")) fprintf(stderr, "  resize %.1f ms\n", std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tr0).count());
    std::atomic<bool> bad{ false };
    auto fill = [&](size_t g0, size_t g1) {
        size_t k = 0;
        while (k + 1 < prims.size() && prims[k + 1].first <= g0) ++k;
        for (size_t g = g0; g < g1; ++g) {
            while (g >= prims[k].first + prims[k].tris) ++k;
            const Prim& pr = prims[k];
            const size_t t = g - pr.first;
            rtc_tri_data d;
            memset(&d, 0, sizeof d);
            d.material = pr.material;
            d.src_tri = (uint32_t)g;
            for (int i = 0; i < 3; ++i) {
                const uint32_t v = read_index(pr.ind, t * 3 + (size_t)i);
                if (v >= pr.pos.count || v >= pr.nrm.count || (pr.has_uv && v >= pr.uv.count)) { bad.store(true); return; }
                const V3 p = transform_position(pr.matrix, read_v3(pr.pos, v));
                const V3 nn = normalize(transform_direction(pr.matrix, read_v3(pr.nrm, v)));     // not the inverse transpose (TODO in the reference)
                float* po = out.positions.data() + 9 * g + 3 * (size_t)i;
                po[0] = p.x; po[1] = p.y; po[2] = p.z;
                d.normal[i][0] = nn.x; d.normal[i][1] = nn.y; d.normal[i][2] = nn.z;
                if (pr.has_uv) memcpy(d.texcoord[i], pr.uv.base + v * pr.uv.stride, 8);
            }
            out.tri_data[g] = d;
        }
    };
    const size_t T = (size_t)(threads < 1 ? 1 : threads);
    const size_t workers = total < 65536 ? 1 : T;
    if (workers == 1) { if (total) fill(0, total); }
    else {
        std::vector<std::thread> pool;
        for (size_t w = 0; w < workers; ++w) pool.emplace_back(fill, total * w / workers, total * (w + 1) / workers);
        for (auto& th : pool) th.join();
    }
    if (bad.load()) fail(RTC_ERR_FORMAT, "glTF: vertex index out of range");
}

// The same walk with the per-triangle loop left to the device (rtc_build.cu: k_assemble): one table row per
// primitive, the buffers the rows point into laid out back to back (256-byte aligned) in one blob.
void plan_geometry(const Asset& a, RtcGltfPlan& out)
{
    out.buffer_base.assign(a.buffers.size(), ~(uint64_t)0);
    uint64_t total = 0;
    auto place = [&](const Accessor& acc) -> uint64_t {
        uint64_t& base = out.buffer_base[acc.buffer];
        if (base == ~(uint64_t)0) { base = out.blob_bytes; out.blob_bytes += (a.buffers[acc.buffer].size() + 255u) & ~(uint64_t)255u; }
        return base + acc.offset;
    };
    auto u32 = [](size_t v, const char* what) -> uint32_t {
        if (v > 0xFFFFFFFFull) fail(RTC_ERR_UNSUPPORTED, std::string("glTF: ") + what + " does not fit 32 bits");
        return (uint32_t)v;
    };
    for_each_primitive(a, [&](const Mat4& matrix, uint32_t material, const Accessor& pos, const Accessor& nrm,
                              const Accessor& uv, bool has_uv, const Accessor& ind) {
        const size_t tris = ind.count / 3;
        if (!tris) return;
        RtcPrimRef r;
        memset(&r, 0, sizeof r);
        for (int c = 0; c < 4; ++c)
            for (int row = 0; row < 4; ++row) r.m[c * 4 + row] = matrix.m[c][row];
        r.off_pos = place(pos); r.stride_pos = u32(pos.stride, "byteStride"); r.count_pos = u32(pos.count, "accessor.count");
        r.off_nrm = place(nrm); r.stride_nrm = u32(nrm.stride, "byteStride"); r.count_nrm = u32(nrm.count, "accessor.count");
        if (has_uv) { r.off_uv = place(uv); r.stride_uv = u32(uv.stride, "byteStride"); r.count_uv = u32(uv.count, "accessor.count"); }
        r.off_ind = place(ind); r.stride_ind = u32(ind.stride, "byteStride");
        r.index_component = (uint32_t)ind.component;
        r.has_uv = has_uv ? 1u : 0u;
        r.material = material;
        if (total + tris > 0xFFFFFFFFull) fail(RTC_ERR_UNSUPPORTED, "glTF: more than 2^32 - 1 triangles");
        r.first_tri = (uint32_t)total; r.num_tris = (uint32_t)tris;
        total += tris;
        out.prims.push_back(r);
    });
    out.num_tris = (uint32_t)total;
}

// loadCamera (src/stage1.zig:282-371).
void load_camera(const Asset& a, const char* camera_name, uint32_t width, uint32_t height, rtc_camera* cam)
{
    const Json& cams = arr_of(a.doc, "cameras");
    if (cams.arr.empty()) fail(RTC_ERR_FORMAT, "NoCamerasAtAll");
    size_t ci = 0;
    if (camera_name) {
        bool found = false;
        for (size_t i = 0; i < cams.arr.size() && !found; ++i)
            if (cams.arr[i].string("name", "Camera_" + std::to_string(i)) == camera_name) { ci = i; found = true; }
        if (!found) fail(RTC_ERR_FORMAT, "CameraNotFound");
    }
    const Json& nodes = arr_of(a.doc, "nodes");
    size_t node = nodes.arr.size();
    for (size_t n = 0; n < nodes.arr.size(); ++n)
        if (nodes.arr[n].has("camera") && (size_t)nodes.arr[n].integer("camera", -1) == ci) { node = n; break; }
    if (node == nodes.arr.size()) fail(RTC_ERR_FORMAT, "CameraNodeNotFound");
    const Json& c = cams.arr[ci];
    if (c.string("type", "") != "perspective" || !c.has("perspective")) fail(RTC_ERR_UNSUPPORTED, "OnlyPerspectiveCamerasSupported");
    const Json& persp = c.at("perspective");
    if (!persp.has("yfov")) fail(RTC_ERR_FORMAT, "glTF: perspective camera without yfov");
    const float yfov = (float)persp.number("yfov", 0.0);
    const bool has_ar = persp.has("aspectRatio");
    uint32_t w = 0, h = 0;
    if (width == 0 && height == 0) fail(RTC_ERR_INVALID_ARGUMENT, "OutputImgSizeIsNotSpecified");
    if (width != 0 && height != 0) {
        if (has_ar) fail(RTC_ERR_INVALID_ARGUMENT, "CameraHasAspectRatio");
        w = width; h = height;
    } else {
        if (!has_ar) fail(RTC_ERR_INVALID_ARGUMENT, "CameraHasntAspectRatio");
        const float ar = (float)persp.number("aspectRatio", 0.0);
        w = width ? width : (uint32_t)((float)height * ar);
        h = height ? height : (uint32_t)((float)width / ar);
    }
    if (w == 0 || h == 0) fail(RTC_ERR_INVALID_ARGUMENT, "output image has a zero dimension");
    const float fw = (float)w, fh = (float)h;
    const Mat4 m = global_transform(a, node);
    const V3 origin = col3(m, 3);
    const V3 fwd = normalize(scale(col3(m, 2), -1.0f));
    const V3 right = normalize(cross(fwd, v3(0.0f, 1.0f, 0.0f)));
    const V3 up = cross(fwd, right);
    const float focal = (fh / 2.0f) / tanf(yfov / 2.0f);
    const V3 llc = (scale(fwd, focal) - scale(right, fw / 2.0f)) - scale(up, fh / 2.0f);
    cam->w = w; cam->h = h;
    cam->origin[0] = origin.x; cam->origin[1] = origin.y; cam->origin[2] = origin.z;
    cam->lower_left_corner[0] = llc.x; cam->lower_left_corner[1] = llc.y; cam->lower_left_corner[2] = llc.z;
    cam->right[0] = right.x; cam->right[1] = right.y; cam->right[2] = right.z;
    cam->up[0] = up.x; cam->up[1] = up.y; cam->up[2] = up.z;
}

}  // namespace

namespace {

// initTexture / initDummyTexture (src/stage1.zig:381-425).
rtc_texture make_texture(const Asset& a, const Json& texture, const Image& image, const float* texels, int channels)
{
    const int32_t lo = std::numeric_limits<int32_t>::min(), hi = std::numeric_limits<int32_t>::max();
    rtc_texture t;
    t.texels = texels; t.w = (int32_t)image.w; t.h = (int32_t)image.h; t.channels = channels;
    t.u_min = lo; t.u_max = hi; t.v_min = lo; t.v_max = hi;
    if (texture.has("sampler")) {
        const Json& s = item(a.doc, "samplers", texture.integer("sampler", -1));
        if (s.integer("wrapS", 10497) == 33071) { t.u_min = 0; t.u_max = (int32_t)image.w - 1; }     // CLAMP_TO_EDGE
        if (s.integer("wrapT", 10497) == 33071) { t.v_min = 0; t.v_max = (int32_t)image.h - 1; }
    }
    return t;
}
rtc_texture dummy_texture(HostScene& out, const float* pixel, int channels)
{
    out.texels.emplace_back(pixel, pixel + channels);
    rtc_texture t;
    t.texels = out.texels.back().data(); t.w = 1; t.h = 1; t.channels = channels;
    t.u_min = 0; t.u_max = 0; t.v_min = 0; t.v_max = 0;
    return t;
}
const Image& image_of(const Asset& a, const Json& texture)
{
    if (!texture.has("source")) fail(RTC_ERR_FORMAT, "glTF: texture without source");
    const long long src = texture.integer("source", -1);
    if (src < 0 || (size_t)src >= a.images.size()) fail(RTC_ERR_FORMAT, "glTF: texture.source out of range");
    return a.images[(size_t)src];
}
// loadColorTexture (src/stage1.zig:427-446): rgb of the image times the factor, or a 1x1 texture of the factor.
// Deferred mode: the texture's texels are left null and a recipe (image, kind, constants) is recorded instead.
struct Recipes {
    std::vector<RtcTexelRecipe>* list = nullptr;
    uint32_t material = 0, slot = 0;
    void add(const Asset& a, const Image& im, uint32_t kind, const float k[3]) const
    {
        RtcTexelRecipe r;
        r.material = material; r.slot = slot; r.image = (uint32_t)(&im - a.images.data()); r.kind = kind;
        r.k[0] = k[0]; r.k[1] = k[1]; r.k[2] = k[2];
        list->push_back(r);
    }
};

rtc_texture color_texture(const Asset& a, HostScene& out, const Json* info, const float factor[3], const Recipes& rec)
{
    if (!info) return dummy_texture(out, factor, 3);
    const Json& texture = item(a.doc, "textures", info->integer("index", -1));
    const Image& im = image_of(a, texture);
    if (rec.list) { rec.add(a, im, RTC_TEXEL_RGB, factor); return make_texture(a, texture, im, nullptr, 3); }
    const size_t n = (size_t)im.w * im.h;
    std::vector<float> px(n * 3);
    for (size_t i = 0; i < n; ++i)
        for (int k = 0; k < 3; ++k) px[3 * i + (size_t)k] = im.data[4 * i + (size_t)k] * factor[k];
    out.texels.push_back(std::move(px));
    return make_texture(a, texture, im, out.texels.back().data(), 3);
}
// loadTransparencyTexture (src/stage1.zig:448-472).
rtc_texture transparency_texture(const Asset& a, HostScene& out, const Json& material, const Recipes& rec)
{
    const float one = 1.0f;
    const std::string mode = material.string("alphaMode", "OPAQUE");
    const Json* pbr = material.find("pbrMetallicRoughness");
    const Json* info = pbr ? pbr->find("baseColorTexture") : nullptr;
    if (mode != "OPAQUE" && info) {
        const Json& texture = item(a.doc, "textures", info->integer("index", -1));
        const Image& im = image_of(a, texture);
        if (im.actual_c == 4 || im.actual_c == 2) {
            const float cutoff = (float)material.number("alphaCutoff", 0.5);
            if (rec.list) {
                const float k[3] = { cutoff, 0.0f, 0.0f };
                rec.add(a, im, mode == "MASK" ? RTC_TEXEL_ALPHA_MASK : RTC_TEXEL_ALPHA, k);
                return make_texture(a, texture, im, nullptr, 1);
            }
            const size_t n = (size_t)im.w * im.h;
            std::vector<float> px(n);
            for (size_t i = 0; i < n; ++i) {
                const float alpha = im.data[4 * i + 3];
                px[i] = mode == "MASK" ? (alpha > cutoff ? 1.0f : 0.0f) : alpha;
            }
            out.texels.push_back(std::move(px));
            return make_texture(a, texture, im, out.texels.back().data(), 1);
        }
    }
    return dummy_texture(out, &one, 1);
}
// loadMaterials (src/stage1.zig:474-496). metallic/roughness, normal maps, emissive_strength: ignored there too.
void load_materials(const Asset& a, HostScene& out, std::vector<RtcTexelRecipe>* recipes = nullptr)
{
    Recipes rec;
    rec.list = recipes;
    for (const Json& m : arr_of(a.doc, "materials").arr) {
        rec.material = (uint32_t)out.materials.size();
        float base[3] = { 1.0f, 1.0f, 1.0f }, emis[3] = { 0.0f, 0.0f, 0.0f };
        const Json* pbr = m.find("pbrMetallicRoughness");
        if (pbr)
            if (const Json* f = pbr->find("baseColorFactor"))
                for (size_t i = 0; i < 3 && i < f->arr.size(); ++i) base[i] = (float)f->arr[i].num;
        if (const Json* f = m.find("emissiveFactor"))
            for (size_t i = 0; i < 3 && i < f->arr.size(); ++i) emis[i] = (float)f->arr[i].num;
        rtc_material mat;
        rec.slot = 0; mat.base_color = color_texture(a, out, pbr ? pbr->find("baseColorTexture") : nullptr, base, rec);
        rec.slot = 1; mat.emissive = color_texture(a, out, m.find("emissiveTexture"), emis, rec);
        rec.slot = 2; mat.transparency = transparency_texture(a, out, m, rec);
        out.materials.push_back(mat);
    }
}

template <typename F>
int guarded(F&& f)
{
    try {
        f();
        return RTC_OK;
    } catch (const HostError& e) {
        return rtc_set_error_(e.code, e.msg.c_str());
    } catch (const std::bad_alloc&) {
        return rtc_set_error_(RTC_ERR_OUT_OF_MEMORY, "host allocation failed");
    } catch (const std::exception& e) {
        return rtc_set_error_(RTC_ERR_FORMAT, e.what());
    }
}

}  // namespace

// rtc_gltf_plan.h
int rtc_gltf_plan_(const char* path, const char* camera_name, uint32_t width, uint32_t height, int num_threads,
                   RtcGltfPlan* out)
{
    return guarded([&] {
        Asset a;
        a.defer_texels = true;
        load_asset(path, num_threads, a);
        HostScene hs;
        load_camera(a, camera_name, width, height, &out->camera);     // order of src/main.zig:109-112
        load_materials(a, hs, &out->recipes);
        plan_geometry(a, *out);
        linear_lut(out->lut);
        out->images8.resize(a.images.size());
        for (size_t i = 0; i < a.images.size(); ++i) {
            out->images8[i].w = a.images[i].w; out->images8[i].h = a.images[i].h;
            out->images8[i].rgba = std::move(a.images[i].rgba8);
        }
        out->materials = std::move(hs.materials);
        out->texels = std::move(hs.texels);          // moving the outer vector keeps the inner data() pointers
        out->num_images = (uint32_t)a.images.size();
        out->buffers = std::move(a.buffers);
    });
}

extern "C" {

int rtc_gltf_load(const char* path, const char* camera_name, uint32_t width, uint32_t height, int num_threads,
                  rtc_host_scene* out)
{
    if (!path || !out) return rtc_set_error_(RTC_ERR_INVALID_ARGUMENT, "rtc_gltf_load: null argument");
    memset(out, 0, sizeof *out);
    return guarded([&] {
        const bool trace = getenv("RTC_TRACE_LOAD") != nullptr;
        auto now = [] { return std::chrono::steady_clock::now(); };
        auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
        const auto t0 = now();
        Asset a;
        load_asset(path, num_threads, a);
        const auto t1 = now();
        std::unique_ptr<HostScene> hs(new HostScene());
        load_camera(a, camera_name, width, height, &out->camera);     // order of src/main.zig:109-112
        load_materials(a, *hs);
        const auto t2 = now();
        load_geometry(a, *hs, num_threads);
        const auto t3 = now();
        if (trace) fprintf(stderr, "rtc_gltf_load: files+images %.1f ms, camera+materials %.1f ms, triangles %.1f ms\n", ms(t0, t1), ms(t1, t2), ms(t2, t3));
        out->positions = hs->positions.data();
        out->tri_data = hs->tri_data.data();
        out->num_tris = (uint32_t)hs->tri_data.size();
        out->materials = hs->materials.data();
        out->num_materials = (uint32_t)hs->materials.size();
        out->num_images = (uint32_t)a.images.size();
        out->owner = hs.release();
    });
}

void rtc_host_scene_free(rtc_host_scene* scene)
{
    if (!scene) return;
    delete static_cast<HostScene*>(scene->owner);
    memset(scene, 0, sizeof *scene);
}

// Config.load (src/main.zig:63-68): std.json.parseFromSlice(Config, ...) with default options — every key
// required (num_threads may be null), unknown keys rejected.
int rtc_config_load(const char* path, rtc_config* out)
{
    if (!path || !out) return rtc_set_error_(RTC_ERR_INVALID_ARGUMENT, "rtc_config_load: null argument");
    return guarded([&] {
        const RtcBytes file = read_file(path);
        const Json doc = rtc_host::JsonParser((const char*)file.data(), file.size()).parse();
        if (doc.kind != Json::Object) fail(RTC_ERR_FORMAT, "config.json: not an object");
        for (const auto& kv : doc.obj)
            if (kv.first != "grid_resolution" && kv.first != "num_threads" && kv.first != "num_samples" && kv.first != "max_bounce")
                fail(RTC_ERR_FORMAT, "config.json: unknown key '" + kv.first + "'");
        for (const char* key : { "grid_resolution", "num_threads", "num_samples", "max_bounce" })
            if (!doc.find(key)) fail(RTC_ERR_FORMAT, std::string("config.json: missing key '") + key + "'");
        const Json& g = doc.at("grid_resolution");
        if (g.kind != Json::Array || g.arr.size() != 3) fail(RTC_ERR_FORMAT, "config.json: grid_resolution needs 3 integers");
        auto uint_of = [](const Json& j, const char* what, double max) -> uint32_t {
            if (j.kind != Json::Number || j.num < 0 || j.num > max || j.num != floor(j.num)) fail(RTC_ERR_FORMAT, std::string("config.json: bad ") + what);
            return (uint32_t)j.num;
        };
        for (int i = 0; i < 3; ++i) out->grid_resolution[i] = uint_of(g.arr[(size_t)i], "grid_resolution", 4294967295.0);
        const Json& nt = doc.at("num_threads");
        out->num_threads = nt.kind == Json::Null ? 0u : uint_of(nt, "num_threads", 255.0);       // ?u8
        out->num_samples = (uint16_t)uint_of(doc.at("num_samples"), "num_samples", 65535.0);      // u16
        out->max_bounce = (uint16_t)uint_of(doc.at("max_bounce"), "max_bounce", 65535.0);         // u16
    });
}

int rtc_png_write(const char* path, uint32_t w, uint32_t h, const uint8_t* rgb)
{
    if (!path || !rgb || w == 0 || h == 0) return rtc_set_error_(RTC_ERR_INVALID_ARGUMENT, "rtc_png_write: bad argument");
    return guarded([&] {
        int threads = (int)std::thread::hardware_concurrency();
        if (const char* e = getenv("RTC_PNG_THREADS")) threads = atoi(e);
        // RTC_PNG_STB=1: the file stbi_write_png would write, byte for byte (single-threaded); default: same pixels, zlib
        // slices on all threads
        const char* stb_mode = getenv("RTC_PNG_STB");
        const std::vector<uint8_t> png = stb_mode && stb_mode[0] == '1' ? rtc_host::png_encode_rgb8_stb(w, h, rgb)
                                                                        : rtc_host::png_encode_rgb8(w, h, rgb, threads);
        FILE* f = fopen(path, "wb");
        if (!f) fail(RTC_ERR_IO, std::string("WritePngFail: cannot open '") + path + "'");
        const size_t n = fwrite(png.data(), 1, png.size(), f);
        if (fclose(f) != 0 || n != png.size()) fail(RTC_ERR_IO, std::string("WritePngFail: short write to '") + path + "'");
    });
}

int rtc_png_read(const char* path, uint32_t* w, uint32_t* h, uint8_t** rgb_out)
{
    if (!path || !w || !h || !rgb_out) return rtc_set_error_(RTC_ERR_INVALID_ARGUMENT, "rtc_png_read: null argument");
    return guarded([&] {
        const RtcBytes file = read_file(path);
        rtc_host::DecodedImage img;
        try { img = rtc_host::png_decode(file.data(), file.size()); }
        catch (const std::runtime_error& e) { fail(RTC_ERR_FORMAT, e.what()); }
        uint8_t* rgb = (uint8_t*)malloc((size_t)img.w * img.h * 3);
        if (!rgb) throw std::bad_alloc();
        for (size_t i = 0; i < (size_t)img.w * img.h; ++i) memcpy(rgb + 3 * i, img.rgba.data() + 4 * i, 3);
        *w = img.w; *h = img.h; *rgb_out = rgb;
    });
}

}  // extern "C"
