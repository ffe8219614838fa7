// rtc_gltf_plan.h — what the glTF loader (rtc_gltf.cpp, host C++) hands to the device assembly
// (rtc_build.cu): everything loadGeometry (src/stage1.zig:217-259) needs, with the per-triangle work left out.
// Internal to librtc.so.
#pragma once
#include "../../include/rtc.h"

#include <cstddef>
#include <cstdint>
#include <memory>
#include <new>
#include <utility>
#include <vector>

// Bytes of a file or a glTF buffer: owned (base64 payloads) or a window into a read-only mapping of the file
// (no read() into a zero-filled vector, no copy of a .glb's binary chunk). Reads like a vector.
struct RtcBytes {
    std::vector<uint8_t> own;
    std::shared_ptr<const void> map;       // keeps the mapping alive; null = `own` holds the bytes
    const uint8_t* p = nullptr;
    size_t n = 0;
    RtcBytes() = default;
    RtcBytes(std::vector<uint8_t>&& v) : own(std::move(v)) {}
    const uint8_t* data() const { return map ? p : own.data(); }
    size_t size() const { return map ? n : own.size(); }
    bool empty() const { return size() == 0; }
    RtcBytes window(size_t off, size_t len) const       // of a mapping: shares it; of owned bytes: a copy
    {
        RtcBytes r;
        if (map) { r.map = map; r.p = p + off; r.n = len; }
        else r.own.assign(own.begin() + (ptrdiff_t)off, own.begin() + (ptrdiff_t)(off + len));
        return r;
    }
};

// One mesh primitive of one node, in the reference's visiting order (nodes in file order, primitives in mesh
// order). Offsets are bytes from the start of the device blob the buffers are uploaded into (RtcGltfPlan::
// buffer_base + bufferView.byteOffset + accessor.byteOffset); the accessors were range-checked by the loader.
// POD: the table is copied to the device as it is.
struct RtcPrimRef {
    float    m[16];                // global node transform, m[column*4 + row] (zgltf Mat4)
    uint64_t off_pos, off_nrm, off_uv, off_ind;
    uint32_t stride_pos, stride_nrm, stride_uv, stride_ind;
    uint32_t count_pos, count_nrm, count_uv;       // vertices the accessors hold (index range check)
    uint32_t index_component;      // 5121 u8 / 5123 u16 / 5125 u32
    uint32_t has_uv;               // TEXCOORD_0 present; absent = texcoords read as zeroes
    uint32_t material;
    uint32_t first_tri, num_tris;  // position in the soup
};

// vector::resize without the zero fill: the soup arrays are hundreds of MB that are overwritten right away, and the
// first touch of their pages then happens in the worker threads that fill them.
template <class T>
struct RtcNoInit : std::allocator<T> {
    template <class U> struct rebind { using other = RtcNoInit<U>; };
    template <class U> void construct(U* p) { ::new ((void*)p) U; }
    template <class U, class... A> void construct(U* p, A&&... a) { ::new ((void*)p) U(std::forward<A>(a)...); }
};

// What an rtc_host_scene's `owner` points at (rtc_host_scene_free deletes it).
struct RtcHostSceneOwner {
    std::vector<float, RtcNoInit<float>> positions;
    std::vector<rtc_tri_data, RtcNoInit<rtc_tri_data>> tri_data;
    std::vector<rtc_material> materials;
    std::vector<std::vector<float>> texels;      // owners of rtc_texture.texels
};

// A texture whose texels are produced on the device from a decoded 8-bit image (rtc_build.cu: k_texels):
//   RGB         rgb = lut[c] * k   (stb's pow(c/255.0f, 2.2f) table, times the material's factor; stage1.zig:427-446)
//   ALPHA       a = c/255          (alphaMode BLEND; stage1.zig:448-472)
//   ALPHA_MASK  a > k[0] ? 1 : 0   (alphaMode MASK, k[0] = alphaCutoff)
// The rtc_material's texture has texels == nullptr; w, h, channels and the clamp range are set.
enum { RTC_TEXEL_RGB = 0, RTC_TEXEL_ALPHA = 1, RTC_TEXEL_ALPHA_MASK = 2 };
struct RtcTexelRecipe {
    uint32_t material, slot;       // slot: 0 base colour, 1 emissive, 2 transparency
    uint32_t image, kind;
    float k[3];
};
struct RtcImage8 {
    uint32_t w = 0, h = 0;
    std::vector<uint8_t> rgba;     // w*h*4
};

struct RtcGltfPlan {
    rtc_camera camera{};
    std::vector<RtcBytes> buffers;                 // the asset's buffers (moved out of the loader)
    std::vector<uint64_t> buffer_base;             // per buffer: offset in the device blob, or ~0 = not referenced
    uint64_t blob_bytes = 0;
    std::vector<RtcPrimRef> prims;                 // num_tris > 0 only
    uint32_t num_tris = 0;
    std::vector<rtc_material> materials;
    std::vector<std::vector<float>> texels;        // owners of materials[].*.texels (1x1 constants)
    std::vector<RtcTexelRecipe> recipes;           // the textures that come from images
    std::vector<RtcImage8> images8;
    float lut[256];                                // stb's ldr -> linear table (computed in double on the host)
    uint32_t num_images = 0;
};

// loadGltfFile + loadCamera + loadMaterials and the per-primitive part of loadGeometry. Returns an RTC_* code
// (error text set).
int rtc_gltf_plan_(const char* path, const char* camera_name, uint32_t width, uint32_t height, int num_threads,
                   RtcGltfPlan* out);
