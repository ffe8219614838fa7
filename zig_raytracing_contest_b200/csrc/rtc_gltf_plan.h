// rtc_gltf_plan.h — what the glTF loader (rtc_gltf.cpp, host C++) hands to the device assembly
// (rtc_build.cu): everything loadGeometry (src/stage1.zig:217-259) needs, with the per-triangle work left out.
// Internal to librtc.so.
#pragma once
#include "../../include/rtc.h"

#include <cstdint>
#include <vector>

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

// What an rtc_host_scene's `owner` points at (rtc_host_scene_free deletes it).
struct RtcHostSceneOwner {
    std::vector<float> positions;
    std::vector<rtc_tri_data> tri_data;
    std::vector<rtc_material> materials;
    std::vector<std::vector<float>> texels;      // owners of rtc_texture.texels
};

struct RtcGltfPlan {
    rtc_camera camera{};
    std::vector<std::vector<uint8_t>> buffers;     // the asset's buffers (moved out of the loader)
    std::vector<uint64_t> buffer_base;             // per buffer: offset in the device blob, or ~0 = not referenced
    uint64_t blob_bytes = 0;
    std::vector<RtcPrimRef> prims;                 // num_tris > 0 only
    uint32_t num_tris = 0;
    std::vector<rtc_material> materials;
    std::vector<std::vector<float>> texels;        // owners of materials[].*.texels
    uint32_t num_images = 0;
};

// loadGltfFile + loadCamera + loadMaterials and the per-primitive part of loadGeometry. Returns an RTC_* code
// (error text set).
int rtc_gltf_plan_(const char* path, const char* camera_name, uint32_t width, uint32_t height, int num_threads,
                   RtcGltfPlan* out);
