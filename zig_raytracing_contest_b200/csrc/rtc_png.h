// rtc_png.h — PNG in and out for the host side, over zlib.
//   png_encode_rgb8 : what the reference asks of stb at src/main.zig:130-135 — stbi_write_png(path, w, h, 3,
//                     img, 3*w): an 8-bit RGB, non-interlaced PNG (pixels identical; the deflate stream is zlib's,
//                     produced in 1 MiB slices on all host threads).
//   png_decode      : what stage1 asks of stb at src/stage1.zig:56-63 (stbi_loadf_from_memory(..., 4)), for
//                     PNG inputs: 8-bit RGBA out + the file's own channel count (`actual_c`). 16-bit samples keep
//                     their high byte (stb's 16 -> 8 conversion); palette + tRNS expand; Adam7 is not supported.
#pragma once
#include <zlib.h>

#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

namespace rtc_host {

inline void be32(std::vector<uint8_t>& v, uint32_t x) { v.push_back(x >> 24); v.push_back(x >> 16); v.push_back(x >> 8); v.push_back(x); }
inline uint32_t rd32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

inline void png_chunk(std::vector<uint8_t>& out, const char type[4], const uint8_t* data, size_t n)
{
    be32(out, (uint32_t)n);
    const size_t at = out.size();
    out.insert(out.end(), type, type + 4);
    if (n) out.insert(out.end(), data, data + n);
    be32(out, (uint32_t)crc32(0L, out.data() + at, (uInt)(n + 4)));
}

// The zlib stream of `raw`, deflated in independent slices of kPngSlice bytes on up to `threads` threads: every slice
// but the last ends in a sync flush (byte-aligned, BFINAL = 0), so the concatenation is one valid deflate stream
// (the pigz construction, without dictionary priming); the Adler-32 trailer is combined from the slices'. The
// slice size is fixed, so the bytes do not depend on the thread count. stb's encoder (stb_image_write.h:1128-1229)
// is a single-threaded deflate of the whole frame — seconds at 4K; the pixels a decoder gets back are the same.
static constexpr size_t kPngSlice = (size_t)1 << 20;
inline std::vector<uint8_t> zlib_parallel(const uint8_t* raw, size_t n, int threads)
{
    const size_t slices = n ? (n + kPngSlice - 1) / kPngSlice : 1;
    std::vector<std::vector<uint8_t>> parts(slices);
    std::vector<uLong> adler(slices, 1);
    std::atomic<size_t> next{ 0 };
    std::atomic<bool> failed{ false };
    auto work = [&] {
        for (size_t i = next.fetch_add(1); i < slices; i = next.fetch_add(1)) {
            const size_t off = i * kPngSlice, len = n - off < kPngSlice ? n - off : kPngSlice;
            z_stream zs;
            memset(&zs, 0, sizeof zs);
            if (deflateInit2(&zs, 6, Z_DEFLATED, -15, 8, Z_DEFAULT_STRATEGY) != Z_OK) { failed.store(true); return; }
            std::vector<uint8_t>& out = parts[i];
            out.resize(deflateBound(&zs, (uLong)len) + 16);
            zs.next_in = const_cast<Bytef*>(raw + off); zs.avail_in = (uInt)len;
            zs.next_out = out.data(); zs.avail_out = (uInt)out.size();
            const int rc = deflate(&zs, i + 1 == slices ? Z_FINISH : Z_SYNC_FLUSH);
            if ((i + 1 == slices && rc != Z_STREAM_END) || (i + 1 != slices && (rc != Z_OK || zs.avail_in != 0 || zs.avail_out == 0))) failed.store(true);
            out.resize(out.size() - zs.avail_out);
            deflateEnd(&zs);
            adler[i] = adler32(1L, raw + off, (uInt)len);
        }
    };
    const size_t T = (size_t)(threads < 1 ? 1 : threads) < slices ? (size_t)(threads < 1 ? 1 : threads) : slices;
    if (T <= 1) work();
    else {
        std::vector<std::thread> pool;
        for (size_t t = 0; t < T; ++t) pool.emplace_back(work);
        for (auto& th : pool) th.join();
    }
    if (failed.load()) throw std::runtime_error("PNG: deflate failed");
    std::vector<uint8_t> z = { 0x78, 0x9C };
    uLong a = 1;
    for (size_t i = 0; i < slices; ++i) {
        z.insert(z.end(), parts[i].begin(), parts[i].end());
        const size_t off = i * kPngSlice, len = n - off < kPngSlice ? n - off : kPngSlice;
        a = i == 0 ? adler[0] : adler32_combine(a, adler[i], (z_off_t)len);
    }
    be32(z, (uint32_t)a);
    return z;
}

inline std::vector<uint8_t> png_encode_rgb8(uint32_t w, uint32_t h, const uint8_t* rgb, int threads = 1)
{
    std::vector<uint8_t> raw((size_t)h * (3 * (size_t)w + 1));
    for (uint32_t y = 0; y < h; ++y) {
        uint8_t* row = raw.data() + (size_t)y * (3 * (size_t)w + 1);
        row[0] = 0;                                               // filter type None
        memcpy(row + 1, rgb + (size_t)y * 3 * w, 3 * (size_t)w);
    }
    const std::vector<uint8_t> z = zlib_parallel(raw.data(), raw.size(), threads);
    std::vector<uint8_t> out = { 0x89, 'P', 'N', 'G', '\r', '\n', 0x1A, '\n' };
    std::vector<uint8_t> ihdr;
    be32(ihdr, w); be32(ihdr, h);
    ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
    png_chunk(out, "IHDR", ihdr.data(), ihdr.size());
    png_chunk(out, "IDAT", z.data(), z.size());
    png_chunk(out, "IEND", nullptr, 0);
    return out;
}

struct DecodedImage {
    uint32_t w = 0, h = 0;
    int actual_c = 0;                 // channels in the file: 1 grey, 2 grey+alpha, 3 rgb, 4 rgba
    std::vector<uint8_t> rgba;        // w*h*4
};

inline int paeth(int a, int b, int c)
{
    const int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

inline DecodedImage png_decode(const uint8_t* data, size_t n)
{
    static const uint8_t sig[8] = { 0x89, 'P', 'N', 'G', '\r', '\n', 0x1A, '\n' };
    if (n < 8 || memcmp(data, sig, 8) != 0) throw std::runtime_error("image is not a PNG (this host decodes PNG textures only)");
    DecodedImage img;
    int depth = 0, ctype = 0, interlace = 0;
    std::vector<uint8_t> idat, plte, trns;
    size_t p = 8;
    while (p + 12 <= n) {
        const uint32_t len = rd32(data + p);
        const char* type = (const char*)data + p + 4;
        const uint8_t* body = data + p + 8;
        if (p + 12 + (size_t)len > n) throw std::runtime_error("PNG: truncated chunk");
        if (!memcmp(type, "IHDR", 4)) {
            if (len < 13) throw std::runtime_error("PNG: bad IHDR");
            img.w = rd32(body); img.h = rd32(body + 4);
            depth = body[8]; ctype = body[9]; interlace = body[12];
        } else if (!memcmp(type, "PLTE", 4)) plte.assign(body, body + len);
        else if (!memcmp(type, "tRNS", 4)) trns.assign(body, body + len);
        else if (!memcmp(type, "IDAT", 4)) idat.insert(idat.end(), body, body + len);
        else if (!memcmp(type, "IEND", 4)) break;
        p += 12 + (size_t)len;
    }
    if (img.w == 0 || img.h == 0) throw std::runtime_error("PNG: missing IHDR");
    if (interlace) throw std::runtime_error("PNG: Adam7-interlaced images are not supported");
    int ch;
    switch (ctype) {
        case 0: ch = 1; break;
        case 2: ch = 3; break;
        case 3: ch = 1; break;
        case 4: ch = 2; break;
        case 6: ch = 4; break;
        default: throw std::runtime_error("PNG: bad colour type");
    }
    if (!(depth == 8 || depth == 16 || ((ctype == 0 || ctype == 3) && (depth == 1 || depth == 2 || depth == 4))))
        throw std::runtime_error("PNG: unsupported bit depth");
    const size_t bpp_bits = (size_t)ch * depth, stride = (img.w * bpp_bits + 7) / 8, bpp = (bpp_bits + 7) / 8;
    std::vector<uint8_t> raw((stride + 1) * img.h);
    uLongf rn = (uLongf)raw.size();
    if (uncompress(raw.data(), &rn, idat.data(), (uLong)idat.size()) != Z_OK || rn != raw.size())
        throw std::runtime_error("PNG: inflate failed");
    // undo the per-row filters in place
    std::vector<uint8_t> prev(stride, 0);
    for (uint32_t y = 0; y < img.h; ++y) {
        uint8_t* row = raw.data() + (size_t)y * (stride + 1);
        const int f = row[0];
        uint8_t* cur = row + 1;
        for (size_t i = 0; i < stride; ++i) {
            const int a = i >= bpp ? cur[i - bpp] : 0, b = prev[i], c = i >= bpp ? prev[i - bpp] : 0;
            int v = cur[i];
            switch (f) {
                case 0: break;
                case 1: v += a; break;
                case 2: v += b; break;
                case 3: v += (a + b) >> 1; break;
                case 4: v += paeth(a, b, c); break;
                default: throw std::runtime_error("PNG: bad filter");
            }
            cur[i] = (uint8_t)v;
        }
        memcpy(prev.data(), cur, stride);
    }
    img.rgba.resize((size_t)img.w * img.h * 4);
    auto sample = [&](const uint8_t* cur, size_t idx) -> int {          // idx-th sample of the row, as 8 bits
        if (depth == 8) return cur[idx];
        if (depth == 16) return cur[2 * idx];
        const int per = 8 / depth, shift = (per - 1 - (int)(idx % per)) * depth;
        const int v = (cur[idx / per] >> shift) & ((1 << depth) - 1);
        return ctype == 3 ? v : v * 255 / ((1 << depth) - 1);
    };
    img.actual_c = ctype == 3 ? (trns.empty() ? 3 : 4) : ch + ((ctype == 0 || ctype == 2) && !trns.empty() ? 1 : 0);
    for (uint32_t y = 0; y < img.h; ++y) {
        const uint8_t* cur = raw.data() + (size_t)y * (stride + 1) + 1;
        uint8_t* o = img.rgba.data() + (size_t)y * img.w * 4;
        for (uint32_t x = 0; x < img.w; ++x, o += 4) {
            if (ctype == 3) {
                const size_t i = (size_t)sample(cur, x);
                if (3 * i + 2 >= plte.size()) throw std::runtime_error("PNG: palette index out of range");
                o[0] = plte[3 * i]; o[1] = plte[3 * i + 1]; o[2] = plte[3 * i + 2];
                o[3] = i < trns.size() ? trns[i] : 255;
            } else if (ch <= 2) {
                const int g = sample(cur, (size_t)x * ch);
                o[0] = o[1] = o[2] = (uint8_t)g;
                o[3] = ch == 2 ? (uint8_t)sample(cur, (size_t)x * 2 + 1) : 255;
                if (ch == 1 && trns.size() >= 2 && depth == 8 && g == trns[1]) o[3] = 0;
            } else {
                o[0] = (uint8_t)sample(cur, (size_t)x * ch); o[1] = (uint8_t)sample(cur, (size_t)x * ch + 1);
                o[2] = (uint8_t)sample(cur, (size_t)x * ch + 2);
                o[3] = ch == 4 ? (uint8_t)sample(cur, (size_t)x * 4 + 3) : 255;
                if (ch == 3 && trns.size() >= 6 && depth == 8 && o[0] == trns[1] && o[1] == trns[3] && o[2] == trns[5]) o[3] = 0;
            }
        }
    }
    return img;
}

}  // namespace rtc_host
