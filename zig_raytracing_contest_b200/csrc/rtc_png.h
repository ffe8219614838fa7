// rtc_png.h — PNG in and out for the host side, over zlib.
//   png_encode_rgb8 : what the reference asks of stb at src/main.zig:130-135 — stbi_write_png(path, w, h, 3,
//                     img, 3*w): an 8-bit RGB, non-interlaced PNG (pixels identical; the deflate stream is zlib's,
//                     produced in 1 MiB slices on all host threads).
//   png_decode      : what stage1 asks of stb at src/stage1.zig:56-63 (stbi_loadf_from_memory(..., 4)), for
//                     PNG inputs: 8-bit RGBA out + the file's own channel count (`actual_c`). 16-bit samples keep
//                     their high byte (stb's 16 -> 8 conversion); palette + tRNS expand; colour keys; Adam7 interlace.
#pragma once
#include <zlib.h>

#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
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

// ---- the reference's own file, byte for byte ---------------------------------------------------------------------
// stbi_write_png (stb_image_write.h:1092-1229) restated: per row the filter (none / sub / up / average / paeth; the
// first row's variants) with the smallest sum of |signed byte|, first winner kept; then stb's own deflate
// (stb_image_write.h:859-1021): one fixed-Huffman block, 16384 hash buckets keyed on 3 bytes, each bucket holding at
// most 2*quality positions (the older half dropped when full), the LAST longest match of at least 3 bytes inside the
// 32 KiB window, abandoned for a literal when the next byte starts a longer one; stored blocks if that came out
// larger. Pinned against the reference's unit in tests/test_ref_stb.py. Single-threaded by construction; the default
// writer (zlib slices on all threads) produces the same pixels in a fraction of the time.
inline std::vector<uint8_t> stb_zlib_compress(const uint8_t* data, int n, int quality)
{
    static const uint16_t len_base[] = { 3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258, 259 };
    static const uint8_t len_extra[] = { 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0 };
    static const uint16_t dist_base[] = { 1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145,
                                          8193, 12289, 16385, 24577, 32768 };
    static const uint8_t dist_extra[] = { 0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13 };
    const int kBuckets = 16384;
    if (quality < 5) quality = 5;
    std::vector<uint8_t> out = { 0x78, 0x5e };
    uint32_t bitbuf = 0;
    int bitcount = 0;
    auto add = [&](uint32_t code, int nbits) {
        bitbuf |= code << bitcount; bitcount += nbits;
        while (bitcount >= 8) { out.push_back((uint8_t)bitbuf); bitbuf >>= 8; bitcount -= 8; }
    };
    auto rev = [](int code, int nbits) { int r = 0; while (nbits--) { r = (r << 1) | (code & 1); code >>= 1; } return r; };
    auto symbol = [&](int v) {                             // the fixed literal/length code of RFC 1951 3.2.6
        if (v <= 143) add((uint32_t)rev(0x30 + v, 8), 8);
        else if (v <= 255) add((uint32_t)rev(0x190 + v - 144, 9), 9);
        else if (v <= 279) add((uint32_t)rev(v - 256, 7), 7);
        else add((uint32_t)rev(0xc0 + v - 280, 8), 8);
    };
    auto hash3 = [&](int i) {
        uint32_t h = (uint32_t)data[i] + ((uint32_t)data[i + 1] << 8) + ((uint32_t)data[i + 2] << 16);
        h ^= h << 3; h += h >> 5; h ^= h << 4; h += h >> 17; h ^= h << 25; h += h >> 6;
        return (int)(h & (uint32_t)(kBuckets - 1));
    };
    auto run = [&](int a, int b, int limit) { int k = 0; while (k < limit && k < 258 && data[a + k] == data[b + k]) ++k; return k; };
    add(1, 1);                                             // BFINAL
    add(1, 2);                                             // BTYPE = fixed Huffman
    std::vector<std::vector<int>> buckets((size_t)kBuckets);
    int i = 0;
    while (i < n - 3) {
        std::vector<int>& bk = buckets[(size_t)hash3(i)];
        int best = 3, where = -1;
        for (int pos : bk)
            if (pos > i - 32768) { const int d = run(pos, i, n - i); if (d >= best) { best = d; where = pos; } }
        if ((int)bk.size() == 2 * quality) bk.erase(bk.begin(), bk.begin() + quality);
        bk.push_back(i);
        if (where >= 0) {
            const std::vector<int>& nx = buckets[(size_t)hash3(i + 1)];
            for (int pos : nx)
                if (pos > i - 32767 && run(pos, i + 1, n - i - 1) > best) { where = -1; break; }
        }
        if (where >= 0) {
            const int d = i - where;
            int j = 0;
            while (best > len_base[j + 1] - 1) ++j;
            symbol(j + 257);
            if (len_extra[j]) add((uint32_t)(best - len_base[j]), len_extra[j]);
            j = 0;
            while (d > dist_base[j + 1] - 1) ++j;
            add((uint32_t)rev(j, 5), 5);
            if (dist_extra[j]) add((uint32_t)(d - dist_base[j]), dist_extra[j]);
            i += best;
        } else {
            symbol(data[i]);
            ++i;
        }
    }
    for (; i < n; ++i) symbol(data[i]);
    symbol(256);
    while (bitcount) add(0, 1);
    if ((long long)out.size() > (long long)n + 2 + ((n + 32766) / 32767) * 5) {          // stored blocks came out smaller
        out.resize(2);
        for (int j = 0; j < n;) {
            int len = n - j;
            if (len > 32767) len = 32767;
            out.push_back((uint8_t)(n - j == len));
            out.push_back((uint8_t)len); out.push_back((uint8_t)(len >> 8));
            out.push_back((uint8_t)~len); out.push_back((uint8_t)(~len >> 8));
            out.insert(out.end(), data + j, data + j + len);
            j += len;
        }
    }
    be32(out, (uint32_t)adler32(1L, data, (uInt)n));
    return out;
}

inline std::vector<uint8_t> png_encode_rgb8_stb(uint32_t w, uint32_t h, const uint8_t* rgb)
{
    const int n = 3, row = (int)w * n;
    if ((uint64_t)h * ((uint64_t)row + 1) > 0x7fffffffull) throw std::runtime_error("PNG: frame too large for the stb-compatible writer");
    std::vector<uint8_t> filt((size_t)h * ((size_t)row + 1));
    std::vector<int8_t> line((size_t)row), keep((size_t)row);
    auto paeth = [](int a, int b, int c) {
        const int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
        return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
    };
    auto encode = [&](uint32_t y, int filter, int8_t* o) {
        static const int first_row[5] = { 0, 1, 0, 5, 6 };
        const int type = y ? filter : first_row[filter];
        const uint8_t* z = rgb + (size_t)y * row;
        const uint8_t* up = z - row;
        if (type == 0) { memcpy(o, z, (size_t)row); return; }
        for (int i = 0; i < n; ++i)
            switch (type) {
                case 2: o[i] = (int8_t)(z[i] - up[i]); break;
                case 3: o[i] = (int8_t)(z[i] - (up[i] >> 1)); break;
                case 4: o[i] = (int8_t)(z[i] - paeth(0, up[i], 0)); break;
                default: o[i] = (int8_t)z[i]; break;
            }
        for (int i = n; i < row; ++i)
            switch (type) {
                case 1: o[i] = (int8_t)(z[i] - z[i - n]); break;
                case 2: o[i] = (int8_t)(z[i] - up[i]); break;
                case 3: o[i] = (int8_t)(z[i] - ((z[i - n] + up[i]) >> 1)); break;
                case 4: o[i] = (int8_t)(z[i] - paeth(z[i - n], up[i], up[i - n])); break;
                case 5: o[i] = (int8_t)(z[i] - (z[i - n] >> 1)); break;
                default: o[i] = (int8_t)(z[i] - paeth(z[i - n], 0, 0)); break;
            }
    };
    for (uint32_t y = 0; y < h; ++y) {
        int best = 0, best_val = 0x7fffffff;
        for (int f = 0; f < 5; ++f) {
            encode(y, f, line.data());
            int est = 0;
            for (int i = 0; i < row; ++i) est += abs((int)line[(size_t)i]);
            if (est < best_val) { best_val = est; best = f; keep.swap(line); }
        }
        uint8_t* dst = filt.data() + (size_t)y * ((size_t)row + 1);
        dst[0] = (uint8_t)best;
        memcpy(dst + 1, keep.data(), (size_t)row);
    }
    const std::vector<uint8_t> z = stb_zlib_compress(filt.data(), (int)filt.size(), 8);
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
    if (interlace > 1) throw std::runtime_error("PNG: bad interlace method");
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
    if ((uint64_t)img.w * img.h > ((uint64_t)1 << 30)) throw std::runtime_error("PNG: image too large");
    const size_t bpp_bits = (size_t)ch * depth, bpp = (bpp_bits + 7) / 8;
    // the sub-images: one for a plain file, the seven Adam7 passes otherwise (each filtered on its own)
    struct Pass { uint32_t x0, y0, dx, dy, w, h; };
    std::vector<Pass> passes;
    if (!interlace) passes.push_back(Pass{ 0, 0, 1, 1, img.w, img.h });
    else {
        static const uint32_t xo[7] = { 0, 4, 0, 2, 0, 1, 0 }, yo[7] = { 0, 0, 4, 0, 2, 0, 1 }, xs[7] = { 8, 8, 4, 4, 2, 2, 1 }, ys[7] = { 8, 8, 8, 4, 4, 2, 2 };
        for (int k = 0; k < 7; ++k) {
            const uint32_t pw = (img.w - xo[k] + xs[k] - 1) / xs[k], ph = (img.h - yo[k] + ys[k] - 1) / ys[k];
            if (img.w > xo[k] && img.h > yo[k] && pw && ph) passes.push_back(Pass{ xo[k], yo[k], xs[k], ys[k], pw, ph });
        }
    }
    size_t raw_size = 0;
    for (const Pass& ps : passes) raw_size += (((size_t)ps.w * bpp_bits + 7) / 8 + 1) * ps.h;
    std::vector<uint8_t> raw(raw_size);
    uLongf rn = (uLongf)raw.size();
    if (uncompress(raw.data(), &rn, idat.data(), (uLong)idat.size()) != Z_OK || rn != raw.size())
        throw std::runtime_error("PNG: inflate failed");
    img.rgba.resize((size_t)img.w * img.h * 4);
    img.actual_c = ctype == 3 ? (trns.empty() ? 3 : 4) : ch + ((ctype == 0 || ctype == 2) && !trns.empty() ? 1 : 0);
    const int scale = depth < 8 ? 255 / ((1 << depth) - 1) : 1;          // stb's depth scale: 0xff, 0x55, 0x11
    auto key16 = [&](size_t k) { return trns.size() >= 2 * k + 2 ? ((int)trns[2 * k] << 8) | trns[2 * k + 1] : -1; };
    uint8_t* base = raw.data();
    for (const Pass& ps : passes) {
        const size_t stride = ((size_t)ps.w * bpp_bits + 7) / 8;
        std::vector<uint8_t> prev(stride, 0);
        for (uint32_t y = 0; y < ps.h; ++y) {
            uint8_t* row = base + (size_t)y * (stride + 1);
            const int f = row[0];
            uint8_t* cur = row + 1;
            for (size_t i = 0; i < stride; ++i) {                       // undo the row filter in place
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
            auto raw_sample = [&](size_t idx) -> int {                  // idx-th sample of the row at the file's depth
                if (depth == 8) return cur[idx];
                if (depth == 16) return ((int)cur[2 * idx] << 8) | cur[2 * idx + 1];
                const int per = 8 / depth, shift = (per - 1 - (int)(idx % per)) * depth;
                return (cur[idx / per] >> shift) & ((1 << depth) - 1);
            };
            auto to8 = [&](int v) -> uint8_t { return (uint8_t)(depth == 16 ? v >> 8 : v * scale); };
            for (uint32_t x = 0; x < ps.w; ++x) {
                uint8_t* o = img.rgba.data() + (((size_t)ps.y0 + (size_t)y * ps.dy) * img.w + ps.x0 + (size_t)x * ps.dx) * 4;
                if (ctype == 3) {
                    const size_t i = (size_t)raw_sample(x);
                    if (3 * i + 2 >= plte.size()) throw std::runtime_error("PNG: palette index out of range");
                    o[0] = plte[3 * i]; o[1] = plte[3 * i + 1]; o[2] = plte[3 * i + 2];
                    o[3] = i < trns.size() ? trns[i] : 255;
                } else if (ch <= 2) {
                    const int g = raw_sample((size_t)x * ch);
                    o[0] = o[1] = o[2] = to8(g);
                    o[3] = ch == 2 ? to8(raw_sample((size_t)x * 2 + 1)) : 255;
                    // colour key (tRNS of a grey image): compared at 16 bits, else on the low byte at the scaled 8-bit value
                    if (ch == 1 && key16(0) >= 0 && (depth == 16 ? g == key16(0) : to8(g) == (uint8_t)((key16(0) & 255) * scale))) o[3] = 0;
                } else {
                    const int r = raw_sample((size_t)x * ch), g = raw_sample((size_t)x * ch + 1), b = raw_sample((size_t)x * ch + 2);
                    o[0] = to8(r); o[1] = to8(g); o[2] = to8(b);
                    o[3] = ch == 4 ? to8(raw_sample((size_t)x * 4 + 3)) : 255;
                    if (ch == 3 && key16(2) >= 0 &&
                        (depth == 16 ? (r == key16(0) && g == key16(1) && b == key16(2))
                                     : (r == (key16(0) & 255) && g == (key16(1) & 255) && b == (key16(2) & 255)))) o[3] = 0;
                }
            }
        }
        base += (stride + 1) * ps.h;
    }
    return img;
}

}  // namespace rtc_host
