// rtc_jpeg.h — JPEG textures for the loader: what stage1 gets from stb at src/stage1.zig:56-63
// (stbi_loadf_from_memory(..., 4)) when a glTF image is a JPEG, as 8-bit RGBA before linearisation.
//
// Entropy decoding is ITU T.81's (Huffman, sequential and progressive, restart intervals, interleaved and
// single-component scans); any conforming decoder produces the same coefficients. What decides the BYTES is the
// arithmetic after it, and that is restated from stb_image.h so that the texels are the reference's, bit for bit
// (pinned against the reference's own unit in tests/test_ref_stb.py):
//   dequantisation to 16-bit            stb_image.h:2208-2262, 3060-3096
//   the 8x8 inverse DCT                 :2427-2516  (jidctint "islow" at 12 fractional bits; 2 extra bits after the
//                                                    column pass, +128 and rounding folded into the row pass)
//   chroma upsampling                   :3448-3653  (h2, v2, h2v2 triangle filters centred as JFIF, else replication)
//   YCbCr -> RGB                        :3655-3682  (20-bit fixed point, the Cb term of green truncated to 16 bits)
// Supported: 8-bit, 1, 3 or 4 components (YCbCr, or RGB when an Adobe APP14 marker says so and there is no JFIF
// header; CMYK / YCCK by the Adobe transform flag). Arithmetic-coded, lossless and hierarchical files are reported
// as unsupported (stb rejects them too).
#pragma once
#include "rtc_png.h"

#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <vector>

namespace rtc_host {

inline bool is_jpeg(const uint8_t* d, size_t n) { return n >= 3 && d[0] == 0xFF && d[1] == 0xD8; }

namespace jpeg_detail {

static const uint8_t kZigzag[64 + 15] = {
    0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
    35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63,
    63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63 };        // the tail absorbs runs that overshoot a corrupt block

[[noreturn]] inline void bad(const char* what) { throw std::runtime_error(std::string("JPEG: ") + what); }

// Canonical Huffman table (T.81 annex C / F.2.2.3).
struct Huff {
    uint8_t values[256];
    int32_t maxcode[18];        // per length: largest code + 1, left-aligned comparisons are avoided by decoding bit by bit
    int32_t mincode[17], valptr[17];
    int16_t lookup[512];        // 9-bit prefix -> (length << 8 | symbol), or -1
    bool present = false;
    void build(const uint8_t counts[16], const uint8_t* syms, int total)
    {
        memcpy(values, syms, (size_t)total);
        int code = 0, k = 0;
        for (int len = 1; len <= 16; ++len) {
            valptr[len] = k; mincode[len] = code;
            code += counts[len - 1]; k += counts[len - 1];
            maxcode[len] = counts[len - 1] ? code - 1 : -1;
            if (code > (1 << len)) bad("bad code lengths");
            code <<= 1;
        }
        for (int i = 0; i < 512; ++i) lookup[i] = -1;
        for (int len = 1; len <= 9; ++len)
            for (int c = mincode[len]; maxcode[len] >= 0 && c <= maxcode[len]; ++c) {
                const int sym = values[valptr[len] + c - mincode[len]];
                for (int f = 0; f < (1 << (9 - len)); ++f) lookup[(c << (9 - len)) | f] = (int16_t)((len << 8) | sym);
            }
        present = true;
    }
};

struct Component {
    int id = 0, h = 1, v = 1, tq = 0, hd = 0, ha = 0, dc_pred = 0;
    int x = 0, y = 0, w2 = 0, h2 = 0, coeff_w = 0, coeff_h = 0;
    std::vector<uint8_t> data;          // w2 * h2 samples
    std::vector<int16_t> coeff;         // progressive: 64 per block
};

struct Decoder {
    const uint8_t* p;
    const uint8_t* end;
    uint32_t bits = 0;                  // left-aligned bit buffer
    int nbits = 0;
    int marker = -1;                    // a marker met inside entropy-coded data
    bool nomore = false;
    uint16_t dequant[4][64];
    Huff hdc[4], hac[4];
    Component comp[4];
    int ncomp = 0, width = 0, height = 0, hmax = 1, vmax = 1, mcu_x = 0, mcu_y = 0;
    bool progressive = false, jfif = false;
    int adobe_transform = -1, restart_interval = 0, todo = 0, eob_run = 0;
    int scan_n = 0, order[4], spec_start = 0, spec_end = 63, succ_high = 0, succ_low = 0;

    int get8() { return p < end ? *p++ : 0; }
    int get16() { const int a = get8(); return (a << 8) | get8(); }

    // ---- bit reader: 0xFF00 is a stuffed 0xFF, any other 0xFFxx ends the entropy-coded segment (zeros follow) ----
    void fill()
    {
        while (nbits <= 24) {
            int b = nomore ? 0 : get8();
            if (b == 0xFF) {
                int c = get8();
                while (c == 0xFF) c = get8();
                if (c != 0) { marker = c; nomore = true; b = 0; }
            }
            bits |= (uint32_t)b << (24 - nbits);
            nbits += 8;
        }
    }
    int getbits(int n)
    {
        if (n == 0) return 0;
        if (nbits < n) fill();
        const int v = (int)(bits >> (32 - n));
        bits <<= n; nbits -= n;
        return v;
    }
    int getbit() { return getbits(1); }
    int decode(const Huff& h)
    {
        if (nbits < 16) fill();
        const int16_t e = h.lookup[bits >> 23];
        if (e >= 0) { const int len = e >> 8; bits <<= len; nbits -= len; return e & 255; }
        int code = 0, len = 10;                           // longer than 9 bits: T.81 F.2.2.3, length by length
        for (; len <= 16; ++len) {
            code = (int)(bits >> (32 - len));
            if (code <= h.maxcode[len]) break;
        }
        if (len > 16) bad("bad huffman code");
        bits <<= len; nbits -= len;
        return h.values[h.valptr[len] + code - h.mincode[len]];
    }
    int extend(int n)                                     // T.81 F.2.2.1 RECEIVE + EXTEND
    {
        if (n == 0) return 0;
        const int v = getbits(n);
        return v < (1 << (n - 1)) ? v - (1 << n) + 1 : v;
    }
    void reset()
    {
        bits = 0; nbits = 0; nomore = false; marker = -1; eob_run = 0;
        for (int i = 0; i < 4; ++i) comp[i].dc_pred = 0;
        todo = restart_interval ? restart_interval : 0x7fffffff;
    }

    // ---- blocks ----
    void block_baseline(int16_t d[64], Component& c)
    {
        const Huff& dc = hdc[c.hd];
        const Huff& ac = hac[c.ha];
        const uint16_t* q = dequant[c.tq];
        memset(d, 0, 64 * sizeof(int16_t));
        const int t = decode(dc);
        if (t > 15) bad("bad DC size");
        c.dc_pred = (int)((uint32_t)c.dc_pred + (uint32_t)extend(t));          // wraps (corrupt files), never overflows
        d[0] = (int16_t)((uint32_t)c.dc_pred * (uint32_t)q[0]);
        for (int k = 1; k < 64;) {
            const int rs = decode(ac), s = rs & 15, r = rs >> 4;
            if (s == 0) {
                if (rs != 0xF0) break;
                k += 16;
            } else {
                k += r;
                const int zig = kZigzag[k++];
                d[zig] = (int16_t)(extend(s) * q[zig]);
            }
        }
    }
    void block_prog_dc(int16_t d[64], Component& c)
    {
        if (spec_end != 0) bad("DC scan with AC coefficients");
        if (succ_high == 0) {
            memset(d, 0, 64 * sizeof(int16_t));
            const int t = decode(hdc[c.hd]);
            if (t > 15) bad("bad DC size");
            c.dc_pred = (int)((uint32_t)c.dc_pred + (uint32_t)extend(t));
            d[0] = (int16_t)((uint32_t)c.dc_pred << succ_low);
        } else if (getbit()) d[0] = (int16_t)(d[0] + (int16_t)(1 << succ_low));
    }
    void block_prog_ac(int16_t d[64], Component& c)
    {
        if (spec_start == 0) bad("AC scan starting at the DC coefficient");
        const Huff& ac = hac[c.ha];
        if (succ_high == 0) {
            const int shift = succ_low;
            if (eob_run) { --eob_run; return; }
            for (int k = spec_start; k <= spec_end;) {
                const int rs = decode(ac), s = rs & 15, r = rs >> 4;
                if (s == 0) {
                    if (r < 15) {
                        eob_run = 1 << r;
                        if (r) eob_run += getbits(r);
                        --eob_run;
                        break;
                    }
                    k += 16;
                } else {
                    k += r;
                    const int zig = kZigzag[k++];
                    d[zig] = (int16_t)(extend(s) * (1 << shift));
                }
            }
        } else {
            const int16_t bit = (int16_t)(1 << succ_low);
            auto refine = [&](int16_t* c0) {
                if (getbit() && (*c0 & bit) == 0) *c0 = (int16_t)(*c0 > 0 ? *c0 + bit : *c0 - bit);
            };
            if (eob_run) {
                --eob_run;
                for (int k = spec_start; k <= spec_end; ++k) {
                    int16_t* c0 = &d[kZigzag[k]];
                    if (*c0 != 0) refine(c0);
                }
                return;
            }
            int k = spec_start;
            do {
                const int rs = decode(ac);
                int s = rs & 15, r = rs >> 4;
                if (s == 0) {
                    if (r < 15) {
                        eob_run = (1 << r) - 1;
                        if (r) eob_run += getbits(r);
                        r = 64;                              // run to the end of the band
                    }
                } else {
                    if (s != 1) bad("bad refinement code");
                    s = getbit() ? bit : -bit;
                }
                while (k <= spec_end) {
                    int16_t* c0 = &d[kZigzag[k++]];
                    if (*c0 != 0) refine(c0);
                    else {
                        if (r == 0) { *c0 = (int16_t)s; break; }
                        --r;
                    }
                }
            } while (k <= spec_end);
        }
    }

    // ---- 8x8 inverse DCT, stb_image.h:2427-2516 ----
    static int fx(double x) { return (int)(x * 4096 + 0.5); }
    // I = int for the column pass (16-bit inputs cannot overflow it); the row pass runs on its results, which a
    // corrupt file can drive beyond what 32-bit products hold, so it uses 64 bits (same values whenever stb's int
    // arithmetic does not overflow, defined behaviour when it would).
    template <class I> struct RowT { I x0, x1, x2, x3, t0, t1, t2, t3; };
    template <class I>
    static RowT<I> idct1d(I s0, I s1, I s2, I s3, I s4, I s5, I s6, I s7)
    {
        static const int k0541 = fx(0.5411961f), k1847 = fx(-1.847759065f), k0765 = fx(0.765366865f), k1175 = fx(1.175875602f),
                         k0298 = fx(0.298631336f), k2053 = fx(2.053119869f), k3072 = fx(3.072711026f), k1501 = fx(1.501321110f),
                         k0899 = fx(-0.899976223f), k2562 = fx(-2.562915447f), k1961 = fx(-1.961570560f), k0390 = fx(-0.390180644f);
        RowT<I> r;
        I p1 = (s2 + s6) * k0541;
        I t2 = p1 + s6 * k1847, t3 = p1 + s2 * k0765;
        I t0 = (s0 + s4) * 4096, t1 = (s0 - s4) * 4096;
        r.x0 = t0 + t3; r.x3 = t0 - t3; r.x1 = t1 + t2; r.x2 = t1 - t2;
        t0 = s7; t1 = s5; t2 = s3; t3 = s1;
        I p3 = t0 + t2, p4 = t1 + t3;
        p1 = t0 + t3;
        I p2 = t1 + t2;
        const I p5 = (p3 + p4) * k1175;
        t0 = t0 * k0298; t1 = t1 * k2053; t2 = t2 * k3072; t3 = t3 * k1501;
        p1 = p5 + p1 * k0899; p2 = p5 + p2 * k2562; p3 = p3 * k1961; p4 = p4 * k0390;
        r.t3 = t3 + p1 + p4; r.t2 = t2 + p2 + p3; r.t1 = t1 + p2 + p4; r.t0 = t0 + p1 + p3;
        return r;
    }
    static uint8_t clamp8(long long x) { return (uint8_t)(x < 0 ? 0 : (x > 255 ? 255 : x)); }
    static void idct(uint8_t* out, int stride, const int16_t d[64])
    {
        int val[64];
        for (int i = 0; i < 8; ++i) {
            const int16_t* c = d + i;
            int* v = val + i;
            if (c[8] == 0 && c[16] == 0 && c[24] == 0 && c[32] == 0 && c[40] == 0 && c[48] == 0 && c[56] == 0) {
                const int dc = c[0] * 4;
                v[0] = v[8] = v[16] = v[24] = v[32] = v[40] = v[48] = v[56] = dc;
            } else {
                RowT<int> r = idct1d<int>(c[0], c[8], c[16], c[24], c[32], c[40], c[48], c[56]);
                r.x0 += 512; r.x1 += 512; r.x2 += 512; r.x3 += 512;
                v[0] = (r.x0 + r.t3) >> 10; v[56] = (r.x0 - r.t3) >> 10;
                v[8] = (r.x1 + r.t2) >> 10; v[48] = (r.x1 - r.t2) >> 10;
                v[16] = (r.x2 + r.t1) >> 10; v[40] = (r.x2 - r.t1) >> 10;
                v[24] = (r.x3 + r.t0) >> 10; v[32] = (r.x3 - r.t0) >> 10;
            }
        }
        for (int i = 0; i < 8; ++i) {
            const int* v = val + 8 * i;
            uint8_t* o = out + (size_t)i * stride;
            RowT<long long> r = idct1d<long long>(v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7]);
            const long long bias = 65536 + (128 << 17);
            r.x0 += bias; r.x1 += bias; r.x2 += bias; r.x3 += bias;
            o[0] = clamp8((r.x0 + r.t3) >> 17); o[7] = clamp8((r.x0 - r.t3) >> 17);
            o[1] = clamp8((r.x1 + r.t2) >> 17); o[6] = clamp8((r.x1 - r.t2) >> 17);
            o[2] = clamp8((r.x2 + r.t1) >> 17); o[5] = clamp8((r.x2 - r.t1) >> 17);
            o[3] = clamp8((r.x3 + r.t0) >> 17); o[4] = clamp8((r.x3 - r.t0) >> 17);
        }
    }

    // ---- segments ----
    void read_dqt()
    {
        int len = get16() - 2;
        while (len > 0) {
            const int q = get8(), prec = q >> 4, t = q & 15;
            if ((prec != 0 && prec != 1) || t > 3) bad("bad DQT");
            for (int i = 0; i < 64; ++i) dequant[t][kZigzag[i]] = (uint16_t)(prec ? get16() : get8());
            len -= prec ? 129 : 65;
        }
        if (len != 0) bad("bad DQT length");
    }
    void read_dht()
    {
        int len = get16() - 2;
        while (len > 0) {
            const int q = get8(), tc = q >> 4, th = q & 15;
            if (tc > 1 || th > 3) bad("bad DHT");
            uint8_t counts[16], syms[256];
            int total = 0;
            for (int i = 0; i < 16; ++i) { counts[i] = (uint8_t)get8(); total += counts[i]; }
            if (total > 256) bad("bad DHT counts");
            for (int i = 0; i < total; ++i) syms[i] = (uint8_t)get8();
            (tc == 0 ? hdc[th] : hac[th]).build(counts, syms, total);
            len -= 17 + total;
        }
        if (len != 0) bad("bad DHT length");
    }
    void skip_segment() { const int len = get16(); if (len < 2) bad("bad segment length"); p = (end - p) < len - 2 ? end : p + (len - 2); }
    void read_app(int m)
    {
        int len = get16();
        if (len < 2) bad("bad APP length");
        len -= 2;
        if (m == 0xE0 && len >= 5) {
            static const char tag[5] = { 'J', 'F', 'I', 'F', 0 };
            bool ok = true;
            for (int i = 0; i < 5; ++i) ok &= get8() == (uint8_t)tag[i];
            len -= 5;
            if (ok) jfif = true;
        } else if (m == 0xEE && len >= 12) {
            static const char tag[6] = { 'A', 'd', 'o', 'b', 'e', 0 };
            bool ok = true;
            for (int i = 0; i < 6; ++i) ok &= get8() == (uint8_t)tag[i];
            len -= 6;
            if (ok) { get8(); get16(); get16(); adobe_transform = get8(); len -= 6; }
        }
        p = (end - p) < len ? end : p + len;
    }
    void read_sof()
    {
        const int len = get16();
        if (get8() != 8) bad("only 8-bit samples are supported");
        height = get16(); width = get16();
        if (height == 0 || width == 0) bad("zero image size");
        if (width > (1 << 24) || height > (1 << 24)) bad("image too large");
        ncomp = get8();
        if (ncomp != 1 && ncomp != 3 && ncomp != 4) bad("bad component count");
        if (len != 8 + 3 * ncomp) bad("bad SOF length");
        for (int i = 0; i < ncomp; ++i) {
            Component& c = comp[i];
            c.id = get8();
            const int q = get8();
            c.h = q >> 4; c.v = q & 15;
            if (!c.h || c.h > 4 || !c.v || c.v > 4) bad("bad sampling factor");
            c.tq = get8();
            if (c.tq > 3) bad("bad quantisation table index");
            if (c.h > hmax) hmax = c.h;
            if (c.v > vmax) vmax = c.v;
        }
        for (int i = 0; i < ncomp; ++i)
            if (hmax % comp[i].h != 0 || vmax % comp[i].v != 0) bad("sampling factors that do not divide the largest");
        mcu_x = (width + hmax * 8 - 1) / (hmax * 8);
        mcu_y = (height + vmax * 8 - 1) / (vmax * 8);
        for (int i = 0; i < ncomp; ++i) {
            Component& c = comp[i];
            c.x = (width * c.h + hmax - 1) / hmax;
            c.y = (height * c.v + vmax - 1) / vmax;
            c.w2 = mcu_x * c.h * 8; c.h2 = mcu_y * c.v * 8;
            c.data.assign((size_t)c.w2 * c.h2, 0);
            if (progressive) {
                c.coeff_w = c.w2 / 8; c.coeff_h = c.h2 / 8;
                c.coeff.assign((size_t)c.w2 * c.h2, 0);
            }
        }
    }
    void read_sos()
    {
        const int len = get16();
        scan_n = get8();
        if (scan_n < 1 || scan_n > ncomp) bad("bad SOS component count");
        if (len != 6 + 2 * scan_n) bad("bad SOS length");
        for (int i = 0; i < scan_n; ++i) {
            const int id = get8(), q = get8();
            int which = 0;
            for (; which < ncomp; ++which)
                if (comp[which].id == id) break;
            if (which == ncomp) bad("SOS names an unknown component");
            comp[which].hd = q >> 4; comp[which].ha = q & 15;
            if (comp[which].hd > 3 || comp[which].ha > 3) bad("bad huffman table index");
            order[i] = which;
        }
        spec_start = get8(); spec_end = get8();
        const int a = get8();
        succ_high = a >> 4; succ_low = a & 15;
        if (progressive) {
            if (spec_start > 63 || spec_end > 63 || spec_start > spec_end || succ_high > 13 || succ_low > 13) bad("bad SOS");
        } else {
            if (spec_start != 0 || succ_high != 0 || succ_low != 0) bad("bad SOS");
            spec_end = 63;
        }
    }

    // After each restart interval: RSTn expected; anything else ends the scan.
    bool restart()
    {
        if (--todo > 0) return true;
        if (nbits < 24) fill();
        if (marker < 0xD0 || marker > 0xD7) return false;
        reset();
        return true;
    }
    void need(const Huff& h) { if (!h.present) bad("missing huffman table"); }
    void scan()
    {
        reset();
        auto one = [&](Component& c, int bx, int by) {
            if (progressive) {
                int16_t* d = c.coeff.data() + 64 * ((size_t)bx + (size_t)by * c.coeff_w);
                if (spec_start == 0) { if (succ_high == 0) need(hdc[c.hd]); block_prog_dc(d, c); }
                else { need(hac[c.ha]); block_prog_ac(d, c); }
            } else {
                int16_t d[64];
                need(hdc[c.hd]); need(hac[c.ha]);
                block_baseline(d, c);
                idct(c.data.data() + (size_t)c.w2 * by * 8 + (size_t)bx * 8, c.w2, d);
            }
        };
        if (scan_n == 1) {                                // a single component: its own block grid, not the MCUs
            Component& c = comp[order[0]];
            const int w = (c.x + 7) >> 3, h = (c.y + 7) >> 3;
            for (int j = 0; j < h; ++j)
                for (int i = 0; i < w; ++i) {
                    one(c, i, j);
                    if (!restart()) return;
                }
        } else {
            for (int j = 0; j < mcu_y; ++j)
                for (int i = 0; i < mcu_x; ++i) {
                    for (int k = 0; k < scan_n; ++k) {
                        Component& c = comp[order[k]];
                        for (int y = 0; y < c.v; ++y)
                            for (int x = 0; x < c.h; ++x) one(c, i * c.h + x, j * c.v + y);
                    }
                    if (!restart()) return;
                }
        }
    }
    void finish_progressive()
    {
        for (int n = 0; n < ncomp; ++n) {
            Component& c = comp[n];
            const int w = (c.x + 7) >> 3, h = (c.y + 7) >> 3;
            const uint16_t* q = dequant[c.tq];
            for (int j = 0; j < h; ++j)
                for (int i = 0; i < w; ++i) {
                    int16_t* d = c.coeff.data() + 64 * ((size_t)i + (size_t)j * c.coeff_w);
                    for (int k = 0; k < 64; ++k) d[k] = (int16_t)(d[k] * q[k]);
                    idct(c.data.data() + (size_t)c.w2 * j * 8 + (size_t)i * 8, c.w2, d);
                }
        }
    }

    void run()
    {
        if (get8() != 0xFF || get8() != 0xD8) bad("no SOI");
        bool have_frame = false;
        for (;;) {
            int m;
            if (marker >= 0) { m = marker; marker = -1; }
            else {
                int b = get8();
                while (b != 0xFF) { if (p >= end) { m = 0xD9; goto got; } b = get8(); }
                do m = get8(); while (m == 0xFF);
            }
        got:
            if (m == 0xD9) break;
            if (m == 0xC0 || m == 0xC1 || m == 0xC2) {
                if (have_frame) bad("more than one frame");
                progressive = m == 0xC2;
                read_sof();
                have_frame = true;
            } else if (m == 0xC4) read_dht();
            else if (m == 0xDB) read_dqt();
            else if (m == 0xDD) { if (get16() != 4) bad("bad DRI"); restart_interval = get16(); }
            else if (m == 0xDA) {
                if (!have_frame) bad("scan before frame header");
                read_sos();
                scan();
                if (marker < 0) {                         // resynchronise on the next marker after the scan's data
                    while (p < end) {
                        if (*p == 0xFF && p + 1 < end && p[1] != 0 && p[1] != 0xFF && !(p[1] >= 0xD0 && p[1] <= 0xD7)) break;
                        ++p;
                    }
                }
                if (p >= end && marker < 0) break;
            } else if (m >= 0xE0 && m <= 0xEF) read_app(m);
            else if (m == 0xFE || (m >= 0xC5 && m <= 0xCF && m != 0xC8 && m != 0xCC)) {
                if (m != 0xFE) bad("arithmetic-coded, lossless and hierarchical files are not supported");
                skip_segment();
            } else if (m == 0x01 || (m >= 0xD0 && m <= 0xD7)) { /* standalone */ }
            else skip_segment();
            if (p >= end && marker < 0) break;
        }
        if (!have_frame) bad("no frame");
        if (progressive) finish_progressive();
    }
};

// stb_image.h:3448-3653
inline const uint8_t* upsample_row(std::vector<uint8_t>& buf, const uint8_t* near_, const uint8_t* far_, int w, int hs, int vs)
{
    uint8_t* out = buf.data();
    if (hs == 1 && vs == 1) return near_;
    if (hs == 1 && vs == 2) {
        for (int i = 0; i < w; ++i) out[i] = (uint8_t)((3 * near_[i] + far_[i] + 2) >> 2);
        return out;
    }
    if (hs == 2 && vs == 1) {
        const uint8_t* in = near_;
        if (w == 1) { out[0] = out[1] = in[0]; return out; }
        out[0] = in[0];
        out[1] = (uint8_t)((in[0] * 3 + in[1] + 2) >> 2);
        int i = 1;
        for (; i < w - 1; ++i) {
            const int n = 3 * in[i] + 2;
            out[i * 2] = (uint8_t)((n + in[i - 1]) >> 2);
            out[i * 2 + 1] = (uint8_t)((n + in[i + 1]) >> 2);
        }
        out[i * 2] = (uint8_t)((in[w - 2] * 3 + in[w - 1] + 2) >> 2);
        out[i * 2 + 1] = in[w - 1];
        return out;
    }
    if (hs == 2 && vs == 2) {
        if (w == 1) { out[0] = out[1] = (uint8_t)((3 * near_[0] + far_[0] + 2) >> 2); return out; }
        int t1 = 3 * near_[0] + far_[0];
        out[0] = (uint8_t)((t1 + 2) >> 2);
        for (int i = 1; i < w; ++i) {
            const int t0 = t1;
            t1 = 3 * near_[i] + far_[i];
            out[i * 2 - 1] = (uint8_t)((3 * t0 + t1 + 8) >> 4);
            out[i * 2] = (uint8_t)((3 * t1 + t0 + 8) >> 4);
        }
        out[w * 2 - 1] = (uint8_t)((t1 + 2) >> 2);
        return out;
    }
    for (int i = 0; i < w; ++i)
        for (int j = 0; j < hs; ++j) out[i * hs + j] = near_[i];
    return out;
}

}  // namespace jpeg_detail

// 8-bit RGBA (alpha 255) + the component count stb reports (3 or 1).
inline DecodedImage jpeg_decode(const uint8_t* data, size_t n)
{
    using namespace jpeg_detail;
    Decoder z;
    z.p = data; z.end = data + n;
    memset(z.dequant, 0, sizeof z.dequant);
    z.run();
    DecodedImage img;
    img.w = (uint32_t)z.width; img.h = (uint32_t)z.height;
    img.actual_c = z.ncomp >= 3 ? 3 : 1;
    img.rgba.resize((size_t)img.w * img.h * 4);
    const bool is_rgb = z.ncomp == 3 && (z.adobe_transform == 0 && !z.jfif);     // stb also honours an 'R','G','B' component-id triple
    bool rgb_ids = z.ncomp == 3 && z.comp[0].id == 'R' && z.comp[1].id == 'G' && z.comp[2].id == 'B';
    struct Res { int hs, vs, ystep, w_lores, ypos; const uint8_t* line0; const uint8_t* line1; std::vector<uint8_t> buf; };
    Res res[4];
    for (int k = 0; k < z.ncomp; ++k) {
        Res& r = res[k];
        r.hs = z.hmax / z.comp[k].h; r.vs = z.vmax / z.comp[k].v;
        r.ystep = r.vs >> 1;
        r.w_lores = (z.width + r.hs - 1) / r.hs;
        r.ypos = 0;
        r.line0 = r.line1 = z.comp[k].data.data();
        r.buf.resize((size_t)z.width + 8);
    }
    const int c_y = (int)(1.40200f * 4096.0f + 0.5f) << 8, c_g1 = -((int)(0.71414f * 4096.0f + 0.5f) << 8),
              c_g2 = -((int)(0.34414f * 4096.0f + 0.5f) << 8), c_b = (int)(1.77200f * 4096.0f + 0.5f) << 8;
    for (int j = 0; j < z.height; ++j) {
        const uint8_t* row[4] = { nullptr, nullptr, nullptr, nullptr };
        for (int k = 0; k < z.ncomp; ++k) {
            Res& r = res[k];
            const bool bot = r.ystep >= (r.vs >> 1);
            row[k] = upsample_row(r.buf, bot ? r.line1 : r.line0, bot ? r.line0 : r.line1, r.w_lores, r.hs, r.vs);
            if (++r.ystep >= r.vs) {
                r.ystep = 0;
                r.line0 = r.line1;
                if (++r.ypos < z.comp[k].y) r.line1 += z.comp[k].w2;
            }
        }
        uint8_t* out = img.rgba.data() + (size_t)j * img.w * 4;
        auto ycc = [&](int i, uint8_t* o) {
            const int yf = (row[0][i] << 20) + (1 << 19), cb = row[1][i] - 128, cr = row[2][i] - 128;
            int r = yf + cr * c_y;
            int g = yf + cr * c_g1 + (int)((uint32_t)(cb * c_g2) & 0xffff0000u);
            int b = yf + cb * c_b;
            r >>= 20; g >>= 20; b >>= 20;
            o[0] = Decoder::clamp8(r); o[1] = Decoder::clamp8(g); o[2] = Decoder::clamp8(b);
        };
        auto mul8 = [](int x, int y) { const unsigned t = (unsigned)(x * y + 128); return (uint8_t)((t + (t >> 8)) >> 8); };   // stb_image.h:3724
        for (int i = 0; i < z.width; ++i, out += 4) {
            if (z.ncomp == 1) { out[0] = out[1] = out[2] = row[0][i]; }
            else if (z.ncomp == 4) {                                   // stb_image.h:3974-3996: CMYK, YCCK, or YCbCr + an ignored plane
                const int k = row[3][i];
                if (z.adobe_transform == 0) { out[0] = mul8(row[0][i], k); out[1] = mul8(row[1][i], k); out[2] = mul8(row[2][i], k); }
                else {
                    ycc(i, out);
                    if (z.adobe_transform == 2) { out[0] = mul8(255 - out[0], k); out[1] = mul8(255 - out[1], k); out[2] = mul8(255 - out[2], k); }
                }
            }
            else if (is_rgb || rgb_ids) { out[0] = row[0][i]; out[1] = row[1][i]; out[2] = row[2][i]; }
            else ycc(i, out);
            out[3] = 255;
        }
    }
    return img;
}

}  // namespace rtc_host
