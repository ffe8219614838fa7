// rtc_bvh.cuh — the opt-in RTC_ACCEL_BVH traversal: an 8-wide BVH with quantised child boxes (80-byte nodes fetched as
// three 256-bit loads) over the scene's source triangles, walked per lane with a short stack. It replaces the reference's
// uniform grid + 3D-DDA (Scene.traceRay / Grid.Iterator, src/stage3.zig:152-186, src/linalg.zig:443-497) as the
// structure that finds the candidate triangles; the triangle test itself is the reference's
// Triangle.rayIntersection (src/linalg.zig:696-722) — rtc_device.cuh tri_ray, the very function the grid path runs —
// so t, u, v of a hit are the same bits whichever structure found it. What differs from the grid:
//   * exact-t ties between two triangles go to the smaller source-triangle id (the grid: first cell along the ray,
//     then source order inside the cell);
//   * a hit whose point the grid's binning (SAT, linalg.zig:500-563) or crossing arithmetic places in a cell the
//     ray does not visit is found here and not there (and vice versa for a box test that fails by rounding).
// Both are the "enumerated near-ties" of SURVEY §8(c); tests/test_gpu_bvh.py counts them against the oracle.
//
// The functions are __host__ __device__: tests/native/bvh_check.cu runs the same node test and step function on the
// CPU against a brute-force loop over all triangles.
//
// NODE (80 B of content = 5 x 16 B, stored at a 96-byte stride so that a node is exactly three 32-byte sectors and is
// fetched with three 256-bit loads (LDG.256, sm_100) instead of five 128-bit loads that touch 3-4 sectors; after
// Ylitie, Karras, Laine, "Efficient Incoherent Ray Traversal on GPUs Through Compressed
// Wide BVHs", HPG 2017 — the published layout, written from the paper):
//   q0: p.x p.y p.z | ex, ey, ez, imask        origin of the quantisation frame, biased exponents of the per-axis
//                                              scale 2^e, bit s of imask = slot s holds an inner node
//   q1: child_base | tri_base | meta[0..3] | meta[4..7]
//        inner child in slot s : meta = 001 11sss (0x20 | 24 + s); its node index = child_base + number of inner
//                                children in the slots below s
//        leaf child            : meta = (unary triangle count: 001, 011, 111) << 5 | offset of its first triangle
//                                from tri_base (0..23)
//        empty slot            : 0
//   q2: qlo.x[8] qlo.y[8]   q3: qlo.z[8] qhi.x[8]   q4: qhi.y[8] qhi.z[8]     child box = p + q * 2^e, 8 bits each
// Slots are assigned by the builder so that slot ^ octant of the ray is roughly front-to-back (slot bit a set =
// the child lies towards +a): the hit mask keeps inner children at bit 24 + (slot ^ octinv), octinv = 7 ^ octant,
// and the highest set bit is the nearest child.
#pragma once
#include "rtc_trav.cuh"

namespace rtc {

struct BvhView {
    uint32_t num_nodes;
    const float4* nodes;       // kBvhNodeQuads float4 per node (5 used + 1 pad: 96 B, sector-aligned); node 0 = root
    const float4* tris;        // kBvhTriQuads float4 per triangle in BVH order: {v0 | src id} {e1} {e2} {pad}: 64 B, two sectors
};
static constexpr uint32_t kBvhNodeQuads = 6;
static constexpr uint32_t kBvhTriQuads = 4;

// 32 bytes in one request: one sector, one L1 tag lookup (the records above are laid out for it)
struct Quad2 { float4 lo, hi; };
RTC_HD Quad2 bvh_ld256(const float4* p)
{
    Quad2 r;
#if defined(__CUDA_ARCH__) && !defined(RTC_BVH_NO_LD256)
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(r.lo.x), "=f"(r.lo.y), "=f"(r.lo.z), "=f"(r.lo.w), "=f"(r.hi.x), "=f"(r.hi.y), "=f"(r.hi.z), "=f"(r.hi.w)
        : "l"(p));
#else
    r.lo = RTC_LDG(p); r.hi = RTC_LDG(p + 1);
#endif
    return r;
}

struct BvhRay {
    float ox, oy, oz;
    float ix, iy, iz;          // 1 / d, with |d| below 2^-80 replaced by +-2^-80 (no 0 * inf in the slab test);
                               // the sign of 1/d is the sign of d (-0.0 included), which is all the slab test needs of d
    uint32_t octinv4;          // (7 ^ octant) replicated into the four bytes
};

static constexpr uint32_t kBvhStack = 48;      // the builder caps the tree depth below this

RTC_HD float bvh_u2f(uint32_t u)
{
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, sizeof f); return f;
#endif
}
RTC_HD uint32_t bvh_clz(uint32_t x)
{
#ifdef __CUDA_ARCH__
    return (uint32_t)__clz((int)x);
#else
    return x ? (uint32_t)__builtin_clz(x) : 32u;
#endif
}
RTC_HD uint32_t bvh_popc(uint32_t x)
{
#ifdef __CUDA_ARCH__
    return (uint32_t)__popc(x);
#else
    return (uint32_t)__builtin_popcount(x);
#endif
}
RTC_HD float bvh_fma(float a, float b, float c)
{
#ifdef __CUDA_ARCH__
    return __fmaf_rn(a, b, c);
#else
    return fmaf(a, b, c);
#endif
}

RTC_HD void bvh_ray_init(BvhRay& r, v3 o, v3 d)
{
    r.ox = o.x; r.oy = o.y; r.oz = o.z;
    const float tiny = bvh_u2f((127u - 80u) << 23);
    const float ax = fabsf(d.x) > tiny ? d.x : copysignf(tiny, d.x);
    const float ay = fabsf(d.y) > tiny ? d.y : copysignf(tiny, d.y);
    const float az = fabsf(d.z) > tiny ? d.z : copysignf(tiny, d.z);
    r.ix = 1.0f / ax; r.iy = 1.0f / ay; r.iz = 1.0f / az;
    const uint32_t oct = (r.ix < 0.0f ? 1u : 0u) | (r.iy < 0.0f ? 2u : 0u) | (r.iz < 0.0f ? 4u : 0u);
    r.octinv4 = (7u ^ oct) * 0x01010101u;
}

// Byte j of w as a float. RTC_BVH_CVT = 0: the conversion instruction (I2F.U8 with its byte selector: one issue slot, on
// the conversion pipe); 1: one byte permute into the mantissa of 2^23 and one FADD (exact: 2^23 + q - 2^23 = q; two
// issue slots on the ALU and FMA pipes); 2: kept for A/B — the far planes by permute, the near planes by conversion.
#ifndef RTC_BVH_CVT
#define RTC_BVH_CVT 0
#endif
RTC_HD float bvh_byte_f(uint32_t w, int j, bool far_plane)
{
#if defined(__CUDA_ARCH__) && RTC_BVH_CVT
    if (RTC_BVH_CVT == 1 || far_plane)
        return __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7650u | (uint32_t)j)) - 8388608.0f;
#endif
    (void)far_plane;
    return (float)((w >> (8 * j)) & 0xFFu);
}

// Slab test of the ray against the (up to) eight quantised child boxes of node ni, clipped to [0, tmax]. Fills the
// node group {child_base, inner hits << 24 | imask} and the triangle group {tri_base, one bit per triangle of the
// leaf children that were hit}.
// `top` (may be null): the first `ntop` nodes — breadth-first order, i.e. the top levels of the tree, which every ray
// visits — as 5 float4 each in shared memory (k_trace_bvh stages them once per CTA): no L1 tag lookups for them.
RTC_HD void bvh_node_hits(const float4* __restrict__ nodes, const float4* top, uint32_t ntop, uint32_t ni, const BvhRay& r, float tmax,
                          uint2& ng, uint2& tg)
{
    float4 q0, q1, q2, q3, q4;
    if (ni < ntop) {
        const float4* n = top + 5u * ni;
        q0 = n[0]; q1 = n[1]; q2 = n[2]; q3 = n[3]; q4 = n[4];
    } else {
        const float4* n = nodes + kBvhNodeQuads * (size_t)ni;
        const Quad2 n01 = bvh_ld256(n), n23 = bvh_ld256(n + 2), n45 = bvh_ld256(n + 4);
        q0 = n01.lo; q1 = n01.hi; q2 = n23.lo; q3 = n23.hi; q4 = n45.lo;
    }
    const uint32_t eim = f_bits(q0.w);
    const float sx = bvh_u2f((eim & 0xFFu) << 23) * r.ix;
    const float sy = bvh_u2f(((eim >> 8) & 0xFFu) << 23) * r.iy;
    const float sz = bvh_u2f(((eim >> 16) & 0xFFu) << 23) * r.iz;
    const float bx = (q0.x - r.ox) * r.ix, by = (q0.y - r.oy) * r.iy, bz = (q0.z - r.oz) * r.iz;
    const bool nx = r.ix < 0.0f, ny = r.iy < 0.0f, nz = r.iz < 0.0f;
    uint32_t hitmask = 0u;
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
    for (int half = 0; half < 2; ++half) {
        const uint32_t meta4 = f_bits(half ? q1.w : q1.z);
        const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
        const uint32_t inner_mask4 = (is_inner4 >> 4) * 0xFFu;
        const uint32_t bit_index4 = (meta4 ^ (r.octinv4 & inner_mask4)) & 0x1F1F1F1Fu;
        const uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;
        const uint32_t lox = f_bits(half ? q2.y : q2.x), loy = f_bits(half ? q2.w : q2.z), loz = f_bits(half ? q3.y : q3.x);
        const uint32_t hix = f_bits(half ? q3.w : q3.z), hiy = f_bits(half ? q4.y : q4.x), hiz = f_bits(half ? q4.w : q4.z);
        const uint32_t nearx = nx ? hix : lox, farx = nx ? lox : hix;
        const uint32_t neary = ny ? hiy : loy, fary = ny ? loy : hiy;
        const uint32_t nearz = nz ? hiz : loz, farz = nz ? loz : hiz;
#ifdef __CUDA_ARCH__
#pragma unroll
#endif
        for (int j = 0; j < 4; ++j) {
            const int sh = 8 * j;
            const float t0x = bvh_fma(bvh_byte_f(nearx, j, false), sx, bx), t1x = bvh_fma(bvh_byte_f(farx, j, true), sx, bx);
            const float t0y = bvh_fma(bvh_byte_f(neary, j, false), sy, by), t1y = bvh_fma(bvh_byte_f(fary, j, true), sy, by);
            const float t0z = bvh_fma(bvh_byte_f(nearz, j, false), sz, bz), t1z = bvh_fma(bvh_byte_f(farz, j, true), sz, bz);
            const float t0 = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, 0.0f));
            const float t1 = fminf(fminf(t1x, t1y), fminf(t1z, tmax));
            if (t0 <= t1) hitmask |= ((child_bits4 >> sh) & 0xFFu) << ((bit_index4 >> sh) & 0xFFu);
        }
    }
    ng.x = f_bits(q1.x); ng.y = (hitmask & 0xFF000000u) | (eim >> 24);
    tg.x = f_bits(q1.y); tg.y = hitmask & 0x00FFFFFFu;
}

// Per-ray traversal state. stack[k * stride] is entry k (the kernel interleaves the lanes of a warp).
struct BvhTrav {
    uint2 ng;                  // current node group
    uint32_t sp;
    float best_t;
    uint32_t best_tri;         // index into BvhView::tris (BVH order) or 0xFFFFFFFF
};
RTC_HD void bvh_trav_begin(BvhTrav& s)
{
    s.ng = make_uint2(0u, 0x80000000u);       // the root as the only child of a virtual group
    s.sp = 0u;
    s.best_t = f_inf();
    s.best_tri = 0xFFFFFFFFu;
}

// The reference's acceptance `nearest.t > hit.t and hit.t > 0` (stage3.zig:174); on an exact tie the smaller source id.
RTC_HD void bvh_test_tri(const BvhView& b, const Ray& ray, uint32_t ti, BvhTrav& s)
{
    const float4* p = b.tris + kBvhTriQuads * (size_t)ti;
    const float4 a = RTC_LDG(p), e1 = RTC_LDG(p + 1), e2 = RTC_LDG(p + 2);
    float t, u, v;
    if (!tri_ray(mk(a.x, a.y, a.z), mk(e1.x, e1.y, e1.z), mk(e2.x, e2.y, e2.z), ray, t, u, v) || !(t > 0.0f)) return;
    if (s.best_t > t) { s.best_t = t; s.best_tri = ti; return; }
    if (s.best_t == t && s.best_tri != 0xFFFFFFFFu) {
        const uint32_t cur = f_bits(RTC_LDG(b.tris + kBvhTriQuads * (size_t)s.best_tri).w);
        if (f_bits(a.w) < cur) s.best_tri = ti;
    }
}

// The traversal stack behind put / get: a plain strided array here; k_trace_bvh may keep its first entries in shared memory.
struct BvhArrayStack {
    uint2* p; uint32_t stride;
    RTC_HD void put(uint32_t k, uint2 v) { p[k * stride] = v; }
    RTC_HD uint2 get(uint32_t k) const { return p[k * stride]; }
};

// One iteration in two halves, so that the kernel can put the triangle tests of a whole warp into one pool between
// them. bvh_step_nodes: pop the nearest pending child (or a postponed triangle group) and test its children; returns
// the triangle group {tri_base, one bit per triangle of the leaves that were hit}. bvh_step_pop: take the next group
// off the stack when the current one is used up; returns true when the ray is done.
// nodes counts what the STATS build reports (80-byte node records fetched).
template <bool STATS, class Stack>
RTC_HD uint2 bvh_step_nodes(const BvhView& b, const BvhRay& r, BvhTrav& s, Stack& stack, uint32_t& nodes,
                            const float4* top = nullptr, uint32_t ntop = 0u)
{
    uint2 tg;
    if (s.ng.y > 0x00FFFFFFu) {
        const uint32_t hits = s.ng.y;
        const uint32_t bit = 31u - bvh_clz(hits);
        s.ng.y &= ~(1u << bit);
        if (s.ng.y > 0x00FFFFFFu) { stack.put(s.sp, s.ng); s.sp += 1u; }
        const uint32_t slot = (bit - 24u) ^ (r.octinv4 & 7u);
        const uint32_t rel = bvh_popc(hits & ~(0xFFFFFFFFu << slot));
        if (STATS) nodes += 1u;
        bvh_node_hits(b.nodes, top, ntop, s.ng.x + rel, r, s.best_t, s.ng, tg);
    } else {
        tg = s.ng;
        s.ng = make_uint2(0u, 0u);
    }
    return tg;
}
template <class Stack>
RTC_HD bool bvh_step_pop(BvhTrav& s, const Stack& stack)
{
    if (s.ng.y <= 0x00FFFFFFu) {
        if (s.sp == 0u) return true;
        s.sp -= 1u;
        s.ng = stack.get(s.sp);
    }
    return false;
}

// The whole iteration for one ray on its own (the host-side checker, tests/native/bvh_check.cu).
template <bool STATS>
RTC_HD bool bvh_step(const BvhView& b, const BvhRay& r, const Ray& ray, BvhTrav& s, uint2* stack_mem, uint32_t stride,
                     uint32_t& nodes, uint32_t& tests)
{
    BvhArrayStack stack{ stack_mem, stride };
    uint2 tg = bvh_step_nodes<STATS>(b, r, s, stack, nodes);
    while (tg.y != 0u) {
        const uint32_t rel = 31u - bvh_clz(tg.y);
        tg.y &= ~(1u << rel);
        if (STATS) tests += 1u;
        bvh_test_tri(b, ray, tg.x + rel, s);
    }
    return bvh_step_pop(s, stack);
}

}  // namespace rtc
