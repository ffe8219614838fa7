// rtc_trav.cuh — Grid.Iterator (src/linalg.zig:471-497) in the register-lean form k_trace walks with, as
// __host__ __device__ code so that the same functions are checked on the CPU against the oracle's iterator
// (tests/native/trav_check.cu, tests/test_trav_twolevel.py).
//
// The cell travels as the packed word of rtc_device.cuh (low 6 bits = cell inside its 4x4x4 block, the rest =
// block index) next to the packed exit cell; t_delta carries the step direction in its sign bit. The
// t_next_crossing / t_delta arithmetic is the reference's, add for add.
//
// TWO LEVELS. Per axis the reference's crossing values are one fixed sequence t, t+d, (t+d)+d, ... (each a
// separate fp32 add), and the cells a ray visits are the 3-way merge of those sequences with the tie order
// z, y, x. Inside a block with no triangle nothing can happen except termination against a hit found earlier,
// and termination at ANY crossing of an empty block implies termination at the block's last crossing (the
// values only grow). So an empty block is taken in ONE step: per axis, the crossing that leaves the block is
// the axis' sequence advanced by the r in-block crossings that come before it (r sequential adds, r = 0..3 from
// the cell's position in the block); the smallest of the three (same tie order) is where the ray leaves; that
// axis' sequence moves on, the other two are left untouched — their in-block crossings that come before the exit
// are consumed lazily, when the ray next enters a block that has triangles (trav_enter). Every value produced is
// an element of the reference's sequences, so the cells with triangles are reached in the reference's order with
// the reference's t values, bit for bit. Requires resolutions that are multiples of 4 (otherwise the grid's
// boundary lies inside a block) and crossing values that are finite or +inf; rays that do not qualify carry
// kTravCellOnly in their exit word and walk cell by cell.
#pragma once
#include "rtc_device.cuh"

#include <string.h>

namespace rtc {

#define RTC_HD __host__ __device__ __forceinline__

#ifdef __CUDA_ARCH__
#define RTC_LDG(p) __ldg(p)
RTC_HD uint32_t f_bits(float f) { return __float_as_uint(f); }
RTC_HD float f_inf() { return __int_as_float(0x7f800000); }
#else
#define RTC_LDG(p) (*(p))
RTC_HD uint32_t f_bits(float f) { uint32_t u; memcpy(&u, &f, sizeof u); return u; }
RTC_HD float f_inf() { const uint32_t u = 0x7f800000u; float f; memcpy(&f, &u, sizeof f); return f; }
#endif

static constexpr uint32_t kTravCellOnly = 0x80000000u;      // flag in Trav::ec: never take an empty block in one step

struct Trav {
    float tx, ty, tz;                   // t_next_crossing
    float dx, dy, dz;                   // |.| = t_delta; sign bit set = stepping towards 0 on that axis
    uint32_t c, ec;                     // packed cell, packed exit cell (| kTravCellOnly)
    uint32_t fine_lo, fine_hi;          // occupancy mask of the block the cell lies in
};

// Occupancy of the block of cell s.c: the coarse bit (shared memory when CSM) decides whether the 8-byte fine
// mask is fetched at all.
template <bool CSM>
RTC_HD void trav_load_block(const DevScene& sc, const uint32_t* __restrict__ coarse, Trav& s)
{
    const uint32_t bi = s.c >> 6;
    const uint32_t w = CSM ? coarse[bi >> 5] : RTC_LDG(sc.coarse + (bi >> 5));
    uint2 f = make_uint2(0u, 0u);
    if ((w >> (bi & 31u)) & 1u) f = RTC_LDG(reinterpret_cast<const uint2*>(sc.fine) + bi);
    s.fine_lo = f.x; s.fine_hi = f.y;
}
RTC_HD bool trav_occupied(const Trav& s)
{
    const uint32_t w = (s.c & 32u) ? s.fine_hi : s.fine_lo;
    return ((w >> (s.c & 31u)) & 1u) != 0u;
}
// The ray is in a block without triangles and may take it in one step.
RTC_HD bool trav_block_mode(const Trav& s) { return (s.fine_lo | s.fine_hi | (s.ec & kTravCellOnly)) == 0u; }

// Iterator.next: axis = {2,1,2,1,2,2,0,0}[(tx<ty)<<2 | (tx<tz)<<1 | (ty<tz)] (linalg.zig:479-484):
// x iff tx<ty && tx<tz; y iff !(tx<ty) && ty<tz; otherwise z (ties prefer z, then y). Returns +inf without
// moving when the cell is the exit cell on that axis (linalg.zig:486-488). The coordinate of the axis is a
// scattered bit field m of the packed cell: with every other bit set, +1 carries across the gap
// ((c | ~m) + 1), and (c | ~m) + m == (c & m) - 1 is the decrement.
RTC_HD float trav_next(const DevScene& sc, Trav& s)
{
    const bool xy = s.tx < s.ty, xz = s.tx < s.tz, yz = s.ty < s.tz;
    const bool ax = xy && xz, ay = !xy && yz;
    const float t = ax ? s.tx : (ay ? s.ty : s.tz);
    const float d = ax ? s.dx : (ay ? s.dy : s.dz);
    const uint32_t m = ax ? sc.mask[0] : (ay ? sc.mask[1] : sc.mask[2]);
    if (((s.c ^ s.ec) & m) == 0u) return f_inf();
    const uint32_t stepped = (s.c | ~m) + ((int32_t)f_bits(d) < 0 ? m : 1u);
    s.c = (stepped & m) | (s.c & ~m);
    const float tn = t + fabsf(d);
    s.tx = ax ? tn : s.tx; s.ty = ay ? tn : s.ty; s.tz = ax ? s.tz : (ay ? s.tz : tn);
    return t;
}

// In-block crossings of one axis that come before the crossing that leaves the block: r = lo when stepping
// towards 0, 3 - lo otherwise (lo = the cell's coordinate inside its block).
RTC_HD uint32_t trav_inner(uint32_t lo, float d) { return (int32_t)f_bits(d) < 0 ? lo : 3u - lo; }
// t advanced by r = trav_inner(lo, d) sequential adds of |d|; the two bits of r = the bits of lo, flipped when
// stepping away from 0. `c` is the packed cell, `sh` the position of the axis' two in-block bits.
RTC_HD float trav_chain(float t, float d, uint32_t c, uint32_t sh)
{
    const bool pos = (int32_t)f_bits(d) >= 0;
    const bool b0 = (((c >> sh) & 1u) != 0u) != pos, b1 = (((c >> (sh + 1u)) & 1u) != 0u) != pos;
    const float a = fabsf(d);
    if (b0) t += a;
    if (b1) { t += a; t += a; }
    return t;
}

// One step of the walk, computed but not yet applied. Cell mode (bm = false): Iterator.next. Block mode: the ray
// leaves its (empty) block through the axis whose block-leaving crossing comes first. Returns false when the ray
// leaves the grid (Iterator.next returning +inf).
struct TravStep {
    float t;            // value of the crossing taken
    float tn;           // that axis' sequence advanced past it
    uint32_t c;         // the cell behind the crossing
    bool ax, ay;        // axis taken: x, y, else z
};
RTC_HD bool trav_peek(const DevScene& sc, const Trav& s, bool bm, TravStep& o)
{
    float X = s.tx, Y = s.ty, Z = s.tz;
    if (bm) {
        X = trav_chain(X, s.dx, s.c, 0u);
        Y = trav_chain(Y, s.dy, s.c, 2u);
        Z = trav_chain(Z, s.dz, s.c, 4u);
    }
    const bool xy = X < Y, xz = X < Z, yz = Y < Z;
    o.ax = xy && xz; o.ay = !xy && yz;
    o.t = o.ax ? X : (o.ay ? Y : Z);
    const float d = o.ax ? s.dx : (o.ay ? s.dy : s.dz);
    const uint32_t m = o.ax ? sc.mask[0] : (o.ay ? sc.mask[1] : sc.mask[2]);
    const uint32_t lo3 = bm ? (o.ax ? 3u : (o.ay ? 12u : 48u)) : 0u;
    // block mode compares block coordinates only: leaving the exit block on this axis is leaving the grid
    if (((s.c ^ s.ec) & m & ~lo3) == 0u) return false;
    const bool neg = (int32_t)f_bits(d) < 0;
    const uint32_t at_face = neg ? (s.c & ~lo3) : (s.c | lo3);      // block mode: the block's last cell on this axis
    const uint32_t stepped = (at_face | ~m) + (neg ? m : 1u);
    o.c = (stepped & m) | (s.c & ~m);
    o.tn = o.t + fabsf(d);
    return true;
}
RTC_HD void trav_commit(Trav& s, const TravStep& o)
{
    s.c = o.c;
    s.tx = o.ax ? o.tn : s.tx; s.ty = o.ay ? o.tn : s.ty; s.tz = o.ax ? s.tz : (o.ay ? s.tz : o.tn);
}

// The ray has just entered a block WITH triangles through `axis` (0/1/2) at crossing value T, coming from a block
// it took in one step. The other two axes still hold the state of when they last crossed a block boundary (or of
// the cell where block mode began): consume their in-block crossings that come before (T, axis) in the merge order
// — smaller value, or equal value and higher priority (z over y over x) — and move the cell accordingly.
RTC_HD void trav_enter_axis(float& t, float d, uint32_t& c, uint32_t shift, float T, bool tie_wins)
{
    const bool neg = (int32_t)f_bits(d) < 0;
    const uint32_t lo = (c >> shift) & 3u;
    const uint32_t r = neg ? lo : 3u - lo;                          // trav_inner
    const float a = fabsf(d);
    const float c0 = t, c1 = c0 + a, c2 = c1 + a, c3 = c2 + a;
    // the values only grow, so the crossings that come before (T, axis) are a prefix: count it, capped by r
    const uint32_t n = ((c0 < T || (tie_wins && c0 == T)) ? 1u : 0u) + ((c1 < T || (tie_wins && c1 == T)) ? 1u : 0u) +
                       ((c2 < T || (tie_wins && c2 == T)) ? 1u : 0u);
    const uint32_t k = n < r ? n : r;
    t = k == 0u ? c0 : (k == 1u ? c1 : (k == 2u ? c2 : c3));
    const uint32_t dk = k << shift;
    c = neg ? c - dk : c + dk;
}
// The two axes other than `axis`, picked with selects so that the arithmetic exists once per axis slot:
//   axis x: (y, wins ties against x), (z, wins);   axis y: (x, loses), (z, wins);   axis z: (x, loses), (y, loses)
RTC_HD void trav_enter(Trav& s, float T, uint32_t axis)
{
    const bool a0 = axis == 0u, a2 = axis == 2u;
    float tA = a0 ? s.ty : s.tx, tB = a2 ? s.ty : s.tz;
    const float dA = a0 ? s.dy : s.dx, dB = a2 ? s.dy : s.dz;
    trav_enter_axis(tA, dA, s.c, a0 ? 2u : 0u, T, a0);
    trav_enter_axis(tB, dB, s.c, a2 ? 2u : 4u, T, !a2);
    s.tx = a0 ? s.tx : tA;
    s.ty = a0 ? tA : (a2 ? tB : s.ty);
    s.tz = a2 ? s.tz : tB;
}

// One iteration of the walk for a ray standing in a cell it has finished with (empty, or its triangles tested):
// take the next crossing; the ray ends when the crossing is at or beyond its nearest hit (stage3.zig:179-182) or
// leads out of the grid. Returns kWalkDone, kWalkMoved, or kWalkEnter + axis: the ray came out of block mode into
// a block with triangles and trav_enter(s, T, axis) must run before its cell is looked at.
enum { kWalkMoved = 0u, kWalkDone = 1u, kWalkEnter = 2u };
template <bool CSM, bool TWO>
RTC_HD uint32_t trav_advance(const DevScene& sc, const uint32_t* __restrict__ coarse, Trav& s, float best_t, float& T)
{
    const bool bm = TWO && trav_block_mode(s);
    TravStep st;
    if (!trav_peek(sc, s, bm, st)) return kWalkDone;
    if (best_t <= st.t) return kWalkDone;
    const bool crossed = ((st.c ^ s.c) >> 6) != 0u;
    trav_commit(s, st);
    if (crossed) {
        trav_load_block<CSM>(sc, coarse, s);
        if (bm && (s.fine_lo | s.fine_hi) != 0u) { T = st.t; return kWalkEnter + (st.ax ? 0u : (st.ay ? 1u : 2u)); }
    }
    return kWalkMoved;
}

}  // namespace rtc
