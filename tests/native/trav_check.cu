// TEST INFRASTRUCTURE (host only, built by tests/test_trav_twolevel.py with nvcc): walks rays through a grid's
// occupancy twice with the product's own traversal functions (csrc/rtc_trav.cuh compiled for the host) — once
// cell by cell (the form pinned to the reference's iterator), once with empty blocks taken in one step — and
// reports the first ray on which the two disagree about the triangle-holding cells reached, the state the ray
// has there, the crossing value that leaves each of them, or how the walk ends.
#include "../../zig_raytracing_contest_b200/csrc/rtc_trav.cuh"

#include <cstdio>
#include <vector>

using namespace rtc;

namespace {
struct Event { uint32_t c; float tx, ty, tz, t_exit; };
struct Tables {
    DevScene sc;
    std::vector<unsigned long long> fine;
    std::vector<uint32_t> coarse;
};

void build(const uint32_t res[3], const uint32_t* cells_lin, Tables& tb)
{
    DevScene& d = tb.sc;
    memset(&d, 0, sizeof d);
    const uint32_t nbx = (res[0] + 3) / 4, nby = (res[1] + 3) / 4, nbz = (res[2] + 3) / 4;
    auto clog2 = [](uint32_t n) { uint32_t l = 0; while ((1u << l) < n) ++l; return l; };
    const uint32_t lx = clog2(nbx), ly = clog2(nby), lz = clog2(nbz);
    d.shift[0] = 6u; d.shift[1] = 6u + lx; d.shift[2] = 6u + lx + ly;
    d.mask[0] = 0x03u | (((1u << lx) - 1u) << d.shift[0]);
    d.mask[1] = 0x0Cu | (((1u << ly) - 1u) << d.shift[1]);
    d.mask[2] = 0x30u | (((1u << lz) - 1u) << d.shift[2]);
    const size_t nb = (size_t)1 << (lx + ly + lz);
    tb.fine.assign(nb, 0ull);
    tb.coarse.assign((nb + 31) / 32, 0u);
    size_t c = 0;
    for (uint32_t z = 0; z < res[2]; ++z)
        for (uint32_t y = 0; y < res[1]; ++y)
            for (uint32_t x = 0; x < res[0]; ++x, ++c)
                if (cells_lin[2 * c + 1] > cells_lin[2 * c]) {
                    const uint32_t pc = pack_cell(d.shift, x, y, z);
                    tb.fine[pc >> 6] |= 1ull << (pc & 63u);
                    tb.coarse[(pc >> 6) >> 5] |= 1u << ((pc >> 6) & 31u);
                }
    d.fine = tb.fine.data(); d.coarse = tb.coarse.data(); d.coarse_words = (uint32_t)tb.coarse.size();
    for (int a = 0; a < 3; ++a) d.res[a] = res[a];
}

// Returns 1 when the ray ended (at a crossing at or beyond best_t, or by leaving the grid: the same outcome for
// k_trace, and a ray that would end inside its last, empty block ends when it leaves it), 0 on the step budget.
// The loop is k_trace's phase A / phase C: occupancy check, then trav_advance.
template <bool TWO>
int walk(const Tables& tb, Trav s, float best_t, std::vector<Event>& ev, uint64_t* steps)
{
    trav_load_block<false>(tb.sc, nullptr, s);
    bool pending = false;
    Event e{ 0u, 0.0f, 0.0f, 0.0f, 0.0f };
    for (int guard = 0; guard < 1 << 20; ++guard) {
        if (trav_occupied(s) && !pending) { e.c = s.c; e.tx = s.tx; e.ty = s.ty; e.tz = s.tz; pending = true; }
        const uint32_t c0 = s.c;
        TravStep st;
        const bool has = trav_peek(tb.sc, s, TWO && trav_block_mode(s), st);     // for the record only
        float T = 0.0f;
        const uint32_t w = trav_advance<false, TWO>(tb.sc, nullptr, s, best_t, T);
        if (w >= kWalkEnter) trav_enter(s, T, w - kWalkEnter);
        const bool done = w == kWalkDone;
        *steps += 1;
        if (pending && (done || s.c != c0)) { e.t_exit = has ? st.t : f_inf(); ev.push_back(e); pending = false; }
        if (done) return 1;
    }
    return 0;
}
}  // namespace

// init_u: per ray cell[3], exit[3], step[3]; init_f: per ray t_delta[3], t_next[3]; best_t: per ray.
// Returns the number of rays whose two walks differ; *first_bad = index of the first (or -1);
// steps[0], steps[1] = loop iterations of the cell-by-cell and of the two-level walk.
extern "C" int trav_check(const uint32_t res[3], const uint32_t* cells_lin, int nrays, const uint32_t* init_u, const float* init_f,
                          const float* best_t, int* first_bad, uint64_t steps[2], uint64_t* events)
{
    Tables tb;
    build(res, cells_lin, tb);
    const bool aligned = res[0] % 4u == 0u && res[1] % 4u == 0u && res[2] % 4u == 0u;
    int bad = 0;
    *first_bad = -1; steps[0] = steps[1] = 0; *events = 0;
    std::vector<Event> a, b;
    for (int i = 0; i < nrays; ++i) {
        const uint32_t* u = init_u + 9 * (size_t)i;
        const float* f = init_f + 6 * (size_t)i;
        Trav s;
        s.tx = f[3]; s.ty = f[4]; s.tz = f[5];
        float d[3];
        for (int k = 0; k < 3; ++k) { uint32_t bits = f_bits(f[k]) | (u[6 + k] == 1u ? 0u : 0x80000000u); memcpy(&d[k], &bits, 4); }
        s.dx = d[0]; s.dy = d[1]; s.dz = d[2];
        s.c = pack_cell(tb.sc.shift, u[0], u[1], u[2]);
        s.ec = pack_cell(tb.sc.shift, u[3], u[4], u[5]);
        // the rule queue_setup applies (rtc_wavefront.cu): cell-only unless the grid is 4-aligned and every
        // crossing value is finite or +inf
        const bool ok = aligned && s.tx > -f_inf() && s.ty > -f_inf() && s.tz > -f_inf();
        if (!ok) s.ec |= kTravCellOnly;
        s.fine_lo = s.fine_hi = 0u;
        a.clear(); b.clear();
        const int ha = walk<false>(tb, s, best_t[i], a, &steps[0]);
        const int hb = walk<true>(tb, s, best_t[i], b, &steps[1]);
        *events += a.size();
        bool same = (ha != 0) == (hb != 0) && ha != 0 && a.size() == b.size();
        for (size_t k = 0; same && k < a.size(); ++k)
            same = a[k].c == b[k].c && f_bits(a[k].tx) == f_bits(b[k].tx) && f_bits(a[k].ty) == f_bits(b[k].ty) &&
                   f_bits(a[k].tz) == f_bits(b[k].tz) && f_bits(a[k].t_exit) == f_bits(b[k].t_exit);
        if (!same) { if (*first_bad < 0) *first_bad = i; ++bad; }
    }
    return bad;
}

