// search_hier.cuh -- K1: the production ring search (helpers.rs:9-54 + depth_image.rs:113-122).
//
// The reference walks ring s = 0, 1, .. around the origin as: row y+s, row y-s (x ascending), column x+s, column x-s
// (y ascending) and returns the FIRST pixel inside the query's colour box.  That is an arg-min of the ring-order key
// (common.cuh) over the matching pixels, so pixels may be visited in any order and whole regions without a match may be
// skipped.  This kernel walks COARSE rings of 8x8-pixel tiles around the origin's tile (tiles.cuh):
//
//   * coarse ring S = the tiles at Chebyshev tile distance S.  Each of its four sides is one 64-tile occupancy window
//     per channel (3 x 8-byte loads + 2 ANDs): a clear bit proves that the tile holds no match for this query's colour;
//   * a flagged tile is tested on the fine planes: the 15 words of the exact match word in one round trip (64 probes),
//     bit-exact without touching the pixels;
//   * the first match inside a tile in reference order is found by bit tricks: the tile is cut by the two diagonals
//     through the origin into the parts that belong to the bottom row / top row / right column / left column of their
//     rings (corners go to the rows, helpers.rs:28-31); inside a part the ring-order key is monotone in (row, column) or
//     (column, row), so the winner is a find-first-set on the row-major word or on its column fold;
//   * every pixel of a tile at coarse ring S+1 lies at ring >= 8S+1, so the walk stops after coarse ring S as soon as the
//     best key so far has ring <= 8S; a tile whose nearest pixel is farther than the best ring so far (or >= max) is skipped.
//
// Work per query is data dependent (most queries stop at ring 0; a few percent walk to the last ring), so the walk is
// split: the NEAR kernel does ring 0 for four patches of main pixels (or slices of the block list), queues the survivors
// in shared memory and walks them 256 at a time through the first coarse rings (full warps, no warp retires early); what
// is still open goes to a work list; the FAR kernel runs twice over the (shrinking) list, 32 consecutive entries per
// warp, pulled with an atomic cursor.  Inside a warp the expensive, rare tile tests are pooled (hier_rings_warp): lanes
// list their flagged tiles as tasks (one shared-memory atomic per lane, tiles out of reach of the best ring so far dropped
// by one bit-range mask per side), the 32 lanes take 32 tasks at a time.  Rings <= 15 have sides of at most 31 tiles and
// use 32-bit flags; longer sides are walked per lane, window by window.  `d as u8` (depth_image.rs:120-121) and the
// per-pair maximum (:124) leave with the result.
// hier_rings() is the same walk for ONE query without the pooling: tests/host_emul compiles it for the host.
#pragma once
#include <algorithm>

#include "ctx.cuh"
#include "search_literal.cuh"
#include "tiles.cuh"

namespace stepth {

constexpr uint32_t KEY21_NONE = 0x1fffffu;  // work-list encoding of KEY_NONE (a real key has ring <= 511: 21 bits)
constexpr uint32_t WL_MAX_PAIRS = 2048;     // pair index travels in the upper 11 bits of a work-list entry

struct HierView {
    const unsigned long long *fine;  // this pair's fine records
    const unsigned long long *crp, *ccp;  // this pair's coarse planes (64-tile windows)
    int ntx, nty, nxw, nyw;
};

STEPTH_HD HierView make_view(const uint8_t *planes, const TileGeom &g, uint32_t pair) {
    const uint8_t *base = planes + (size_t)pair * g.pair_bytes;
    HierView v;
    v.fine = reinterpret_cast<const unsigned long long *>(base);
    v.crp = reinterpret_cast<const unsigned long long *>(base + g.off_crp);
    v.ccp = reinterpret_cast<const unsigned long long *>(base + g.off_ccp);
    v.ntx = (int)g.ntx; v.nty = (int)g.nty; v.nxw = (int)g.nxw; v.nyw = (int)g.nyw;
    return v;
}

// ---- diagonal masks of an 8x8 tile (bit = ly * 8 + lx) ------------------------------------------------------------
// diag_le(a) = { lx - ly <= a },  diag_ge(b) = { lx + ly >= b };  any integer a, b
STEPTH_HD unsigned long long diag_le(int a) {
    if (a >= 7) return ~0ull;
    if (a < -7) return 0ull;
    // row ly keeps lx <= a + ly: the low (a + ly + 1) bits of its byte
    unsigned long long m = 0;
#pragma unroll
    for (int ly = 0; ly < 8; ++ly) {
        int n = a + ly + 1;
        n = n < 0 ? 0 : (n > 8 ? 8 : n);
        m |= (unsigned long long)((1u << n) - 1u) << (8 * ly);
    }
    return m;
}
STEPTH_HD unsigned long long diag_ge(int b) {
    if (b <= 0) return ~0ull;
    if (b > 14) return 0ull;
    // row ly keeps lx >= b - ly
    unsigned long long m = 0;
#pragma unroll
    for (int ly = 0; ly < 8; ++ly) {
        int n = b - ly;
        n = n < 0 ? 0 : (n > 8 ? 8 : n);
        m |= (unsigned long long)((0xffu << n) & 0xffu) << (8 * ly);
    }
    return m;
}

// tables of the two families for a in [-9, 7] and b in [0, 16]; the kernels keep them in shared memory
constexpr int DIAG_N = 17;
STEPTH_HD int diag_le_index(int a) { return (a < -9 ? -9 : (a > 7 ? 7 : a)) + 9; }
STEPTH_HD int diag_ge_index(int b) { return b < 0 ? 0 : (b > 16 ? 16 : b); }

// The first match (smallest ring-order key) among the set bits of tile mask m; (bx, by) = offset of the tile's pixel
// (0,0) from the origin.  tab_le / tab_ge: the diagonal tables above.
STEPTH_HD uint32_t tile_min_key(unsigned long long m, int bx, int by, const unsigned long long *tab_le,
                                const unsigned long long *tab_ge) {
    // pixel (lx, ly): dx = bx + lx, dy = by + ly.   dy >= dx <=> lx - ly <= by - bx;   dy >= -dx <=> lx + ly >= -(bx + by)
    const int a = by - bx, b = -(bx + by);
    const unsigned long long le = tab_le[diag_le_index(a)], lt = tab_le[diag_le_index(a - 1)];
    const unsigned long long ge = tab_ge[diag_ge_index(b)], gt = tab_ge[diag_ge_index(b + 1)];
    uint32_t best = KEY_NONE;
    const unsigned long long mb = m & le & ge;       // bottom row of its ring: dy = s,  |dx| <= s   (origin included)
    if (mb) {
        const int i = ffs64(mb) - 1, ly = i >> 3, lx = i & 7;
        const int s = by + ly, dx = bx + lx;
        best = ((uint32_t)s << 12) | (uint32_t)(dx + s);
    }
    const unsigned long long mt = m & ~lt & ~gt;     // top row: dy = -s, |dx| <= s
    if (mt) {
        const int ly = (63 - clz64(mt)) >> 3;
        const uint32_t row = (uint32_t)(mt >> (8 * ly)) & 0xffu;
        const int lx = ffs32(row) - 1;
        const int s = -(by + ly), dx = bx + lx;
        const uint32_t k = ((uint32_t)s << 12) | (uint32_t)((2 * s + 1) + dx + s);
        best = k < best ? k : best;
    }
    const unsigned long long mr = m & ~le & gt;      // right column: dx = s, |dy| < s
    const unsigned long long ml = m & lt & ~ge;      // left column: dx = -s, |dy| < s
    if (mr) {
        uint32_t f = (uint32_t)mr | (uint32_t)(mr >> 32);
        f |= f >> 16;
        f |= f >> 8;
        const int lx = ffs32(f & 0xffu) - 1;
        const unsigned long long colbits = (mr >> lx) & 0x0101010101010101ull;
        const int ly = (ffs64(colbits) - 1) >> 3;
        const int s = bx + lx, dy = by + ly;
        const uint32_t k = ((uint32_t)s << 12) | (uint32_t)(2 * (2 * s + 1) + (dy + s - 1));
        best = k < best ? k : best;
    }
    if (ml) {
        uint32_t f = (uint32_t)ml | (uint32_t)(ml >> 32);
        f |= f >> 16;
        f |= f >> 8;
        const int lx = 31 - clz32(f & 0xffu);
        const unsigned long long colbits = (ml >> lx) & 0x0101010101010101ull;
        const int ly = (ffs64(colbits) - 1) >> 3;
        const int s = -(bx + lx), dy = by + ly;
        const uint32_t k = ((uint32_t)s << 12) | (uint32_t)(2 * (2 * s + 1) + (2 * s - 1) + (dy + s - 1));
        best = k < best ? k : best;
    }
    return best;
}

// One query's walk over coarse rings [S0, S1).  Returns true when the answer is final (`best` = the first match in
// reference order, or KEY_NONE); false when rings >= S1 still have to be walked.
struct HierQuery {
    int ox, oy;            // origin in the coordinates of the plane coverage (oy may lie outside it in a row band)
    uint32_t col;          // r | g << 8 | b << 16
    uint32_t jr, jg, jb;   // fine level: slot / bin of each channel
    uint32_t cr, cg, cb;   // coarse level: word offset of each channel's exact-value plane inside a cell
};

STEPTH_HD HierQuery make_query(int ox, int oy, uint32_t col) {
    HierQuery q;
    q.ox = ox; q.oy = oy; q.col = col;
    q.jr = w_u(col & 0xffu);
    q.jg = w_u((col >> 8) & 0xffu);
    q.jb = w_u((col >> 16) & 0xffu);
    q.cr = col & 0xffu;
    q.cg = C_Q + ((col >> 8) & 0xffu);
    q.cb = 2 * C_Q + ((col >> 16) & 0xffu);
    return q;
}

// Tests tile (TX, TY) for query q and lowers `best` to the tile's first match if that precedes it.  All 15 plane words
// (3 candidate + 12 for the exact word) are requested at once: the walk is latency bound, one round trip beats two.
STEPTH_HD void fine_test(const HierView &v, const HierQuery &q, int TX, int TY, int D, const unsigned long long *tab_le,
                         const unsigned long long *tab_ge, uint32_t &best) {
    const int bx = TX * 8 - q.ox, by = TY * 8 - q.oy;
    const int dxm = bx > 0 ? bx : (bx + 7 < 0 ? -(bx + 7) : 0);
    const int dym = by > 0 ? by : (by + 7 < 0 ? -(by + 7) : 0);
    const int smin = dxm > dym ? dxm : dym;              // ring of the tile's nearest pixel
    if (smin >= D || smin > (int)(best >> 12)) return;   // KEY_NONE >> 12 is larger than any ring
    const unsigned long long *rec = v.fine + ((size_t)TY * v.ntx + TX) * T_REC;
    unsigned long long m = ~0ull;
#if defined(STEPTH_FINE_CANDIDATE) && defined(__CUDA_ARCH__)
    // the candidate word first (3 loads: the pixels inside the union of the bin's ranges in all three channels); the 12 words
    // that make it exact only when it is not empty.  Off by default: 16 pairs 1080p `nat` 2.10 -> 2.15 ms (a second dependent
    // round trip for the two tests out of three that pass), `geo` 5.09 -> 4.93 ms (profiles/recipes/r2_run25.sh)
#pragma unroll
    for (int c = 0; c < 3; ++c) m &= ldg(rec + c * T_SLOTS + w_u((q.col >> (8 * c)) & 0xffu));
    if (!m) return;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const uint32_t qc = (q.col >> (8 * c)) & 0xffu;
        const unsigned long long *ch = rec + c * T_SLOTS;
        m &= (ldg(ch + w_up1(qc)) | ldg(ch + w_nta(qc))) & (ldg(ch + w_um1(qc)) | ldg(ch + w_ntb(qc)));
    }
#else
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const uint32_t qc = (q.col >> (8 * c)) & 0xffu;
        const unsigned long long *ch = rec + c * T_SLOTS;
        m &= ldg(ch + w_u(qc)) & (ldg(ch + w_up1(qc)) | ldg(ch + w_nta(qc))) & (ldg(ch + w_um1(qc)) | ldg(ch + w_ntb(qc)));
    }
#endif
    if (!m) return;
    const uint32_t k = tile_min_key(m, bx, by, tab_le, tab_ge);
    if ((int)(k >> 12) < D && k < best) best = k;
}

// occupancy window `cell` (= line * windows_per_line + w) of a coarse plane for query q: bit i <-> tile 32 w + i
STEPTH_HD unsigned long long coarse_window(const unsigned long long *plane, uint32_t cell, const HierQuery &q) {
    const unsigned long long *p = plane + (size_t)cell * C_CELL;
    return ldg(p + q.cr) & ldg(p + q.cg) & ldg(p + q.cb);
}

// flags of the tile run [lo, hi] of plane line `line` as one word (bit i <-> tile lo + i); needs (lo & 31) + hi - lo < 64
STEPTH_HD unsigned long long side_flags(const unsigned long long *plane, int wpl, int line, int n_lines, int lo, int hi,
                                        const HierQuery &q) {
    if ((unsigned)line >= (unsigned)n_lines || lo > hi) return 0ull;
    unsigned long long m = coarse_window(plane, (uint32_t)(line * wpl + (lo >> 5)), q) >> (lo & 31);
    const int len = hi - lo + 1;
    if (len < 64) m &= (1ull << len) - 1ull;
    return m;
}

// ---- the first step: coarse rings 0 and 1 together -------------------------------------------------------------------
// The 3x3 tile block around the origin's tile is three runs of (at most) three tiles in three consecutive tile rows: THREE
// row-major windows give all nine occupancy flags -- the origin's own tile included, which ring 0 alone would test blind
// (on a natural pair two out of three own tiles hold no match), and the two column sides of ring 1, which are single
// tiles of the middle row -- where rings 0 and 1 walked separately cost one un-flagged tile test and four windows.
struct Block3 {
    uint32_t fa, fb, fc;  // flags of tile rows T0y-1, T0y, T0y+1; bit i <-> tile rlo + i
    int rlo, rhi;         // the tile run, clipped to the grid
};

STEPTH_HD uint32_t row_flags3(const unsigned long long *plane, int wpl, int line, int n_lines, int lo, int hi, const HierQuery &q) {
    if ((unsigned)line >= (unsigned)n_lines || lo > hi) return 0u;
    const unsigned long long w = coarse_window(plane, (uint32_t)(line * wpl + (lo >> 5)), q);
    return (uint32_t)(w >> (lo & 31)) & ((2u << (hi - lo)) - 1u);   // hi - lo <= 2
}

STEPTH_HD Block3 block3_flags(const HierView &v, const HierQuery &q, int T0x, int T0y) {
    Block3 b;
    b.rlo = T0x - 1 < 0 ? 0 : T0x - 1;
    b.rhi = T0x + 1 > v.ntx - 1 ? v.ntx - 1 : T0x + 1;
    b.fa = row_flags3(v.crp, v.nxw, T0y - 1, v.nty, b.rlo, b.rhi, q);
    b.fb = row_flags3(v.crp, v.nxw, T0y, v.nty, b.rlo, b.rhi, q);
    b.fc = row_flags3(v.crp, v.nxw, T0y + 1, v.nty, b.rlo, b.rhi, q);
    return b;
}
// flag of tile TX of the middle row (0 when TX lies outside the run)
STEPTH_HD uint32_t block3_mid(const Block3 &b, int TX) { return (TX < b.rlo || TX > b.rhi) ? 0u : (b.fb >> (TX - b.rlo)) & 1u; }

// Every pixel outside the (2S+1)^2 tile block around the origin's tile lies at ring >= 8S + slack + 1, slack = the distance
// from the origin to the nearest border of its own tile: after coarse ring S the walk is over once the best ring <= 8S + slack.
STEPTH_HD int stop_slack(const HierQuery &q) {
    const int lx = q.ox & 7, ly = q.oy & 7;
    const int mx = lx < 7 - lx ? lx : 7 - lx, my = ly < 7 - ly ? ly : 7 - ly;
    return mx < my ? mx : my;
}

// coarse rings [S0, S1): true when the answer is final, false when rings >= S1 still have to be walked.
// Ring S: side 0 = tile row T0y+S, 1 = row T0y-S (tiles T0x-S .. T0x+S), 2 = tile column T0x+S, 3 = column T0x-S
// (tiles T0y-S+1 .. T0y+S-1), all clipped to the grid.
STEPTH_HD bool hier_rings(const HierView &v, const HierQuery &q, int D, int S0, int S1, const unsigned long long *tab_le,
                          const unsigned long long *tab_ge, uint32_t &best) {
    const int T0x = q.ox >> 3, T0y = q.oy >> 3;  // arithmetic shift: floor for a negative oy
    const int Smax = (D + 6) >> 3;               // tiles of coarse ring S start at ring >= 8S - 7
    const int S_end = S1 < Smax + 1 ? S1 : Smax + 1;
    const int xmax = v.ntx - 1, ymax = v.nty - 1;
    const int slack = stop_slack(q);
    for (int S = S0; S < S_end; ++S) {
        // the coarse square has left the tile grid on all four sides: nothing further out exists
        if (T0x - S < 0 && T0x + S > xmax && T0y - S < 0 && T0y + S > ymax) return true;
        if (S == 0 && S_end >= 2) {
            // rings 0 and 1 from three row windows (see Block3)
            const Block3 b = block3_flags(v, q, T0x, T0y);
            if (block3_mid(b, T0x)) fine_test(v, q, T0x, T0y, D, tab_le, tab_ge, best);
            if (best != KEY_NONE && (int)(best >> 12) <= slack) return true;
            S = 1;
            if (T0x - 1 < 0 && T0x + 1 > xmax && T0y - 1 < 0 && T0y + 1 > ymax) return true;
            for (int side = 0; side < 2; ++side) {            // rows T0y+1, T0y-1
                uint32_t m = side == 0 ? b.fc : b.fa;
                while (m) {
                    const int p = b.rlo + ffs32(m) - 1;
                    m &= m - 1;
                    fine_test(v, q, p, side == 0 ? T0y + 1 : T0y - 1, D, tab_le, tab_ge, best);
                }
            }
            if (block3_mid(b, T0x + 1)) fine_test(v, q, T0x + 1, T0y, D, tab_le, tab_ge, best);
            if (block3_mid(b, T0x - 1)) fine_test(v, q, T0x - 1, T0y, D, tab_le, tab_ge, best);
        } else if (S == 0) {
            if ((unsigned)T0x < (unsigned)v.ntx && (unsigned)T0y < (unsigned)v.nty)
                fine_test(v, q, T0x, T0y, D, tab_le, tab_ge, best);
        } else {
            const int rlo = T0x - S < 0 ? 0 : T0x - S, rhi = T0x + S > xmax ? xmax : T0x + S;               // row sides
            const int clo = T0y - S + 1 < 0 ? 0 : T0y - S + 1, chi = T0y + S - 1 > ymax ? ymax : T0y + S - 1;  // column sides
            if (S <= 16) {
                // a side is at most 33 tiles: one window each; all four are requested before any is used
                const unsigned long long f0 = side_flags(v.crp, v.nxw, T0y + S, v.nty, rlo, rhi, q);
                const unsigned long long f1 = side_flags(v.crp, v.nxw, T0y - S, v.nty, rlo, rhi, q);
                const unsigned long long f2 = side_flags(v.ccp, v.nyw, T0x + S, v.ntx, clo, chi, q);
                const unsigned long long f3 = side_flags(v.ccp, v.nyw, T0x - S, v.ntx, clo, chi, q);
                if (f0 | f1 | f2 | f3) {
#pragma unroll 1
                    for (int side = 0; side < 4; ++side) {
                        unsigned long long m = side == 0 ? f0 : side == 1 ? f1 : side == 2 ? f2 : f3;
                        const int line = (side >= 2 ? T0x : T0y) + ((side & 1) ? -S : S);
                        const int lo = side >= 2 ? clo : rlo;
                        while (m) {
                            const int p = lo + ffs64(m) - 1;
                            m &= m - 1;
                            fine_test(v, q, side >= 2 ? line : p, side >= 2 ? p : line, D, tab_le, tab_ge, best);
                        }
                    }
                }
            } else {
                // longer sides (max_rings > 134): window by window
#pragma unroll 1
                for (int side = 0; side < 4; ++side) {
                    const bool is_col = side >= 2;
                    const int line = (is_col ? T0x : T0y) + ((side & 1) ? -S : S);
                    if ((unsigned)line >= (unsigned)(is_col ? v.ntx : v.nty)) continue;
                    const int lo = is_col ? clo : rlo, hi = is_col ? chi : rhi;
                    for (int w = lo >> 5; 32 * w <= hi; w += 2) {
                        unsigned long long m = coarse_window(is_col ? v.ccp : v.crp, (uint32_t)(line * (is_col ? v.nyw : v.nxw) + w), q);
                        if (32 * w < lo) m &= ~0ull << (lo - 32 * w);
                        if (32 * w + 63 > hi) m &= ~0ull >> (32 * w + 63 - hi);
                        while (m) {
                            const int p = 32 * w + ffs64(m) - 1;
                            m &= m - 1;
                            fine_test(v, q, is_col ? line : p, is_col ? p : line, D, tab_le, tab_ge, best);
                        }
                    }
                }
            }
        }
        if (best != KEY_NONE && (int)(best >> 12) <= 8 * S + slack) return true;
    }
    return S1 > Smax;
}

#if defined(__CUDACC__)

struct HierArgs {
    SearchArgs s;
    const uint8_t *planes;
    TileGeom g;
    // work lists: entry.x = query index inside its pair, entry.y = pair << 21 | key21
    uint2 *wl_out;
    uint32_t *wl_out_count;
    const uint2 *wl_in;
    const uint32_t *wl_in_count;
    uint32_t *wl_in_cursor;
    uint32_t wl_cap;   // entries wl_out can hold
    int S0, S1;        // coarse rings this launch walks
    int tail_split;    // last stage only: lists of at most this many entries go to search_tail_kernel, longer ones to search_far_kernel
    uint32_t *dbg_err; // -DSTEPTH_DEBUG_CHECKS builds only: bit mask of the index checks that failed (null otherwise)
};

// compute-sanitizer is closed on this pool; a -DSTEPTH_DEBUG_CHECKS build (tests run it through STEPTH_LIB) checks every index the
// search kernels form into shared-memory lists, the tile grid and the work lists, and reports a mask through HierArgs::dbg_err
#ifdef STEPTH_DEBUG_CHECKS
#define STEPTH_CHECK(a, cond, bit)                                  \
    do {                                                            \
        if (!(cond) && (a).dbg_err) atomicOr((a).dbg_err, 1u << (bit)); \
    } while (0)
#else
#define STEPTH_CHECK(a, cond, bit) \
    do {                           \
    } while (0)
#endif

__device__ __forceinline__ void load_diag_tables(unsigned long long *s_le, unsigned long long *s_ge) {
    for (int t = threadIdx.x; t < 2 * DIAG_N; t += blockDim.x) {
        if (t < DIAG_N) s_le[t] = diag_le(t - 9);
        else s_ge[t - DIAG_N] = diag_ge(t - DIAG_N);
    }
    __syncthreads();
}

// origin, colour and output slot of query `qi` of pair `pair` (depth_image.rs:114-118)
template <bool BLOCKS>
__device__ __forceinline__ void query_of(const SearchArgs &s, uint32_t pair, uint32_t qi, int &ox, int &oyg, uint32_t &col) {
    if (BLOCKS) {
        const stepth_block bl = s.blocks[qi];
        ox = (int)((bl.x + bl.w) / 2);  // literal (pos + size) / 2, not the centre
        oyg = (int)((bl.y + bl.h) / 2);
        col = bl.value[0] | (bl.value[1] << 8) | (bl.value[2] << 16);
    } else {
        const uint32_t x = qi % s.w, y = qi / s.w;
        ox = (int)((x + 1) >> 1);
        oyg = (int)((y + s.y0 + 1) >> 1);
        col = reinterpret_cast<const uint32_t *>(s.main_rgba)[(size_t)pair * s.n_queries + qi] & 0xffffffu;
    }
}

template <bool ONE_PAIR = false>  // ONE_PAIR: every lane of the warp works on the same pair (near kernel)
__device__ __forceinline__ void publish_result(const SearchArgs &s, uint32_t pair, uint32_t qi, uint32_t key, bool active) {
    uint32_t d = 0;
    if (active) {
        d = key == KEY_NONE ? 0u : (uint32_t)key_to_d8(key);
        s.d8[(size_t)pair * s.n_queries + qi] = (uint8_t)d;
    }
    // per-pair maximum: lanes of a warp may belong to different pairs only in the far kernel; match on the pair there
    const uint32_t mask = __activemask();
    uint32_t peers;
    if (ONE_PAIR) {  // two groups: the lanes with a result, and the rest (a lane must be a member of the mask it names)
        const uint32_t act = __ballot_sync(mask, active);
        peers = active ? act : (mask & ~act);
    } else {
        peers = __match_any_sync(mask, active ? pair : 0xffffffffu);
    }
    const uint32_t m = __reduce_max_sync(peers, d);
    if (active && m && (threadIdx.x & 31) == (uint32_t)(__ffs(peers) - 1)) {
        if (m > *(volatile uint32_t *)(s.max_out + pair)) atomicMax(s.max_out + pair, m);
    }
}

// appends the still open queries of a warp to the work list (one atomic per warp)
__device__ __forceinline__ void export_open(const HierArgs &a, bool open, uint32_t pair, uint32_t qi, uint32_t best) {
    const uint32_t mask = __activemask();
    const uint32_t bal = __ballot_sync(mask, open);
    if (!bal) return;
    const int lane = threadIdx.x & 31;
    uint32_t base = 0;
    if (lane == __ffs(bal) - 1) base = atomicAdd(a.wl_out_count, (uint32_t)__popc(bal));
    base = __shfl_sync(mask, base, __ffs(bal) - 1);
    if (open) {
        const uint32_t slot = base + __popc(bal & ((1u << lane) - 1u));
        if (slot < a.wl_cap)
            a.wl_out[slot] = make_uint2(qi, (pair << 21) | (best == KEY_NONE ? KEY21_NONE : best));
    }
}

// ---- warp-cooperative walk ---------------------------------------------------------------------------------------------
// Reading the occupancy windows of a ring is cheap and every lane does it for its own query.  The expensive part -- the
// tile tests -- is rare per lane (a few percent of the tiles are flagged), so lane-private tile tests would run with one
// or two active lanes at a time.  Instead the warp pools them: every lane lists the flagged tiles of its query as tasks
// (owner lane, tile) in shared memory, the 32 lanes take 32 tasks at a time -- fetching the owner's query with shuffles --
// and post the key they find with an atomicMin on the owner's slot.  Pruning (`smin > best ring`) uses the best key
// known when the task is listed; the arg-min does not depend on the order in which tiles are tested.
constexpr int WTASKS = 96;  // task slots of a warp

struct WarpScratch {
    uint32_t task[WTASKS];  // owner << 26 | TX << 13 | TY
    uint32_t best[32];      // per owner lane: best key so far
    uint32_t n;             // tasks listed in this pass (may run past WTASKS: the surplus is listed again next pass)
};

// ring of the nearest pixel of tile (TX, TY) as seen from the origin of q
STEPTH_HD int tile_smin(const HierQuery &q, int TX, int TY) {
    const int bx = TX * 8 - q.ox, by = TY * 8 - q.oy;
    const int dxm = bx > 0 ? bx : (bx + 7 < 0 ? -(bx + 7) : 0);
    const int dym = by > 0 ? by : (by + 7 < 0 ? -(by + 7) : 0);
    return dxm > dym ? dxm : dym;
}

// flags of the tile run [lo, hi] of plane line `line` as one 32-bit word (bit i <-> tile lo + i); needs hi - lo < 31, which holds
// for the sides of coarse rings <= 15 (at most 31 tiles): the run lies inside the 64-tile window that starts at lo & ~31
__device__ __forceinline__ uint32_t side_flags32(const unsigned long long *plane, int wpl, int line, int n_lines, int lo, int hi,
                                                 const HierQuery &q) {
    if ((unsigned)line >= (unsigned)n_lines || lo > hi) return 0u;
    const unsigned long long w = coarse_window(plane, (uint32_t)(line * wpl + (lo >> 5)), q);
    const uint32_t m = __funnelshift_r((uint32_t)w, (uint32_t)(w >> 32), lo & 31);
    return m & ((2u << (hi - lo)) - 1u);
}

// keeps the bits i of m with tmin <= lo + i <= tmax
__device__ __forceinline__ uint32_t keep_range(uint32_t m, int lo, int tmin, int tmax) {
    int a = tmin - lo, b = tmax - lo;
    a = a < 0 ? 0 : a;
    if (b < a || a > 31) return 0u;
    b = b > 31 ? 31 : b;
    return m & (~0u << a) & (~0u >> (31 - b));
}

// Drops the flagged tiles of ring S that cannot hold the answer: a tile whose nearest pixel lies beyond `limit` (the best
// ring so far, or max_rings - 1).  A whole side goes when its line is too far; along the side the tiles within reach are a
// run, so the pruning is one bit-range mask per side instead of a distance per tile.
__device__ __forceinline__ void prune_sides(const HierQuery &q, int S, int limit, int rlo, int clo, uint32_t &f0, uint32_t &f1,
                                            uint32_t &f2, uint32_t &f3) {
    const int T0x = q.ox >> 3, T0y = q.oy >> 3;
    const int xa = (q.ox - limit) >> 3, xb = (q.ox + limit) >> 3;   // tile columns with a pixel at |dx| <= limit
    const int ya = (q.oy - limit) >> 3, yb = (q.oy + limit) >> 3;
    f0 = (T0y + S) * 8 - q.oy > limit ? 0u : keep_range(f0, rlo, xa, xb);
    f1 = q.oy - ((T0y - S) * 8 + 7) > limit ? 0u : keep_range(f1, rlo, xa, xb);
    f2 = (T0x + S) * 8 - q.ox > limit ? 0u : keep_range(f2, clo, ya, yb);
    f3 = q.ox - ((T0x - S) * 8 + 7) > limit ? 0u : keep_range(f3, clo, ya, yb);
}

// Lists the flagged tiles of one side of a ring as tasks of the warp from slot `slot` on (the lane reserved its slots with
// one atomic for all four sides); the bits that do not fit below WTASKS come back and are listed again in the next pass.
__device__ __forceinline__ uint32_t emit_side(uint32_t m, int lo, int line, bool is_col, uint32_t lane, uint32_t &slot,
                                              WarpScratch *ws) {
    // task = owner << 26 | TX << 13 | TY: row sides step TX (bit 13), column sides step TY (bit 0)
    const uint32_t code = lane << 26 | (is_col ? (uint32_t)line << 13 | (uint32_t)lo : (uint32_t)lo << 13 | (uint32_t)line);
    const int sh = is_col ? 0 : 13;
    while (m && slot < (uint32_t)WTASKS) {
        ws->task[slot++] = code + ((uint32_t)(__ffs((int)m) - 1) << sh);
        m &= m - 1;
    }
    return m;
}

// All 32 lanes call this together.  `walking` lanes carry a query (v, q, best); the others only help with tile tests.
// Walks coarse rings [S0, S1); returns true when this lane's answer is final.
__device__ __forceinline__ bool hier_rings_warp(bool walking, const uint8_t *planes, const TileGeom &g, uint32_t pair,
                                                const HierQuery &q, int D, int S0, int S1, const unsigned long long *tab_le,
                                                const unsigned long long *tab_ge, uint32_t &best, WarpScratch *ws, uint32_t *dbg = nullptr) {
    const uint32_t lane = threadIdx.x & 31;
    const HierView v = make_view(planes, g, pair);
    const int T0x = q.ox >> 3, T0y = q.oy >> 3;
    const int Smax = (D + 6) >> 3;
    const int S_end = S1 < Smax + 1 ? S1 : Smax + 1;
    const int xmax = v.ntx - 1, ymax = v.nty - 1;
    const int slack = walking ? stop_slack(q) : 0;
    bool final_ = !walking;
    for (int S = S0; S < S_end; ++S) {
        uint32_t f0 = 0, f1 = 0, f2 = 0, f3 = 0;
        int rlo = 0, clo = 0;
        bool go = walking && !final_;
        if (S == 0 && S_end >= 2) {   // warp-uniform: rings 0 and 1 from three row windows (see Block3)
            if (go) {
                const Block3 b = block3_flags(v, q, T0x, T0y);
                // the origin's own tile: every walking lane has at most this one test, nothing to pool
                if (block3_mid(b, T0x)) fine_test(v, q, T0x, T0y, D, tab_le, tab_ge, best);
                if ((best != KEY_NONE && (int)(best >> 12) <= slack) ||
                    (T0x - 1 < 0 && T0x + 1 > xmax && T0y - 1 < 0 && T0y + 1 > ymax)) {
                    final_ = true;
                    go = false;
                } else {
                    rlo = b.rlo;
                    clo = T0y;
                    f0 = b.fc;
                    f1 = b.fa;
                    f2 = block3_mid(b, T0x + 1);
                    f3 = block3_mid(b, T0x - 1);
                }
            }
            S = 1;
        } else if (go) {
            if (T0x - S < 0 && T0x + S > xmax && T0y - S < 0 && T0y + S > ymax) {
                final_ = true;  // the coarse square has left the tile grid on all four sides
            } else if (S == 0) {
                // the origin's own tile: every walking lane has exactly this one test, nothing to pool
                if ((unsigned)T0x < (unsigned)v.ntx && (unsigned)T0y < (unsigned)v.nty)
                    fine_test(v, q, T0x, T0y, D, tab_le, tab_ge, best);
            } else {
                rlo = T0x - S < 0 ? 0 : T0x - S;
                clo = T0y - S + 1 < 0 ? 0 : T0y - S + 1;
                const int rhi = T0x + S > xmax ? xmax : T0x + S, chi = T0y + S - 1 > ymax ? ymax : T0y + S - 1;
                if (S <= 15) {
                    // a side is at most 31 tiles: 32-bit flags, the warp pools the tile tests below
                    f0 = side_flags32(v.crp, v.nxw, T0y + S, v.nty, rlo, rhi, q);
                    f1 = side_flags32(v.crp, v.nxw, T0y - S, v.nty, rlo, rhi, q);
                    f2 = side_flags32(v.ccp, v.nyw, T0x + S, v.ntx, clo, chi, q);
                    f3 = side_flags32(v.ccp, v.nyw, T0x - S, v.ntx, clo, chi, q);
                } else {
                    // longer sides (max_rings > 120): this lane tests its tiles itself, window by window
#pragma unroll 1
                    for (int side = 0; side < 4; ++side) {
                        const bool is_col = side >= 2;
                        const int line = (is_col ? T0x : T0y) + ((side & 1) ? -S : S);
                        if ((unsigned)line >= (unsigned)(is_col ? v.ntx : v.nty)) continue;
                        const int lo = is_col ? clo : rlo, hi = is_col ? chi : rhi;
                        for (int w = lo >> 5; 32 * w <= hi; w += 2) {
                            unsigned long long m = coarse_window(is_col ? v.ccp : v.crp, (uint32_t)(line * (is_col ? v.nyw : v.nxw) + w), q);
                            if (32 * w < lo) m &= ~0ull << (lo - 32 * w);
                            if (32 * w + 63 > hi) m &= ~0ull >> (32 * w + 63 - hi);
                            while (m) {
                                const int p = 32 * w + __ffsll((long long)m) - 1;
                                m &= m - 1;
                                fine_test(v, q, is_col ? line : p, is_col ? p : line, D, tab_le, tab_ge, best);
                            }
                        }
                    }
                }
            }
        }
        // ---- pool the tile tests of the warp -----------------------------------------------------------------------------
        while (__any_sync(0xffffffffu, (f0 | f1 | f2 | f3) != 0u)) {
            if (lane == 0) ws->n = 0;
            ws->best[lane] = best;
            __syncwarp();
            if (f0 | f1 | f2 | f3) {
                // with a match in hand (or at the last rings) most flagged tiles are out of reach and are not listed
                const int limit = min(D - 1, (int)(best >> 12));
                if (limit < 8 * S + 7) prune_sides(q, S, limit, rlo, clo, f0, f1, f2, f3);  // else every tile of the ring is in reach
                const uint32_t cnt = __popc(f0) + __popc(f1) + __popc(f2) + __popc(f3);
                if (cnt) {
                    uint32_t slot = atomicAdd(&ws->n, cnt);  // slots [slot, slot + cnt); those beyond WTASKS wait for the next pass
                    f0 = emit_side(f0, rlo, T0y + S, false, lane, slot, ws);
                    f1 = emit_side(f1, rlo, T0y - S, false, lane, slot, ws);
                    f2 = emit_side(f2, clo, T0x + S, true, lane, slot, ws);
                    f3 = emit_side(f3, clo, T0x - S, true, lane, slot, ws);
                }
            }
            __syncwarp();
            const int n_pass = (int)min(ws->n, (uint32_t)WTASKS);
            for (int b = 0; b < n_pass; b += 32) {
                const bool has = b + (int)lane < n_pass;
                const uint32_t task = has ? ws->task[b + lane] : 0u;
                const int owner = (int)(task >> 26);
                const int o_ox = __shfl_sync(0xffffffffu, q.ox, owner), o_oy = __shfl_sync(0xffffffffu, q.oy, owner);
                const uint32_t o_col = __shfl_sync(0xffffffffu, q.col, owner);
                const uint32_t o_pair = __shfl_sync(0xffffffffu, pair, owner);
                if (has) {
                    const HierView ov = make_view(planes, g, o_pair);
                    const HierQuery oq = make_query(o_ox, o_oy, o_col);
#ifdef STEPTH_DEBUG_CHECKS
                    if (dbg && ((task >> 13 & 0x1fffu) >= (uint32_t)ov.ntx || (task & 0x1fffu) >= (uint32_t)ov.nty)) atomicOr(dbg, 1u << 8);
#endif
                    uint32_t kb = ws->best[owner];
                    const uint32_t before = kb;
                    fine_test(ov, oq, (int)((task >> 13) & 0x1fffu), (int)(task & 0x1fffu), D, tab_le, tab_ge, kb);
                    if (kb < before) atomicMin(&ws->best[owner], kb);
                }
            }
            __syncwarp();
            best = ws->best[lane];
        }
        if (go && best != KEY_NONE && (int)(best >> 12) <= 8 * S + slack) final_ = true;
        if (__all_sync(0xffffffffu, final_)) return true;
    }
    return final_ || S1 > Smax;
}

// NEAR kernel: ring 0 (one pixel compare) and coarse rings [0, S1) for every query.
// Dense mode: a warp owns an 8x4 patch of main pixels (= 4x2 origins: neighbouring queries share almost all loads), a
// CTA 32x8 pixels.  Block mode: one thread per DiscretePixel, list order.
#ifndef STEPTH_NEAR_MINB
#define STEPTH_NEAR_MINB 5   // resident CTAs per SM the register allocation must allow: 5 x 256 threads -> 48 registers
#endif
#ifndef STEPTH_FAR_MINB
#define STEPTH_FAR_MINB 10
#endif
#ifndef STEPTH_NEAR_KP
#define STEPTH_NEAR_KP 4     // most 32x8-pixel patches (or 256-block slices) one CTA settles before it walks: 4, 2 or 1
#endif

// ring 0 of one query (helpers.rs:26-47 with step = 0): `open` = the walk has to go on
template <bool BLOCKS>
__device__ __forceinline__ void ring0_of(const SearchArgs &s, uint32_t pair, uint32_t qi, uint32_t X, uint32_t Y, int &ox, int &oyl,
                                         uint32_t &col, uint32_t &key, bool &open) {
    int oyg;
    if (BLOCKS) {
        query_of<true>(s, pair, qi, ox, oyg, col);
    } else {  // query_of<false> without the division: X, Y are at hand
        ox = (int)((X + 1) >> 1);
        oyg = (int)((Y + s.y0 + 1) >> 1);
        col = reinterpret_cast<const uint32_t *>(s.main_rgba)[(size_t)pair * s.n_queries + qi] & 0xffffffu;
    }
    const int src_wa = (int)s.wa, src_ha = (int)s.ha;
    // ring 0 out of bounds => `still_in_bounds` stays false => break => None   (helpers.rs:49-51)
    if (ox < src_wa && oyg < (int)s.add_img_h) {
        oyl = oyg - (int)s.add_y0;  // origin row inside the sub-image held on this device
        bool hit0 = false;
        if (oyl >= 0 && oyl < src_ha) {
            const uint8_t *px = s.add_rgb + (size_t)pair * src_ha * src_wa * 3 + ((size_t)oyl * src_wa + ox) * 3;
            hit0 = pixel_in_box(px[0], px[1], px[2], col, s.pr, s.pg, s.pb, s.sat != 0);
        }
        if (hit0) key = 0;
        else open = s.max_rings > 1;
    }
}

// KPT = patches per CTA (0: no survivor queue, a thread keeps its own query)
template <bool BLOCKS, int KPT>
__global__ void __launch_bounds__(256, STEPTH_NEAR_MINB) search_near_kernel(HierArgs a) {
    constexpr bool COMPACT = KPT > 0;
    constexpr int KP = COMPACT ? KPT : 1;
    __shared__ unsigned long long s_le[DIAG_N], s_ge[DIAG_N];
    __shared__ WarpScratch s_ws[8];
    __shared__ uint4 s_open[COMPACT ? KP * 256 : 1];  // queries that survive ring 0: {qi, ox, oy, colour}
    __shared__ uint32_t s_n;
    if (threadIdx.x == 0) s_n = 0;
    load_diag_tables(s_le, s_ge);  // ends with __syncthreads()
    const SearchArgs &s = a.s;
    const uint32_t pair = BLOCKS ? 0u : blockIdx.z;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (COMPACT) {
        // Ring 0 settles about half of the queries of a natural pair.  The CTA runs ring 0 for KP patches, queues the
        // survivors in shared memory and then walks them 256 at a time: full warps, and every warp of the CTA stays busy.
#pragma unroll 1
        for (int k = 0; k < KP; ++k) {
            bool valid;
            uint32_t qi, X = 0, Y = 0;
            if (BLOCKS) {
                qi = (blockIdx.x * KP + k) * 256u + tid;
                valid = qi < s.n_queries;
            } else {
                X = (blockIdx.x * KP + k) * 32u + (warp & 3) * 8u + (lane & 7);
                Y = blockIdx.y * 8u + (warp >> 2) * 4u + (lane >> 3);
                valid = X < s.w && Y < s.h;
                qi = Y * s.w + X;
            }
            uint32_t key = KEY_NONE, col = 0;
            bool open = false;
            int ox = 0, oyl = 0;
            if (valid) ring0_of<BLOCKS>(s, pair, qi, X, Y, ox, oyl, col, key, open);
            // settled at ring 0: a hit has distance 0 and `None` maps to 0 too (depth_image.rs:120), neither moves the maximum
            if (valid && !open) s.d8[(size_t)pair * s.n_queries + qi] = 0;
            const uint32_t bal = __ballot_sync(0xffffffffu, open);
            if (bal) {
                uint32_t base = 0;
                if (lane == __ffs(bal) - 1) base = atomicAdd(&s_n, (uint32_t)__popc(bal));
                base = __shfl_sync(0xffffffffu, base, __ffs(bal) - 1);
                if (open) s_open[base + __popc(bal & ((1u << lane) - 1u))] = make_uint4(qi, (uint32_t)ox, (uint32_t)oyl, col);
            }
        }
        __syncthreads();
        const uint32_t n_open = s_n;
#pragma unroll 1
        for (uint32_t i0 = (uint32_t)warp * 32u; i0 < n_open; i0 += 256u) {  // warp-uniform
            const bool has = i0 + lane < n_open;
            const uint4 e = has ? s_open[i0 + lane] : make_uint4(0, 0, 0, 0);
            const HierQuery q = make_query((int)e.y, (int)e.z, e.w);
            uint32_t key = KEY_NONE;
            const bool still = !hier_rings_warp(has, a.planes, a.g, pair, q, (int)s.max_rings, 0, a.S1, s_le, s_ge, key, &s_ws[warp], a.dbg_err);
            export_open(a, has && still, pair, e.x, key);
            publish_result<true>(s, pair, e.x, key, has && !still);
        }
        return;
    }
    bool valid;
    uint32_t qi = 0, X = 0, Y = 0;
    if (BLOCKS) {
        qi = blockIdx.x * 256u + tid;
        valid = qi < s.n_queries;
    } else {
        X = blockIdx.x * 32u + (warp & 3) * 8u + (lane & 7);
        Y = blockIdx.y * 8u + (warp >> 2) * 4u + (lane >> 3);
        valid = X < s.w && Y < s.h;
        qi = Y * s.w + X;
    }
    uint32_t key = KEY_NONE, col = 0;
    bool open = false;
    int ox = 0, oyl = 0;
    if (valid) ring0_of<BLOCKS>(s, pair, qi, X, Y, ox, oyl, col, key, open);
    if (__any_sync(0xffffffffu, open)) {
        const HierQuery q = make_query(ox, oyl, col);
        open = !hier_rings_warp(open, a.planes, a.g, pair, q, (int)s.max_rings, 0, a.S1, s_le, s_ge, key, &s_ws[warp], a.dbg_err);
    }
    export_open(a, open, pair, qi, key);
    publish_result<true>(s, pair, qi, key, valid && !open);
}

// A work list this short is a matter of latency, not of throughput: the last stage hands it to search_tail_kernel (one WARP per
// query) instead of walking it with one lane per query.  Both kernels are launched; the list's length, known on the device only,
// decides which of them works.
// (HierArgs::tail_split = the longest list the tail kernel takes; 0 = none.  Measured on one README pair, 255 rings: block list
// 0.210 -> 0.183 ms, dense 0.150 -> 0.131 ms; one 1080p pair at 128 rings, four rings left for this stage: no gain)

// FAR kernel: resumes the listed queries at coarse ring S0; a warp pulls 32 consecutive entries at a time.
template <bool BLOCKS>
__global__ void __launch_bounds__(128, STEPTH_FAR_MINB) search_far_kernel(HierArgs a) {
    __shared__ unsigned long long s_le[DIAG_N], s_ge[DIAG_N];
    __shared__ WarpScratch s_ws[4];
    load_diag_tables(s_le, s_ge);
    const SearchArgs &s = a.s;
    const int lane = threadIdx.x & 31;
    const uint32_t n_entries = min(*a.wl_in_count, a.wl_cap);
    if (n_entries <= (uint32_t)a.tail_split) return;   // search_tail_kernel takes this list
    for (;;) {
        uint32_t e0 = 0;
        if (lane == 0) e0 = atomicAdd(a.wl_in_cursor, 32u);
        e0 = __shfl_sync(0xffffffffu, e0, 0);
        if (e0 >= n_entries) break;
        const uint32_t e = e0 + lane;
        const bool valid = e < n_entries;
        uint32_t pair = 0, qi = 0, key = KEY_NONE, col = 0;
        int ox = 0, oyl = 0;
        if (valid) {
            const uint2 ent = a.wl_in[e];
            qi = ent.x;
            pair = ent.y >> 21;
            const uint32_t k21 = ent.y & KEY21_NONE;
            key = k21 == KEY21_NONE ? KEY_NONE : k21;
            int oyg;
            query_of<BLOCKS>(s, pair, qi, ox, oyg, col);
            oyl = oyg - (int)s.add_y0;
        }
        const HierQuery q = make_query(ox, oyl, col);
        const bool open = !hier_rings_warp(valid, a.planes, a.g, pair, q, (int)s.max_rings, a.S0, a.S1, s_le, s_ge, key,
                                           &s_ws[threadIdx.x >> 5], a.dbg_err);
        if (a.wl_out) export_open(a, open, pair, qi, key);
        publish_result(s, pair, qi, key, valid && !open);
    }
}

// TAIL kernel: the same rings for a SHORT list, one warp per query.  A ring has four sides of at most 65 / 63 tiles = ten runs of at
// most 32 tiles; the lanes take the thirty runs of three consecutive rings, each reads its occupancy window (one 64-tile window
// holds any run of 32) and tests its flagged tiles itself; a warp minimum folds the keys.  A query that walks twenty empty rings
// pays seven window round trips instead of twenty times four (the README pair at 255 rings: 62 -> about 15 us for the stage).
template <bool BLOCKS>
__global__ void __launch_bounds__(128) search_tail_kernel(HierArgs a) {
    __shared__ unsigned long long s_le[DIAG_N], s_ge[DIAG_N];
    load_diag_tables(s_le, s_ge);
    const SearchArgs &s = a.s;
    const int lane = threadIdx.x & 31;
    const uint32_t n_entries = min(*a.wl_in_count, a.wl_cap);
    if (n_entries > (uint32_t)a.tail_split) return;   // search_far_kernel takes this list
    const int D = (int)s.max_rings, Smax = (D + 6) >> 3;
    const uint32_t warps = gridDim.x * (blockDim.x >> 5);
    const int g = lane / 10, item = lane - 10 * g;   // ring of the group, run of the ring (lanes 30, 31: none)
#pragma unroll 1
    for (uint32_t e = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); e < n_entries; e += warps) {
        const uint2 ent = a.wl_in[e];
        const uint32_t qi = ent.x, pair = ent.y >> 21, k21 = ent.y & KEY21_NONE;
        uint32_t best = k21 == KEY21_NONE ? KEY_NONE : k21;
        int ox, oyg;
        uint32_t col;
        query_of<BLOCKS>(s, pair, qi, ox, oyg, col);
        const HierQuery q = make_query(ox, oyg - (int)s.add_y0, col);
        const HierView v = make_view(a.planes, a.g, pair);
        const int T0x = q.ox >> 3, T0y = q.oy >> 3, xmax = v.ntx - 1, ymax = v.nty - 1;
        const int slack = stop_slack(q);
#pragma unroll 1
        for (int S = a.S0; S <= Smax; S += 3) {
            if (T0x - S < 0 && T0x + S > xmax && T0y - S < 0 && T0y + S > ymax) break;   // the coarse square has left the tile grid
            const int R = S + g;
            uint32_t kb = best;
            if (g < 3 && R <= Smax) {
                // runs: 0-2 row T0y+R, 3-5 row T0y-R (tiles T0x-R .. T0x+R); 6-7 column T0x+R, 8-9 column T0x-R (tiles T0y-R+1 .. T0y+R-1)
                const bool is_col = item >= 6;
                const int side = is_col ? 2 + (item - 6) / 2 : item / 3, seg = is_col ? (item - 6) & 1 : item % 3;
                const int line = (is_col ? T0x : T0y) + ((side & 1) ? -R : R);
                const int first = is_col ? T0y - R + 1 : T0x - R, last = is_col ? T0y + R - 1 : T0x + R;
                int lo = first + 32 * seg, hi = lo + 31 < last ? lo + 31 : last;
                const int lim = is_col ? ymax : xmax;
                lo = lo < 0 ? 0 : lo;
                hi = hi > lim ? lim : hi;
                if ((unsigned)line < (unsigned)(is_col ? v.ntx : v.nty) && lo <= hi) {
                    const unsigned long long w = coarse_window(is_col ? v.ccp : v.crp, (uint32_t)(line * (is_col ? v.nyw : v.nxw) + (lo >> 5)), q);
                    uint32_t m = __funnelshift_r((uint32_t)w, (uint32_t)(w >> 32), lo & 31) & ((2u << (hi - lo)) - 1u);
                    while (m) {
                        const int p = lo + __ffs((int)m) - 1;
                        m &= m - 1;
                        fine_test(v, q, is_col ? line : p, is_col ? p : line, D, s_le, s_ge, kb);
                    }
                }
            }
            best = __reduce_min_sync(0xffffffffu, kb);
            const int R_last = S + 2 < Smax ? S + 2 : Smax;
            if (best != KEY_NONE && (int)(best >> 12) <= 8 * R_last + slack) break;
        }
        if (lane == 0) {
            const uint32_t d = best == KEY_NONE ? 0u : (uint32_t)key_to_d8(best);
            s.d8[(size_t)pair * s.n_queries + qi] = (uint8_t)d;
            if (d && d > *(volatile uint32_t *)(s.max_out + pair)) atomicMax(s.max_out + pair, d);
        }
    }
}

#endif  // __CUDACC__

}  // namespace stepth
