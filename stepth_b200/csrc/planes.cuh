// planes.cuh -- K0: candidate bit-planes of the additional image.
//
// The match predicate of the reference, `v.substract(what).lt(precision)` (helpers.rs:35), is a per-channel interval
// test  lo_c(what) <= v_c <= hi_c(what)  (common.cuh).  `precision` is the same for every query of a call, so the
// interval only depends on the query colour.  K0 bins the QUERY colour per channel into NB bins of QSTEP levels and,
// for every bin j of every channel c, stores one bit per additional pixel:
//
//     U[c][j](pixel) = 1  iff  v_c lies in the UNION of the intervals of all query values q in [j*QSTEP, j*QSTEP+QSTEP)
//                    = 1  iff  max(0, j*QSTEP - p_c + 1) <= v_c <= min(255, j*QSTEP + QSTEP - 1 + p_c - 1)      (abs/all)
//
// For a query (r,g,b) the AND of U[0][r/QSTEP], U[1][g/QSTEP], U[2][b/QSTEP] is a SUPERSET of its matches with at most
// QSTEP-1 foreign levels per channel: the search (search_tiles.cuh) tests 32 pixels with three loads and one AND, and
// verifies the (rare) set bits against the real pixel, in the reference's probe order -- so the result stays exact.
// Out-of-image bits are 0: the reference skips out-of-bounds probes (helpers.rs:19-25,44), a 0 bit is never a candidate.
//
// Layout (u32 words).  A "cell" is the 3*NB words describing 32 consecutive pixels; all NB bins of one channel of one
// cell share one 128-byte line, so the lanes of a warp that look at the same 32 pixels with different query colours are
// served by a single L1 wavefront.
//   row-major    RP[(y * nxw + xw) * CELL + c * NB + j]   bit b <-> pixel (xw*32 + b, y)      (row passes of a ring)
//   transposed   CP[(x * nyw + yw) * CELL + c * NB + j]   bit b <-> pixel (x, yw*32 + b)      (column passes)
// Algorithmic bytes per additional pixel: 3 read, 2 * 3 * NB / 8 = 24 written.
#pragma once
#include "common.cuh"

namespace stepth {

constexpr int NB = 32;               // query-colour bins per channel
constexpr int QSHIFT = 3;            // log2(QSTEP)
constexpr int QSTEP = 1 << QSHIFT;   // 8 levels per bin
constexpr int CELL = 3 * NB;         // u32 words per cell (96)
static_assert(NB * QSTEP == 256, "bins must cover the u8 range");

struct PlaneGeom {
    uint32_t wa, ha;    // additional (sub-)image held on the device
    uint32_t nxw, nyw;  // ceil(wa/32), ceil(ha/32)
    size_t rp_words;    // ha * nxw * CELL
    size_t cp_words;    // wa * nyw * CELL
    size_t pair_words;  // rp_words + cp_words
};

inline PlaneGeom make_plane_geom(uint32_t wa, uint32_t ha) {
    PlaneGeom g;
    g.wa = wa; g.ha = ha;
    g.nxw = (wa + 31) / 32;
    g.nyw = (ha + 31) / 32;
    g.rp_words = (size_t)ha * g.nxw * CELL;
    g.cp_words = (size_t)wa * g.nyw * CELL;
    g.pair_words = g.rp_words + g.cp_words;
    return g;
}

struct PlaneArgs {
    const uint8_t *add_rgb;  // [n_pairs, ha, wa, 3]
    uint32_t *planes;        // [n_pairs, pair_words]
    PlaneGeom g;
    uint32_t pr, pg, pb, sat;
};

// union interval of query bin j (see header comment)
__device__ __forceinline__ void bin_interval(uint32_t j, uint32_t p, bool sat, int &lo, int &hi) {
    int qlo = (int)(j << QSHIFT), qhi = qlo + QSTEP - 1;
    lo = sat ? 0 : qlo - (int)p + 1;
    hi = qhi + (int)p - 1;
    if (lo < 0) lo = 0;
    if (hi > 255) hi = 255;
}

// One CTA per 32x32 pixel tile: stage it in shared memory, then every warp turns (row, channel) and (column, channel)
// lines of 32 pixels into NB plane words with one ballot per bin; lane j keeps bin j, so the store is one 128 B line.
__global__ void __launch_bounds__(256) planes_kernel(PlaneArgs a) {
    __shared__ uint8_t s_px[3][32][36];
    const uint32_t pair = blockIdx.z, tx = blockIdx.x, ty = blockIdx.y;
    const uint32_t wa = a.g.wa, ha = a.g.ha;
    const uint8_t *add = a.add_rgb + (size_t)pair * ha * wa * 3;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < 32 * 96; i += 256) {
        int r = i / 96, k = i - r * 96;
        uint32_t x = tx * 32 + k / 3, y = ty * 32 + r;
        uint8_t v = 0;
        if (x < wa && y < ha) v = add[((size_t)y * wa + tx * 32) * 3 + k];
        s_px[k % 3][r][k / 3] = v;
    }
    __syncthreads();
    uint32_t *rp = a.planes + (size_t)pair * a.g.pair_words;
    uint32_t *cp = rp + a.g.rp_words;
    const uint32_t prec[3] = {a.pr, a.pg, a.pb};
    const bool sat = a.sat != 0;
    // row-major cells (y, xw = tx)
    for (int rc = warp; rc < 96; rc += 8) {
        const int r = rc / 3, c = rc - r * 3;
        const uint32_t y = ty * 32 + r;
        if (y >= ha) continue;  // warp-uniform
        const bool valid = tx * 32 + lane < wa;
        const int v = s_px[c][r][lane];
        uint32_t mine = 0;
#pragma unroll 8
        for (int j = 0; j < NB; ++j) {
            int lo, hi;
            bin_interval((uint32_t)j, prec[c], sat, lo, hi);
            uint32_t m = __ballot_sync(0xffffffffu, valid && v >= lo && v <= hi);
            if (lane == j) mine = m;
        }
        rp[((size_t)y * a.g.nxw + tx) * CELL + c * NB + lane] = mine;
    }
    // transposed cells (x, yw = ty): lane <-> row inside the tile
    for (int cc = warp; cc < 96; cc += 8) {
        const int cx = cc / 3, c = cc - cx * 3;
        const uint32_t x = tx * 32 + cx;
        if (x >= wa) continue;  // warp-uniform
        const bool valid = ty * 32 + lane < ha;
        const int v = s_px[c][lane][cx];
        uint32_t mine = 0;
#pragma unroll 8
        for (int j = 0; j < NB; ++j) {
            int lo, hi;
            bin_interval((uint32_t)j, prec[c], sat, lo, hi);
            uint32_t m = __ballot_sync(0xffffffffu, valid && v >= lo && v <= hi);
            if (lane == j) mine = m;
        }
        cp[((size_t)x * a.g.nyw + ty) * CELL + c * NB + lane] = mine;
    }
}

}  // namespace stepth
