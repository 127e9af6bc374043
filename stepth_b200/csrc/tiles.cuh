// tiles.cuh -- K0: two-level match bit-planes of the additional image (8x8-pixel tiles + a tile-occupancy level).
//
// The match predicate of the reference, `v.substract(what).lt(precision)` (helpers.rs:35), is a per-channel interval
// test  T(q) = [q - p + 1, q + p - 1]  (abs/all; sat/all: [0, q + p - 1]) on the pixel value v, q = the query's channel
// value, p = `precision` -- the same p for every query of a call.  K0 turns the additional image into bit-planes from
// which the search evaluates the predicate for 64 probes (one 8x8 tile) with a few 8-byte loads and ANDs.  Per channel,
// with the query value split as q = 8 j + r:
//
//   U[j]    v in the UNION of T(q) over the bin q in [8j, 8j+7]:  max(0, 8j-p+1) <= v <= min(255, 8j+p+6)   j = -1 .. 32
//   NTA[r]  ((v + p - 1) & 7) >= r        (all ones in sat mode)                                          r = 0 .. 7
//   NTB[r]  ((v - p + 1) & 7) <= r                                                                        r = 0 .. 7
//
//   candidate(q) = U[j]                                    a superset of the matches: 3 loads + 2 ANDs per 64 probes
//   exact(q)     = U[j] & (U[j+1] | NTA[r]) & (U[j-1] | NTB[r])
//
// Why exact: inside U[j] the pixels below T(q) are exactly those not in U[j+1] whose offset (v+p-1)&7 is < r, and the
// pixels above T(q) are exactly those not in U[j-1] whose offset (v-p+1)&7 is > r (derivation in DESIGN.md).  Out-of-image
// bits are 0 in every plane: the reference skips out-of-bounds probes (helpers.rs:19-25,44).
//
// FINE level (u64 words).  One record per 8x8 tile, 3 channels x 64 slots:
//   slot 0 U[-1] | 1..32 U[0..31] | 33 U[32] | 34..41 NTA[0..7] | 42..49 NTB[0..7] | 50..63 unused (never written)
//   (U[j-1], U[j], U[j+1] of a query are three consecutive words: one address, no special cases at j = 0 and j = 31)
//   F[((TY * ntx + TX) * 3 + c) * 64 + slot]      bit (ly * 8 + lx) <-> pixel (8 TX + lx, 8 TY + ly)
// The 34 U words of one channel of one tile are contiguous: lanes that look at the same tile with different query colours
// hit two or three 128-byte lines.
//
// COARSE level: E_c[q] bit = "some pixel of this tile has channel c inside T(q)", one plane per exact query value
// q = 0..255.  A cell is a 64-tile WINDOW that starts every 32 tiles (u64 = {tiles 32w .. 32w+31, tiles 32w+32 .. 32w+63}),
// so one 8-byte load holds any run of up to 33 tiles -- a whole side of a coarse ring up to ring 16 -- wherever it starts:
//   row-major   CR[((TY * nxw + xw) * 3 + c) * 256 + q]   bit b <-> tile (32 xw + b, TY)     (row sides of a coarse ring)
//   transposed  CC[((TX * nyw + yw) * 3 + c) * 256 + q]   bit b <-> tile (TX, 32 yw + b)     (column sides)
// A tile whose three coarse bits are set for a query MAY hold a match (the three channels may be satisfied by different
// pixels); a tile with a clear bit cannot.  Per tile and channel E is the set of values present in the tile, dilated by
// p - 1 on both sides (abs) / smeared upwards from min - (p - 1) (sat).
//
// Algorithmic bytes per additional pixel: 3 read; 3 * 50 * 8 / 64 = 18.75 written (fine) + 2 * 2 * 3 * 256 / (8 * 64) = 6 (coarse).
#pragma once
#include <cuda.h>   // CUtensorMap (the encode function is resolved at run time through cudaGetDriverEntryPoint: no libcuda link)

#include "common.cuh"

namespace stepth {

constexpr int NB = 32;               // query-value bins per channel
constexpr int QSHIFT = 3;            // log2(QSTEP)
constexpr int QSTEP = 1 << QSHIFT;   // 8 levels per bin
constexpr int T_SLOTS = 64;          // u64 slots per channel of a tile record
constexpr int T_REC = 3 * T_SLOTS;   // u64 words per tile record (192 = 1536 bytes)
constexpr int C_Q = 256;             // coarse planes per channel: one per exact query value
constexpr int C_CELL = 3 * C_Q;      // u64 windows per coarse cell (768 = 6144 bytes)
constexpr int W_UM1 = 0, W_U0 = 1, W_UP32 = 33, W_NTA = 34, W_NTB = 42, W_USED = 50;
static_assert(NB * QSTEP == 256 && NB == 32, "the layouts assume 32 bins of 8 levels");

struct TileGeom {
    uint32_t wa, ha;      // the part of the additional (sub-)image the planes cover: columns [0, wa), rows [0, ha)
    uint32_t src_wa;      // row pitch (pixels) of the source image
    uint32_t ntx, nty;    // tiles: ceil(wa/8), ceil(ha/8)
    uint32_t nxw, nyw;    // coarse words per tile row / tile column: ceil(ntx/32), ceil(nty/32)
    size_t fine_words;    // u64: ntx * nty * T_REC
    size_t crp_words;     // u64: nty * nxw * C_CELL
    size_t ccp_words;     // u64: ntx * nyw * C_CELL
    size_t off_crp, off_ccp;  // byte offsets of the coarse sections inside one pair's block
    size_t pair_bytes;    // bytes per pair (sections 256-byte aligned)
};

inline TileGeom make_tile_geom(uint32_t wa, uint32_t ha, uint32_t src_wa) {
    TileGeom g;
    g.wa = wa; g.ha = ha; g.src_wa = src_wa;
    g.ntx = (wa + 7) / 8;
    g.nty = (ha + 7) / 8;
    g.nxw = (g.ntx + 31) / 32;
    g.nyw = (g.nty + 31) / 32;
    g.fine_words = (size_t)g.ntx * g.nty * T_REC;
    g.crp_words = (size_t)g.nty * g.nxw * C_CELL;
    g.ccp_words = (size_t)g.ntx * g.nyw * C_CELL;
    auto up = [](size_t b) { return (b + 255) & ~(size_t)255; };
    g.off_crp = up(g.fine_words * 8);
    g.off_ccp = g.off_crp + up(g.crp_words * 8);
    g.pair_bytes = g.off_ccp + up(g.ccp_words * 8);
    return g;
}

// slot offsets (inside one channel of a tile record) a query value q needs
STEPTH_HD uint32_t w_um1(uint32_t q) { return q >> QSHIFT; }          // U[j - 1]
STEPTH_HD uint32_t w_u(uint32_t q) { return (q >> QSHIFT) + 1; }      // U[j]
STEPTH_HD uint32_t w_up1(uint32_t q) { return (q >> QSHIFT) + 2; }    // U[j + 1]
STEPTH_HD uint32_t w_nta(uint32_t q) { return W_NTA + (q & (QSTEP - 1)); }
STEPTH_HD uint32_t w_ntb(uint32_t q) { return W_NTB + (q & (QSTEP - 1)); }

// membership of one pixel value in the 50 planes of its channel: lo = U[0..31], hi = bits 0..17 <-> slots 32..49
STEPTH_HD void plane_bits(int v, int p, bool sat, bool valid, uint32_t &lo, uint32_t &hi) {
    if (!valid) { lo = 0; hi = 0; return; }
    // U[j] contains v  <=>  8j - p + 1 <= v (abs only)  and  v <= 8j + p + 6
    int jb = sat ? NB : (v + p - 1) >> 3;            // largest j with 8j <= v + p - 1
    const int ja = (v - p - 6 + 7) >> 3;              // smallest j with 8j >= v - p - 6 (arithmetic shift = floor)
    // bits for j = 0 .. 31
    uint32_t m_hi_side = jb >= 31 ? 0xffffffffu : (jb < 0 ? 0u : (0xffffffffu >> (31 - jb)));   // j <= jb
    uint32_t m_lo_side = ja <= 0 ? 0xffffffffu : (ja > 31 ? 0u : (0xffffffffu << ja));          // j >= ja
    lo = m_hi_side & m_lo_side;
    uint32_t h = 0;
    if (ja <= -1 && jb >= -1) h |= 1u;                // U[-1]:  v <= p - 2 (and, abs, -p - 7 <= v: always)
    if (ja <= 32 && jb >= 32) h |= 2u;                // U[32]:  v >= 257 - p (abs) / always (sat); v <= 262 + p always
    const uint32_t a = sat ? 7u : (uint32_t)((v + p - 1) & 7), b = (uint32_t)((v - p + 1) & 7);
    h |= ((1u << (a + 1)) - 1u) << 2;                 // NTA[r]: r <= a
    h |= ((0xffu << b) & 0xffu) << 10;                // NTB[r]: r >= b
    hi = h;
}

#if defined(__CUDACC__)

struct TileBuildArgs {
    const uint8_t *add_rgb;   // [n_pairs, src_ha, src_wa, 3]
    size_t src_pair_bytes;    // src_ha * src_wa * 3
    uint8_t *planes;          // [n_pairs, pair_bytes]
    TileGeom g;
    uint32_t pr, pg, pb, sat;
};

// 32x32 bit-matrix transpose across the lanes of a warp: in: lane i holds row i; out: lane i holds column i
__device__ __forceinline__ uint32_t warp_bit_transpose(uint32_t x, int lane) {
    {   // the 16- and 8-bit steps move whole bytes: one byte permute each
        uint32_t y = __shfl_xor_sync(0xffffffffu, x, 16);
        x = __byte_perm(x, y, (lane & 16) ? 0x3276u : 0x5410u);   // hi: {y.b2, y.b3 | x.b2, x.b3}   lo: {x.b0, x.b1 | y.b0, y.b1}
        y = __shfl_xor_sync(0xffffffffu, x, 8);
        x = __byte_perm(x, y, (lane & 8) ? 0x3715u : 0x6240u);    // hi: bytes {y1, x1, y3, x3}      lo: bytes {x0, y0, x2, y2}
    }
    // 4-, 2-, 1-bit steps: a lane keeps the half of every 2d-bit group that stays (mask km), sends the other half already
    // moved to where the partner wants it (a rotate: the bits that would wrap are zero after masking) and ORs in what comes
#pragma unroll
    for (int s = 2; s < 5; ++s) {
        const int d = 16 >> s;
        const uint32_t m = d == 4 ? 0x0f0f0f0fu : d == 2 ? 0x33333333u : 0x55555555u;
        const uint32_t km = (lane & d) ? ~m : m;             // lane-constant: hoisted out of the callers' loops
        const int rot = (lane & d) ? d : 32 - d;             // upper lane sends its low halves up (<< d), lower lane its high halves down (>> d)
        const uint32_t give = x & ~km;
        const uint32_t y = __shfl_xor_sync(0xffffffffu, __funnelshift_l(give, give, rot), d);
        x = (x & km) | y;
    }
    return x;
}

// ---- 512-bit shifts spread over the 16 lanes of a half warp (lane16 = word index, word 0 = bits 0..31) ------------------
__device__ __forceinline__ uint32_t shl512(uint32_t x, int lane16, int nbits) {  // nbits warp-uniform, 0 <= nbits < 512
    if (nbits < 32) {  // the doubling steps of a dilation by a typical precision: the neighbour word is all that moves
        uint32_t lo = __shfl_up_sync(0xffffffffu, x, 1, 16);
        if (lane16 == 0) lo = 0;
        return __funnelshift_l(lo, x, nbits);
    }
    const int wsh = nbits >> 5, bsh = nbits & 31;
    uint32_t hi = __shfl_up_sync(0xffffffffu, x, wsh, 16);
    uint32_t lo = __shfl_up_sync(0xffffffffu, x, (wsh + 1) & 15, 16);
    if (lane16 < wsh) hi = 0;
    if (lane16 < wsh + 1) lo = 0;
    return __funnelshift_l(lo, hi, bsh);
}
__device__ __forceinline__ uint32_t shr512(uint32_t x, int lane16, int nbits) {
    const int wsh = nbits >> 5, bsh = nbits & 31;
    uint32_t lo = __shfl_down_sync(0xffffffffu, x, wsh, 16);
    uint32_t hi = __shfl_down_sync(0xffffffffu, x, (wsh + 1) & 15, 16);
    if (lane16 + wsh > 15) lo = 0;
    if (lane16 + wsh + 1 > 15) hi = 0;
    return __funnelshift_r(lo, hi, bsh);
}

// E = { q : some value v of the set P satisfies v in T(q) }, P and E as 256-bit sets in words 0..7 of a half warp.
//   abs: T(q) = [q - R, q + R]  ->  E[q] = OR_{s = 0..2R} P[q + R - s]   (window-OR by doubling, then a right shift by R)
//   sat: T(q) = [0, q + R]      ->  the same with an unbounded window
__device__ __forceinline__ uint32_t dilate_values(uint32_t x, int lane16, int R, bool sat) {
    const int W = sat ? 512 : 2 * R + 1;  // window width, 1 <= W <= 512
    int span = 1;                         // x covers the window [0, span)
    while (2 * span <= W) {
        x |= shl512(x, lane16, span);
        span *= 2;
    }
    if (W > span) x |= shl512(x, lane16, W - span);
    return shr512(x, lane16, R);
}

// One CTA per coarse row-major cell = 32 tiles along x of one tile row = a 256x8 pixel strip.  The strip is staged in
// shared memory; every warp takes 4 tiles x 3 channels: each lane computes the plane membership of its pixel (rows 0-3,
// then rows 4-7 of the tile), two warp bit transposes per half make lane j the owner of plane word j, and the stores are
// 256 contiguous bytes (U) + 144 bytes (the 18 extra planes).  The coarse level: the set of values present in the tile
// (warp OR-reductions) is dilated into E, 256 bits per (tile, channel); the CTA transposes the 32 tiles x 256 values of
// a channel into 256 words of 32 tile bits and stores the cell with coalesced lines.
// `tmap` (use_tmap != 0): a rank-3 tensor map over the additional images as u32 words {row words, rows, pairs}; the CTA's strip
// is ONE 2-D box of 192 words x 8 rows brought in by the TMA unit (cp.async.bulk.tensor), rows and words beyond the image
// zero-filled by the unit.  A tensor map needs what the per-row bulk copies need (rows that start on 16 bytes), so images
// whose pitch is not a multiple of 16 bytes -- the README pair's 1800-byte rows -- stay on the plain-load paths below.
__global__ void __launch_bounds__(256) tile_planes_kernel(TileBuildArgs a, const __grid_constant__ CUtensorMap tmap, int use_tmap) {
    __shared__ __align__(128) uint32_t s_raw[8][192];   // [row][interleaved RGB of 256 pixels]
    __shared__ __align__(8) unsigned long long s_bar;   // mbarrier of the bulk copies
    __shared__ uint2 s_tab[3][256];       // plane membership (plane_bits) of every value of a channel
    __shared__ uint32_t s_E[3][8][32];  // [channel][value word g][tile]: bit b <-> q = 32 g + b
    const uint32_t pair = blockIdx.z, xw = blockIdx.x, TY = blockIdx.y;
    const uint32_t wa = a.g.wa, ha = a.g.ha;
    const uint8_t *add = a.add_rgb + (size_t)pair * a.src_pair_bytes;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, lane16 = lane & 15;
    const uint32_t x_base = xw * 256u, y_base = TY * 8u;
    // the strip as it lies in memory (interleaved RGB, 768 bytes per row).  When the rows start on 16 bytes the TMA unit
    // brings them in (one bulk copy per row, completion counted on an mbarrier) while the CTA builds its tables
    const size_t pitch = (size_t)a.g.src_wa * 3;
    const uint32_t row_bytes = min(768u, (a.g.src_wa - x_base) * 3u);   // x_base < wa <= src_wa
    const uint32_t rows = min(8u, ha - y_base);
    const bool bulk = (pitch & 15) == 0 && (reinterpret_cast<uintptr_t>(add) & 15) == 0 && (row_bytes & 15) == 0;
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_bar);
    if (use_tmap) {
        if (tid == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(1));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(8u * 768u) : "memory");   // the whole box, zero fill included
            const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&s_raw[0][0]);
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst),
                         "l"(&tmap), "r"((int)(xw * 192u)), "r"((int)y_base), "r"((int)pair), "r"(bar)
                         : "memory");
        }
    } else if (bulk) {
        if (tid == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(1));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(rows * row_bytes) : "memory");
            for (uint32_t r = 0; r < rows; ++r) {
                const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&s_raw[r][0]);
                const uint8_t *src = add + (size_t)(y_base + r) * pitch + (size_t)x_base * 3;
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                             "l"(src), "r"(row_bytes), "r"(bar)
                             : "memory");
            }
        }
    } else if ((pitch & 3) == 0 && (reinterpret_cast<uintptr_t>(add) & 3) == 0) {
        const uint32_t row_words = (uint32_t)(pitch >> 2);            // words of one image row
        for (int i = tid; i < 8 * 192; i += 256) {
            const int r = i / 192, k = i - r * 192;
            const uint32_t y = y_base + r, wi = xw * 192u + k;         // word inside the row
            uint32_t v = 0;
            if (y < ha && wi < row_words) v = reinterpret_cast<const uint32_t *>(add + (size_t)y * pitch)[wi];
            s_raw[r][k] = v;
        }
    } else {
        uint8_t *raw = reinterpret_cast<uint8_t *>(&s_raw[0][0]);
        for (int i = tid; i < 8 * 768; i += 256) {
            const int r = i / 768, k = i - r * 768;
            const uint32_t x = x_base + k / 3, y = y_base + r;
            uint8_t v = 0;
            if (x < a.g.src_wa && y < ha) v = add[(size_t)y * pitch + (size_t)x_base * 3 + k];
            raw[r * 768 + k] = v;
        }
    }
    for (int i = tid; i < 3 * 8 * 32; i += 256) (&s_E[0][0][0])[i] = 0;
    const int prec[3] = {(int)a.pr, (int)a.pg, (int)a.pb};
    const bool sat = a.sat != 0;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        uint32_t lo, hi;
        plane_bits(tid, prec[c], sat, true, lo, hi);
        s_tab[c][tid] = make_uint2(lo, hi);
    }
    __syncthreads();   // tables, s_E, the staged strip of the non-bulk paths -- and the mbarrier's initialisation
    if (use_tmap || bulk) {   // rows and columns beyond the image: zero (tensor map) or whatever shared memory held (bulk rows); masked either way (ok0 / ok1)
        uint32_t done;
        do {
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done)
                         : "r"(bar), "r"(0)
                         : "memory");
        } while (!done);
    }
    uint8_t *pair_base = a.planes + (size_t)pair * a.g.pair_bytes;
    unsigned long long *fine = reinterpret_cast<unsigned long long *>(pair_base);
    const uint8_t *raw = reinterpret_cast<const uint8_t *>(&s_raw[0][0]);
    const bool same_prec = prec[0] == prec[1] && prec[1] == prec[2];
#pragma unroll 1
    for (int t = 0; t < 4; ++t) {
        const int tl = warp * 4 + t;                 // tile inside the cell
        const uint32_t TX = xw * 32u + tl;
        if (TX >= a.g.ntx) break;                    // warp-uniform
        const int lx = lane & 7, ly = lane >> 3;
        const bool ok0 = TX * 8u + lx < wa && y_base + ly < ha, ok1 = TX * 8u + lx < wa && y_base + ly + 4 < ha;
        uint32_t pw01 = 0, pw2 = 0;  // value-presence words: lanes 0-15 channel 0 | lanes 16-31 channel 1;  channel 2
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const int v0 = raw[ly * 768 + (tl * 8 + lx) * 3 + c], v1 = raw[(ly + 4) * 768 + (tl * 8 + lx) * 3 + c];
            const uint2 t0 = s_tab[c][v0], t1 = s_tab[c][v1];
            uint32_t lo0 = ok0 ? t0.x : 0u, hi0 = ok0 ? t0.y : 0u, lo1 = ok1 ? t1.x : 0u, hi1 = ok1 ? t1.y : 0u;
            lo0 = warp_bit_transpose(lo0, lane);
            lo1 = warp_bit_transpose(lo1, lane);
            hi0 = warp_bit_transpose(hi0, lane);
            hi1 = warp_bit_transpose(hi1, lane);
            unsigned long long *rec = fine + (((size_t)TY * a.g.ntx + TX) * 3 + c) * T_SLOTS;
            rec[W_U0 + lane] = (unsigned long long)lo0 | ((unsigned long long)lo1 << 32);
            // hi word of lane 0 = U[-1] -> slot 0; lane 1 = U[32] -> slot 33; lanes 2..17 = NTA / NTB -> slots 34..49
            if (lane < W_USED - 32) rec[lane ? 32 + lane : W_UM1] = (unsigned long long)hi0 | ((unsigned long long)hi1 << 32);
            // the set of values present in the tile, one 32-value word at a time: only the words some pixel falls into
            // are visited (a natural tile spans one to three of the eight)
            uint32_t need = (ok0 ? 1u << (v0 >> 5) : 0u) | (ok1 ? 1u << (v1 >> 5) : 0u);
            need = __reduce_or_sync(0xffffffffu, need);
            while (need) {
                const int g = __ffs((int)need) - 1;
                need &= need - 1;
                uint32_t contrib = 0;
                if (ok0 && (v0 >> 5) == g) contrib |= 1u << (v0 & 31);
                if (ok1 && (v1 >> 5) == g) contrib |= 1u << (v1 & 31);
                const uint32_t pg = __reduce_or_sync(0xffffffffu, contrib);
                if (lane16 == g) {
                    if (c == 0 && lane < 16) pw01 = pg;
                    if (c == 1 && lane >= 16) pw01 = pg;
                    if (c == 2) pw2 = pg;
                }
            }
        }
        // the precision differs per channel, so the three dilations are separate calls (warp-uniform shift counts)
        const uint32_t e0 = dilate_values(pw01, lane16, prec[0] - 1, sat);   // lanes 0-15 carry channel 0
        const uint32_t e1 = same_prec ? e0 : dilate_values(pw01, lane16, prec[1] - 1, sat);   // lanes 16-31 carry channel 1
        const uint32_t e2 = dilate_values(pw2, lane16, prec[2] - 1, sat);    // both halves carry channel 2
        if (lane16 < 8) {
            if (lane < 16) { s_E[0][lane16][tl] = e0; s_E[2][lane16][tl] = e2; }
            else s_E[1][lane16][tl] = e1;
        }
    }
    __syncthreads();
    // 24 (channel, value word) blocks of 32 tiles x 32 values -> 32 values x 32 tiles
    // window xw gets these 32 tiles as its low half, window xw - 1 as its high half; the last window's high half is empty
    uint32_t *crp = reinterpret_cast<uint32_t *>(pair_base + a.g.off_crp) + ((size_t)TY * a.g.nxw + xw) * C_CELL * 2;
    for (int cg = warp; cg < 24; cg += 8) {
        const int c = cg >> 3, g = cg & 7;
        const uint32_t x = warp_bit_transpose(s_E[c][g][lane], lane);  // lane b: tile bits of q = 32 g + b
        const int slot = (c * C_Q + g * 32 + lane) * 2;
        crp[slot] = x;
        if (xw > 0) crp[slot + 1 - C_CELL * 2] = x;
        if (xw + 1 == a.g.nxw) crp[slot + 1] = 0u;
    }
}

// coarse row-major -> coarse transposed: one warp per (32x32 tile block, channel*256 + value); lane r reads the word of tile
// row 32 by + r, the warp transposes the 32x32 bit block, lane k writes the word of tile column 32 bx + k.
constexpr int CT_PARTS = 8;  // CTAs per (32x32 tile block, pair): each transposes C_CELL / CT_PARTS planes

__global__ void __launch_bounds__(256) coarse_transpose_kernel(TileBuildArgs a) {
    const uint32_t pair = blockIdx.z / CT_PARTS, part = blockIdx.z % CT_PARTS, bx = blockIdx.x, by = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint8_t *pair_base = a.planes + (size_t)pair * a.g.pair_bytes;
    const uint32_t *crp = reinterpret_cast<const uint32_t *>(pair_base + a.g.off_crp);
    uint32_t *ccp = reinterpret_cast<uint32_t *>(pair_base + a.g.off_ccp);
    constexpr int PER_WARP = C_CELL / CT_PARTS / 8;  // 12 planes per warp, four loads in flight at a time
    const uint32_t TY = by * 32u + lane;
    const uint32_t *src = crp + ((size_t)TY * a.g.nxw + bx) * C_CELL * 2;
    const int cj0 = (int)part * (C_CELL / CT_PARTS) + warp * PER_WARP;
    // all twelve low halves are requested before any is used (the kernel is latency bound): six 16-byte loads per lane, each
    // bringing two (low, high) window words of which the low ones belong to this 32x32 block
    static_assert(PER_WARP % 2 == 0, "two planes per 16-byte load");
    uint32_t x[PER_WARP];
    const uint4 *src4 = reinterpret_cast<const uint4 *>(src + (size_t)cj0 * 2);  // cj0 is a multiple of 12: 32-byte aligned
#pragma unroll
    for (int i = 0; i < PER_WARP / 2; ++i) {
        const uint4 t = TY < a.g.nty ? __ldg(src4 + i) : make_uint4(0, 0, 0, 0);
        x[2 * i] = t.x;
        x[2 * i + 1] = t.z;
    }
    // transposed words go through shared memory so that a warp stores one tile column at a time: the 96 planes of this part
    // are 96 consecutive (low, high) words of the column's cell -> runs of 8-byte-strided words instead of 32 scattered ones
    constexpr int PART = C_CELL / CT_PARTS;            // 96 planes
    __shared__ uint32_t s_t[32][PART + 1];             // [tile column][plane]
#pragma unroll
    for (int i = 0; i < PER_WARP; ++i) s_t[lane][warp * PER_WARP + i] = warp_bit_transpose(x[i], lane);
    __syncthreads();
    const int p0 = (int)part * PART;
    for (int k = warp; k < 32; k += 8) {               // tile column 32 bx + k
        const uint32_t TXk = bx * 32u + k;
        if (TXk >= a.g.ntx) break;                     // warp-uniform; later columns of this warp are beyond the grid too
        uint32_t *d = ccp + ((size_t)TXk * a.g.nyw + by) * C_CELL * 2 + p0 * 2;
#pragma unroll
        for (int j = lane; j < PART; j += 32) {
            const uint32_t t = s_t[k][j];
            d[j * 2] = t;
            if (by > 0) d[j * 2 + 1 - C_CELL * 2] = t;
            if (by + 1 == a.g.nyw) d[j * 2 + 1] = 0u;
        }
    }
}

#endif  // __CUDACC__

}  // namespace stepth
