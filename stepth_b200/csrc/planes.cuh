// planes.cuh -- K0: match bit-planes of the additional image.
//
// The match predicate of the reference, `v.substract(what).lt(precision)` (helpers.rs:35), is a per-channel interval
// test  T(q) = [q - p + 1, q + p - 1]  (abs/all; sat/all: [0, q + p - 1]) on the pixel value v, q = the query's channel
// value, p = `precision` -- the same p for every query of a call.  K0 turns the additional image into bit-planes (one bit
// per pixel, 32 pixels per word) from which the search evaluates the predicate for 32 probes with a few loads and ANDs.
// Per channel, with the query value split as q = 8 j + r:
//
//   U[j]    v in the UNION of T(q) over the bin q in [8j, 8j+7]:  max(0, 8j-p+1) <= v <= min(255, 8j+p+6)   j = -1 .. 32
//   NTA[r]  ((v + p - 1) & 7) >= r        (all ones in sat mode)                                          r = 0 .. 7
//   NTB[r]  ((v - p + 1) & 7) <= r                                                                        r = 0 .. 7
//
//   candidate(q) = U[j]                                    a superset of the matches: 3 loads + 1 AND per 32 probes
//   exact(q)     = U[j] & (U[j+1] | NTA[r]) & (U[j-1] | NTB[r])
//
// Why: inside U[j] the pixels below T(q) are exactly those not in U[j+1] whose offset (v+p-1)&7 is < r, and the pixels
// above T(q) are exactly those not in U[j-1] whose offset (v-p+1)&7 is > r (planes.cuh derivation in DESIGN.md).  The
// search tests `candidate` first and refines a non-zero word with `exact` (12 more loads), so its result is bit-exact
// without ever touching the pixels again.  Out-of-image bits are 0 in every U plane: the reference skips out-of-bounds
// probes (helpers.rs:19-25,44).
//
// Layout (u32 words).  A "cell" describes 32 consecutive pixels: 3 channels x 64 words
//   [0..31] U[0..31] | [32] U[-1] | [33] U[32] | [34..41] NTA[0..7] | [42..49] NTB[0..7] | [50..63] unused
// so the U words of one channel of one cell share one 128-byte line: lanes of a warp that look at the same 32 pixels with
// different query colours are served by a single L1 wavefront.
//   row-major    RP[(y * nxw + xw) * CELL + c * 64 + word]   bit b <-> pixel (xw*32 + b, y)      (row passes of a ring)
//   transposed   CP[(x * nyw + yw) * CELL + c * 64 + word]   bit b <-> pixel (x, yw*32 + b)      (column passes)
// Algorithmic bytes per additional pixel: 3 read, 2 * 24 = 48 written.
#pragma once
#include "common.cuh"

namespace stepth {

constexpr int NB = 32;               // query-value bins per channel
constexpr int QSHIFT = 3;            // log2(QSTEP)
constexpr int QSTEP = 1 << QSHIFT;   // 8 levels per bin
constexpr int CH_WORDS = 64;         // u32 words per channel of a cell
constexpr int CELL = 3 * CH_WORDS;   // u32 words per cell (192)
constexpr int W_UM1 = 32, W_UP32 = 33, W_NTA = 34, W_NTB = 42;
static_assert(NB * QSTEP == 256 && NB == 32, "the cell layout assumes 32 bins of 8 levels");

struct PlaneGeom {
    uint32_t wa, ha;    // the part of the additional (sub-)image the planes cover: columns [0, wa), rows [0, ha)
    uint32_t src_wa;    // row pitch (pixels) of the source image
    uint32_t nxw, nyw;  // ceil(wa/32), ceil(ha/32)
    size_t rp_words;    // ha * nxw * CELL
    size_t cp_words;    // wa * nyw * CELL
    size_t pair_words;  // rp_words + cp_words
};

inline PlaneGeom make_plane_geom(uint32_t wa, uint32_t ha, uint32_t src_wa) {
    PlaneGeom g;
    g.wa = wa; g.ha = ha; g.src_wa = src_wa;
    g.nxw = (wa + 31) / 32;
    g.nyw = (ha + 31) / 32;
    g.rp_words = (size_t)ha * g.nxw * CELL;
    g.cp_words = (size_t)wa * g.nyw * CELL;
    g.pair_words = g.rp_words + g.cp_words;
    return g;
}

struct PlaneArgs {
    const uint8_t *add_rgb;   // [n_pairs, src_ha, src_wa, 3]
    size_t src_pair_bytes;    // src_ha * src_wa * 3
    uint32_t *planes;         // [n_pairs, pair_words]
    PlaneGeom g;
    uint32_t pr, pg, pb, sat;
};

// word offsets (inside one channel of a cell) a query value q needs
__device__ __forceinline__ uint32_t w_u(uint32_t q) { return q >> QSHIFT; }
__device__ __forceinline__ uint32_t w_um1(uint32_t q) { uint32_t j = q >> QSHIFT; return j ? j - 1 : (uint32_t)W_UM1; }
__device__ __forceinline__ uint32_t w_up1(uint32_t q) { uint32_t j = q >> QSHIFT; return j < NB - 1 ? j + 1 : (uint32_t)W_UP32; }
__device__ __forceinline__ uint32_t w_nta(uint32_t q) { return W_NTA + (q & (QSTEP - 1)); }
__device__ __forceinline__ uint32_t w_ntb(uint32_t q) { return W_NTB + (q & (QSTEP - 1)); }

// exact(q) for the three channels of colour `col`, given the candidate word m of the same cell
__device__ __forceinline__ uint32_t refine_exact(const uint32_t *__restrict__ cell, uint32_t col, uint32_t m) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const uint32_t q = (col >> (8 * c)) & 0xffu;
        const uint32_t *ch = cell + c * CH_WORDS;
        m &= (__ldg(ch + w_up1(q)) | __ldg(ch + w_nta(q))) & (__ldg(ch + w_um1(q)) | __ldg(ch + w_ntb(q)));
    }
    return m;
}

// membership of one pixel value in the 50 planes of its channel: lo = U[0..31], hi = bits 0..17 <-> words 32..49
__device__ __forceinline__ void plane_bits(int v, int p, bool sat, bool valid, uint32_t &lo, uint32_t &hi) {
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

// 32x32 bit-matrix transpose across the lanes of a warp: in: lane i holds row i; out: lane i holds column i
__device__ __forceinline__ uint32_t warp_bit_transpose(uint32_t x, int lane) {
#pragma unroll
    for (int s = 0; s < 5; ++s) {
        const int d = 16 >> s;
        const uint32_t m = d == 16 ? 0x0000ffffu : d == 8 ? 0x00ff00ffu : d == 4 ? 0x0f0f0f0fu : d == 2 ? 0x33333333u : 0x55555555u;
        const uint32_t y = __shfl_xor_sync(0xffffffffu, x, d);
        x = (lane & d) ? (((y >> d) & m) | (x & ~m)) : ((x & m) | ((y << d) & ~m));
    }
    return x;
}

// One CTA per 32x32 pixel tile: stage it in shared memory; every warp turns (row, channel) and (column, channel) lines of
// 32 pixels into their 50 plane words: each lane computes the plane membership of its pixel (two words), two warp bit
// transposes make lane j the owner of plane word j, and the store is one coalesced 128-byte line per half.
__global__ void __launch_bounds__(256) planes_kernel(PlaneArgs a) {
    __shared__ uint8_t s_px[3][32][36];
    const uint32_t pair = blockIdx.z, tx = blockIdx.x, ty = blockIdx.y;
    const uint32_t wa = a.g.wa, ha = a.g.ha;
    const uint8_t *add = a.add_rgb + (size_t)pair * a.src_pair_bytes;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < 32 * 96; i += 256) {
        int r = i / 96, k = i - r * 96;
        uint32_t x = tx * 32 + k / 3, y = ty * 32 + r;
        uint8_t v = 0;
        if (x < wa && y < ha) v = add[((size_t)y * a.g.src_wa + tx * 32) * 3 + k];
        s_px[k % 3][r][k / 3] = v;
    }
    __syncthreads();
    uint32_t *rp = a.planes + (size_t)pair * a.g.pair_words;
    uint32_t *cp = rp + a.g.rp_words;
    const int prec[3] = {(int)a.pr, (int)a.pg, (int)a.pb};
    const bool sat = a.sat != 0;
    // row-major cells (y, xw = tx): lane <-> pixel x
    for (int rc = warp; rc < 96; rc += 8) {
        const int r = rc / 3, c = rc - r * 3;
        const uint32_t y = ty * 32 + r;
        if (y >= ha) continue;  // warp-uniform
        uint32_t lo, hi;
        plane_bits(s_px[c][r][lane], prec[c], sat, tx * 32 + lane < wa, lo, hi);
        lo = warp_bit_transpose(lo, lane);
        hi = warp_bit_transpose(hi, lane);
        uint32_t *dst = rp + ((size_t)y * a.g.nxw + tx) * CELL + c * CH_WORDS;
        dst[lane] = lo;
        dst[32 + lane] = hi;
    }
    // transposed cells (x, yw = ty): lane <-> pixel y
    for (int cc = warp; cc < 96; cc += 8) {
        const int cx = cc / 3, c = cc - cx * 3;
        const uint32_t x = tx * 32 + cx;
        if (x >= wa) continue;  // warp-uniform
        uint32_t lo, hi;
        plane_bits(s_px[c][lane][cx], prec[c], sat, ty * 32 + lane < ha, lo, hi);
        lo = warp_bit_transpose(lo, lane);
        hi = warp_bit_transpose(hi, lane);
        uint32_t *dst = cp + ((size_t)x * a.g.nyw + ty) * CELL + c * CH_WORDS;
        dst[lane] = lo;
        dst[32 + lane] = hi;
    }
}

}  // namespace stepth
