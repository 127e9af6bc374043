// search_origins.cuh -- K1 (dense mode): the production ring search, one thread per ORIGIN.
//
// depth_image.rs:114-117 puts the search origin of main pixel (x, y) at ((x+1)/2, (y+1)/2): the four main pixels
// (2k-1 .. 2k, 2l-1 .. 2l) share origin (k, l), hence every ring, every row/column segment, every plane cell and every
// clip mask.  A thread therefore walks the rings of ONE origin for its (up to) four queries: addresses, ranges and masks
// are computed once, each query adds three plane loads and one AND per 32 probes (planes.cuh).
//
//   * a CTA owns 16x16 origins (32x32 main pixels); ring 0 is one RGBX load per origin;
//   * rings are walked in chunks [s0, s1) of growing thickness; per chunk the origins that still have an unresolved
//     query are compacted in shared memory into a row-ordered and a column-ordered list.  ROW pass: a warp holds
//     consecutive origins of (at most two) origin rows and reads rows y+s, y-s of the row-major planes; COLUMN pass:
//     threads are re-assigned so that a warp holds origins of one or two origin columns and reads columns x+s, x-s of the
//     transposed planes -- in both passes the lanes of a warp read the same 128-byte plane lines;
//   * the reference's probe order (helpers.rs:26-31: row y+s, row y-s, column x+s, column x-s, ascending inside each)
//     is the order of the ring key (common.cuh); a query's answer is min(row-pass key, column-pass key) and the column
//     pass of a chunk only walks rings below that chunk's row hit;
//   * plane bits are a superset of the matches: every set bit is verified, in probe order, against the pixel's RGBX
//     word (one load, byte-SIMD |v - what| < precision); out-of-image probes are 0 bits / clipped ranges.
#pragma once
#include "search_tiles.cuh"

namespace stepth {

// Walks the segment [lo, hi] of plane line `line` for the queries in `need` (bit k <-> query k of the origin).
// candidate word (3 loads + AND) -> exact word (planes.cuh refine_exact) -> first set bit = first match of the segment.
// A hit at position p stores key[k] = keybase + p and clears bit k.  Returns the remaining `need`.
__device__ __forceinline__ uint32_t scan4(const uint32_t *__restrict__ planes, uint32_t words_per_line, int line,
                                          int n_lines, int lo, int hi, int n, uint32_t need, const uint32_t (&jr)[4],
                                          const uint32_t (&jg)[4], const uint32_t (&jb)[4], const uint32_t (&col)[4],
                                          uint32_t keybase, uint32_t (&key)[4]) {
    if ((unsigned)line >= (unsigned)n_lines) return need;
    if (lo < 0) lo = 0;
    if (hi > n - 1) hi = n - 1;
    if (lo > hi) return need;
    const int w0 = lo >> 5, w1 = hi >> 5;
    const uint32_t *cell = planes + ((size_t)line * words_per_line + w0) * CELL;
    for (int w = w0; w <= w1; ++w, cell += CELL) {
        uint32_t mask = 0xffffffffu;
        if (w == w0) mask <<= (lo & 31);
        if (w == w1) mask &= 0xffffffffu >> (31 - (hi & 31));
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (need & (1u << k)) {
                uint32_t m = __ldg(cell + jr[k]) & __ldg(cell + jg[k]) & __ldg(cell + jb[k]) & mask;
                if (m) {
                    m = refine_exact(cell, col[k], m);
                    if (m) {
                        key[k] = keybase + (uint32_t)((w << 5) + __ffs(m) - 1);
                        need &= ~(1u << k);
                    }
                }
            }
        }
        if (!need) break;
    }
    return need;
}

// `v.substract(what).lt(precision)` (helpers.rs:35) on the bytes of one pixel (ring 0 only)
__device__ __forceinline__ bool px_match_rgb(const uint8_t *v, uint32_t col, uint32_t pr, uint32_t pg, uint32_t pb, bool sat) {
    const ColourBox bx = make_box(col & 0xff, (col >> 8) & 0xff, (col >> 16) & 0xff, pr, pg, pb, sat);
    return box_contains(bx, v[0], v[1], v[2]);
}

__device__ __forceinline__ bool key_is_dead_from(uint32_t k) { return k >= KEY_DEAD_FROM && k < KEY_DEAD_FROM + 0x10000u; }

constexpr int ORG_NT = 64;            // threads per CTA: small CTAs keep many independent barrier groups on an SM
constexpr int ORG_NW = ORG_NT / 32;   // warps per CTA
constexpr int ORG_ROUNDS = 256 / ORG_NT;

__global__ void __launch_bounds__(ORG_NT, 1024 / ORG_NT) search_origins_kernel(TileArgs a) {
    __shared__ uint32_t q_col[4][256], q_key[4][256];
    __shared__ uint16_t list_r[256], list_c[256];
    __shared__ uint32_t s_cnt[2][ORG_ROUNDS * ORG_NW];
    __shared__ uint32_t s_max, s_base;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t pair = blockIdx.z;
    const int wa = (int)a.g.wa, ha = (int)a.g.ha;          // plane coverage (cropped to what a query can reach)
    const int src_wa = (int)a.s.wa, src_ha = (int)a.s.ha;  // the additional (sub-)image itself
    const int add_y0 = (int)a.s.add_y0, img_h = (int)a.s.add_img_h;
    const uint32_t *rp = a.planes + (size_t)pair * a.g.pair_words;
    const uint32_t *cp = rp + a.g.rp_words;
    const uint8_t *add = a.s.add_rgb + (size_t)pair * src_ha * src_wa * 3;
    const bool sat = a.s.sat != 0;
    const int ox_base = 16 * (int)blockIdx.x, oy_base = 16 * (int)(a.by0 + blockIdx.y);
    const int W = (int)a.s.w, Hb = (int)a.s.h, y0 = (int)a.s.y0;
    const uint32_t *main_px = reinterpret_cast<const uint32_t *>(a.s.main_rgba) + (size_t)pair * a.s.n_queries;

    // ---- the origin of this thread and its four queries; ring 0 -----------------------------------------------------
    for (int o = tid; o < 256; o += ORG_NT) {
        const int ox = ox_base + (o & 15), oyg = oy_base + (o >> 4), oyl = oyg - add_y0;
        const bool origin_ok = ox < src_wa && oyg < img_h;  // else ring 0 has no in-bounds probe -> None (helpers.rs:49-51)
        const bool have0 = origin_ok && oyl >= 0 && oyl < src_ha;
        const uint8_t *px0 = add + ((size_t)(have0 ? oyl : 0) * src_wa + (have0 ? ox : 0)) * 3;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int X = 2 * ox - 1 + (k & 1), Yl = 2 * oyg - 1 + (k >> 1) - y0;
            const bool valid = X >= 0 && X < W && Yl >= 0 && Yl < Hb;
            uint32_t c = 0, key = KEY_DEAD;
            if (valid) {
                c = main_px[(size_t)Yl * W + X] & 0xffffffu;
                if (origin_ok) key = (have0 && px_match_rgb(px0, c, a.s.pr, a.s.pg, a.s.pb, sat)) ? 0u : KEY_NONE;
            }
            q_col[k][o] = c;
            q_key[k][o] = key;
        }
    }
    if (tid == 0) s_max = 0;
    __syncthreads();

    const int D = (int)a.s.max_rings;
    bool exported = false;
    for (int s0 = 1; s0 < D;) {
        const int s1 = min(D, s0 + max(1, s0 >> 1));
        // ---- compact the origins with an unresolved query into a row-ordered and a column-ordered list ---------------
        bool alive_r[ORG_ROUNDS], alive_c[ORG_ROUNDS];
        uint32_t br[ORG_ROUNDS], bc[ORG_ROUNDS];
#pragma unroll
        for (int r = 0; r < ORG_ROUNDS; ++r) {
            const int o = r * ORG_NT + tid;
            const int oc = ((o & 15) << 4) | (o >> 4);  // column-major enumeration of the 16x16 origins
            alive_r[r] = q_key[0][o] == KEY_NONE || q_key[1][o] == KEY_NONE || q_key[2][o] == KEY_NONE || q_key[3][o] == KEY_NONE;
            alive_c[r] = q_key[0][oc] == KEY_NONE || q_key[1][oc] == KEY_NONE || q_key[2][oc] == KEY_NONE || q_key[3][oc] == KEY_NONE;
            br[r] = __ballot_sync(0xffffffffu, alive_r[r]);
            bc[r] = __ballot_sync(0xffffffffu, alive_c[r]);
            if (lane == 0) { s_cnt[0][r * ORG_NW + warp] = __popc(br[r]); s_cnt[1][r * ORG_NW + warp] = __popc(bc[r]); }
        }
        __syncthreads();
        uint32_t n_alive = 0;
        {
            uint32_t run_r = 0, run_c = 0;
            const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
            for (int r = 0; r < ORG_ROUNDS; ++r) {
#pragma unroll
                for (int w = 0; w < ORG_NW; ++w) {
                    if (w == warp) {
                        const int o = r * ORG_NT + tid;
                        if (alive_r[r]) list_r[run_r + __popc(br[r] & lt)] = (uint16_t)o;
                        if (alive_c[r]) list_c[run_c + __popc(bc[r] & lt)] = (uint16_t)(((o & 15) << 4) | (o >> 4));
                    }
                    run_r += s_cnt[0][r * ORG_NW + w];
                    run_c += s_cnt[1][r * ORG_NW + w];
                }
            }
            n_alive = run_r;
        }
        if (n_alive == 0) break;  // uniform
        __syncthreads();
        // ---- sparse tile: hand the unresolved origins to the warp-per-origin kernel (search_tail.cuh) and retire -----
        if (n_alive < a.export_below) {  // uniform
            if (tid == 0) s_base = atomicAdd(a.wl_count, n_alive);
            __syncthreads();
            for (uint32_t e = tid; e < n_alive; e += ORG_NT) {
                const int o = list_r[e];
                uint32_t need = 0;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (q_key[k][o] == KEY_NONE) need |= 1u << k;
                a.wl[s_base + e] = make_uint2((uint32_t)(ox_base + (o & 15)) | ((uint32_t)(oy_base + (o >> 4)) << 16),
                                              pair | ((uint32_t)s0 << 16) | (need << 28));
            }
            exported = true;
            break;
        }

        // ---- row pass: rows y+s then y-s of rings s0 .. s1-1 ----------------------------------------------------------
        for (uint32_t e = tid; e < n_alive; e += ORG_NT) {
            const int o = list_r[e];
            const int qx = ox_base + (o & 15), qyg = oy_base + (o >> 4), qy = qyg - add_y0;
            uint32_t jr[4], jg[4], jb[4], col[4], key[4], need = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t c = q_col[k][o];
                key[k] = q_key[k][o];
                if (key[k] == KEY_NONE) need |= 1u << k;
                jr[k] = w_u(c & 0xff);
                jg[k] = CH_WORDS + w_u((c >> 8) & 0xff);
                jb[k] = 2 * CH_WORDS + w_u((c >> 16) & 0xff);
                col[k] = c;
            }
            const uint32_t need0 = need;
            for (int s = s0; s < s1 && need; ++s) {
                // no in-bounds probe from ring s on (helpers.rs:49-51): the column pass may still own rings < s
                if (qx - s < 0 && qx + s >= src_wa && qyg - s < 0 && qyg + s >= img_h) {
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        if (need & (1u << k)) key[k] = KEY_DEAD_FROM | (uint32_t)s;
                    need = 0;
                    break;
                }
                const uint32_t kb = ((uint32_t)s << 12) + (uint32_t)(s - qx);  // key of (dx, +s) = s<<12 | dx + s
                need = scan4(rp, a.g.nxw, qy + s, ha, qx - s, qx + s, wa, need, jr, jg, jb, col, kb, key);
                if (!need) break;
                need = scan4(rp, a.g.nxw, qy - s, ha, qx - s, qx + s, wa, need, jr, jg, jb, col, kb + (uint32_t)(2 * s + 1), key);
            }
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if ((need0 & (1u << k)) && key[k] != KEY_NONE) q_key[k][o] = key[k];
        }
        __syncthreads();
        // ---- column pass: columns x+s then x-s, per query only the rings below its row hit of this chunk --------------
        for (uint32_t e = tid; e < n_alive; e += ORG_NT) {
            const int o = list_c[e];
            const int qx = ox_base + (o & 15), qyg = oy_base + (o >> 4), qy = qyg - add_y0;
            uint32_t jr[4], jg[4], jb[4], col[4], key[4], krow[4], fallback[4];
            int lim[4], smax = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t c = q_col[k][o];
                const uint32_t kr = q_key[k][o];
                krow[k] = kr;
                fallback[k] = kr;
                if (kr == KEY_NONE) lim[k] = s1;
                else if (key_is_dead_from(kr)) { lim[k] = (int)(kr & 0xffffu); fallback[k] = KEY_DEAD; }
                else if (kr == KEY_DEAD) lim[k] = 0;
                else lim[k] = (int)(kr >> 12);  // found: this chunk's row hit (ring >= s0) or an older result (ring < s0)
                smax = max(smax, lim[k]);
                key[k] = KEY_NONE;
                jr[k] = w_u(c & 0xff);
                jg[k] = CH_WORDS + w_u((c >> 8) & 0xff);
                jb[k] = 2 * CH_WORDS + w_u((c >> 16) & 0xff);
                col[k] = c;
            }
            uint32_t found = 0;
            for (int s = s0; s < smax; ++s) {
                uint32_t need = 0;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (s < lim[k]) need |= 1u << k;
                need &= ~found;
                if (!need) break;
                const uint32_t kb = ((uint32_t)s << 12) + (uint32_t)(5 * s + 1 - qy);  // key of (+s, dy) = s<<12 | 2(2s+1) + dy+s-1
                uint32_t rest = scan4(cp, a.g.nyw, qx + s, wa, qy - s + 1, qy + s - 1, ha, need, jr, jg, jb, col, kb, key);
                found |= need & ~rest;
                if (rest) {
                    const uint32_t rest2 = scan4(cp, a.g.nyw, qx - s, wa, qy - s + 1, qy + s - 1, ha, rest, jr, jg, jb, col,
                                                 kb + (uint32_t)(2 * s - 1), key);
                    found |= rest & ~rest2;
                }
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t nk = (found & (1u << k)) ? key[k] : fallback[k];
                if (nk != krow[k]) q_key[k][o] = nk;
            }
        }
        __syncthreads();
        s0 = s1;
    }

    // ---- `d as u8` per query, per-pair maximum (depth_image.rs:120-124) ---------------------------------------------
    uint32_t dmax = 0;
    for (int o = tid; o < 256; o += ORG_NT) {
        const int ox = ox_base + (o & 15), oyg = oy_base + (o >> 4);
        uint8_t *d8 = a.s.d8 + (size_t)pair * a.s.n_queries;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int X = 2 * ox - 1 + (k & 1), Yl = 2 * oyg - 1 + (k >> 1) - y0;
            if (X >= 0 && X < W && Yl >= 0 && Yl < Hb) {
                const uint32_t key = q_key[k][o];
                if (exported && key == KEY_NONE) continue;  // the tail kernel writes this one
                const uint32_t d = key >= KEY_DEAD_FROM ? 0u : (uint32_t)key_to_d8(key);
                d8[(size_t)Yl * W + X] = (uint8_t)d;
                dmax = max(dmax, d);
            }
        }
    }
    const uint32_t m = __reduce_max_sync(0xffffffffu, dmax);
    if (lane == 0 && m) atomicMax(&s_max, m);
    __syncthreads();
    if (tid == 0 && s_max) atomicMax(a.s.max_out + pair, s_max);
}

}  // namespace stepth
