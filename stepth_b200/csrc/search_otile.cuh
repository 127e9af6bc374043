// search_otile.cuh -- K1a (dense mode): ring 0 and the first coarse rings, ONE WARP PER ORIGIN TILE.
//
// Origins are half-coordinates (depth_image.rs:114-117): main pixel (x, y) starts its search at ((x+1)/2, (y+1)/2), so the
// 16 x 16 main pixels x in [16 TX - 1, 16 TX + 14], y likewise, all start inside tile (TX, TY) of the additional image.  Their
// walks over coarse rings (search_hier.cuh) therefore read the SAME occupancy cells and test the SAME tiles -- only the
// colour, the position of the origin inside the tile and the best key so far differ.  A warp that owns the 256 queries of one
// origin tile keeps that coherence for the whole walk, which the walking near kernel (queries of four patches packed into
// warps once, then walked to the last ring with whatever lanes are still open) gives up:
//
//   ring 0    one pixel compare per query (helpers.rs:26-47 with step = 0), 32 queries at a time; the survivors -- about
//             half of a natural pair -- are packed into a list in shared memory: {colour | local index, best key};
//   pass S    (S = 0, 1, 2, ..)  the list is read 32 entries at a time.  An entry whose best key cannot be beaten any more
//             (ring <= 8 (S-1) + slack, search_hier.cuh: stop_slack) is published (`d as u8`, depth_image.rs:120-121) and
//             leaves; the others are packed again IN PLACE -- every pass runs on full warps however few queries are left --
//             read the occupancy windows of coarse ring S (for S = 0 the three row windows that also hold ring 1) and list
//             their flagged tiles as tasks {entry, tile};
//   tasks     the warp tests the listed tiles 32 at a time on the fine records (fine_test: 15 plane words, exact) and lowers
//             the entry's best key with a shared-memory atomicMin: the arg-min does not care who tests which tile, or when;
//   the end   what is still open after coarse ring S1 - 1 goes to the work list of the far kernel (one atomic per warp).
//
// No barrier, no shuffle of query state, no per-lane bookkeeping of "still walking": the list is the state.
#pragma once
#include "search_hier.cuh"

namespace stepth {

#if defined(__CUDACC__)

#ifndef STEPTH_OT_MINB
#define STEPTH_OT_MINB 18   // resident CTAs per SM the register allocation must allow: 18 x 64 threads -> 56 registers (64: 2.16 ms, 56: 2.10, 48: 2.12, 40: 2.16)
#endif
#ifndef STEPTH_OT_WARPS
#define STEPTH_OT_WARPS 2   // warps per CTA; measured (16 pairs 1080p, search stage): 2 warps / 48 registers 2.13 ms, 4 / 48 2.15, 4 / 64 2.21, 8 / 48 2.45
#endif
constexpr int OT_WARPS = STEPTH_OT_WARPS;   // warps (origin tiles, or half tiles) per CTA, side by side
#ifndef STEPTH_OT_TASKS
#define STEPTH_OT_TASKS 256   // measured (16 pairs 1080p, search stage): 128 2.18 ms, 192 2.11, 256 2.10, 384 2.11, 512 2.10
#endif
constexpr int OT_TASKS = STEPTH_OT_TASKS;    // task slots of a warp
constexpr int OT_MAX_S1 = 16;    // rings <= 15 have sides of at most 31 tiles: 32-bit flags

struct OtWarp {
    uint32_t ent[256];        // colour | local query index << 24:  x = 16 TX - 1 + (lq & 15),  global row = 2 (8 TY + add_y0) - 1 + (lq >> 4)
    uint32_t best[256];       // best key so far
    uint32_t task[OT_TASKS];  // entry | (dx + 16) << 8 | (dy + 16) << 13:  tile (TX + dx, TY + dy)
    uint8_t fl[256];          // flags of coarse ring 1, kept from the three row windows read in pass 0
    uint32_t n_task;
};

// the warp tests the listed tiles, 32 at a time
__device__ __forceinline__ void ot_run_tasks(OtWarp *ws, const HierView &v, int T0x, int T0y, int D, const unsigned long long *tab_le,
                                             const unsigned long long *tab_ge, uint32_t lane, uint32_t *dbg = nullptr, int n_list = 256) {
    __syncwarp();
    const int nt = (int)min(ws->n_task, (uint32_t)OT_TASKS);
    for (int t0 = 0; t0 < nt; t0 += 32) {
        if (t0 + (int)lane < nt) {
            const uint32_t task = ws->task[t0 + lane];
            const uint32_t e = task & 0xffu;
            const int dx = (int)((task >> 8) & 31u) - 16, dy = (int)((task >> 13) & 31u) - 16;
#ifdef STEPTH_DEBUG_CHECKS
            if (dbg && ((int)e >= n_list || (unsigned)(T0x + dx) >= (unsigned)v.ntx || (unsigned)(T0y + dy) >= (unsigned)v.nty || dx < -15 || dx > 15 ||
                        dy < -15 || dy > 15))
                atomicOr(dbg, 1u << 0);
#endif
            const uint32_t ent = ws->ent[e];
            const uint32_t lq = ent >> 24;
            const HierQuery q = make_query(8 * T0x + (int)((lq & 15u) >> 1), 8 * T0y + (int)(lq >> 5), ent & 0xffffffu);
            uint32_t kb = ws->best[e];   // may have been lowered by another lane meanwhile: that only costs a test
            const uint32_t before = kb;
            fine_test(v, q, T0x + dx, T0y + dy, D, tab_le, tab_ge, kb);
            if (kb < before) atomicMin(&ws->best[e], kb);
        }
    }
    __syncwarp();
    if (lane == 0) ws->n_task = 0;
    __syncwarp();
}

// lists the flagged tiles of one side from slot `slot` on; the bits that do not fit below OT_TASKS come back
__device__ __forceinline__ uint32_t ot_emit_side(uint32_t m, uint32_t code, int sh, uint32_t &slot, OtWarp *ws) {
    while (m && slot < (uint32_t)OT_TASKS) {
        ws->task[slot++] = code + ((uint32_t)(__ffs((int)m) - 1) << sh);
        m &= m - 1;
    }
    return m;
}

// One origin tile (NIT = 8) or NIT of its eight origin rows, start to finish, by one warp.
// NIT = origin rows per warp: 8 = a whole origin tile (256 queries); 4, 2, 1 = a half, a quarter, an eighth (small launches)
template <int NIT>
__device__ __forceinline__ void ot_tile(const HierArgs &a, const uint32_t pair, const int T0x, const int T0y, const int it0, OtWarp *ws,
                                        const unsigned long long *s_le, const unsigned long long *s_ge, const uint32_t lane) {
    const SearchArgs &s = a.s;
    if (T0x > (int)((s.w >> 1) >> 3)) return;   // beyond the last origin tile column (warp-uniform)
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t *main32 = reinterpret_cast<const uint32_t *>(s.main_rgba) + (size_t)pair * s.n_queries;
    uint8_t *d8 = s.d8 + (size_t)pair * s.n_queries;
    const int row_base = 2 * (8 * T0y + (int)s.add_y0) - 1 - (int)s.y0;   // band row of local row 0 of the tile's 16 x 16 main pixels
    const int x_base = 16 * T0x - 1;

    // ---- ring 0: one pixel compare per query; the survivors are listed ------------------------------------------------------
    int n = 0;
    {
        const uint8_t *addp = s.add_rgb + (size_t)pair * s.ha * s.wa * 3;
        const int x = x_base + (int)(lane & 15u), ox = 8 * T0x + (int)((lane & 15u) >> 1);
#pragma unroll 2
        for (int it = it0; it < it0 + NIT; ++it) {
            const int y = row_base + 2 * it + (int)(lane >> 4);
            const bool valid = x >= 0 && x < (int)s.w && y >= 0 && y < (int)s.h;
            bool open = false;
            uint32_t col = 0;
            if (valid) {
                const uint32_t qi = (uint32_t)y * s.w + (uint32_t)x;
                col = main32[qi] & 0xffffffu;
                const int oyl = 8 * T0y + it, oyg = oyl + (int)s.add_y0;   // = (global row + 1) / 2
                // ring 0 out of bounds => `still_in_bounds` stays false => break => None   (helpers.rs:49-51)
                if (ox < (int)s.wa && oyg < (int)s.add_img_h) {
                    bool hit0 = false;
                    if (oyl >= 0 && oyl < (int)s.ha) {
                        const uint8_t *px = addp + ((size_t)oyl * s.wa + ox) * 3;
                        hit0 = pixel_in_box(px[0], px[1], px[2], col, s.pr, s.pg, s.pb, s.sat != 0);
                    }
                    open = !hit0 && s.max_rings > 1;
                }
                // settled at ring 0: a hit has distance 0 and `None` maps to 0 too (depth_image.rs:120), neither moves the maximum
                if (!open) d8[qi] = 0;
            }
            const uint32_t bal = __ballot_sync(0xffffffffu, open);
            if (open) {
                const int pos = n + __popc(bal & lt_mask);
                STEPTH_CHECK(a, pos >= 0 && pos < 256, 1);
                ws->ent[pos] = col | ((uint32_t)(it * 32) + lane) << 24;
                ws->best[pos] = KEY_NONE;
            }
            n += __popc(bal);
        }
    }
    if (n == 0) return;   // warp-uniform

    // ---- the walk --------------------------------------------------------------------------------------------------------------
    const HierView v = make_view(a.planes, a.g, pair);
    const int D = (int)s.max_rings;
    const int Smax = (D + 6) >> 3;                           // tiles of coarse ring S start at ring >= 8S - 7
    const int S_end = a.S1 < Smax + 1 ? a.S1 : Smax + 1;     // rings [0, S_end) are walked here; the host keeps S1 in [2, OT_MAX_S1]
    const int xmax = v.ntx - 1, ymax = v.nty - 1;
    uint32_t mx = 0;                                         // largest distance this lane published
#pragma unroll 1
    for (int S = 0; S <= S_end; ++S) {
        // pass S: settles what ring S - 1 finished, packs the rest, walks ring S (when S < S_end)
        const bool walk = S < S_end;
        // the coarse square has left the tile grid on all four sides (nothing further out exists), or every ring below
        // max_rings has been walked: whatever the entries hold is final
        const bool all_final = (T0x - S < 0 && T0x + S > xmax && T0y - S < 0 && T0y + S > ymax) || (!walk && a.S1 > Smax);
        const int rlo = T0x - S < 0 ? 0 : T0x - S, rhi = T0x + S > xmax ? xmax : T0x + S;                  // row sides (S >= 2)
        const int clo = T0y - S + 1 < 0 ? 0 : T0y - S + 1, chi = T0y + S - 1 > ymax ? ymax : T0y + S - 1;  // column sides
        int n_new = 0;
#pragma unroll 1
        for (int b0 = 0; b0 < n; b0 += 32) {
            const int e = b0 + (int)lane;
            const bool has = e < n;
            uint32_t ent = 0, best = KEY_NONE, fl = 0;
            if (has) {
                ent = ws->ent[e];
                best = ws->best[e];
                if (S == 1) fl = ws->fl[e];
            }
            const uint32_t lq = ent >> 24;
            const int lx = (int)((lq & 15u) >> 1), ly = (int)(lq >> 5);
            const HierQuery q = make_query(8 * T0x + lx, 8 * T0y + ly, ent & 0xffffffu);
            // every pixel outside the (2S-1)^2 tile block lies at ring >= 8 (S-1) + slack + 1
            const bool fin = has && (all_final || (S > 0 && best != KEY_NONE && (int)(best >> 12) <= 8 * (S - 1) + stop_slack(q)));
            if (fin) {
                STEPTH_CHECK(a, row_base + (int)(lq >> 4) >= 0 && row_base + (int)(lq >> 4) < (int)s.h && x_base + (int)(lq & 15u) >= 0 &&
                                    x_base + (int)(lq & 15u) < (int)s.w, 3);
                const uint32_t d = best == KEY_NONE ? 0u : (uint32_t)key_to_d8(best);
                d8[(uint32_t)(row_base + (int)(lq >> 4)) * s.w + (uint32_t)(x_base + (int)(lq & 15u))] = (uint8_t)d;
                mx = max(mx, d);
            }
            const bool keep = has && !fin;
            const uint32_t bal = __ballot_sync(0xffffffffu, keep);
            const int pos = n_new + __popc(bal & lt_mask);
            __syncwarp();   // every lane has read its entry of this batch: the packed ones may overwrite them
            if (keep) {
                STEPTH_CHECK(a, pos >= 0 && pos <= e && e < 256, 2);   // packing in place never overtakes the read position
                ws->ent[pos] = ent;
                ws->best[pos] = best;
            }
            n_new += __popc(bal);
            if (!walk) continue;   // warp-uniform
            uint32_t f0 = 0, f1 = 0, f2 = 0, f3 = 0;   // flagged tiles of rows T0y+S, T0y-S (from rlo), columns T0x+S, T0x-S (from clo)
            int r_lo = rlo, c_lo = clo;
            if (keep) {
                if (S == 0) {
                    // three row windows give the nine flags of coarse rings 0 and 1 (search_hier.cuh: Block3)
                    const Block3 b = block3_flags(v, q, T0x, T0y);
                    f2 = block3_mid(b, T0x);   // the origin's own tile: "column T0x + 0, tile T0y"
                    c_lo = T0y;
                    const int sh = T0x - b.rlo;   // 0 or 1: bit of tile T0x - 1 + i in a row's flags is i - 1 + sh
                    const uint32_t fa = sh ? b.fa : b.fa << 1, fc = sh ? b.fc : b.fc << 1;   // bit i <-> tile T0x - 1 + i
                    ws->fl[pos] = (uint8_t)((fa & 7u) | (fc & 7u) << 3 | block3_mid(b, T0x + 1) << 6 | block3_mid(b, T0x - 1) << 7);
                } else if (S == 1) {
                    r_lo = T0x - 1;   // may be -1: bit 0 is clear then
                    c_lo = T0y;
                    f1 = fl & 7u;
                    f0 = (fl >> 3) & 7u;
                    f2 = (fl >> 6) & 1u;
                    f3 = (fl >> 7) & 1u;
                } else {
                    f0 = side_flags32(v.crp, v.nxw, T0y + S, v.nty, rlo, rhi, q);
                    f1 = side_flags32(v.crp, v.nxw, T0y - S, v.nty, rlo, rhi, q);
                    f2 = side_flags32(v.ccp, v.nyw, T0x + S, v.ntx, clo, chi, q);
                    f3 = side_flags32(v.ccp, v.nyw, T0x - S, v.ntx, clo, chi, q);
                }
                if (S > 0 && (f0 | f1 | f2 | f3)) {
                    // with a match in hand (or at the last rings) most flagged tiles are out of reach and are not listed
                    const int limit = min(D - 1, (int)(best >> 12));
                    if (limit < 8 * S + 7) prune_sides(q, S, limit, r_lo, c_lo, f0, f1, f2, f3);
                }
            }
            // list the tiles; lanes whose tasks do not fit keep their flags and list them after the warp has tested a full list
            for (;;) {
                const uint32_t cnt = __popc(f0) + __popc(f1) + __popc(f2) + __popc(f3);
                if (cnt) {
                    uint32_t slot = atomicAdd(&ws->n_task, cnt);
                    const uint32_t code = (uint32_t)pos | 16u << 8 | 16u << 13;
                    f0 = ot_emit_side(f0, code + ((uint32_t)(r_lo - T0x) << 8) + ((uint32_t)S << 13), 8, slot, ws);
                    f1 = ot_emit_side(f1, code + ((uint32_t)(r_lo - T0x) << 8) - ((uint32_t)S << 13), 8, slot, ws);
                    f2 = ot_emit_side(f2, code + ((uint32_t)S << 8) + ((uint32_t)(c_lo - T0y) << 13), 13, slot, ws);
                    f3 = ot_emit_side(f3, code - ((uint32_t)S << 8) + ((uint32_t)(c_lo - T0y) << 13), 13, slot, ws);
                }
                __syncwarp();
                const bool more = __any_sync(0xffffffffu, (f0 | f1 | f2 | f3) != 0u);
                if (more || ws->n_task > (uint32_t)(OT_TASKS - 96)) ot_run_tasks(ws, v, T0x, T0y, D, s_le, s_ge, lane, a.dbg_err, n_new > n ? n_new : n);
                if (!more) break;
            }
        }
        n = n_new;
        if (n == 0) break;
        if (walk) ot_run_tasks(ws, v, T0x, T0y, D, s_le, s_ge, lane, a.dbg_err, n_new > n ? n_new : n);
    }

    // ---- what is still open goes to the far kernel ------------------------------------------------------------------------------
    if (n > 0) {
        __syncwarp();
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(a.wl_out_count, (uint32_t)n);
        base = __shfl_sync(0xffffffffu, base, 0);
        for (int e = (int)lane; e < n; e += 32) {
            const uint32_t ent = ws->ent[e], best = ws->best[e], lq = ent >> 24;
            const uint32_t qi = (uint32_t)(row_base + (int)(lq >> 4)) * s.w + (uint32_t)(x_base + (int)(lq & 15u));
            if (base + (uint32_t)e < a.wl_cap)
                a.wl_out[base + e] = make_uint2(qi, (pair << 21) | (best == KEY_NONE ? KEY21_NONE : best));
        }
    }
    const uint32_t m = __reduce_max_sync(0xffffffffu, mx);
    if (lane == 0 && m && m > *(volatile uint32_t *)(s.max_out + pair)) atomicMax(s.max_out + pair, m);
}

#ifndef STEPTH_OT_SPAN
#define STEPTH_OT_SPAN 0   // > 0: a CTA owns this many consecutive origin tiles of a tile row and its warps pull them one by one
#endif

template <int NIT>
__global__ void __launch_bounds__(OT_WARPS * 32, STEPTH_OT_MINB) search_otile_kernel(HierArgs a, int ty_first) {
    constexpr int HALVES = 8 / NIT;
    __shared__ unsigned long long s_le[DIAG_N], s_ge[DIAG_N];
    __shared__ OtWarp s_w[OT_WARPS];
#if STEPTH_OT_SPAN > 0
    __shared__ int s_next;
    if (threadIdx.x == 0) s_next = 0;
#endif
    load_diag_tables(s_le, s_ge);  // ends with __syncthreads()
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const int T0y = ty_first + (int)(blockIdx.y / HALVES);
    const int it0 = (int)(blockIdx.y % HALVES) * NIT;
    OtWarp *ws = &s_w[warp];
    if (lane == 0) ws->n_task = 0;
#if STEPTH_OT_SPAN > 0
    // neighbouring tiles read the same occupancy cells and tile records: the warps of a CTA stay next to each other (L1) and
    // none of them idles while a slower one finishes its tile, as it would with one tile per warp in a wide CTA
    for (;;) {
        int t = 0;
        if (lane == 0) t = atomicAdd(&s_next, 1);
        t = __shfl_sync(0xffffffffu, t, 0);
        if (t >= STEPTH_OT_SPAN) break;
        ot_tile<NIT>(a, blockIdx.z, (int)blockIdx.x * STEPTH_OT_SPAN + t, T0y, it0, ws, s_le, s_ge, lane);
        __syncwarp();
    }
#else
    ot_tile<NIT>(a, blockIdx.z, (int)blockIdx.x * OT_WARPS + (int)warp, T0y, it0, ws, s_le, s_ge, lane);
#endif
}

#endif  // __CUDACC__

}  // namespace stepth
