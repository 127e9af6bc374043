// search_tail.cuh -- K1b: warp-per-origin ring search for the sparse tail.
//
// search_origins.cuh (thread-per-origin) is efficient while a 16x16-origin tile still has many unresolved origins.  The
// cost of a query grows with the square of its first-match ring, so most of the work of a natural pair sits in the few
// percent of queries that go far -- and those are scattered: a tile with a handful of unresolved origins would keep
// 2 warps busy with 3 active lanes.  Such tiles EXPORT their unresolved origins (origin, resume ring, open queries) to a
// global work list; this kernel gives every exported origin a whole warp:
//
//   * the 32 lanes take the plane words of the next ring(s) in the reference's probe order -- item = (ring, side, word)
//     with side = row y+s, row y-s, column x+s, column x-s (helpers.rs:26-31).  Small rings have few words, so one step
//     covers several rings; a ring above ~112 takes more than one step;
//   * every lane tests its 32 probes for each open query of the origin (candidate word = three plane loads + AND, exact
//     word = planes.cuh refine_exact, first set bit), then one ballot per query finds the lowest item with a hit = the first match in
//     reference order.  All lanes read the plane cells of ONE origin, i.e. a few 128-byte lines per step;
//   * warps pull origins from the list with an atomic cursor, so the load balance is dynamic and there is no barrier.
#pragma once
#include "search_origins.cuh"

namespace stepth {

struct TailArgs {
    TileArgs t;
    const uint2 *list;      // exported origins: x = ox | oyg << 16, y = pair | s_resume << 16 | need << 28
    const uint32_t *count;  // number of entries (written by the head kernel)
    uint32_t *cursor;       // atomic work cursor (zeroed before the launch)
};

__global__ void __launch_bounds__(128) search_tail_kernel(TailArgs a) {
    const int lane = threadIdx.x & 31;
    const TileArgs &t = a.t;
    const int wa = (int)t.g.wa, ha = (int)t.g.ha, src_wa = (int)t.s.wa;
    const int add_y0 = (int)t.s.add_y0, img_h = (int)t.s.add_img_h;
    const int W = (int)t.s.w, y0 = (int)t.s.y0, D = (int)t.s.max_rings;
    const uint32_t n_entries = *a.count;

    for (;;) {
        uint32_t e = 0;
        if (lane == 0) e = atomicAdd(a.cursor, 1u);
        e = __shfl_sync(0xffffffffu, e, 0);
        if (e >= n_entries) break;
        const uint2 ent = a.list[e];
        const int qx = (int)(ent.x & 0xffffu), qyg = (int)(ent.x >> 16), qy = qyg - add_y0;
        const uint32_t pair = ent.y & 0xffffu;
        const int s_resume = (int)((ent.y >> 16) & 0xfffu);
        uint32_t need = ent.y >> 28;
        const uint32_t *rp = t.planes + (size_t)pair * t.g.pair_words;
        const uint32_t *cp = rp + t.g.rp_words;
        const uint32_t *main_px = reinterpret_cast<const uint32_t *>(t.s.main_rgba) + (size_t)pair * t.s.n_queries;

        // the four queries of the origin: main pixels (2qx-1 .. 2qx, 2qyg-1 .. 2qyg)
        uint32_t jr[4], jg[4], jb[4], col[4], key[4];
        {
            uint32_t c = 0;
            if (lane < 4 && (need & (1u << lane))) {
                const int X = 2 * qx - 1 + (lane & 1), Yl = 2 * qyg - 1 + (lane >> 1) - y0;
                c = main_px[(size_t)Yl * W + X] & 0xffffffu;
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t ck = __shfl_sync(0xffffffffu, c, k);
                jr[k] = w_u(ck & 0xff);
                jg[k] = CH_WORDS + w_u((ck >> 8) & 0xff);
                jb[k] = 2 * CH_WORDS + w_u((ck >> 16) & 0xff);
                col[k] = ck;
                key[k] = KEY_NONE;
            }
        }

        int s = s_resume;
        while (s < D && need) {
            // no in-bounds probe from ring s on -> None for what is still open (helpers.rs:49-51)
            if (qx - s < 0 && qx + s >= src_wa && qyg - s < 0 && qyg + s >= img_h) break;
            // slots per side: an upper bound on the words a (2s+1)-pixel segment can touch; R rings per step
            int S = ((2 * s + 31) >> 5) + 1;
            int R = max(1, min(32 / (4 * S), D - s));
            const int S_last = ((2 * (s + R - 1) + 31) >> 5) + 1;  // the last ring of the step may touch one word more
            if (S_last > S) { S = S_last; R = max(1, min(32 / (4 * S), D - s)); }
            const int items = R * 4 * S;
            for (int base = 0; base < items && need; base += 32) {
                const int i = base + lane;
                uint32_t hit[4] = {KEY_NONE, KEY_NONE, KEY_NONE, KEY_NONE};
                if (i < items) {
                    const int r = i / (4 * S), rem = i - r * 4 * S, side = rem / S, wi = rem - side * S;
                    const int sr = s + r;
                    const bool is_col = side >= 2, neg = (side & 1) != 0;
                    const int line = is_col ? (neg ? qx - sr : qx + sr) : (neg ? qy - sr : qy + sr);
                    const int n_lines = is_col ? wa : ha, n = is_col ? ha : wa;
                    int lo = is_col ? qy - sr + 1 : qx - sr, hi = is_col ? qy + sr - 1 : qx + sr;
                    lo = max(lo, 0);
                    hi = min(hi, n - 1);
                    const int w = (lo >> 5) + wi;
                    if ((unsigned)line < (unsigned)n_lines && lo <= hi && w <= (hi >> 5)) {
                        uint32_t mask = 0xffffffffu;
                        if (w == (lo >> 5)) mask <<= (lo & 31);
                        if (w == (hi >> 5)) mask &= 0xffffffffu >> (31 - (hi & 31));
                        const uint32_t *cell = (is_col ? cp : rp) + ((size_t)line * (is_col ? t.g.nyw : t.g.nxw) + w) * CELL;
                        // key of position p on this side (common.cuh ring_key)
                        const uint32_t s12 = (uint32_t)sr << 12;
                        const uint32_t kb = is_col ? s12 + (uint32_t)(5 * sr + 1 - qy) + (neg ? (uint32_t)(2 * sr - 1) : 0u)
                                                : s12 + (uint32_t)(sr - qx) + (neg ? (uint32_t)(2 * sr + 1) : 0u);
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            if (need & (1u << k)) {
                                uint32_t m = __ldg(cell + jr[k]) & __ldg(cell + jg[k]) & __ldg(cell + jb[k]) & mask;
                                if (m) m = refine_exact(cell, col[k], m);
                                if (m) hit[k] = kb + (uint32_t)((w << 5) + __ffs(m) - 1);
                            }
                        }
                    }
                }
                // lowest item with a hit = first match in probe order (items ascend with ring, side, word)
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (need & (1u << k)) {
                        const uint32_t bal = __ballot_sync(0xffffffffu, hit[k] != KEY_NONE);
                        if (bal) {
                            key[k] = __shfl_sync(0xffffffffu, hit[k], __ffs(bal) - 1);
                            need &= ~(1u << k);
                        }
                    }
                }
            }
            s += R;
        }

        // `d as u8` of the exported queries (None -> 0) and the per-pair maximum
        if (lane < 4 && ((ent.y >> 28) & (1u << lane))) {
            uint32_t kk = KEY_NONE;
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (k == lane) kk = key[k];
            const uint32_t d = kk == KEY_NONE ? 0u : (uint32_t)key_to_d8(kk);
            const int X = 2 * qx - 1 + (lane & 1), Yl = 2 * qyg - 1 + (lane >> 1) - y0;
            t.s.d8[(size_t)pair * t.s.n_queries + (size_t)Yl * W + X] = (uint8_t)d;
            if (d && d > *(volatile uint32_t *)(t.s.max_out + pair)) atomicMax(t.s.max_out + pair, d);
        }
    }
}

inline void launch_search_tail(stepth_ctx *ctx, cudaStream_t st, const TileArgs &t) {
    TailArgs a;
    a.t = t;
    a.list = t.wl;
    a.count = t.wl_count;
    a.cursor = t.wl_count + 1;
    search_tail_kernel<<<ctx->sm_count * 8, 128, 0, st>>>(a);
}

}  // namespace stepth
