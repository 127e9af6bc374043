// search_tile.cuh -- K1a (dense mode): ring 0 and coarse rings 0..1 for ALL the queries whose origin lies in two neighbouring
// 8x8 tiles of the additional image, with the tile records they need staged in shared memory by the TMA unit.
//
// Origins are half-coordinates (depth_image.rs:114-117): the origin of main pixel (x, y) is ((x+1)/2, (y+1)/2), so the 256
// main pixels x in [16 TX - 1, 16 TX + 14], y likewise, all start their search inside tile (TX, TY) of the additional image
// and -- for the first two coarse rings, where three out of four walks of a natural pair end -- look at the same 3 x 3 tile
// block.  A CTA takes two such tiles side by side: the 3 x 4 fine records around them (3 bulk copies of up to 6 KB, one per
// tile row, completion counted on an mbarrier: the records of a tile row are contiguous in memory) arrive while the CTA runs
// ring 0 (one pixel compare per query, helpers.rs:26-47 with step = 0) for its 512 queries and packs the survivors -- about
// half -- into full warps.  A survivor then settles coarse rings 0 and 1 from shared memory alone: the candidate word of a
// tile (3 loads, 2 ANDs: 64 probes) says whether the exact word (12 more loads) is worth building; tiles out of reach of the
// best ring so far are skipped; the first match inside a tile in reference order comes from tile_min_key.  No global load,
// no 64-bit address arithmetic, no occupancy windows in this stage.  What is still open after ring 1 goes to the work list of
// the walk kernels (search_hier.cuh), which resume at coarse ring 2.
#pragma once
#include "search_hier.cuh"

namespace stepth {

#if defined(__CUDACC__)

#ifndef STEPTH_TILE_MINB
#define STEPTH_TILE_MINB 4
#endif

constexpr int ST_COLS = 4;   // tile columns staged per CTA: TXa - 1 .. TXa + 2 for the two origin tiles TXa, TXa + 1

// the exact match word of a staged tile for a query (tiles.cuh): u_* = slot of U[j] inside each channel's 64 slots, r_* = q & 7
__device__ __forceinline__ unsigned long long staged_exact(const unsigned long long *rec, uint32_t ur, uint32_t ug, uint32_t ub, uint32_t rr,
                                                           uint32_t rg, uint32_t rb) {
    const unsigned long long *c0 = rec, *c1 = rec + T_SLOTS, *c2 = rec + 2 * T_SLOTS;
    unsigned long long m = c0[ur] & (c0[ur + 1] | c0[W_NTA + rr]) & (c0[ur - 1] | c0[W_NTB + rr]);
    m &= c1[ug] & (c1[ug + 1] | c1[W_NTA + rg]) & (c1[ug - 1] | c1[W_NTB + rg]);
    m &= c2[ub] & (c2[ub + 1] | c2[W_NTA + rb]) & (c2[ub - 1] | c2[W_NTB + rb]);
    return m;
}

__global__ void __launch_bounds__(256, STEPTH_TILE_MINB) search_tile_kernel(HierArgs a, int ty_first) {
    __shared__ __align__(128) unsigned long long s_rec[3][ST_COLS][T_REC];   // 3 x 4 tile records: 18 KB
    __shared__ __align__(8) unsigned long long s_bar;
    __shared__ unsigned long long s_le[DIAG_N], s_ge[DIAG_N];
    __shared__ uint2 s_open[512];   // survivors of ring 0: {k << 8 | thread, colour}
    __shared__ uint32_t s_n;
    const SearchArgs &s = a.s;
    const TileGeom &g = a.g;
    const uint32_t pair = blockIdx.z;
    const int tid = threadIdx.x, lane = tid & 31;
    const int TXa = (int)blockIdx.x * 2, TY = ty_first + (int)blockIdx.y;
    const int ntx = (int)g.ntx, nty = (int)g.nty;

    // ---- stage the 3 x 4 records: one bulk copy per tile row that exists; what does not exist reads as "no match" ----------
    const int clo = TXa - 1 < 0 ? 0 : TXa - 1, chi = TXa + 2 > ntx - 1 ? ntx - 1 : TXa + 2;
    const uint32_t run_bytes = clo <= chi ? (uint32_t)(chi - clo + 1) * (uint32_t)(T_REC * 8) : 0u;
    int n_rows = 0;
#pragma unroll
    for (int r = 0; r < 3; ++r) n_rows += (TY - 1 + r >= 0 && TY - 1 + r < nty && run_bytes) ? 1 : 0;
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_bar);
    if (tid == 0) {
        s_n = 0;
        if (n_rows) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(1));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(run_bytes * (uint32_t)n_rows) : "memory");
            const unsigned long long *fine = reinterpret_cast<const unsigned long long *>(a.planes + (size_t)pair * g.pair_bytes);
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                const int ty = TY - 1 + r;
                if (ty < 0 || ty >= nty) continue;
                const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&s_rec[r][clo - (TXa - 1)][0]);
                const unsigned long long *src = fine + ((size_t)ty * g.ntx + clo) * T_REC;
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
                             "r"(run_bytes), "r"(bar)
                             : "memory");
            }
        }
    }
    // records outside the grid: zero (the bulk copies write the others; the two sets are disjoint)
    for (int i = tid; i < 3 * ST_COLS * T_REC; i += 256) {
        const int r = i / (ST_COLS * T_REC), c = (i / T_REC) % ST_COLS;
        const int ty = TY - 1 + r, tx = TXa - 1 + c;
        if (ty < 0 || ty >= nty || tx < clo || tx > chi) (&s_rec[0][0][0])[i] = 0ull;
    }
    if (tid < DIAG_N) s_le[tid] = diag_le(tid - 9);
    else if (tid < 2 * DIAG_N) s_ge[tid - DIAG_N] = diag_ge(tid - DIAG_N);
    __syncthreads();   // s_n, the tables, the zeroed records and the mbarrier's initialisation

    // ---- ring 0 for the 2 x 256 queries of the two origin tiles; survivors are queued ---------------------------------------
    const int lxq = tid & 15, lyq = tid >> 4;
    const int oyl = 8 * TY + (lyq >> 1);                              // origin row inside the sub-image held here
    const int Yg = 2 * (8 * TY + (int)s.add_y0) - 1 + lyq;            // global main row
    const int y = Yg - (int)s.y0;                                     // row inside the band
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const int x = 16 * (TXa + k) - 1 + lxq;
        const bool valid = x >= 0 && x < (int)s.w && y >= 0 && y < (int)s.h;
        bool open = false;
        uint32_t col = 0;
        if (valid) {
            const uint32_t qi = (uint32_t)y * s.w + (uint32_t)x;
            const int ox = (x + 1) >> 1, oyg = (Yg + 1) >> 1;
            col = reinterpret_cast<const uint32_t *>(s.main_rgba)[(size_t)pair * s.n_queries + qi] & 0xffffffu;
            // ring 0 out of bounds => `still_in_bounds` stays false => break => None   (helpers.rs:49-51)
            if (ox < (int)s.wa && oyg < (int)s.add_img_h) {
                bool hit0 = false;
                if (oyl >= 0 && oyl < (int)s.ha) {
                    const uint8_t *px = s.add_rgb + (size_t)pair * s.ha * s.wa * 3 + ((size_t)oyl * s.wa + ox) * 3;
                    hit0 = pixel_in_box(px[0], px[1], px[2], col, s.pr, s.pg, s.pb, s.sat != 0);
                }
                open = !hit0 && s.max_rings > 1;
            }
            // settled at ring 0: a hit has distance 0 and `None` maps to 0 too (depth_image.rs:120), neither moves the maximum
            if (!open) s.d8[(size_t)pair * s.n_queries + qi] = 0;
        }
        const uint32_t bal = __ballot_sync(0xffffffffu, open);
        if (bal) {
            uint32_t base = 0;
            if (lane == __ffs(bal) - 1) base = atomicAdd(&s_n, (uint32_t)__popc(bal));
            base = __shfl_sync(0xffffffffu, base, __ffs(bal) - 1);
            if (open) s_open[base + __popc(bal & ((1u << lane) - 1u))] = make_uint2((uint32_t)(k << 8 | tid), col);
        }
    }
    __syncthreads();
    if (n_rows) {   // every thread waits: a CTA must not retire with a bulk copy into its shared memory in flight
        uint32_t done;
        do {
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done)
                         : "r"(bar), "r"(0)
                         : "memory");
        } while (!done);
    }
    const uint32_t n_open = s_n;
    const int D = (int)s.max_rings;
    const int Smax = (D + 6) >> 3;
#pragma unroll 1
    for (uint32_t i0 = (uint32_t)(tid & ~31); i0 < n_open; i0 += 256u) {   // warp-uniform
        const bool has = i0 + lane < n_open;
        const uint2 e = has ? s_open[i0 + lane] : make_uint2(0, 0);
        const int k = (int)(e.x >> 8), t = (int)(e.x & 255u);
        const int qlx = t & 15, qly = t >> 4;
        const int qx = 16 * (TXa + k) - 1 + qlx, qy = 2 * (8 * TY + (int)s.add_y0) - 1 + qly - (int)s.y0;
        const uint32_t qi = (uint32_t)qy * s.w + (uint32_t)qx;
        uint32_t best = KEY_NONE;
        bool final_ = true;
        if (has) {
            const int lx = qlx >> 1, ly = qly >> 1;   // the origin inside its own tile
            const uint32_t qr = e.y & 0xffu, qg = (e.y >> 8) & 0xffu, qb = (e.y >> 16) & 0xffu;
            const uint32_t ur = (qr >> QSHIFT) + W_U0, ug = (qg >> QSHIFT) + W_U0, ub = (qb >> QSHIFT) + W_U0;
            const uint32_t rr = qr & (QSTEP - 1), rg = qg & (QSTEP - 1), rb = qb & (QSTEP - 1);
            const int T0x = TXa + k;
            // own tile first, then the four edge neighbours, then the corners: nearer tiles first, so that a match in hand
            // puts the farther ones out of reach.  Offsets packed two bits each: dx + 1, dy + 1.
            const uint32_t order = (1u | 1u << 2) | (1u | 2u << 2) << 4 | (1u | 0u << 2) << 8 | (2u | 1u << 2) << 12 | (0u | 1u << 2) << 16 |
                                   (2u | 2u << 2) << 20 | (0u | 2u << 2) << 24 | (2u | 0u << 2) << 28;
            const uint32_t order8 = 0u | 0u << 2;   // the ninth: (-1, -1)
#pragma unroll 1
            for (int j = 0; j < 9; ++j) {
                const uint32_t o = j < 8 ? (order >> (4 * j)) & 15u : order8;
                const int dx = (int)(o & 3u) - 1, dy = (int)(o >> 2) - 1;
                const int bx = 8 * dx - lx, by = 8 * dy - ly;   // offset of the tile's pixel (0,0) from the origin
                const int dxm = dx > 0 ? 8 - lx : (dx < 0 ? lx + 1 : 0), dym = dy > 0 ? 8 - ly : (dy < 0 ? ly + 1 : 0);
                const int smin = dxm > dym ? dxm : dym;          // ring of the tile's nearest pixel
                if (smin >= D || smin > (int)(best >> 12)) continue;
                const unsigned long long *rec = &s_rec[dy + 1][k + 1 + dx][0];
                if (!(rec[ur] & rec[T_SLOTS + ug] & rec[2 * T_SLOTS + ub])) continue;   // candidate word: a superset of the matches
                const unsigned long long m = staged_exact(rec, ur, ug, ub, rr, rg, rb);
                if (!m) continue;
                const uint32_t key = tile_min_key(m, bx, by, s_le, s_ge);
                if ((int)(key >> 12) < D && key < best) best = key;
            }
            // every pixel outside the 3 x 3 block lies at ring >= 8 + slack + 1 (search_hier.cuh: stop_slack)
            const int mx = lx < 7 - lx ? lx : 7 - lx, my = ly < 7 - ly ? ly : 7 - ly;
            const int slack = mx < my ? mx : my;
            const int T0y = TY;
            const bool exhausted = T0x - 1 < 0 && T0x + 1 > ntx - 1 && T0y - 1 < 0 && T0y + 1 > nty - 1;
            final_ = (best != KEY_NONE && (int)(best >> 12) <= 8 + slack) || exhausted || Smax < 2;
        }
        export_open(a, has && !final_, pair, qi, best);
        publish_result<true>(s, pair, qi, best, has && final_);
    }
}

#endif  // __CUDACC__

}  // namespace stepth
