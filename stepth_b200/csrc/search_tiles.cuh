// search_tiles.cuh -- K1: the production ring search (helpers.rs:9-54 + depth_image.rs:113-122).
//
// The reference walks ring s = 0, 1, .. around the origin as: row y+s, row y-s (x ascending), column x+s, column x-s
// (y ascending) and returns the FIRST pixel inside the query's colour box.  That is an arg-min of the ring-order key
// (common.cuh); this kernel finds the same pixel, but 32 probes at a time on the candidate bit-planes of planes.cuh:
//
//   * a CTA owns a 16x16 tile of queries = 8x8 origins (the literal `(pos + size) / 2` origin of depth_image.rs:114-117
//     makes main pixels (2k-1, 2k) share origin k, so tiles start at odd coordinates); ring 0 is tested directly;
//   * rings are walked in chunks [s0, s1) of growing thickness.  Per chunk the still unresolved queries are compacted
//     in shared memory into a row-ordered and a column-ordered list; the ROW pass gives a warp 32 queries of (mostly) one
//     origin row -> every lane reads the same 128 B plane line (lanes differ only in the bin word j); the COLUMN pass
//     re-assigns threads so that a warp holds queries of one origin column and reads the transposed planes the same
//     way.  A query's answer is min(row key, column key); the column pass of a chunk only walks rings below the row hit;
//   * candidates (set bits) are verified against the real pixel in probe order (__ffs), so the superset planes never
//     change the result; out-of-image probes are 0 bits / clipped ranges (helpers.rs:19-25,44);
//   * `d as u8` (depth_image.rs:120-121) and the per-pair maximum (:124) leave the kernel: one byte per query, one
//     atomicMax per CTA.
//
// The same kernel serves the block list (one CTA per 256 consecutive DiscretePixels; list order is kept).
#pragma once
#include <algorithm>

#include "ctx.cuh"
#include "planes.cuh"
#include "search_literal.cuh"

namespace stepth {

constexpr uint32_t KEY_DEAD = 0xfffffffeu;       // no further in-bounds ring / origin outside the image: None
constexpr uint32_t KEY_DEAD_FROM = 0x80000000u;  // row pass -> column pass: "rings >= (key & 0xffff) are out of bounds"

struct TileArgs {
    SearchArgs s;
    const uint32_t *planes;
    PlaneGeom g;
    uint32_t by0;  // first tile row (global tile coordinates) of a row band
    // sparse-tail export (search_origins.cuh -> search_tail.cuh)
    uint2 *wl;              // exported origins
    uint32_t *wl_count;     // [0] = number of entries, [1] = tail cursor
    uint32_t export_below;  // a tile exports its unresolved origins once fewer than this many are left (0 = never)
};

__device__ __forceinline__ bool verify_px(const uint8_t *add, uint32_t wa, int x, int y, const ColourBox &box) {
    const uint8_t *v = add + ((size_t)y * wa + x) * 3;
    return box_contains(box, v[0], v[1], v[2]);
}

// first candidate-and-verified pixel of the segment [lo, hi] (clipped to [0, n-1]) of line `line` of a plane set.
// `planes` is RP (line = y, pixels along x) or CP (line = x, pixels along y); `swap` tells verify which is which.
template <bool COLS>
__device__ __forceinline__ int scan_line(const uint32_t *__restrict__ planes, uint32_t words_per_line, int line,
                                         int n_lines, int lo, int hi, int n, uint32_t jr, uint32_t jg, uint32_t jb,
                                         uint32_t col) {
    if ((unsigned)line >= (unsigned)n_lines) return -1;
    if (lo < 0) lo = 0;
    if (hi > n - 1) hi = n - 1;
    if (lo > hi) return -1;
    const int w0 = lo >> 5, w1 = hi >> 5;
    const uint32_t *cell = planes + ((size_t)line * words_per_line + w0) * CELL;
    for (int w = w0; w <= w1; ++w, cell += CELL) {
        uint32_t m = __ldg(cell + jr) & __ldg(cell + jg) & __ldg(cell + jb);
        if (w == w0) m &= 0xffffffffu << (lo & 31);
        if (w == w1) m &= 0xffffffffu >> (31 - (hi & 31));
        if (m) m = refine_exact(cell, col, m);
        if (m) return (w << 5) + __ffs(m) - 1;
    }
    return -1;
}

template <bool BLOCKS>
__global__ void __launch_bounds__(256) search_tiles_kernel(TileArgs a) {
    __shared__ uint32_t q_pos[256], q_col[256], q_key[256];
    __shared__ uint16_t list_r[256], list_c[256];
    __shared__ uint32_t s_cnt[2][8];
    __shared__ uint32_t s_max;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t pair = BLOCKS ? 0u : blockIdx.z;
    const int wa = (int)a.g.wa, ha = (int)a.g.ha;
    const int src_wa = (int)a.s.wa, src_ha = (int)a.s.ha;
    const int add_y0 = (int)a.s.add_y0, img_h = (int)a.s.add_img_h;
    const uint8_t *add = a.s.add_rgb + (size_t)pair * src_ha * src_wa * 3;
    const uint32_t *rp = a.planes + (size_t)pair * a.g.pair_words;
    const uint32_t *cp = rp + a.g.rp_words;
    const bool sat = a.s.sat != 0;

    // ---- this thread's query: origin (depth_image.rs:114-117), colour, output slot --------------------------------
    bool valid;
    int ox = 0, oyg = 0;
    uint32_t col = 0;
    size_t out_idx = 0;
    if (BLOCKS) {
        const uint32_t i = blockIdx.x * 256u + tid;
        valid = i < a.s.n_queries;
        if (valid) {
            const stepth_block bl = a.s.blocks[i];
            ox = (int)((bl.x + bl.w) / 2);
            oyg = (int)((bl.y + bl.h) / 2);
            col = bl.value[0] | (bl.value[1] << 8) | (bl.value[2] << 16);
            out_idx = i;
        }
    } else {
        const int X = 16 * (int)blockIdx.x - 1 + (tid & 15);
        const int Yg = 16 * (int)(a.by0 + blockIdx.y) - 1 + (tid >> 4);
        const int Yl = Yg - (int)a.s.y0;
        valid = X >= 0 && X < (int)a.s.w && Yl >= 0 && Yl < (int)a.s.h;
        if (valid) {
            const size_t q = (size_t)Yl * a.s.w + X;
            col = reinterpret_cast<const uint32_t *>(a.s.main_rgba)[(size_t)pair * a.s.n_queries + q] & 0xffffffu;
            ox = (X + 1) >> 1;
            oyg = (Yg + 1) >> 1;
            out_idx = (size_t)pair * a.s.n_queries + q;
        }
    }
    {
        uint32_t key = KEY_NONE;
        if (!valid || ox >= src_wa || oyg >= img_h || ox > 0xffff || oyg > 0xffff) {
            key = KEY_DEAD;  // ring 0 has no in-bounds probe: `still_in_bounds` stays false -> None (helpers.rs:49-51)
        } else {
            const int oyl = oyg - add_y0;
            const ColourBox box = make_box(col & 0xff, (col >> 8) & 0xff, (col >> 16) & 0xff, a.s.pr, a.s.pg, a.s.pb, sat);
            if (oyl >= 0 && oyl < src_ha && verify_px(add, (uint32_t)src_wa, ox, oyl, box)) key = 0;  // ring 0
        }
        q_pos[tid] = (uint32_t)ox | ((uint32_t)oyg << 16);
        q_col[tid] = col;
        q_key[tid] = key;
    }
    if (tid == 0) s_max = 0;
    __syncthreads();

    const int D = (int)a.s.max_rings;
    // column-major enumeration of the tile (dense) / list order (blocks)
    const int qc = BLOCKS ? tid : (((tid & 15) << 4) | (tid >> 4));
    for (int s0 = 1; s0 < D;) {
        const int s1 = min(D, s0 + max(1, s0 >> 1));
        // ---- compact the unresolved queries into a row-ordered and a column-ordered list -----------------------
        const bool alive_r = q_key[tid] == KEY_NONE, alive_c = q_key[qc] == KEY_NONE;
        const uint32_t br = __ballot_sync(0xffffffffu, alive_r), bc = __ballot_sync(0xffffffffu, alive_c);
        if (lane == 0) { s_cnt[0][warp] = __popc(br); s_cnt[1][warp] = __popc(bc); }
        __syncthreads();
        uint32_t pre_r = 0, pre_c = 0, n_alive = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w) {
            const uint32_t c0 = s_cnt[0][w], c1 = s_cnt[1][w];
            if (w < warp) { pre_r += c0; pre_c += c1; }
            n_alive += c0;
        }
        if (n_alive == 0) break;  // uniform
        const uint32_t lt = (1u << lane) - 1u;
        if (alive_r) list_r[pre_r + __popc(br & lt)] = (uint16_t)tid;
        if (alive_c) list_c[pre_c + __popc(bc & lt)] = (uint16_t)qc;
        __syncthreads();

        // ---- row pass: rows y+s then y-s of rings s0 .. s1-1 ------------------------------------------------------
        if ((uint32_t)tid < n_alive) {
            const int q = list_r[tid];
            const uint32_t pos = q_pos[q], c = q_col[q];
            const int qx = (int)(pos & 0xffff), qyg = (int)(pos >> 16), qy = qyg - add_y0;
            const uint32_t r = c & 0xff, g = (c >> 8) & 0xff, b = (c >> 16) & 0xff;
            const uint32_t jr = w_u(r), jg = CH_WORDS + w_u(g), jb = 2 * CH_WORDS + w_u(b);
            uint32_t key = KEY_NONE;
            for (int s = s0; s < s1; ++s) {
                // no in-bounds probe from ring s on (helpers.rs:49-51): the column pass may still own rings < s
                if (qx - s < 0 && qx + s >= src_wa && qyg - s < 0 && qyg + s >= img_h) { key = KEY_DEAD_FROM | (uint32_t)s; break; }
                int x = scan_line<false>(rp, a.g.nxw, qy + s, ha, qx - s, qx + s, wa, jr, jg, jb, c);
                if (x >= 0) { key = ring_key(x - qx, s); break; }
                x = scan_line<false>(rp, a.g.nxw, qy - s, ha, qx - s, qx + s, wa, jr, jg, jb, c);
                if (x >= 0) { key = ring_key(x - qx, -s); break; }
            }
            if (key != KEY_NONE) q_key[q] = key;
        }
        __syncthreads();
        // ---- column pass: columns x+s then x-s, only rings below the row hit of this chunk -------------------------
        if ((uint32_t)tid < n_alive) {
            const int q = list_c[tid];
            const uint32_t krow = q_key[q];
            {
                const uint32_t pos = q_pos[q], c = q_col[q];
                const int qx = (int)(pos & 0xffff), qyg = (int)(pos >> 16), qy = qyg - add_y0;
                const uint32_t r = c & 0xff, g = (c >> 8) & 0xff, b = (c >> 16) & 0xff;
                    const uint32_t jr = w_u(r), jg = CH_WORDS + w_u(g), jb = 2 * CH_WORDS + w_u(b);
                const bool dead_from = krow != KEY_NONE && (krow & KEY_DEAD_FROM) != 0;
                const int lim = krow == KEY_NONE ? s1 : dead_from ? (int)(krow & 0xffffu) : (int)(krow >> 12);
                uint32_t key = dead_from ? KEY_DEAD : krow;
                for (int s = s0; s < lim; ++s) {
                    int y = scan_line<true>(cp, a.g.nyw, qx + s, wa, qy - s + 1, qy + s - 1, ha, jr, jg, jb, c);
                    if (y >= 0) { key = ring_key(s, y - qy); break; }
                    y = scan_line<true>(cp, a.g.nyw, qx - s, wa, qy - s + 1, qy + s - 1, ha, jr, jg, jb, c);
                    if (y >= 0) { key = ring_key(-s, y - qy); break; }
                }
                if (key != krow) q_key[q] = key;
            }
        }
        __syncthreads();
        s0 = s1;
    }

    // ---- `d as u8` per query, per-pair maximum (depth_image.rs:120-124) -------------------------------------------
    const uint32_t key = q_key[tid];
    const uint32_t d = key >= KEY_DEAD ? 0u : (uint32_t)key_to_d8(key);
    if (valid) a.s.d8[out_idx] = (uint8_t)d;
    const uint32_t m = __reduce_max_sync(0xffffffffu, d);
    if (lane == 0 && m) atomicMax(&s_max, m);
    __syncthreads();
    if (tid == 0 && s_max) atomicMax(a.s.max_out + pair, s_max);
}

// forward: the dense-mode kernel lives in search_origins.cuh
constexpr int ORG_NT_FWD = 64;
__global__ void search_origins_kernel(TileArgs a);
void launch_search_tail(stepth_ctx *ctx, cudaStream_t st, const TileArgs &t);  // search_tail.cuh

// planes (K0) + search (K1) for n_pairs dense pairs, or for one block list (a.blocks != nullptr)
inline int32_t launch_search_tiles(stepth_ctx *ctx, cudaStream_t st, const SearchArgs &a, uint32_t n_pairs) {
    if (a.wa > 0xffffu || a.add_img_h > 0xffffu)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "additional image larger than 65535 in one dimension");
    TileArgs t;
    memset(&t, 0, sizeof t);
    t.s = a;
    // planes only cover what a query can reach: dense origins are half-coordinates (depth_image.rs:114-117), so columns
    // [0, w/2 + max_rings) and local rows [0, (y0+h)/2 + max_rings - add_y0) -- the rest of the image is never probed
    uint32_t cover_w = a.wa, cover_h = a.ha;
    if (!a.blocks) {
        cover_w = (uint32_t)std::min<uint64_t>(a.wa, (uint64_t)a.w / 2 + a.max_rings);
        const int64_t last_row = (int64_t)(a.y0 + a.h) / 2 + a.max_rings - 1 - (int64_t)a.add_y0;  // local, inclusive
        cover_h = (uint32_t)std::max<int64_t>(1, std::min<int64_t>(a.ha, last_row + 1));
    }
    t.g = make_plane_geom(cover_w, cover_h, a.wa);
    const size_t plane_bytes = t.g.pair_words * sizeof(uint32_t) * n_pairs;
    int32_t rc = ensure(ctx, ctx->planes, plane_bytes);
    if (rc != STEPTH_OK) return rc;
    t.planes = (const uint32_t *)ctx->planes.p;
    PlaneArgs p{a.add_rgb, (size_t)a.ha * a.wa * 3, (uint32_t *)ctx->planes.p, t.g, a.pr, a.pg, a.pb, a.sat};
    if (t.g.nyw > 65535u || n_pairs > 65535u)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "too many plane tile rows or pairs for one launch");
    planes_kernel<<<dim3(t.g.nxw, t.g.nyw, n_pairs), 256, 0, st>>>(p);
    ctx->launches++;
    STEPTH_CUDA_TRY(ctx, cudaGetLastError());
    if (a.blocks) {
        t.by0 = 0;
        search_tiles_kernel<true><<<dim3((a.n_queries + 255) / 256, 1, 1), 256, 0, st>>>(t);
    } else {
        // 16x16 origins per CTA; origin (k, l) serves main pixels (2k-1 .. 2k, 2l-1 .. 2l)
        t.by0 = ((a.y0 + 1) / 2) / 16;
        const uint32_t by1 = ((a.y0 + a.h) / 2) / 16;  // origin row of the last band row y0 + h - 1
        const uint32_t tiles_y = by1 - t.by0 + 1, tiles_x = (a.w / 2) / 16 + 1;
        if (tiles_y > 65535u) return set_err(ctx, STEPTH_ERR_INVALID_ARG, "band taller than 2M rows");
        // work list of the sparse tail: at most one entry per origin
        const size_t max_entries = (size_t)tiles_x * tiles_y * 256 * n_pairs;
        rc = ensure(ctx, ctx->worklist, max_entries * sizeof(uint2) + 64);
        if (rc != STEPTH_OK) return rc;
        t.wl_count = (uint32_t *)ctx->worklist.p;
        t.wl = (uint2 *)((char *)ctx->worklist.p + 64);
        t.export_below = ctx->export_below;
        STEPTH_CUDA_TRY(ctx, cudaMemsetAsync(t.wl_count, 0, 64, st));
        search_origins_kernel<<<dim3(tiles_x, tiles_y, n_pairs), ORG_NT_FWD, 0, st>>>(t);
        if (t.export_below) {
            ctx->launches++;
            STEPTH_CUDA_TRY(ctx, cudaGetLastError());
            launch_search_tail(ctx, st, t);
        }
    }
    ctx->launches++;
    STEPTH_CUDA_TRY(ctx, cudaGetLastError());
    return STEPTH_OK;
}

}  // namespace stepth
