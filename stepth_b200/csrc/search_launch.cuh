// search_launch.cuh -- host side of the ring search: K0 (planes) + K1 (near stage, far stage(s)) on one stream.
#pragma once
#include "search_hier.cuh"
#include "search_otile.cuh"
#include "search_tile.cuh"

namespace stepth {

#if defined(__CUDACC__)

// planes (K0) + search (K1) for n_pairs dense pairs, or for one block list (a.blocks != nullptr)
inline int32_t launch_search_hier(stepth_ctx *ctx, cudaStream_t st, const SearchArgs &a, uint32_t n_pairs) {
    if (a.wa > 0xffffu || a.add_img_h > 0xffffu)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "additional image larger than 65535 in one dimension");
    static_assert(WL_MAX_PAIRS * CT_PARTS <= 65535, "grid.z of the transpose kernel");
    if (n_pairs > WL_MAX_PAIRS)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "more than 2048 pairs in one launch");
    HierArgs h;
    memset(&h, 0, sizeof h);
    h.s = a;
    // planes only cover what a query can reach: dense origins are half-coordinates (depth_image.rs:114-117), so columns
    // [0, w/2 + max_rings) and local rows [0, (y0+h)/2 + max_rings - add_y0) -- the rest of the image is never probed
    uint32_t cover_w = a.wa, cover_h = a.ha;
    if (!a.blocks) {
        cover_w = (uint32_t)std::min<uint64_t>(a.wa, (uint64_t)a.w / 2 + a.max_rings);
        const int64_t last_row = (int64_t)(a.y0 + a.h) / 2 + a.max_rings - 1 - (int64_t)a.add_y0;  // local, inclusive
        cover_h = (uint32_t)std::max<int64_t>(1, std::min<int64_t>(a.ha, last_row + 1));
    }
    h.g = make_tile_geom(cover_w, cover_h, a.wa);
    stepth_ctx::Buf &planes = ctx->slot ? ctx->planes2 : ctx->planes, &worklist = ctx->slot ? ctx->worklist2 : ctx->worklist;
    int32_t rc = ensure(ctx, planes, h.g.pair_bytes * n_pairs);
    if (rc != STEPTH_OK) return rc;
    h.planes = (const uint8_t *)planes.p;
    if (h.g.nty > 65535u) return set_err(ctx, STEPTH_ERR_INVALID_ARG, "too many tile rows for one launch");
    TileBuildArgs p{a.add_rgb, (size_t)a.ha * a.wa * 3, (uint8_t *)planes.p, h.g, a.pr, a.pg, a.pb, a.sat};
    // the strips of K0 come in through a tensor map when the images allow one: rows that start on 16 bytes (the same condition
    // the per-row bulk copies have), at least one full 16-byte word per row
    CUtensorMap tmap;
    memset(&tmap, 0, sizeof tmap);
    int use_tmap = 0;
    const size_t pitch = (size_t)a.wa * 3;
    if (ctx->k0_tensor_map && (pitch & 15) == 0 && (reinterpret_cast<uintptr_t>(a.add_rgb) & 15) == 0 && pitch / 4 <= 0xffffffffull) {
        static const auto encode = [] {
            void *fn = nullptr;
            cudaDriverEntryPointQueryResult q;
            if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) fn = nullptr;
            (void)cudaGetLastError();
            return (decltype(&cuTensorMapEncodeTiled))fn;
        }();
        if (encode) {
            const cuuint64_t dims[3] = {(cuuint64_t)(pitch / 4), (cuuint64_t)a.ha, (cuuint64_t)n_pairs};
            const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)p.src_pair_bytes};   // bytes, dimensions 1 and 2
            const cuuint32_t box[3] = {192u, 8u, 1u}, estr[3] = {1u, 1u, 1u};
            use_tmap = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, (void *)a.add_rgb, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
        }
    }
    tile_planes_kernel<<<dim3(h.g.nxw, h.g.nty, n_pairs), 256, 0, st>>>(p, tmap, use_tmap);
    ctx->launches++;
    STEPTH_CUDA_TRY(ctx, cudaGetLastError());
    coarse_transpose_kernel<<<dim3(h.g.nxw, h.g.nyw, n_pairs * CT_PARTS), 256, 0, st>>>(p);
    ctx->launches++;
    STEPTH_CUDA_TRY(ctx, cudaGetLastError());
    if (ctx->main_ready) {   // K0 needed the additional images only; the queries come from the main ones
        cudaEvent_t ev = ctx->main_ready;
        ctx->main_ready = nullptr;
        STEPTH_CUDA_TRY(ctx, cudaStreamWaitEvent(st, ev, 0));
    }

    // work lists: [count, cursor, pad ..][entries]; at most one entry per query
    const size_t cap = (size_t)a.n_queries * n_pairs;
    if (cap > 0xfffffff0ull) return set_err(ctx, STEPTH_ERR_INVALID_ARG, "more than 2^32 queries in one launch");
    const size_t wl_bytes = 64 + cap * sizeof(uint2);
    rc = ensure(ctx, worklist, 2 * wl_bytes);
    if (rc != STEPTH_OK) return rc;
    uint8_t *wl0 = (uint8_t *)worklist.p, *wl1 = wl0 + wl_bytes;
    STEPTH_CUDA_TRY(ctx, cudaMemsetAsync(wl0, 0, 64, st));
    STEPTH_CUDA_TRY(ctx, cudaMemsetAsync(wl1, 0, 64, st));
    h.wl_cap = (uint32_t)cap;
#ifdef STEPTH_DEBUG_CHECKS
    rc = ensure(ctx, ctx->dbg, 256);
    if (rc != STEPTH_OK) return rc;
    STEPTH_CUDA_TRY(ctx, cudaMemsetAsync(ctx->dbg.p, 0, 4, st));
    h.dbg_err = (uint32_t *)ctx->dbg.p;
#endif

    // NEAR: ring 0 + the first coarse rings
    // dense mode: 2 (default) = one warp per origin tile (search_otile.cuh), 1 = rings 0..1 from TMA-staged tile records
    // (search_tile.cuh), 0 = the walking near kernel (also what block mode runs)
    const bool tile_stage = !a.blocks && ctx->near_tile == 1;
    // 3 (default): origin tiles, except for launches of 4K .. 28K origin tiles (one to three 1080p pairs), where the walking kernel's
    // shorter per-CTA work leaves a shorter tail (one 1080p pair 0.221 vs 0.237 ms, three 0.531 vs 0.578; four 0.663 vs 0.660; one
    // README pair 0.132 vs 0.127; two 4K pairs 1.236 vs 1.174: profiles/recipes/r2_run21.sh)
    const uint64_t origin_tiles = (uint64_t)(((a.w >> 1) >> 3) + 1) * (uint64_t)((a.h >> 4) + 2) * n_pairs;
    const bool otile_stage = !a.blocks && (ctx->near_tile == 2 || (ctx->near_tile == 3 && (origin_tiles < 4096 || origin_tiles >= 28000)));
    h.S0 = 0;
    h.S1 = tile_stage ? 2 : otile_stage ? std::min(OT_MAX_S1, std::max(2, ctx->near_rings_otile)) : ctx->near_rings;
    h.wl_out = (uint2 *)(wl0 + 64);
    h.wl_out_count = (uint32_t *)wl0;
    // A CTA settles ring 0 for up to STEPTH_NEAR_KP patches before it walks; a launch that would not fill the SMs a few
    // times over takes fewer (one 1080p pair: 4 patches 0.257 ms, 1 patch 0.238 ms; 16 pairs: the other way round)
    const uint32_t units = a.blocks ? (uint32_t)((a.n_queries + 255) / 256) : ((a.w + 31) / 32) * ((a.h + 7) / 8) * n_pairs;
    const uint32_t per_sm = units / (uint32_t)std::max(1, ctx->sm_count);
    int kp = !ctx->near_compact ? 0 : (per_sm >= 160 ? 4 : per_sm >= 80 ? 2 : 1);
    if (kp > STEPTH_NEAR_KP) kp = STEPTH_NEAR_KP;
    if (tile_stage || otile_stage) {
        // origin tiles: origins are (x+1)/2, (y+y0+1)/2 - add_y0 (depth_image.rs:114-117)
        auto floor8 = [](int64_t v) { return (int)(v >= 0 ? v / 8 : -((-v + 7) / 8)); };
        const int64_t oy_min = (int64_t)((a.y0 + 1) >> 1) - (int64_t)a.add_y0, oy_max = (int64_t)((a.y0 + a.h) >> 1) - (int64_t)a.add_y0;
        const int ty_first = floor8(oy_min), n_ty = floor8(oy_max) - ty_first + 1;
        const uint32_t n_tx = ((a.w >> 1) >> 3) + 1;
        if (n_ty > 8191) return set_err(ctx, STEPTH_ERR_INVALID_ARG, "band taller than 128K rows");
        if (tile_stage) {
            search_tile_kernel<<<dim3((n_tx + 1) / 2, (uint32_t)n_ty, n_pairs), 256, 0, st>>>(h, ty_first);   // one CTA per two origin tiles
        } else {
            // one warp per origin tile (256 queries); half, quarter or eighth tiles when whole ones would not fill the SMs a few
            // times over (the list a warp packs shrinks with it, and so does the share of full warps in the late passes)
            const uint64_t tiles = (uint64_t)n_tx * (uint64_t)n_ty * n_pairs, want = (uint64_t)ctx->sm_count * ctx->otile_warps_per_sm;
            int nit = 8;
            while (nit > 1 && tiles * (uint64_t)(8 / nit) < want) nit >>= 1;
            if (ctx->otile_nit == 8 || ctx->otile_nit == 4 || ctx->otile_nit == 2 || ctx->otile_nit == 1) nit = ctx->otile_nit;
            constexpr uint32_t per_cta = STEPTH_OT_SPAN > 0 ? STEPTH_OT_SPAN : OT_WARPS;   // origin tiles of a tile row per CTA
            const dim3 grid((n_tx + per_cta - 1) / per_cta, (uint32_t)n_ty * (uint32_t)(8 / nit), n_pairs);
            if (nit == 8) search_otile_kernel<8><<<grid, OT_WARPS * 32, 0, st>>>(h, ty_first);
            else if (nit == 4) search_otile_kernel<4><<<grid, OT_WARPS * 32, 0, st>>>(h, ty_first);
            else if (nit == 2) search_otile_kernel<2><<<grid, OT_WARPS * 32, 0, st>>>(h, ty_first);
            else search_otile_kernel<1><<<grid, OT_WARPS * 32, 0, st>>>(h, ty_first);
        }
    } else if (a.blocks) {
        const uint32_t gx = (uint32_t)((a.n_queries + 255) / 256);
        if (kp == 4) search_near_kernel<true, 4><<<dim3((gx + 3) / 4, 1, 1), 256, 0, st>>>(h);
        else if (kp == 2) search_near_kernel<true, 2><<<dim3((gx + 1) / 2, 1, 1), 256, 0, st>>>(h);
        else if (kp == 1) search_near_kernel<true, 1><<<dim3(gx, 1, 1), 256, 0, st>>>(h);
        else search_near_kernel<true, 0><<<dim3(gx, 1, 1), 256, 0, st>>>(h);
    } else {
        const uint32_t gx = (a.w + 31) / 32, gy = (a.h + 7) / 8;
        if (gy > 65535u) return set_err(ctx, STEPTH_ERR_INVALID_ARG, "band taller than 524280 rows");
        if (kp == 4) search_near_kernel<false, 4><<<dim3((gx + 3) / 4, gy, n_pairs), 256, 0, st>>>(h);
        else if (kp == 2) search_near_kernel<false, 2><<<dim3((gx + 1) / 2, gy, n_pairs), 256, 0, st>>>(h);
        else if (kp == 1) search_near_kernel<false, 1><<<dim3(gx, gy, n_pairs), 256, 0, st>>>(h);
        else search_near_kernel<false, 0><<<dim3(gx, gy, n_pairs), 256, 0, st>>>(h);
    }
    ctx->launches++;
    STEPTH_CUDA_TRY(ctx, cudaGetLastError());

    // FAR: one or two resumptions; the last one walks to the end
    // one resumption by default: with the pooled tile tests a second stage boundary no longer pays (16 pairs 1080p: nat 2.318 ->
    // 2.308 ms, worst case 6.59 -> 6.00 ms; one pair 0.260 -> 0.241 ms); STEPTH_MID_RINGS brings it back
    const uint32_t far_grid = (uint32_t)ctx->sm_count * 16;
    uint8_t *in = wl0, *out = wl1;
    // resumption points between the first stage and the last: after the tile stage the walk resumes at coarse ring 2 and may be cut
    // again at near_rings and at mid_rings; after the walking near kernel only at mid_rings
    const int cuts[2] = {tile_stage ? ctx->near_rings : 0, ctx->mid_rings};
    bool used_out = false;
    for (int ci = 0; ci < 2; ++ci) {
        const int cut = cuts[ci];
        if (cut <= h.S1) continue;
        if (used_out) STEPTH_CUDA_TRY(ctx, cudaMemsetAsync(out, 0, 64, st));   // a list consumed two stages ago is the output again
        h.S0 = h.S1;
        h.S1 = cut;
        h.wl_in = (const uint2 *)(in + 64);
        h.wl_in_count = (const uint32_t *)in;
        h.wl_in_cursor = (uint32_t *)in + 1;
        h.wl_out = (uint2 *)(out + 64);
        h.wl_out_count = (uint32_t *)out;
        if (a.blocks) search_far_kernel<true><<<far_grid, 128, 0, st>>>(h);
        else search_far_kernel<false><<<far_grid, 128, 0, st>>>(h);
        ctx->launches++;
        STEPTH_CUDA_TRY(ctx, cudaGetLastError());
        std::swap(in, out);
        used_out = true;
    }
    h.S0 = h.S1;
    h.S1 = 1 << 20;
    h.wl_in = (const uint2 *)(in + 64);
    h.wl_in_count = (const uint32_t *)in;
    h.wl_in_cursor = (uint32_t *)in + 1;
    h.wl_out = nullptr;
    h.wl_out_count = nullptr;
    // a short list (one README pair, one 1080p pair) is walked by one warp per query (search_tail_kernel), a long one by one lane
    // per query: both kernels are launched, the list's length decides on the device which of them works.  The tail kernel's runs of
    // 32 tiles cover sides of up to 96 tiles: rings <= 32, max_rings <= 257
    {
        const int smax = ((int)a.max_rings + 6) / 8, left = smax - h.S0;   // coarse rings this stage may have to walk
        h.tail_split = (ctx->tail_kernel && smax <= 32 && left > 0) ? (left >= 12 ? 24576 : 6144) : 0;
    }
    if (a.blocks) search_far_kernel<true><<<far_grid, 128, 0, st>>>(h);
    else search_far_kernel<false><<<far_grid, 128, 0, st>>>(h);
    ctx->launches++;
    STEPTH_CUDA_TRY(ctx, cudaGetLastError());
    if (h.tail_split) {
        const uint32_t tail_grid = (uint32_t)ctx->sm_count * 8;
        if (a.blocks) search_tail_kernel<true><<<tail_grid, 128, 0, st>>>(h);
        else search_tail_kernel<false><<<tail_grid, 128, 0, st>>>(h);
        ctx->launches++;
        STEPTH_CUDA_TRY(ctx, cudaGetLastError());
    }
#ifdef STEPTH_DEBUG_CHECKS
    {   // debug builds wait for the search and report the index checks that failed
        uint32_t bad = 0;
        STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(&bad, ctx->dbg.p, 4, cudaMemcpyDeviceToHost, st));
        STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(st));
        if (bad) return set_err(ctx, STEPTH_ERR_CUDA, "STEPTH_DEBUG_CHECKS: index checks failed, mask 0x%x (search_hier.cuh / search_otile.cuh)", bad);
    }
#endif
    return STEPTH_OK;
}

#endif  // __CUDACC__

}  // namespace stepth
