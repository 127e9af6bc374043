// multi_api.cuh -- section 5 of libstepth_cuda.so: the multi-GPU entry points of include/stepth_cuda.h (see multi.cuh).
// Included at the end of stepth_cuda.cu: it drives the single-GPU entry points and stage calls defined there.
#pragma once
#include "multi.cuh"

using namespace stepth;

extern "C" void stepth_host_band_plan(uint32_t h, uint32_t add_h, uint32_t max_rings, uint32_t n_ranks, uint32_t rank,
                                      uint32_t out[4]) {
    if (!out) return;
    out[0] = out[1] = out[2] = out[3] = 0;
    if (n_ranks == 0 || rank >= n_ranks || max_rings == 0) return;
    band_plan(h, add_h, max_rings, n_ranks, rank, out);
}

extern "C" int32_t stepth_cuda_comm_unique_id(uint8_t *id_out) {
    if (!id_out) return STEPTH_ERR_INVALID_ARG;
    NcclApi &n = nccl_api();
    if (!n.lib) {
        g_create_err = n.why;
        return STEPTH_ERR_NCCL;
    }
    ncclUniqueId id;
    ncclResult_t r = n.GetUniqueId(&id);
    if (r != ncclSuccess) {
        g_create_err = std::string("ncclGetUniqueId failed: ") + n.GetErrorString(r);
        return STEPTH_ERR_NCCL;
    }
    static_assert(sizeof(ncclUniqueId) == 128, "the header promises 128 bytes");
    memcpy(id_out, &id, sizeof id);
    return STEPTH_OK;
}

static int32_t comm_alloc(stepth_ctx *ctx, int32_t n_ranks, int32_t rank, ncclComm_t nccl, stepth_comm **out) {
    stepth_comm *c = new stepth_comm();
    c->ctx = ctx;
    c->n_ranks = n_ranks;
    c->rank = rank;
    c->nccl = nccl;
    *out = c;
    return STEPTH_OK;
}

extern "C" void stepth_cuda_comm_destroy(stepth_comm *c) {
    if (!c) return;
    if (c->ctx) cudaSetDevice(c->ctx->device);
    if (c->nccl && c->owns_nccl && nccl_api().lib) nccl_api().CommDestroy(c->nccl);
    if (c->d8h.p) cudaFree(c->d8h.p);
    if (c->small.p) cudaFree(c->small.p);
    if (c->h_small) cudaFreeHost(c->h_small);
    delete c;
}

extern "C" int32_t stepth_cuda_comm_create(stepth_ctx *ctx, int32_t n_ranks, int32_t rank, const uint8_t *id,
                                           stepth_comm **comm_out) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!comm_out || n_ranks < 1 || rank < 0 || rank >= n_ranks || (n_ranks > 1 && !id))
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "comm_create: null or invalid argument");
    *comm_out = nullptr;
    STEPTH_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    ncclComm_t nc = nullptr;
    if (n_ranks > 1) {
        NcclApi &n = nccl_api();
        if (!n.lib) return set_err(ctx, STEPTH_ERR_NCCL, "%s", n.why.c_str());
        ncclUniqueId uid;
        memcpy(&uid, id, sizeof uid);
        STEPTH_NCCL_TRY(ctx, n.CommInitRank(&nc, n_ranks, uid, rank));
    }
    return comm_alloc(ctx, n_ranks, rank, nc, comm_out);
}

extern "C" int32_t stepth_cuda_dev_depth_dense_band(stepth_comm *c, void *stream, uint32_t w, uint32_t h, uint32_t wa,
                                                    uint32_t add_h, const stepth_params *params, const uint8_t *main_band,
                                                    const uint8_t *add_band, uint8_t *depth_band_out, uint8_t *depth_full_out,
                                                    uint8_t *mask_band_out, uint8_t *rgba_band_out) {
    if (!c || !c->ctx) return STEPTH_ERR_INVALID_ARG;
    stepth_ctx *ctx = c->ctx;
    if (!params_ok(params) || !main_band || !depth_band_out || w == 0 || h == 0 || wa == 0 || add_h == 0)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "dev_depth_dense_band: null or invalid argument");
    const uint32_t flags = params->flags;
    const bool want_fg = (flags & STEPTH_FLAG_FOREGROUND) != 0, want_apply = (flags & STEPTH_FLAG_APPLY) != 0;
    if (want_apply && (!want_fg || !rgba_band_out))
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "APPLY needs FOREGROUND and rgba_band_out");
    const int n = c->n_ranks, r = c->rank;
    uint32_t pl[4];
    band_plan(h, add_h, params->max_rings, (uint32_t)n, (uint32_t)r, pl);
    const uint32_t y0 = pl[0], y1 = pl[1], ay0 = pl[2], ay1 = pl[3], rows = y1 - y0;
    if (n > 1 && rows < BAND_HALO)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "row band of %u rows is shorter than the Gaussian's 3-row halo: use fewer ranks", rows);
    if (ay1 > ay0 && !add_band) return set_err(ctx, STEPTH_ERR_INVALID_ARG, "dev_depth_dense_band: add_band is null");
    NcclApi &nc = nccl_api();
    if (n > 1 && !nc.lib) return set_err(ctx, STEPTH_ERR_NCCL, "%s", nc.why.c_str());
    cudaStream_t st = (cudaStream_t)stream;
    STEPTH_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    int32_t rc;
    if ((rc = ensure(ctx, c->d8h, (size_t)(rows + 2 * BAND_HALO) * w)) != STEPTH_OK) return rc;
    if ((rc = ensure(ctx, c->small, stepth_comm::SMALL_BYTES)) != STEPTH_OK) return rc;
    if (!c->h_small) STEPTH_CUDA_TRY(ctx, cudaMallocHost(&c->h_small, stepth_comm::SMALL_BYTES));
    uint32_t *sm = (uint32_t *)c->small.p;
    uint32_t *d_max = sm, *d_nout = sm + 2, *d_hist = sm + 4;
    int32_t *d_status = (int32_t *)(sm + 1);
    uint8_t *d_ranges = (uint8_t *)(sm + 260);
    STEPTH_CUDA_TRY(ctx, cudaMemsetAsync(c->small.p, 0, stepth_comm::SMALL_BYTES, st));
    uint8_t *d8 = (uint8_t *)c->d8h.p + (size_t)BAND_HALO * w;

    // 1. ring search of this band (helpers.rs:9-54, depth_image.rs:113-122); the per-band maximum lands in d_max
    stepth_dev_shape s;
    memset(&s, 0, sizeof s);
    s.n_pairs = 1; s.w = w; s.h = rows; s.img_h = h; s.y0 = y0;
    s.wa = wa; s.ha = ay1 - ay0; s.add_y0 = ay0; s.add_img_h = add_h;
    if ((rc = stepth_cuda_dev_search_dense(ctx, st, &s, params, main_band, add_band, d8, d_max)) != STEPTH_OK) return rc;
    const uint32_t above = r > 0 ? std::min(BAND_HALO, y0) : 0u, below = r < n - 1 ? std::min(BAND_HALO, h - y1) : 0u;
    if (n > 1) {
        // 2. exchange 1: the normaliser is the maximum over the WHOLE image (depth_image.rs:124)
        STEPTH_NCCL_TRY(ctx, nc.AllReduce(d_max, d_max, 1, ncclUint32, ncclMax, c->nccl, st));
        // 3. exchange 2: the unit-scale Gaussian (:133) reads 3 rows across each seam
        const size_t hb = (size_t)BAND_HALO * w;
        STEPTH_NCCL_TRY(ctx, nc.GroupStart());
        if (r > 0) {
            STEPTH_NCCL_TRY(ctx, nc.Send(d8, hb, ncclUint8, r - 1, c->nccl, st));
            STEPTH_NCCL_TRY(ctx, nc.Recv(c->d8h.p, hb, ncclUint8, r - 1, c->nccl, st));
        }
        if (r < n - 1) {
            STEPTH_NCCL_TRY(ctx, nc.Send(d8 + (size_t)(rows - BAND_HALO) * w, hb, ncclUint8, r + 1, c->nccl, st));
            STEPTH_NCCL_TRY(ctx, nc.Recv(d8 + (size_t)rows * w, hb, ncclUint8, r + 1, c->nccl, st));
        }
        STEPTH_NCCL_TRY(ctx, nc.GroupEnd());
    }
    // 4. normalise + Gaussian (+ invert, + histogram) of the band, halo rows as source only
    const uint32_t fin_flags = flags & (STEPTH_RESAMPLE_MASK | STEPTH_FLAG_INVERT);
    if ((rc = stepth_cuda_dev_finish(ctx, st, 1, w, h, d8 - (size_t)above * w, y0 - above, above + rows + below, d_max, fin_flags,
                                     depth_band_out, y0, rows, want_fg ? d_hist : nullptr, d_status)) != STEPTH_OK)
        return rc;
    if (want_fg) {
        // 5. exchange 3: depth_split runs on the histogram of the whole image (depth_image.rs:162-218); every rank then
        // computes the same clusters
        if (n > 1) STEPTH_NCCL_TRY(ctx, nc.AllReduce(d_hist, d_hist, 256, ncclUint32, ncclSum, c->nccl, st));
        if ((rc = stepth_cuda_dev_depth_split(ctx, st, 1, d_hist, 2, d_ranges, d_nout, d_status)) != STEPTH_OK) return rc;
        if (mask_band_out || want_apply)
            if ((rc = stepth_cuda_dev_mask_apply(ctx, st, 1, (size_t)rows * w, depth_band_out, d_ranges, 512, d_status,
                                                 want_apply ? main_band : nullptr, mask_band_out,
                                                 want_apply ? rgba_band_out : nullptr)) != STEPTH_OK)
                return rc;
    }
    if (depth_full_out) {
        // 6. exchange 4 (optional): the assembled map on every GPU -- one broadcast per band, grouped (bands may differ in height)
        if (n > 1) {
            STEPTH_NCCL_TRY(ctx, nc.GroupStart());
            for (int q = 0; q < n; ++q) {
                uint32_t pq[4];
                band_plan(h, add_h, params->max_rings, (uint32_t)n, (uint32_t)q, pq);
                if (pq[1] > pq[0])
                    STEPTH_NCCL_TRY(ctx, nc.Broadcast(depth_band_out, depth_full_out + (size_t)pq[0] * w, (size_t)(pq[1] - pq[0]) * w,
                                                      ncclUint8, q, c->nccl, st));
            }
            STEPTH_NCCL_TRY(ctx, nc.GroupEnd());
        } else if (depth_full_out != depth_band_out) {
            STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(depth_full_out, depth_band_out, (size_t)rows * w, cudaMemcpyDeviceToDevice, st));
        }
    }
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(c->h_small, c->small.p, stepth_comm::SMALL_BYTES, cudaMemcpyDeviceToHost, st));
    return STEPTH_OK;
}

extern "C" int32_t stepth_cuda_comm_band_result(stepth_comm *c, void *stream, int32_t *status_out, uint8_t *from_to_out) {
    if (!c || !c->ctx) return STEPTH_ERR_INVALID_ARG;
    if (!c->h_small) return set_err(c->ctx, STEPTH_ERR_INVALID_ARG, "comm_band_result: no band call has been enqueued");
    STEPTH_CUDA_TRY(c->ctx, cudaSetDevice(c->ctx->device));
    STEPTH_CUDA_TRY(c->ctx, cudaStreamSynchronize((cudaStream_t)stream));
    const uint32_t *sm = (const uint32_t *)c->h_small;
    // max == 0 is a property of the whole image: every rank holds the same all-reduced value
    const int32_t st = sm[0] == 0 ? STEPTH_ERR_DIV_BY_ZERO_MAX : (int32_t)sm[1];
    if (status_out) *status_out = st;
    if (from_to_out) {
        from_to_out[0] = ((const uint8_t *)(sm + 260))[0];
        from_to_out[1] = ((const uint8_t *)(sm + 260))[1];
    }
    return STEPTH_OK;
}
