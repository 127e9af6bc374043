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

// ---- peer windows (peer.cuh) ---------------------------------------------------------------------------------------------------
namespace {
struct PeerInfo {   // what the ranks tell each other about their windows (bootstrap, over NCCL)
    uint64_t pid, ptr;
    int32_t device, ok;
    cudaIpcMemHandle_t handle;
};
static_assert(sizeof(PeerInfo) <= 128, "bootstrap record");

void comm_close_peers(stepth_comm *c) {
    for (int q = 0; q < 8; ++q) {
        if (c->peer_opened[q] && c->peer_base[q]) cudaIpcCloseMemHandle(c->peer_base[q]);
        c->peer_opened[q] = false;
        c->peer_base[q] = nullptr;
    }
}

// window geometry for an image of w x h in n row bands (identical on every rank)
struct WinPlan {
    size_t halo_zone, off_d8h, off_full, bytes;
};
WinPlan win_plan(uint32_t w, uint32_t h, uint32_t add_h, uint32_t max_rings, int n, bool full) {
    auto up = [](size_t b) { return (b + 255) & ~(size_t)255; };
    uint32_t rows_max = 0;
    for (int q = 0; q < n; ++q) {
        uint32_t pq[4];
        band_plan(h, add_h, max_rings, (uint32_t)n, (uint32_t)q, pq);
        rows_max = std::max(rows_max, pq[1] - pq[0]);
    }
    WinPlan p;
    p.halo_zone = ((size_t)BAND_HALO * w + 15) & ~(size_t)15;
    p.off_d8h = up(PW_OFF_HALO + 4 * p.halo_zone);
    p.off_full = up(p.off_d8h + (size_t)(rows_max + 2 * BAND_HALO) * w);
    p.bytes = p.off_full + (full ? up((size_t)w * h) : 0);
    return p;
}

// Makes sure every rank owns (and has mapped everybody's) window of at least plan.bytes.  COLLECTIVE: every rank of the group
// calls it with the same plan.  On a box without peer access between the GPUs peer_mode becomes 0 and the caller uses NCCL.
int32_t comm_ensure_window(stepth_comm *c, cudaStream_t st, const WinPlan &plan, bool full) {
    stepth_ctx *ctx = c->ctx;
    const int n = c->n_ranks, r = c->rank;
    if (c->peer_mode == 0) return STEPTH_OK;
    if (c->win && c->win_bytes >= plan.bytes && (c->win_has_full || !full) && c->off_d8h == plan.off_d8h &&
        (!full || c->off_full == plan.off_full)) {
        c->halo_zone = plan.halo_zone;
        return STEPTH_OK;
    }
    NcclApi &nc = nccl_api();
    int32_t rc;
    if ((rc = ensure(ctx, c->xchg, 128 * 9 + 64)) != STEPTH_OK) return rc;
    uint8_t *x_send = (uint8_t *)c->xchg.p, *x_recv = x_send + 128;
    int *x_flag = (int *)(x_send + 128 * 9);
    // nobody frees a window a peer still has mapped: close, meet, free
    STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(st));
    comm_close_peers(c);
    STEPTH_CUDA_TRY(ctx, cudaMemsetAsync(x_flag, 0, sizeof(int), st));
    STEPTH_NCCL_TRY(ctx, nc.AllReduce(x_flag, x_flag, 1, ncclInt32, ncclSum, c->nccl, st));
    STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(st));
    if (c->win) STEPTH_CUDA_TRY(ctx, cudaFree(c->win));
    c->win = nullptr;
    c->win_bytes = 0;
    STEPTH_CUDA_TRY(ctx, cudaMalloc(&c->win, plan.bytes));
    STEPTH_CUDA_TRY(ctx, cudaMemsetAsync(c->win, 0, PW_OFF_HALO, st));
    PeerInfo me;
    memset(&me, 0, sizeof me);
    me.pid = (uint64_t)getpid();
    me.ptr = (uint64_t)(uintptr_t)c->win;
    me.device = ctx->device;
    me.ok = cudaIpcGetMemHandle(&me.handle, c->win) == cudaSuccess ? 1 : 0;
    (void)cudaGetLastError();
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(x_send, &me, sizeof me, cudaMemcpyHostToDevice, st));
    STEPTH_NCCL_TRY(ctx, nc.AllGather(x_send, x_recv, 128, ncclUint8, c->nccl, st));
    std::vector<uint8_t> all((size_t)128 * n);
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(all.data(), x_recv, all.size(), cudaMemcpyDeviceToHost, st));
    STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(st));
    int ok = 1;
    for (int q = 0; q < n; ++q) {
        PeerInfo pi;
        memcpy(&pi, all.data() + (size_t)128 * q, sizeof pi);
        if (q == r) {
            c->peer_base[q] = (uint8_t *)c->win;
        } else if (pi.pid == me.pid) {   // a thread of this process drives that GPU (stepth_multi): plain peer access
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, ctx->device, pi.device) != cudaSuccess || !can) ok = 0;
            else {
                const cudaError_t e = cudaDeviceEnablePeerAccess(pi.device, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) ok = 0;
                (void)cudaGetLastError();
                c->peer_base[q] = (uint8_t *)(uintptr_t)pi.ptr;
            }
        } else {
            void *p = nullptr;
            if (!pi.ok || cudaIpcOpenMemHandle(&p, pi.handle, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                ok = 0;
                (void)cudaGetLastError();
            } else {
                c->peer_base[q] = (uint8_t *)p;
                c->peer_opened[q] = true;
            }
        }
    }
    // all or nobody
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(x_flag, &ok, sizeof(int), cudaMemcpyHostToDevice, st));
    STEPTH_NCCL_TRY(ctx, nc.AllReduce(x_flag, x_flag, 1, ncclInt32, ncclMin, c->nccl, st));
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(&ok, x_flag, sizeof(int), cudaMemcpyDeviceToHost, st));
    STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(st));
    if (!ok) {
        comm_close_peers(c);
        STEPTH_CUDA_TRY(ctx, cudaMemsetAsync(x_flag, 0, sizeof(int), st));   // the owners free only after everybody has let go
        STEPTH_NCCL_TRY(ctx, nc.AllReduce(x_flag, x_flag, 1, ncclInt32, ncclSum, c->nccl, st));
        STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(st));
        cudaFree(c->win);
        c->win = nullptr;
        c->peer_mode = 0;
        return STEPTH_OK;
    }
    c->peer_mode = 1;
    c->win_bytes = plan.bytes;
    c->win_has_full = full;
    c->off_d8h = plan.off_d8h;
    c->off_full = plan.off_full;
    c->halo_zone = plan.halo_zone;
    c->seq = 0;   // the new windows' flags start at zero on every rank
    return STEPTH_OK;
}

PeerView comm_view(const stepth_comm *c) {
    PeerView v;
    memset(&v, 0, sizeof v);
    for (int q = 0; q < c->n_ranks && q < 8; ++q) v.base[q] = c->peer_base[q];
    v.n = c->n_ranks;
    v.rank = c->rank;
    return v;
}
}  // namespace

extern "C" void stepth_cuda_comm_destroy(stepth_comm *c) {
    if (!c) return;
    if (c->ctx) cudaSetDevice(c->ctx->device);
    comm_close_peers(c);
    if (c->win) cudaFree(c->win);
    if (c->xchg.p) cudaFree(c->xchg.p);
    if (c->nccl && c->owns_nccl && nccl_api().lib) nccl_api().CommDestroy(c->nccl);
    if (c->d8h.p) cudaFree(c->d8h.p);
    if (c->small.p) cudaFree(c->small.p);
    if (c->h_small) cudaFreeHost(c->h_small);
    for (cudaEvent_t e : c->marks)
        if (e) cudaEventDestroy(e);
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
    // the exchanges: stores into the peers' windows over NVLink (peer.cuh) when every GPU of the group reaches every other, NCCL
    // otherwise (and with STEPTH_PEER=0)
    bool peer = n > 1 && n <= PW_MAX_RANKS && ctx->peer_enabled && c->peer_mode != 0;
    if (peer) {
        const WinPlan plan = win_plan(w, h, add_h, params->max_rings, n, depth_full_out != nullptr || c->win_has_full);
        if ((rc = comm_ensure_window(c, st, plan, depth_full_out != nullptr || c->win_has_full)) != STEPTH_OK) return rc;
        peer = c->peer_mode == 1;
    }
    if (!peer && (rc = ensure(ctx, c->d8h, (size_t)(rows + 2 * BAND_HALO) * w)) != STEPTH_OK) return rc;
    if ((rc = ensure(ctx, c->small, stepth_comm::SMALL_BYTES)) != STEPTH_OK) return rc;
    if (!c->h_small) STEPTH_CUDA_TRY(ctx, cudaMallocHost(&c->h_small, stepth_comm::SMALL_BYTES));
    uint32_t *sm = (uint32_t *)c->small.p;
    uint32_t *d_max = sm, *d_nout = sm + 2, *d_hist = sm + 4;
    int32_t *d_status = (int32_t *)(sm + 1);
    uint8_t *d_ranges = (uint8_t *)(sm + 260);
    for (cudaEvent_t &e : c->marks)
        if (!e) STEPTH_CUDA_TRY(ctx, cudaEventCreate(&e));
    STEPTH_CUDA_TRY(ctx, cudaEventRecord(c->marks[0], st));
    STEPTH_CUDA_TRY(ctx, cudaMemsetAsync(c->small.p, 0, stepth_comm::SMALL_BYTES, st));
    uint8_t *d8h = peer ? (uint8_t *)c->win + c->off_d8h : (uint8_t *)c->d8h.p;
    uint8_t *d8 = d8h + (size_t)BAND_HALO * w;
    const PeerView pv = comm_view(c);
    if (peer) ++c->seq;

    // 1. ring search of this band (helpers.rs:9-54, depth_image.rs:113-122); the per-band maximum lands in d_max
    stepth_dev_shape s;
    memset(&s, 0, sizeof s);
    s.n_pairs = 1; s.w = w; s.h = rows; s.img_h = h; s.y0 = y0;
    s.wa = wa; s.ha = ay1 - ay0; s.add_y0 = ay0; s.add_img_h = add_h;
    if ((rc = stepth_cuda_dev_search_dense(ctx, st, &s, params, main_band, add_band, d8, d_max)) != STEPTH_OK) return rc;
    STEPTH_CUDA_TRY(ctx, cudaEventRecord(c->marks[1], st));
    const uint32_t above = r > 0 ? std::min(BAND_HALO, y0) : 0u, below = r < n - 1 ? std::min(BAND_HALO, h - y1) : 0u;
    if (peer) {
        // 2 + 3. one kernel: my maximum into every mailbox, my first / last three rows into the neighbours' halo zones, flags,
        // the maximum over all bands (depth_image.rs:124), the neighbours' rows next to my band (:133)
        PeerExchangeArgs xa;
        xa.pv = pv; xa.seq = c->seq; xa.halo_zone = c->halo_zone; xa.halo_bytes = (size_t)BAND_HALO * w;
        xa.band = d8; xa.rows = rows; xa.w = w; xa.max_io = d_max; xa.timed_out = sm + 3;
        peer_exchange_kernel<<<8, 256, 0, st>>>(xa);
        ctx->launches++;
        STEPTH_CUDA_TRY(ctx, cudaGetLastError());
    } else if (n > 1) {
        // 2. exchange 1: the normaliser is the maximum over the WHOLE image (depth_image.rs:124)
        STEPTH_NCCL_TRY(ctx, nc.AllReduce(d_max, d_max, 1, ncclUint32, ncclMax, c->nccl, st));
        // 3. exchange 2: the unit-scale Gaussian (:133) reads 3 rows across each seam
        const size_t hb = (size_t)BAND_HALO * w;
        STEPTH_NCCL_TRY(ctx, nc.GroupStart());
        if (r > 0) {
            STEPTH_NCCL_TRY(ctx, nc.Send(d8, hb, ncclUint8, r - 1, c->nccl, st));
            STEPTH_NCCL_TRY(ctx, nc.Recv(d8h, hb, ncclUint8, r - 1, c->nccl, st));
        }
        if (r < n - 1) {
            STEPTH_NCCL_TRY(ctx, nc.Send(d8 + (size_t)(rows - BAND_HALO) * w, hb, ncclUint8, r + 1, c->nccl, st));
            STEPTH_NCCL_TRY(ctx, nc.Recv(d8 + (size_t)rows * w, hb, ncclUint8, r + 1, c->nccl, st));
        }
        STEPTH_NCCL_TRY(ctx, nc.GroupEnd());
    }
    STEPTH_CUDA_TRY(ctx, cudaEventRecord(c->marks[2], st));
    // 4. normalise + Gaussian (+ invert, + histogram) of the band, halo rows as source only
    const uint32_t fin_flags = flags & (STEPTH_RESAMPLE_MASK | STEPTH_FLAG_INVERT);
    const bool peer_gather = peer && depth_full_out != nullptr;
    if (peer_gather) {   // 6 rides on 4: the pass that produces the depth band stores it into every rank's assembled map
        for (int q = 0; q < n; ++q) ctx->finish_mirror[q] = pv.base[q] + c->off_full + (size_t)y0 * w;
        ctx->finish_n_mirror = (uint32_t)n;
    }
    ctx->finish_mirrored = 0;
    rc = stepth_cuda_dev_finish(ctx, st, 1, w, h, d8 - (size_t)above * w, y0 - above, above + rows + below, d_max, fin_flags,
                                depth_band_out, y0, rows, want_fg ? d_hist : nullptr, d_status);
    ctx->finish_n_mirror = 0;
    const bool mirrored = ctx->finish_mirrored != 0;
    ctx->finish_mirrored = 0;
    if (rc != STEPTH_OK) return rc;
    STEPTH_CUDA_TRY(ctx, cudaEventRecord(c->marks[3], st));
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
    if (peer_gather) {
        // 6. exchange 4: the bands are already on their way (mirror stores of the finish pass; a copy kernel when that pass could
        // not mirror); one flag round later every band has landed in every rank's window
        if (!mirrored) {
            peer_scatter_kernel<<<(unsigned)ctx->sm_count, 256, 0, st>>>(pv, c->off_full + (size_t)y0 * w, depth_band_out, (size_t)rows * w);
            ctx->launches++;
            STEPTH_CUDA_TRY(ctx, cudaGetLastError());
        }
        peer_barrier_kernel<<<1, 32, 0, st>>>(pv, 1, c->seq, sm + 3);
        ctx->launches++;
        STEPTH_CUDA_TRY(ctx, cudaGetLastError());
        uint8_t *own_full = (uint8_t *)c->win + c->off_full;
        if (depth_full_out != own_full)
            STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(depth_full_out, own_full, (size_t)w * h, cudaMemcpyDeviceToDevice, st));
    } else if (depth_full_out) {
        // 6. exchange 4 (optional): the assembled map on every GPU -- one broadcast per band, grouped (bands may differ in height)
        bool equal = true;   // equal bands are consecutive slabs of the assembled map: one all-gather straight into it
        for (int q = 0; q < n; ++q) {
            uint32_t pq[4];
            band_plan(h, add_h, params->max_rings, (uint32_t)n, (uint32_t)q, pq);
            equal = equal && pq[1] - pq[0] == rows;
        }
        if (n > 1 && equal) {
            STEPTH_NCCL_TRY(ctx, nc.AllGather(depth_band_out, depth_full_out, (size_t)rows * w, ncclUint8, c->nccl, st));
        } else if (n > 1) {
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
    STEPTH_CUDA_TRY(ctx, cudaEventRecord(c->marks[4], st));
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(c->h_small, c->small.p, stepth_comm::SMALL_BYTES, cudaMemcpyDeviceToHost, st));
    return STEPTH_OK;
}

// 1 when the band exchanges of this group go through peer windows over NVLink (peer.cuh), 0 when through NCCL, -1 before the
// first multi-rank band call has decided
extern "C" int32_t stepth_cuda_comm_peer_mode(const stepth_comm *c) { return c ? c->peer_mode : -1; }

// The assembled depth map inside this rank's window (device pointer, w * h bytes, valid until the group is destroyed or a larger
// image is banded): a caller that hands it to stepth_cuda_dev_depth_dense_band as depth_full_out saves the copy out of the
// window.  COLLECTIVE (every rank calls it with the same arguments); *ptr_out = NULL when the group runs over NCCL.
extern "C" int32_t stepth_cuda_comm_full_map(stepth_comm *c, void *stream, uint32_t w, uint32_t h, uint32_t add_h, uint32_t max_rings,
                                             uint8_t **ptr_out) {
    if (!c || !c->ctx) return STEPTH_ERR_INVALID_ARG;
    if (!ptr_out || w == 0 || h == 0 || add_h == 0 || max_rings == 0) return set_err(c->ctx, STEPTH_ERR_INVALID_ARG, "comm_full_map: invalid argument");
    *ptr_out = nullptr;
    stepth_ctx *ctx = c->ctx;
    STEPTH_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    if (c->n_ranks < 2 || c->n_ranks > PW_MAX_RANKS || !ctx->peer_enabled || c->peer_mode == 0) return STEPTH_OK;
    NcclApi &nc = nccl_api();
    if (!nc.lib) return set_err(ctx, STEPTH_ERR_NCCL, "%s", nc.why.c_str());
    const WinPlan plan = win_plan(w, h, add_h, max_rings, c->n_ranks, true);
    const int32_t rc = comm_ensure_window(c, (cudaStream_t)stream, plan, true);
    if (rc != STEPTH_OK) return rc;
    if (c->peer_mode == 1) *ptr_out = (uint8_t *)c->win + c->off_full;
    return STEPTH_OK;
}

// milliseconds the last band call spent in: search | MAX all-reduce + halo exchange (the wait for the slowest band shows up
// here) | finish | tail + gather.  Call after stepth_cuda_comm_band_result.
extern "C" int32_t stepth_cuda_comm_band_stage_ms(stepth_comm *c, float *ms_out) {
    if (!c || !c->ctx) return STEPTH_ERR_INVALID_ARG;
    if (!ms_out || !c->marks[0]) return set_err(c->ctx, STEPTH_ERR_INVALID_ARG, "comm_band_stage_ms: no band call has been made");
    for (int i = 0; i < 4; ++i) STEPTH_CUDA_TRY(c->ctx, cudaEventElapsedTime(&ms_out[i], c->marks[i], c->marks[i + 1]));
    return STEPTH_OK;
}

extern "C" int32_t stepth_cuda_comm_band_result(stepth_comm *c, void *stream, int32_t *status_out, uint8_t *from_to_out) {
    if (!c || !c->ctx) return STEPTH_ERR_INVALID_ARG;
    if (!c->h_small) return set_err(c->ctx, STEPTH_ERR_INVALID_ARG, "comm_band_result: no band call has been enqueued");
    STEPTH_CUDA_TRY(c->ctx, cudaSetDevice(c->ctx->device));
    STEPTH_CUDA_TRY(c->ctx, cudaStreamSynchronize((cudaStream_t)stream));
    const uint32_t *sm = (const uint32_t *)c->h_small;
    if (sm[3]) return set_err(c->ctx, STEPTH_ERR_NCCL, "row-band exchange: a peer's flag did not arrive within 20 s (peer.cuh)");
    // max == 0 is a property of the whole image: every rank holds the same all-reduced value
    const int32_t st = sm[0] == 0 ? STEPTH_ERR_DIV_BY_ZERO_MAX : (int32_t)sm[1];
    if (status_out) *status_out = st;
    if (from_to_out) {
        from_to_out[0] = ((const uint8_t *)(sm + 260))[0];
        from_to_out[1] = ((const uint8_t *)(sm + 260))[1];
    }
    return STEPTH_OK;
}

// ============================================================================================================
// stepth_multi: one process, several GPUs (a context per device, a host thread per device per call)
// ============================================================================================================
struct stepth_multi {
    std::vector<int> devices;
    std::vector<stepth_ctx *> ctxs;
    std::vector<stepth_comm *> comms;   // row-band group over all the devices
    std::string err;
};

static thread_local std::string g_multi_err;

extern "C" void stepth_cuda_multi_destroy(stepth_multi *m) {
    if (!m) return;
    for (auto *c : m->comms) stepth_cuda_comm_destroy(c);
    for (auto *c : m->ctxs) stepth_cuda_ctx_destroy(c);
    delete m;
}

extern "C" const char *stepth_cuda_multi_last_error(const stepth_multi *m) { return m ? m->err.c_str() : g_multi_err.c_str(); }

extern "C" int32_t stepth_cuda_multi_create(const int32_t *devices, uint32_t n_devices, stepth_multi **multi_out) {
    if (!multi_out) return STEPTH_ERR_INVALID_ARG;
    *multi_out = nullptr;
    if (!devices || n_devices == 0 || n_devices > 64) {
        g_multi_err = "multi_create: null or empty device list";
        return STEPTH_ERR_INVALID_ARG;
    }
    for (uint32_t i = 0; i < n_devices; ++i)
        for (uint32_t j = 0; j < i; ++j)
            if (devices[i] == devices[j]) {
                g_multi_err = "multi_create: a device is listed twice";
                return STEPTH_ERR_INVALID_ARG;
            }
    stepth_multi *m = new stepth_multi();
    for (uint32_t i = 0; i < n_devices; ++i) {
        stepth_ctx *c = nullptr;
        const int32_t rc = stepth_cuda_ctx_create(devices[i], &c);
        if (rc != STEPTH_OK) {
            g_multi_err = stepth_cuda_last_error(nullptr);
            stepth_cuda_multi_destroy(m);
            return rc;
        }
        m->devices.push_back(devices[i]);
        m->ctxs.push_back(c);
    }
    std::vector<ncclComm_t> nc(n_devices, nullptr);
    if (n_devices > 1) {
        NcclApi &n = nccl_api();
        if (!n.lib) {
            g_multi_err = n.why;
            stepth_cuda_multi_destroy(m);
            return STEPTH_ERR_NCCL;
        }
        const ncclResult_t r = n.CommInitAll(nc.data(), (int)n_devices, m->devices.data());
        if (r != ncclSuccess) {
            g_multi_err = std::string("ncclCommInitAll failed: ") + n.GetErrorString(r);
            stepth_cuda_multi_destroy(m);
            return STEPTH_ERR_NCCL;
        }
    }
    for (uint32_t i = 0; i < n_devices; ++i) {
        stepth_comm *c = nullptr;
        comm_alloc(m->ctxs[i], (int32_t)n_devices, (int32_t)i, nc[i], &c);
        m->comms.push_back(c);
    }
    *multi_out = m;
    return STEPTH_OK;
}

extern "C" uint64_t stepth_cuda_multi_launch_count(const stepth_multi *m) {
    uint64_t n = 0;
    if (m)
        for (auto *c : m->ctxs) n += c->launches;
    return n;
}

// configs[3]: independent pairs dealt to the devices (pair i -> device i mod n), each device's share through the single-GPU
// batch call on its own context and host thread: no collective anywhere on the data path
extern "C" int32_t stepth_cuda_multi_depth_dense_batch(stepth_multi *m, uint32_t n_pairs, const uint8_t *const *main_rgba,
                                                       uint32_t w, uint32_t h, const uint8_t *const *add_rgb, uint32_t wa,
                                                       uint32_t ha, const stepth_params *params, uint8_t *const *depth_out,
                                                       uint8_t *const *rgba_out, uint8_t *const *mask_out, uint8_t *from_to_out,
                                                       int32_t *status_out) {
    if (!m) return STEPTH_ERR_INVALID_ARG;
    if (!main_rgba || !add_rgb || !depth_out || !status_out || n_pairs == 0) {
        m->err = "multi_depth_dense_batch: null or invalid argument";
        return STEPTH_ERR_INVALID_ARG;
    }
    const size_t nd = m->ctxs.size();
    std::vector<int32_t> rcs(nd, STEPTH_OK);
    std::vector<std::thread> th;
    for (size_t d = 0; d < nd; ++d) {
        th.emplace_back([&, d] {
            std::vector<const uint8_t *> pm, pa;
            std::vector<uint8_t *> pd, pr, pk;
            std::vector<uint32_t> idx;
            for (uint32_t i = (uint32_t)d; i < n_pairs; i += (uint32_t)nd) {
                idx.push_back(i);
                pm.push_back(main_rgba[i]);
                pa.push_back(add_rgb[i]);
                pd.push_back(depth_out[i]);
                if (rgba_out) pr.push_back(rgba_out[i]);
                if (mask_out) pk.push_back(mask_out[i]);
            }
            if (idx.empty()) return;
            std::vector<int32_t> st(idx.size(), 0);
            std::vector<uint8_t> ft(2 * idx.size(), 0);
            rcs[d] = stepth_cuda_depth_dense_batch(m->ctxs[d], (uint32_t)idx.size(), pm.data(), w, h, pa.data(), wa, ha, params,
                                                   pd.data(), rgba_out ? pr.data() : nullptr, mask_out ? pk.data() : nullptr,
                                                   ft.data(), st.data());
            for (size_t k = 0; k < idx.size(); ++k) {
                status_out[idx[k]] = st[k];
                if (from_to_out) {
                    from_to_out[2 * idx[k]] = ft[2 * k];
                    from_to_out[2 * idx[k] + 1] = ft[2 * k + 1];
                }
            }
        });
    }
    for (auto &t : th) t.join();
    for (size_t d = 0; d < nd; ++d)
        if (rcs[d] != STEPTH_OK) {
            m->err = std::string("device ") + std::to_string(m->devices[d]) + ": " + stepth_cuda_last_error(m->ctxs[d]);
            return rcs[d];
        }
    return STEPTH_OK;
}

// configs[4]: ONE pair, main rows banded over the devices.  Every device uploads its band and the additional rows it can
// reach, runs stepth_cuda_dev_depth_dense_band (NCCL for the three small exchanges) and writes its rows of the results
// straight into the caller's buffers.
extern "C" int32_t stepth_cuda_multi_depth_dense_banded(stepth_multi *m, const uint8_t *main_rgba, uint32_t w, uint32_t h,
                                                        const uint8_t *add_rgb, uint32_t wa, uint32_t ha,
                                                        const stepth_params *params, uint8_t *depth_out, uint8_t *rgba_out,
                                                        uint8_t *mask_out, uint8_t *from_to_out) {
    if (!m) return STEPTH_ERR_INVALID_ARG;
    if (!main_rgba || !add_rgb || !depth_out || !params_ok(params) || w == 0 || h == 0 || wa == 0 || ha == 0) {
        m->err = "multi_depth_dense_banded: null or invalid argument";
        return STEPTH_ERR_INVALID_ARG;
    }
    if ((uint64_t)w * h < 8) {
        m->err = "fewer than 8 pixels: par_chunks_mut(0) panics";
        return STEPTH_ERR_TOO_FEW_BLOCKS;
    }
    const uint32_t flags = params->flags;
    const bool want_fg = (flags & STEPTH_FLAG_FOREGROUND) != 0, want_apply = (flags & STEPTH_FLAG_APPLY) != 0;
    if (want_apply && (!want_fg || !rgba_out)) {
        m->err = "APPLY needs FOREGROUND and rgba_out";
        return STEPTH_ERR_INVALID_ARG;
    }
    const size_t nd = m->ctxs.size();
    // only the columns a query can reach are uploaded (origins are half-coordinates: depth_image.rs:114-117)
    const uint32_t wa_d = std::min(wa, w / 2 + params->max_rings);
    std::vector<int32_t> rcs(nd, STEPTH_OK), status(nd, 0);
    std::vector<uint8_t> ft(2 * nd, 0);
    std::vector<std::thread> th;
    for (size_t d = 0; d < nd; ++d) {
        th.emplace_back([&, d] {
            stepth_ctx *ctx = m->ctxs[d];
            stepth_comm *c = m->comms[d];
            auto run = [&]() -> int32_t {
                STEPTH_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
                cudaStream_t st = ctx->stream;
                uint32_t pl[4];
                band_plan(h, ha, params->max_rings, (uint32_t)nd, (uint32_t)d, pl);
                const uint32_t y0 = pl[0], rows = pl[1] - pl[0], ay0 = pl[2], arows = pl[3] - pl[2];
                const size_t px = (size_t)rows * w;
                int32_t rc;
                if ((rc = ensure(ctx, ctx->main_rgba, px * 4 + 16)) != STEPTH_OK) return rc;
                if ((rc = ensure(ctx, ctx->add_rgb, (size_t)arows * wa_d * 3 + 16)) != STEPTH_OK) return rc;
                if ((rc = ensure(ctx, ctx->depth, px + 16)) != STEPTH_OK) return rc;
                if (want_fg && mask_out && (rc = ensure(ctx, ctx->mask, px + 16)) != STEPTH_OK) return rc;
                if (want_apply && (rc = ensure(ctx, ctx->rgba_out, px * 4 + 16)) != STEPTH_OK) return rc;
                if (arows) {
                    if (wa_d == wa)
                        STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->add_rgb.p, add_rgb + (size_t)ay0 * wa * 3, (size_t)arows * wa * 3,
                                                             cudaMemcpyHostToDevice, st));
                    else
                        STEPTH_CUDA_TRY(ctx, cudaMemcpy2DAsync(ctx->add_rgb.p, (size_t)wa_d * 3, add_rgb + (size_t)ay0 * wa * 3,
                                                               (size_t)wa * 3, (size_t)wa_d * 3, arows, cudaMemcpyHostToDevice, st));
                }
                if (rows) STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->main_rgba.p, main_rgba + (size_t)y0 * w * 4, px * 4, cudaMemcpyHostToDevice, st));
                if ((rc = stepth_cuda_dev_depth_dense_band(c, st, w, h, wa_d, ha, params, (const uint8_t *)ctx->main_rgba.p,
                                                           (const uint8_t *)ctx->add_rgb.p, (uint8_t *)ctx->depth.p, nullptr,
                                                           want_fg && mask_out ? (uint8_t *)ctx->mask.p : nullptr,
                                                           want_apply ? (uint8_t *)ctx->rgba_out.p : nullptr)) != STEPTH_OK)
                    return rc;
                if (rows) {
                    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(depth_out + (size_t)y0 * w, ctx->depth.p, px, cudaMemcpyDeviceToHost, st));
                    if (want_fg && mask_out)
                        STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(mask_out + (size_t)y0 * w, ctx->mask.p, px, cudaMemcpyDeviceToHost, st));
                    if (want_apply)
                        STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(rgba_out + (size_t)y0 * w * 4, ctx->rgba_out.p, px * 4, cudaMemcpyDeviceToHost, st));
                }
                return stepth_cuda_comm_band_result(c, st, &status[d], &ft[2 * d]);
            };
            rcs[d] = run();
        });
    }
    for (auto &t : th) t.join();
    for (size_t d = 0; d < nd; ++d)
        if (rcs[d] != STEPTH_OK) {
            m->err = std::string("device ") + std::to_string(m->devices[d]) + ": " + stepth_cuda_last_error(m->ctxs[d]);
            return rcs[d];
        }
    // the status is a property of the whole image (max == 0, degenerate depth range): every device reports the same one
    if (status[0] != STEPTH_OK) {
        m->err = status[0] == STEPTH_ERR_DIV_BY_ZERO_MAX ? "max distance is 0: attempt to divide by zero (depth_image.rs:127)"
                                                          : "depth_split failed on the depth range (depth_image.rs:212-216)";
        return status[0];
    }
    if (want_fg && from_to_out) {
        from_to_out[0] = ft[0];
        from_to_out[1] = ft[1];
    }
    return STEPTH_OK;
}
