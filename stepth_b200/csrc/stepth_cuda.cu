// stepth_cuda.cu -- libstepth_cuda.so: the C ABI declared in include/stepth_cuda.h.
//
// Layout of this translation unit:
//   1. context                                     stepth_cuda_ctx_* / last_error / version / launch_count
//   2. device-resident stage API (stepth_cuda_dev_*): enqueue kernels on a caller stream, device pointers
//   3. host-buffer entry points (what the patched `stepth` crate binds): H2D -> stages -> D2H, synchronous
//   4. host-side stand-in for the disage discretiser
// There is no CPU fallback anywhere: without a usable sm_100 device every entry point returns STEPTH_ERR_CUDA.
#include "ctx.cuh"
#include "search_literal.cuh"
#include "search_launch.cuh"
#include "tail.cuh"
#include "brightness.cuh"
#include "intpeak.cuh"
#include "edit.cuh"
#include "disc.cuh"
#include "codec.cuh"

#include <sched.h>

#include <algorithm>

using namespace stepth;

static thread_local std::string g_create_err;

#define LAUNCH_CHECK(ctx)                                                                                   \
    do {                                                                                                    \
        (ctx)->launches++;                                                                                  \
        STEPTH_CUDA_TRY((ctx), cudaGetLastError());                                                         \
    } while (0)

// ============================================================================================================
// 1. context
// ============================================================================================================
extern "C" int32_t stepth_cuda_device_count(int32_t *count_out) {
    if (!count_out) return STEPTH_ERR_INVALID_ARG;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        g_create_err = std::string("cudaGetDeviceCount failed: ") + cudaGetErrorString(e);
        *count_out = 0;
        return STEPTH_ERR_CUDA;
    }
    *count_out = n;
    return STEPTH_OK;
}

extern "C" void stepth_cuda_ctx_destroy(stepth_ctx *ctx);

extern "C" int32_t stepth_cuda_ctx_create(int32_t device, stepth_ctx **ctx_out) {
    if (!ctx_out) return STEPTH_ERR_INVALID_ARG;
    *ctx_out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        g_create_err = std::string("no CUDA device (there is no CPU fallback): ") + cudaGetErrorString(e);
        return STEPTH_ERR_CUDA;
    }
    if (device < 0 || device >= n) {
        g_create_err = "device index out of range";
        return STEPTH_ERR_INVALID_ARG;
    }
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) {
        g_create_err = std::string("cudaGetDeviceProperties failed: ") + cudaGetErrorString(e);
        return STEPTH_ERR_CUDA;
    }
    if (prop.major != 10) {
        char buf[160];
        snprintf(buf, sizeof buf, "device %d is sm_%d%d; this library only carries sm_100a code", device, prop.major,
                 prop.minor);
        g_create_err = buf;
        return STEPTH_ERR_CUDA;
    }
    if ((e = cudaSetDevice(device)) != cudaSuccess) {
        g_create_err = std::string("cudaSetDevice failed: ") + cudaGetErrorString(e);
        return STEPTH_ERR_CUDA;
    }
    stepth_ctx *ctx = new stepth_ctx();
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    if (const char *ev = getenv("STEPTH_NEAR_RINGS")) ctx->near_rings = std::max(1, atoi(ev));  // tuning knobs
    if (const char *ev = getenv("STEPTH_MID_RINGS")) ctx->mid_rings = atoi(ev);
    if (const char *ev = getenv("STEPTH_NEAR_COMPACT")) ctx->near_compact = atoi(ev);
    if (const char *ev = getenv("STEPTH_NEAR_TILE")) ctx->near_tile = atoi(ev);
    if (const char *ev = getenv("STEPTH_OT_RINGS")) ctx->near_rings_otile = atoi(ev);
    if (const char *ev = getenv("STEPTH_OT_NIT")) ctx->otile_nit = atoi(ev);
    if (const char *ev = getenv("STEPTH_OT_WARPS_PER_SM")) ctx->otile_warps_per_sm = std::max(1, atoi(ev));
    if (const char *ev = getenv("STEPTH_BATCH_CHUNKS")) ctx->batch_chunks = std::max(0, atoi(ev));
    if (const char *ev = getenv("STEPTH_TWO_STREAMS")) ctx->two_streams = atoi(ev);
    if (const char *ev = getenv("STEPTH_FINISH_STREAMING")) ctx->finish_streaming = atoi(ev);
    if (const char *ev = getenv("STEPTH_TRACE")) ctx->trace = atoi(ev);
    if (const char *ev = getenv("STEPTH_PEER")) ctx->peer_enabled = atoi(ev);
    if (const char *ev = getenv("STEPTH_K0_TMAP")) ctx->k0_tensor_map = atoi(ev);
    if (const char *ev = getenv("STEPTH_TAIL")) ctx->tail_kernel = atoi(ev);
    if (const char *ev = getenv("STEPTH_UPLOAD_FULL_ROWS")) ctx->upload_full_rows = atoi(ev);
    {   // copy threads behind pageable host buffers: half of the CPUs this process may run on (its affinity mask -- a launcher
        // that binds one process per GPU to the GPU's NUMA neighbourhood narrows it), at most 16 (64 x 4K, pageable: 8 threads
        // 33.5 ms per 16 pairs, 12: 33.3, 16: 29.2)
        unsigned cpus = std::thread::hardware_concurrency();
        cpu_set_t set;
        if (sched_getaffinity(0, sizeof set, &set) == 0 && CPU_COUNT(&set) > 0) cpus = (unsigned)CPU_COUNT(&set);
        // a launcher that runs one process per GPU says so (torchrun: LOCAL_WORLD_SIZE): the processes share the CPUs
        unsigned procs = 1;
        if (const char *ev = getenv("LOCAL_WORLD_SIZE")) procs = (unsigned)std::max(1, atoi(ev));
        ctx->stage_threads = (int)std::max(2u, std::min(16u, cpus / (2 * procs)));
    }
    if (const char *ev = getenv("STEPTH_STAGE_THREADS")) ctx->stage_threads = std::max(0, std::min(64, atoi(ev)));
    if ((e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) {
        g_create_err = std::string("cudaStreamCreate failed: ") + cudaGetErrorString(e);
        delete ctx;
        return STEPTH_ERR_CUDA;
    }
    if (cudaStreamCreateWithFlags(&ctx->copy_in, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&ctx->stream2, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&ctx->copy_out, cudaStreamNonBlocking) != cudaSuccess) {
        g_create_err = "cudaStreamCreate (copy streams) failed";
        stepth_cuda_ctx_destroy(ctx);
        return STEPTH_ERR_CUDA;
    }
    *ctx_out = ctx;
    return STEPTH_OK;
}

extern "C" void stepth_cuda_ctx_destroy(stepth_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    stepth_ctx::Buf *bufs[] = {&ctx->main_rgba, &ctx->add_rgb, &ctx->d8,   &ctx->depth, &ctx->mask, &ctx->rgba_out,
                               &ctx->blocks,    &ctx->plane,   &ctx->small, &ctx->u16a,  &ctx->u16b, &ctx->u16o,
                               &ctx->work,      &ctx->planes, &ctx->worklist, &ctx->planes2, &ctx->worklist2, &ctx->work2,
                               &ctx->main_rgb, &ctx->disc_counts, &ctx->codec_a, &ctx->codec_b, &ctx->dbg};
    for (auto *b : bufs)
        if (b->p) cudaFree(b->p);
    for (auto &kv : ctx->taps)
        if (kv.second.dev) cudaFree(kv.second.dev);
    if (ctx->h_small) cudaFreeHost(ctx->h_small);
    codec_destroy(ctx);
    ctx->stager.destroy();
    for (cudaEvent_t e : ctx->events) cudaEventDestroy(e);
    if (ctx->copy_in) cudaStreamDestroy(ctx->copy_in);
    if (ctx->copy_out) cudaStreamDestroy(ctx->copy_out);
    if (ctx->stream2) cudaStreamDestroy(ctx->stream2);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

extern "C" const char *stepth_cuda_last_error(const stepth_ctx *ctx) {
    return ctx ? ctx->err.c_str() : g_create_err.c_str();
}

extern "C" const char *stepth_cuda_version(void) { return "stepth-b200 0.1.0 (sm_100a)"; }

extern "C" uint64_t stepth_cuda_launch_count(const stepth_ctx *ctx) { return ctx ? ctx->launches : 0; }

// ============================================================================================================
// 2. device-resident stage API
// ============================================================================================================
static bool params_ok(const stepth_params *p) {
    if (!p) return false;
    if (p->max_rings == 0 || p->max_rings > 512) return false;  // ring keys hold s <= 511 (common.cuh)
    if ((p->flags & STEPTH_RESAMPLE_MASK) == 3u) return false;
    return true;
}

static SearchArgs make_search_args(const stepth_params *p) {
    SearchArgs a;
    memset(&a, 0, sizeof a);
    a.max_rings = p->max_rings;
    a.pr = p->precision[0];
    a.pg = p->precision[1];
    a.pb = p->precision[2];
    a.sat = (p->flags & STEPTH_FLAG_PIXELOPPS_SAT) ? 1u : 0u;
    return a;
}

extern "C" int32_t stepth_cuda_dev_search_dense(stepth_ctx *ctx, void *stream, const stepth_dev_shape *s,
                                                const stepth_params *params, const uint8_t *main_rgba,
                                                const uint8_t *add_rgb, uint8_t *d8_out, uint32_t *max_out) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!s || !params_ok(params) || !main_rgba || (!add_rgb && s->ha) || !d8_out || !max_out)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "dev_search_dense: null or invalid argument");
    if (s->n_pairs == 0 || s->w == 0 || s->h == 0 || s->wa == 0 || (s->ha == 0 && s->add_img_h == 0) ||
        (uint64_t)s->w * s->h > 0xffffffffull || ((uintptr_t)main_rgba & 3))
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "dev_search_dense: bad shape or unaligned RGBA plane");
    cudaStream_t st = (cudaStream_t)stream;
    if (s->ha == 0) {
        // a row band whose reachable additional rows lie below the additional image: every origin is out of bounds, ring 0
        // has no in-bounds probe -> None -> 0 (helpers.rs:49-51, depth_image.rs:120), as the unbanded call answers
        if (ctx->main_ready) {
            cudaEvent_t ev = ctx->main_ready;
            ctx->main_ready = nullptr;
            STEPTH_CUDA_TRY(ctx, cudaStreamWaitEvent(st, ev, 0));
        }
        STEPTH_CUDA_TRY(ctx, cudaMemsetAsync(max_out, 0, sizeof(uint32_t) * s->n_pairs, st));
        STEPTH_CUDA_TRY(ctx, cudaMemsetAsync(d8_out, 0, (size_t)s->n_pairs * s->w * s->h, st));
        return STEPTH_OK;
    }
    const bool hier = !(params->flags & STEPTH_FLAG_LITERAL_SEARCH) && params->precision[0] && params->precision[1] && params->precision[2];
    if (ctx->main_ready && !hier) {   // only the tile search has work to do before it needs the main images
        cudaEvent_t ev = ctx->main_ready;
        ctx->main_ready = nullptr;
        STEPTH_CUDA_TRY(ctx, cudaStreamWaitEvent(st, ev, 0));
    }
    STEPTH_CUDA_TRY(ctx, cudaMemsetAsync(max_out, 0, sizeof(uint32_t) * s->n_pairs, st));
    if (params->precision[0] == 0 || params->precision[1] == 0 || params->precision[2] == 0) {
        // `x < 0` is false for every u8: nothing ever matches, every distance is None -> 0   (helpers.rs:35)
        STEPTH_CUDA_TRY(ctx, cudaMemsetAsync(d8_out, 0, (size_t)s->n_pairs * s->w * s->h, st));
        return STEPTH_OK;
    }
    SearchArgs a = make_search_args(params);
    a.main_rgba = main_rgba;
    a.n_queries = s->w * s->h;
    a.w = s->w; a.h = s->h; a.y0 = s->y0;
    a.add_rgb = add_rgb;
    a.wa = s->wa; a.ha = s->ha; a.add_y0 = s->add_y0;
    a.add_img_h = s->add_img_h ? s->add_img_h : s->ha;
    a.d8 = d8_out;
    a.max_out = max_out;
    if (params->flags & STEPTH_FLAG_LITERAL_SEARCH) {
        dim3 grid((a.n_queries + 255) / 256, s->n_pairs);
        search_literal_kernel<<<grid, 256, 0, st>>>(a);
        LAUNCH_CHECK(ctx);
        return STEPTH_OK;
    }
    const int32_t rc = launch_search_hier(ctx, st, a, s->n_pairs);
    if (ctx->main_ready) {   // an error before K1: the wait is still owed to the caller's stream
        cudaStreamWaitEvent(st, ctx->main_ready, 0);
        ctx->main_ready = nullptr;
    }
    return rc;
}

// measurement aid: the in-order kernel with a counter of the probes the reference's scan order needs (every in-bounds pixel
// of a ring once, up to and including the first match) -- what orc_last_unique_probes counts on the CPU
extern "C" int32_t stepth_cuda_dev_count_probes(stepth_ctx *ctx, void *stream, const stepth_dev_shape *s,
                                                const stepth_params *params, const uint8_t *main_rgba,
                                                const uint8_t *add_rgb, uint8_t *d8_out, uint32_t *max_out,
                                                uint64_t *probes_out) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!s || !params_ok(params) || !main_rgba || !add_rgb || !d8_out || !max_out || !probes_out || s->n_pairs == 0 ||
        s->w == 0 || s->h == 0 || s->wa == 0 || s->ha == 0 || (uint64_t)s->w * s->h > 0xffffffffull || ((uintptr_t)main_rgba & 3))
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "dev_count_probes: null or invalid argument");
    cudaStream_t st = (cudaStream_t)stream;
    STEPTH_CUDA_TRY(ctx, cudaMemsetAsync(max_out, 0, sizeof(uint32_t) * s->n_pairs, st));
    STEPTH_CUDA_TRY(ctx, cudaMemsetAsync(probes_out, 0, sizeof(uint64_t) * s->n_pairs, st));
    if (params->precision[0] == 0 || params->precision[1] == 0 || params->precision[2] == 0)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "dev_count_probes: a zero precision has no probes to count here");
    SearchArgs a = make_search_args(params);
    a.main_rgba = main_rgba;
    a.n_queries = s->w * s->h;
    a.w = s->w; a.h = s->h; a.y0 = s->y0;
    a.add_rgb = add_rgb;
    a.wa = s->wa; a.ha = s->ha; a.add_y0 = s->add_y0;
    a.add_img_h = s->add_img_h ? s->add_img_h : s->ha;
    a.d8 = d8_out;
    a.max_out = max_out;
    a.probes_out = reinterpret_cast<unsigned long long *>(probes_out);
    dim3 grid((a.n_queries + 255) / 256, s->n_pairs);
    search_literal_kernel<<<grid, 256, 0, st>>>(a);
    LAUNCH_CHECK(ctx);
    return STEPTH_OK;
}

extern "C" int32_t stepth_cuda_dev_search_blocks(stepth_ctx *ctx, void *stream, const stepth_params *params,
                                                 const uint8_t *add_rgb, uint32_t wa, uint32_t ha,
                                                 const stepth_block *blocks, size_t n_blocks, uint8_t *d8_out,
                                                 uint32_t *max_out) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!params_ok(params) || !add_rgb || !blocks || !d8_out || !max_out || wa == 0 || ha == 0 ||
        n_blocks > 0xffffffffull || ((uintptr_t)blocks & 3))
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "dev_search_blocks: null or invalid argument");
    if (n_blocks < 8) return set_err(ctx, STEPTH_ERR_TOO_FEW_BLOCKS, "fewer than 8 blocks: par_chunks_mut(0) panics");
    cudaStream_t st = (cudaStream_t)stream;
    STEPTH_CUDA_TRY(ctx, cudaMemsetAsync(max_out, 0, sizeof(uint32_t), st));
    if (params->precision[0] == 0 || params->precision[1] == 0 || params->precision[2] == 0) {
        STEPTH_CUDA_TRY(ctx, cudaMemsetAsync(d8_out, 0, n_blocks, st));
        return STEPTH_OK;
    }
    SearchArgs a = make_search_args(params);
    a.blocks = blocks;
    a.n_queries = (uint32_t)n_blocks;
    a.add_rgb = add_rgb;
    a.wa = wa; a.ha = ha; a.add_y0 = 0; a.add_img_h = ha;
    a.d8 = d8_out;
    a.max_out = max_out;
    if (params->flags & STEPTH_FLAG_LITERAL_SEARCH) {
        dim3 grid((a.n_queries + 255) / 256, 1);
        search_literal_kernel<<<grid, 256, 0, st>>>(a);
        LAUNCH_CHECK(ctx);
        return STEPTH_OK;
    }
    return launch_search_hier(ctx, st, a, 1);
}

extern "C" int32_t stepth_cuda_dev_paint_blocks(stepth_ctx *ctx, void *stream, const stepth_block *blocks,
                                                const uint8_t *d8, size_t n_blocks, uint32_t w, uint32_t h,
                                                uint8_t *plane_out) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!blocks || !d8 || !plane_out || w == 0 || h == 0 || n_blocks == 0 || n_blocks > 0x7ffffffull)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "dev_paint_blocks: null or invalid argument");
    cudaStream_t st = (cudaStream_t)stream;
    STEPTH_CUDA_TRY(ctx, cudaMemsetAsync(plane_out, 0, (size_t)w * h, st));
    PaintArgs a{blocks, d8, plane_out, (uint32_t)n_blocks, w, h};
    uint32_t grid = (uint32_t)((n_blocks * 32 + 255) / 256);
    paint_kernel<<<grid, 256, 0, st>>>(a);
    LAUNCH_CHECK(ctx);
    return STEPTH_OK;
}

extern "C" int32_t stepth_cuda_dev_finish(stepth_ctx *ctx, void *stream, uint32_t n_pairs, uint32_t w,
                                          uint32_t img_h, const uint8_t *src, uint32_t src_y0, uint32_t src_rows,
                                          const uint32_t *max_in, uint32_t flags, uint8_t *depth_out,
                                          uint32_t out_y0, uint32_t out_rows, uint32_t *hist_out,
                                          int32_t *status_out) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!src || !max_in || !depth_out || n_pairs == 0 || w == 0 || img_h == 0 || out_rows == 0 ||
        (flags & STEPTH_RESAMPLE_MASK) == 3u || out_y0 + out_rows > img_h || n_pairs > 65535)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "dev_finish: null or invalid argument");
    cudaStream_t st = (cudaStream_t)stream;
    FinishArgs a;
    memset(&a, 0, sizeof a);
    a.src = src; a.dst = depth_out; a.max_in = max_in; a.hist = hist_out; a.status = status_out;
    a.w = w; a.img_h = img_h; a.src_y0 = src_y0; a.src_rows = src_rows; a.out_y0 = out_y0; a.out_rows = out_rows;
    a.mode = flags & STEPTH_RESAMPLE_MASK;
    a.invert = (flags & STEPTH_FLAG_INVERT) ? 1u : 0u;
    if (a.mode != STEPTH_RESAMPLE_OFF) {
        int32_t rc = get_taps(ctx, st, img_h, &a.taps_y, a.wy_int);
        if (rc != STEPTH_OK) return rc;
        rc = get_taps(ctx, st, w, &a.taps_x, a.wx_int);
        if (rc != STEPTH_OK) return rc;
    }
    if (a.mode == STEPTH_RESAMPLE_U8 && (w & 3) == 0 && w >= 8 && ctx->finish_streaming &&
        ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(depth_out)) & 3) == 0) {
        // the default mode on whole-word rows: two streaming kernels with a 1 B/px plane between them (finish.cuh)
        stepth_ctx::Buf &tmp = ctx->slot ? ctx->work2 : ctx->work;
        edge_taps(ctx, w, a.wx_edge);
        int32_t rc = ensure(ctx, tmp, (size_t)n_pairs * out_rows * w);
        if (rc != STEPTH_OK) return rc;
        a.seg_rows = FV_ROWS;   // a small launch takes short segments: more threads, shorter dependent chains (one 1080p pair: 16 -> 9 us)
        if ((uint64_t)((w / 4 + 127) / 128) * ((out_rows + FV_ROWS - 1) / FV_ROWS) * n_pairs < (uint64_t)ctx->sm_count * 9) a.seg_rows = 7;
        dim3 gv((w / 4 + 127) / 128, (out_rows + a.seg_rows - 1) / a.seg_rows, n_pairs);
        dim3 gh((w + 255) / 256, (out_rows + 31) / 32, n_pairs);
        if (gv.y > 65535) return set_err(ctx, STEPTH_ERR_INVALID_ARG, "dev_finish: more than 1M rows in one band");
        finish_vert_kernel<<<gv, 128, 0, st>>>(a, (uint8_t *)tmp.p);
        LAUNCH_CHECK(ctx);
        if (ctx->finish_n_mirror && n_pairs == 1) {   // set by the row-band call (multi_api.cuh) around this one
            for (uint32_t q = 0; q < ctx->finish_n_mirror && q < 8; ++q) a.mirror[q] = ctx->finish_mirror[q];
            a.n_mirror = std::min(ctx->finish_n_mirror, 8u);
            ctx->finish_mirrored = 1;
        }
        finish_horz_kernel<<<gh, 256, 0, st>>>(a, (const uint8_t *)tmp.p);
        LAUNCH_CHECK(ctx);
        return STEPTH_OK;
    }
    dim3 grid((w + FIN_TILE_W - 1) / FIN_TILE_W, (out_rows + FIN_TILE_H - 1) / FIN_TILE_H, n_pairs);
    if (grid.y > 65535) return set_err(ctx, STEPTH_ERR_INVALID_ARG, "dev_finish: more than 2M rows in one band");
    finish_kernel<<<grid, 256, 0, st>>>(a);
    LAUNCH_CHECK(ctx);
    return STEPTH_OK;
}

extern "C" int32_t stepth_cuda_dev_depth_split(stepth_ctx *ctx, void *stream, uint32_t n_pairs,
                                               const uint32_t *hist, uint8_t zones, uint8_t *ranges_out,
                                               uint32_t *n_out, int32_t *status_out) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!hist || !ranges_out || !n_out || !status_out || n_pairs == 0 || zones < 2)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "dev_depth_split: null or invalid argument (zones >= 2)");
    SplitArgs a{hist, ranges_out, n_out, status_out, zones, 100000u};
    depth_split_kernel<<<n_pairs, 256, 0, (cudaStream_t)stream>>>(a);
    LAUNCH_CHECK(ctx);
    return STEPTH_OK;
}

extern "C" int32_t stepth_cuda_dev_mask_apply(stepth_ctx *ctx, void *stream, uint32_t n_pairs, size_t px,
                                              const uint8_t *depth, const uint8_t *from_to, uint32_t stride_ft,
                                              const int32_t *status_in, const uint8_t *rgba_in, uint8_t *mask_out,
                                              uint8_t *rgba_out) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!depth || !from_to || n_pairs == 0 || px == 0 || (rgba_out && !rgba_in) || (!mask_out && !rgba_out) ||
        n_pairs > 65535 || ((uintptr_t)rgba_in & 3) || ((uintptr_t)rgba_out & 3))
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "dev_mask_apply: null or invalid argument");
    MaskArgs a{depth, from_to, status_in, rgba_in, mask_out, rgba_out, px, stride_ft, 0u};
    uintptr_t al = (uintptr_t)depth | (uintptr_t)rgba_in | (uintptr_t)mask_out | (uintptr_t)rgba_out;
    a.vec_ok = ((al & 15) == 0 && px % 16 == 0) ? 1u : 0u;
    size_t items = a.vec_ok ? px / 16 : px;
    uint32_t gx = (uint32_t)std::min<size_t>((items + 255) / 256, (size_t)ctx->sm_count * 16);
    mask_apply_kernel<<<dim3(gx ? gx : 1, n_pairs), 256, 0, (cudaStream_t)stream>>>(a);
    LAUNCH_CHECK(ctx);
    return STEPTH_OK;
}

extern "C" int32_t stepth_cuda_dev_histogram(stepth_ctx *ctx, void *stream, const uint8_t *data, size_t n,
                                             uint32_t *hist_out) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!data || !hist_out) return set_err(ctx, STEPTH_ERR_INVALID_ARG, "dev_histogram: null argument");
    if (n == 0) return STEPTH_OK;
    uint32_t grid = (uint32_t)std::min<size_t>((n + 256 * 64 - 1) / (256 * 64), (size_t)ctx->sm_count * 8);
    histogram_kernel<<<grid ? grid : 1, 256, 0, (cudaStream_t)stream>>>(data, n, hist_out);
    LAUNCH_CHECK(ctx);
    return STEPTH_OK;
}

// ============================================================================================================
// 3. host-buffer entry points
// ============================================================================================================
namespace {

// per-pair small state kept together so that one copy brings it back: [max][hist][status][n_out][ranges]
struct Small {
    uint32_t n;
    size_t off_max, off_hist, off_status, off_nout, off_ranges, bytes;
    explicit Small(uint32_t n_) : n(n_) {
        off_max = 0;
        off_hist = off_max + sizeof(uint32_t) * n;
        off_status = off_hist + sizeof(uint32_t) * 256 * n;
        off_nout = off_status + sizeof(int32_t) * n;
        off_ranges = off_nout + sizeof(uint32_t) * n;
        bytes = off_ranges + 512 * (size_t)n;
    }
    uint32_t *max(void *b) const { return (uint32_t *)((char *)b + off_max); }
    uint32_t *hist(void *b) const { return (uint32_t *)((char *)b + off_hist); }
    int32_t *status(void *b) const { return (int32_t *)((char *)b + off_status); }
    uint32_t *nout(void *b) const { return (uint32_t *)((char *)b + off_nout); }
    uint8_t *ranges(void *b) const { return (uint8_t *)((char *)b + off_ranges); }
};

int32_t ensure_host_small(stepth_ctx *ctx, size_t bytes) {
    if (bytes <= ctx->h_small_cap) return STEPTH_OK;
    if (ctx->h_small) STEPTH_CUDA_TRY(ctx, cudaFreeHost(ctx->h_small));
    ctx->h_small = nullptr;
    ctx->h_small_cap = 0;
    STEPTH_CUDA_TRY(ctx, cudaMallocHost(&ctx->h_small, bytes));
    ctx->h_small_cap = bytes;
    return STEPTH_OK;
}

#define TRY_RC(expr)                      \
    do {                                  \
        int32_t rc__ = (expr);            \
        if (rc__ != STEPTH_OK) return rc__; \
    } while (0)

// depth_image.rs:110-134 (+ tail) for n pairs of one shape; `blocks` selects block mode (n == 1 only).
int32_t run_pairs_impl(stepth_ctx *ctx, uint32_t n, const uint8_t *const *main_rgba, uint32_t w, uint32_t h,
                       const uint8_t *const *add_rgb, uint32_t wa, uint32_t ha, const stepth_block *blocks,
                       size_t n_blocks, const stepth_params *params, uint8_t *const *depth_out, uint8_t *raw_d8_out,
                       const uint8_t *const *rgba_in, uint8_t *const *rgba_out, uint8_t *const *mask_out,
                       uint8_t *from_to_out, int32_t *status_out, std::vector<cudaEvent_t> &t_out, uint32_t main_stride) {
    const uint32_t flags = params->flags;
    const bool want_fg = (flags & STEPTH_FLAG_FOREGROUND) != 0;
    const bool want_apply = (flags & STEPTH_FLAG_APPLY) != 0;
    if (want_apply && (!want_fg || !rgba_out || !rgba_in))
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "APPLY needs FOREGROUND, rgba_in and rgba_out");
    if (want_apply)   // per element: the single-pair entry points hand in the addresses of their own arguments
        for (uint32_t i = 0; i < n; ++i)
            if (!rgba_out[i] || !rgba_in[i])
                return set_err(ctx, STEPTH_ERR_INVALID_ARG, "APPLY: pair %u has no rgba_in or no rgba_out buffer", i);
    STEPTH_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const size_t px = (size_t)w * h;
    // only the part of the additional image a query can reach is uploaded (origins are half-coordinates:
    // depth_image.rs:114-117): columns [0, w/2 + max), rows [0, h/2 + max)
    uint32_t reach_w = w / 2 + params->max_rings, reach_h = h / 2 + params->max_rings;
    if (blocks) { reach_w = (w + 1) / 2 + params->max_rings + 1; reach_h = (h + 1) / 2 + params->max_rings + 1; }
    // full_rows (off; measured slower, 4.98 vs 4.79 ms per 16 pairs): whole rows as one linear copy per image instead of a
    // pitched 2-D copy of the reachable columns; the planes cover only the reachable columns either way
    const uint32_t wa_d = ctx->upload_full_rows ? wa : std::min(wa, reach_w), ha_d = std::min(ha, reach_h);
    const size_t add_bytes = (size_t)wa_d * ha_d * 3;
    const size_t n_q = blocks ? n_blocks : px;

    Small sm(n);
    TRY_RC(ensure(ctx, ctx->small, sm.bytes));
    TRY_RC(ensure_host_small(ctx, sm.bytes));
    TRY_RC(ensure(ctx, ctx->add_rgb, add_bytes * n));
    TRY_RC(ensure(ctx, ctx->d8, n_q * n));
    TRY_RC(ensure(ctx, ctx->depth, px * n));
    if (blocks) {
        if (!ctx->blocks_on_device) TRY_RC(ensure(ctx, ctx->blocks, sizeof(stepth_block) * n_blocks));
        TRY_RC(ensure(ctx, ctx->plane, px));
    }
    if (!blocks || want_apply) TRY_RC(ensure(ctx, ctx->main_rgba, px * 4 * n));
    if (main_stride == 3) TRY_RC(ensure(ctx, ctx->main_rgb, px * 3 * n));
    // Host buffers that are not page-locked (a Rust Vec<u8>, a numpy array) go through the library's pinned ring, filled and
    // emptied by its copy threads (staging.cuh); page-locked ones are read and written by the DMA engines directly.
    auto upload = [&](void *dev, const void *host, size_t src_pitch, size_t row_bytes, size_t rows, cudaStream_t s) -> int32_t {
        if (ctx->stage_threads > 0 && row_bytes * rows >= (64u << 10) && (src_pitch == row_bytes || row_bytes <= STAGE_IN_RING / 4) &&
            host_ptr_is_pageable(host)) {
            STEPTH_CUDA_TRY(ctx, ctx->stager.ensure(ctx->stage_threads));
            STEPTH_CUDA_TRY(ctx, ctx->stager.h2d(dev, host, src_pitch, row_bytes, rows, s));
        } else if (src_pitch == row_bytes || rows == 1) {
            STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(dev, host, row_bytes * rows, cudaMemcpyHostToDevice, s));
        } else {
            STEPTH_CUDA_TRY(ctx, cudaMemcpy2DAsync(dev, row_bytes, host, src_pitch, row_bytes, rows, cudaMemcpyHostToDevice, s));
        }
        return STEPTH_OK;
    };
    auto download = [&](void *host, const void *dev, size_t bytes, cudaStream_t s) -> int32_t {
        if (ctx->stage_threads > 0 && bytes >= (64u << 10) && host_ptr_is_pageable(host)) {
            STEPTH_CUDA_TRY(ctx, ctx->stager.ensure(ctx->stage_threads));
            STEPTH_CUDA_TRY(ctx, ctx->stager.d2h(host, dev, bytes, s));
        } else {
            STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, s));
        }
        return STEPTH_OK;
    };
    if (want_fg && mask_out) TRY_RC(ensure(ctx, ctx->mask, px * n));
    if (want_apply) TRY_RC(ensure(ctx, ctx->rgba_out, px * 4 * n));

    STEPTH_CUDA_TRY(ctx, cudaMemsetAsync(ctx->small.p, 0, sm.bytes, st));
    // The batch is cut into chunks of pairs that move through three streams: copy_in uploads chunk k+1 while the compute
    // stream works on chunk k and copy_out brings chunk k-1 back, so the PCIe transfers hide behind the kernels.
    // The first chunk is half a chunk (the compute stream starts sooner) and so, then, is the last (a shorter copy-out tail).
    // 8 chunks, or chunks of about 4 pairs for larger batches (64 x 4K fused in 8 / 16 / 32 chunks: 76.1 / 72.4 / 74.2 ms per call; the pipeline's un-overlapped
    // first upload and last download shrink with the chunk)
    const uint32_t parts = (uint32_t)std::max(1, ctx->batch_chunks > 0 ? ctx->batch_chunks : (int)std::min(32u, std::max(8u, n / 4)));
    const uint32_t chunk = (blocks || n <= 2) ? n : (n + parts - 1) / parts;
    std::vector<uint32_t> bounds{0};
    if (chunk >= 2 && chunk < n) bounds.push_back(chunk / 2);
    while (bounds.back() < n) bounds.push_back(std::min(n, bounds.back() + chunk));
    const uint32_t n_chunks = (uint32_t)bounds.size() - 1;
    while (ctx->events.size() < 3 * (size_t)n_chunks + 2) {
        cudaEvent_t e;
        STEPTH_CUDA_TRY(ctx, cudaEventCreateWithFlags(&e, ctx->trace ? cudaEventDefault : cudaEventDisableTiming));
        ctx->events.push_back(e);
    }
    // dense pairs always go through the copy streams: the additional images go up first and K0 runs on them while the main
    // images follow (`split`)
    const bool split = !blocks;
    const bool piped = n_chunks > 1 || split;
    // t_out (trace only): end of each chunk's copy-out; owned and destroyed by run_pairs
    if (ctx->trace && piped)
        for (uint32_t k = 0; k < n_chunks; ++k) {
            cudaEvent_t e;
            STEPTH_CUDA_TRY(ctx, cudaEventCreate(&e));
            t_out.push_back(e);
        }
    cudaStream_t s_in = piped ? ctx->copy_in : st, s_out = piped ? ctx->copy_out : st;
    // two alternating compute streams pay off for small chunks (their poorly filled last kernels overlap the next chunk's
    // first ones); a chunk that fills the GPU by itself only finishes later that way, and so does its copy-out (measured:
    // 1080p pairs 4.31 -> 4.03 ms per 16, one 4K pair per chunk 5.54 -> 5.92 ms per 4)
    const bool two = n_chunks > 1 && ctx->two_streams && ctx->stream2 && (uint64_t)chunk * px <= 6000000ull;
    cudaStream_t st0 = st;
    if (piped) {  // the copy streams must not overtake the memset / earlier work on the compute stream
        STEPTH_CUDA_TRY(ctx, cudaEventRecord(ctx->events[2 * n_chunks], st));
        STEPTH_CUDA_TRY(ctx, cudaStreamWaitEvent(s_in, ctx->events[2 * n_chunks], 0));
        STEPTH_CUDA_TRY(ctx, cudaStreamWaitEvent(s_out, ctx->events[2 * n_chunks], 0));
        if (two) STEPTH_CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream2, ctx->events[2 * n_chunks], 0));
    }
    uint8_t *d_main = (uint8_t *)ctx->main_rgba.p, *d_add = (uint8_t *)ctx->add_rgb.p, *d_d8 = (uint8_t *)ctx->d8.p;
    uint8_t *d_depth = (uint8_t *)ctx->depth.p, *d_mask = (uint8_t *)ctx->mask.p, *d_rgba = (uint8_t *)ctx->rgba_out.p;
    for (uint32_t k = 0; k < n_chunks; ++k) {
        const uint32_t i0 = bounds[k], i1 = bounds[k + 1], nc = i1 - i0;
        ctx->slot = two ? (int)(k & 1) : 0;   // chunks alternate between two compute streams, each with its own scratch
        st = ctx->slot ? ctx->stream2 : st0;
        for (uint32_t i = i0; i < i1; ++i)   // the reachable window of each additional image (whole rows when it spans them)
            TRY_RC(upload(d_add + add_bytes * i, add_rgb[i], (size_t)wa * 3, (size_t)wa_d * 3, ha_d, s_in));
        if (ctx->stager.pool) STEPTH_CUDA_TRY(ctx, ctx->stager.flush_in());   // staged pieces still filling: their DMAs go first
        if (split) {   // the compute stream may start K0 now; K1 waits for the second event (ctx->main_ready)
            STEPTH_CUDA_TRY(ctx, cudaEventRecord(ctx->events[2 * n_chunks + 2 + k], s_in));
            STEPTH_CUDA_TRY(ctx, cudaStreamWaitEvent(st, ctx->events[2 * n_chunks + 2 + k], 0));
        }
        for (uint32_t i = i0; i < i1; ++i) {
            const uint8_t *rgba_src = blocks ? (want_apply ? rgba_in[i] : nullptr) : main_rgba[i];
            if (!rgba_src || ctx->blocks_on_device) continue;   // (the device discretiser's caller left the main image in d_main)
            if (main_stride == 3) TRY_RC(upload((uint8_t *)ctx->main_rgb.p + px * 3 * i, rgba_src, px * 3, px * 3, 1, s_in));
            else TRY_RC(upload(d_main + px * 4 * i, rgba_src, px * 4, px * 4, 1, s_in));
        }
        if (ctx->stager.pool) STEPTH_CUDA_TRY(ctx, ctx->stager.flush_in());
        if (main_stride == 3 && !blocks) {   // RGB8 mains: widened to the RGBA words the search reads, behind their upload
            const size_t n_px = px * (size_t)nc;
            const uint32_t grid = (uint32_t)std::min<size_t>((n_px / 4 + 255) / 256 + 1, (size_t)ctx->sm_count * 8);
            rgb_to_rgba_kernel<<<grid, 256, 0, s_in>>>((const uint8_t *)ctx->main_rgb.p + px * 3 * i0,
                                                       (uint32_t *)(d_main + px * 4 * i0), n_px);
            LAUNCH_CHECK(ctx);
        }
        if (piped) {
            STEPTH_CUDA_TRY(ctx, cudaEventRecord(ctx->events[2 * k], s_in));
            if (split) ctx->main_ready = ctx->events[2 * k];
            else STEPTH_CUDA_TRY(ctx, cudaStreamWaitEvent(st, ctx->events[2 * k], 0));
        }
        uint32_t *c_max = sm.max(ctx->small.p) + i0, *c_hist = sm.hist(ctx->small.p) + 256 * (size_t)i0;
        int32_t *c_status = sm.status(ctx->small.p) + i0;
        uint32_t *c_nout = sm.nout(ctx->small.p) + i0;
        uint8_t *c_ranges = sm.ranges(ctx->small.p) + 512 * (size_t)i0;
        const uint8_t *plane = nullptr;  // the raw-distance plane handed to finish
        if (blocks) {
            // the list is the caller's (host) or the device discretiser's (already in HBM: stepth_cuda_depth_from_additional)
            const stepth_block *d_blocks = ctx->blocks_on_device ? blocks : (const stepth_block *)ctx->blocks.p;
            if (!ctx->blocks_on_device)
                STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->blocks.p, blocks, sizeof(stepth_block) * n_blocks,
                                                     cudaMemcpyHostToDevice, st));
            TRY_RC(stepth_cuda_dev_search_blocks(ctx, st, params, d_add, wa_d, ha_d, d_blocks, n_blocks, d_d8, c_max));
            TRY_RC(stepth_cuda_dev_paint_blocks(ctx, st, d_blocks, d_d8, n_blocks, w, h, (uint8_t *)ctx->plane.p));
            plane = (const uint8_t *)ctx->plane.p;
        } else {
            stepth_dev_shape s;
            memset(&s, 0, sizeof s);
            s.n_pairs = nc; s.w = w; s.h = h; s.img_h = h; s.y0 = 0;
            s.wa = wa_d; s.ha = ha_d; s.add_y0 = 0; s.add_img_h = ha;
            TRY_RC(stepth_cuda_dev_search_dense(ctx, st, &s, params, d_main + px * 4 * i0, d_add + add_bytes * i0,
                                                d_d8 + px * i0, c_max));
            plane = d_d8 + px * i0;
        }
        TRY_RC(stepth_cuda_dev_finish(ctx, st, nc, w, h, plane, 0, h, c_max, flags, d_depth + px * i0, 0, h,
                                      want_fg ? c_hist : nullptr, c_status));
        if (want_fg) {
            TRY_RC(stepth_cuda_dev_depth_split(ctx, st, nc, c_hist, 2, c_ranges, c_nout, c_status));
            if (mask_out || want_apply)
                TRY_RC(stepth_cuda_dev_mask_apply(ctx, st, nc, px, d_depth + px * i0, c_ranges, 512, c_status,
                                                  want_apply ? d_main + px * 4 * i0 : nullptr,
                                                  mask_out ? d_mask + px * i0 : nullptr,
                                                  want_apply ? d_rgba + px * 4 * i0 : nullptr));
        }
        if (piped) {
            STEPTH_CUDA_TRY(ctx, cudaEventRecord(ctx->events[2 * k + 1], st));
            STEPTH_CUDA_TRY(ctx, cudaStreamWaitEvent(s_out, ctx->events[2 * k + 1], 0));
        }
        for (uint32_t i = i0; i < i1; ++i) {
            TRY_RC(download(depth_out[i], d_depth + px * i, px, s_out));
            if (want_fg && mask_out && mask_out[i]) TRY_RC(download(mask_out[i], d_mask + px * i, px, s_out));
            if (want_apply && rgba_out[i]) TRY_RC(download(rgba_out[i], d_rgba + px * 4 * i, px * 4, s_out));
        }
        if (!t_out.empty()) STEPTH_CUDA_TRY(ctx, cudaEventRecord(t_out[k], s_out));
    }
    ctx->slot = 0;
    ctx->main_ready = nullptr;
    st = st0;
    if (two) {  // the small state of the odd chunks is written on stream2
        STEPTH_CUDA_TRY(ctx, cudaEventRecord(ctx->events[2 * n_chunks + 1], ctx->stream2));
        STEPTH_CUDA_TRY(ctx, cudaStreamWaitEvent(st, ctx->events[2 * n_chunks + 1], 0));
    }
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->h_small, ctx->small.p, sm.bytes, cudaMemcpyDeviceToHost, st));
    if (raw_d8_out) STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(raw_d8_out, ctx->d8.p, n_q * n, cudaMemcpyDeviceToHost, st));
    // everything is enqueued: the staged results reach the caller's buffers piece by piece as their DMAs complete
    if (ctx->stager.pool) STEPTH_CUDA_TRY(ctx, ctx->stager.finish());
    STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(st));
    if (piped) {
        STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(s_in));
        STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(s_out));
    }
    if (!t_out.empty()) {
        fprintf(stderr, "stepth trace: %u pairs in %u chunks%s (ms after the start: copy-in, compute, copy-out)\n", n, n_chunks,
                two ? ", two compute streams" : "");
        for (uint32_t k = 0; k < n_chunks; ++k) {
            float a = 0, b = 0, c = 0;
            cudaEventElapsedTime(&a, ctx->events[2 * n_chunks], ctx->events[2 * k]);
            cudaEventElapsedTime(&b, ctx->events[2 * n_chunks], ctx->events[2 * k + 1]);
            cudaEventElapsedTime(&c, ctx->events[2 * n_chunks], t_out[k]);
            fprintf(stderr, "  chunk %u [%u,%u): %.3f %.3f %.3f\n", k, bounds[k], bounds[k + 1], a, b, c);
        }
    }

    int32_t first_bad = STEPTH_OK;
    for (uint32_t i = 0; i < n; ++i) {
        int32_t s = sm.status(ctx->h_small)[i];
        if (status_out) status_out[i] = s;
        if (s != STEPTH_OK && first_bad == STEPTH_OK) first_bad = s;
        if (want_fg && from_to_out) {
            from_to_out[2 * i] = sm.ranges(ctx->h_small)[512 * (size_t)i];
            from_to_out[2 * i + 1] = sm.ranges(ctx->h_small)[512 * (size_t)i + 1];
        }
    }
    if (first_bad != STEPTH_OK && !status_out) {
        const char *why = first_bad == STEPTH_ERR_DIV_BY_ZERO_MAX ? "max distance is 0: attempt to divide by zero (depth_image.rs:127)"
                          : first_bad == STEPTH_ERR_DEGENERATE_DEPTH ? "depth_split on a degenerate depth range (depth_image.rs:212-216)"
                                                                     : "depth_split did not converge";
        return set_err(ctx, first_bad, "%s", why);
    }
    return STEPTH_OK;
}

// Every exit of the batch path goes through here: an error inside the chunk loop returns at once, possibly with device->host
// copies into the caller's buffers still in flight and the context's chunk state (slot, main_ready) mid-batch.  The caller
// owns those buffers again the moment this returns, so all four streams are drained and the state reset first.
int32_t run_pairs(stepth_ctx *ctx, uint32_t n, const uint8_t *const *main_rgba, uint32_t w, uint32_t h,
                  const uint8_t *const *add_rgb, uint32_t wa, uint32_t ha, const stepth_block *blocks,
                  size_t n_blocks, const stepth_params *params, uint8_t *const *depth_out, uint8_t *raw_d8_out,
                  const uint8_t *const *rgba_in, uint8_t *const *rgba_out, uint8_t *const *mask_out,
                  uint8_t *from_to_out, int32_t *status_out, uint32_t main_stride = 4) {
    std::vector<cudaEvent_t> t_out;
    const int32_t rc = run_pairs_impl(ctx, n, main_rgba, w, h, add_rgb, wa, ha, blocks, n_blocks, params, depth_out, raw_d8_out,
                                      rgba_in, rgba_out, mask_out, from_to_out, status_out, t_out, main_stride);
    if (rc != STEPTH_OK) {
        const std::string why = ctx->err;   // the drain must not replace the first error's text
        for (cudaStream_t s : {ctx->copy_in, ctx->stream, ctx->stream2, ctx->copy_out})
            if (s) cudaStreamSynchronize(s);
        ctx->stager.abandon();
        ctx->err = why;
    }
    ctx->slot = 0;
    ctx->main_ready = nullptr;
    for (cudaEvent_t e : t_out) cudaEventDestroy(e);
    return rc;
}

}  // namespace

extern "C" int32_t stepth_cuda_depth_from_blocks(stepth_ctx *ctx, const uint8_t *add_rgb, uint32_t wa, uint32_t ha,
                                                 const stepth_block *blocks, size_t n_blocks, uint32_t w, uint32_t h,
                                                 const stepth_params *params, uint8_t *depth_out, uint8_t *raw_d8_out,
                                                 const uint8_t *rgba_in, uint8_t *rgba_out, uint8_t *mask_out,
                                                 uint8_t *from_to_out) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!add_rgb || !blocks || !depth_out || !params_ok(params) || w == 0 || h == 0 || wa == 0 || ha == 0)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "depth_from_blocks: null or invalid argument");
    if (n_blocks < 8) return set_err(ctx, STEPTH_ERR_TOO_FEW_BLOCKS, "fewer than 8 blocks: par_chunks_mut(0) panics (depth_image.rs:111-112)");
    for (size_t b = 0; b < n_blocks; ++b)
        if ((uint64_t)blocks[b].x + blocks[b].w > w || (uint64_t)blocks[b].y + blocks[b].h > h)
            return set_err(ctx, STEPTH_ERR_INVALID_ARG, "block %zu lies outside the %ux%u image", b, w, h);
    return run_pairs(ctx, 1, nullptr, w, h, &add_rgb, wa, ha, blocks, n_blocks, params, &depth_out, raw_d8_out,
                     &rgba_in, &rgba_out, mask_out ? &mask_out : nullptr, from_to_out, nullptr);
}

extern "C" int32_t stepth_cuda_depth_dense(stepth_ctx *ctx, const uint8_t *main_rgba, uint32_t w, uint32_t h,
                                           const uint8_t *add_rgb, uint32_t wa, uint32_t ha,
                                           const stepth_params *params, uint8_t *depth_out, uint8_t *raw_d8_out,
                                           uint8_t *rgba_out, uint8_t *mask_out, uint8_t *from_to_out) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!main_rgba || !add_rgb || !depth_out || !params_ok(params) || w == 0 || h == 0 || wa == 0 || ha == 0)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "depth_dense: null or invalid argument");
    if ((uint64_t)w * h < 8) return set_err(ctx, STEPTH_ERR_TOO_FEW_BLOCKS, "fewer than 8 pixels: par_chunks_mut(0) panics");
    return run_pairs(ctx, 1, &main_rgba, w, h, &add_rgb, wa, ha, nullptr, 0, params, &depth_out, raw_d8_out,
                     &main_rgba, &rgba_out, mask_out ? &mask_out : nullptr, from_to_out, nullptr);
}

extern "C" int32_t stepth_cuda_depth_dense_batch(stepth_ctx *ctx, uint32_t n_pairs, const uint8_t *const *main_rgba,
                                                 uint32_t w, uint32_t h, const uint8_t *const *add_rgb, uint32_t wa,
                                                 uint32_t ha, const stepth_params *params, uint8_t *const *depth_out,
                                                 uint8_t *const *rgba_out, uint8_t *const *mask_out,
                                                 uint8_t *from_to_out, int32_t *status_out) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!main_rgba || !add_rgb || !depth_out || !status_out || !params_ok(params) || n_pairs == 0 || w == 0 || h == 0 ||
        wa == 0 || ha == 0)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "depth_dense_batch: null or invalid argument");
    if ((uint64_t)w * h < 8) return set_err(ctx, STEPTH_ERR_TOO_FEW_BLOCKS, "fewer than 8 pixels: par_chunks_mut(0) panics");
    for (uint32_t i = 0; i < n_pairs; ++i)
        if (!main_rgba[i] || !add_rgb[i] || !depth_out[i])
            return set_err(ctx, STEPTH_ERR_INVALID_ARG, "depth_dense_batch: pair %u has a null buffer", i);
    return run_pairs(ctx, n_pairs, main_rgba, w, h, add_rgb, wa, ha, nullptr, 0, params, depth_out, nullptr, main_rgba,
                     rgba_out, mask_out, from_to_out, status_out);
}

extern "C" int32_t stepth_cuda_depth_dense_batch_rgb(stepth_ctx *ctx, uint32_t n_pairs, const uint8_t *const *main_rgb,
                                                     uint32_t w, uint32_t h, const uint8_t *const *add_rgb, uint32_t wa,
                                                     uint32_t ha, const stepth_params *params, uint8_t *const *depth_out,
                                                     uint8_t *const *mask_out, uint8_t *from_to_out, int32_t *status_out) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!main_rgb || !add_rgb || !depth_out || !status_out || !params_ok(params) || n_pairs == 0 || w == 0 || h == 0 ||
        wa == 0 || ha == 0)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "depth_dense_batch_rgb: null or invalid argument");
    if (params->flags & STEPTH_FLAG_APPLY)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "depth_dense_batch_rgb: APPLY returns RGBA pixels and needs the RGBA entry point");
    if ((uint64_t)w * h < 8) return set_err(ctx, STEPTH_ERR_TOO_FEW_BLOCKS, "fewer than 8 pixels: par_chunks_mut(0) panics");
    for (uint32_t i = 0; i < n_pairs; ++i)
        if (!main_rgb[i] || !add_rgb[i] || !depth_out[i])
            return set_err(ctx, STEPTH_ERR_INVALID_ARG, "depth_dense_batch_rgb: pair %u has a null buffer", i);
    return run_pairs(ctx, n_pairs, main_rgb, w, h, add_rgb, wa, ha, nullptr, 0, params, depth_out, nullptr, main_rgb, nullptr,
                     mask_out, from_to_out, status_out, 3);
}

// ---- tail operations on host buffers -----------------------------------------------------------------------
extern "C" int32_t stepth_cuda_invert_depth(stepth_ctx *ctx, uint8_t *depth_inout, size_t n) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!depth_inout) return set_err(ctx, STEPTH_ERR_INVALID_ARG, "invert_depth: null buffer");
    if (n == 0) return STEPTH_OK;
    STEPTH_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    TRY_RC(ensure(ctx, ctx->depth, n));
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->depth.p, depth_inout, n, cudaMemcpyHostToDevice, st));
    uint32_t grid = (uint32_t)std::min<size_t>((n / 4 + 255) / 256 + 1, (size_t)ctx->sm_count * 16);
    invert_kernel<<<grid, 256, 0, st>>>((uint8_t *)ctx->depth.p, n);
    LAUNCH_CHECK(ctx);
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(depth_inout, ctx->depth.p, n, cudaMemcpyDeviceToHost, st));
    STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(st));
    return STEPTH_OK;
}

// uploads `depth`, histograms it, runs depth_split(zones); leaves the plane in ctx->depth and the small state
// in ctx->h_small (Small(1) layout).  Shared by depth_split / select_foreground.
static int32_t split_on_device(stepth_ctx *ctx, const uint8_t *depth, size_t n, uint8_t zones) {
    STEPTH_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    Small sm(1);
    TRY_RC(ensure(ctx, ctx->small, sm.bytes));
    TRY_RC(ensure_host_small(ctx, sm.bytes));
    TRY_RC(ensure(ctx, ctx->depth, n ? n : 1));
    STEPTH_CUDA_TRY(ctx, cudaMemsetAsync(ctx->small.p, 0, sm.bytes, st));
    if (n) STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->depth.p, depth, n, cudaMemcpyHostToDevice, st));
    TRY_RC(stepth_cuda_dev_histogram(ctx, st, (const uint8_t *)ctx->depth.p, n, sm.hist(ctx->small.p)));
    TRY_RC(stepth_cuda_dev_depth_split(ctx, st, 1, sm.hist(ctx->small.p), zones, sm.ranges(ctx->small.p),
                                       sm.nout(ctx->small.p), sm.status(ctx->small.p)));
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->h_small, ctx->small.p, sm.bytes, cudaMemcpyDeviceToHost, st));
    return STEPTH_OK;
}

extern "C" int32_t stepth_cuda_depth_split(stepth_ctx *ctx, const uint8_t *depth, size_t n, uint8_t zones,
                                           uint8_t *ranges_out, uint32_t *n_out, int32_t *has_values_out) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!depth || !ranges_out || !n_out || !has_values_out)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "depth_split: null argument");
    if (zones < 2) {  // depth_image.rs:163-165  vec![(None, None)]
        *n_out = 1;
        *has_values_out = 0;
        return STEPTH_OK;
    }
    *has_values_out = 1;
    TRY_RC(split_on_device(ctx, depth, n, zones));
    STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    Small sm(1);
    int32_t s = sm.status(ctx->h_small)[0];
    if (s != STEPTH_OK) return set_err(ctx, s, "depth_split failed on the depth range (depth_image.rs:212-216)");
    *n_out = sm.nout(ctx->h_small)[0];
    memcpy(ranges_out, sm.ranges(ctx->h_small), 2 * (size_t)*n_out);
    return STEPTH_OK;
}

extern "C" int32_t stepth_cuda_slice(stepth_ctx *ctx, const uint8_t *depth, size_t n, int32_t from, int32_t to,
                                     uint8_t *mask_out) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!depth || !mask_out || from > 255 || to > 255)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "slice: null or invalid argument");
    if (n == 0) return STEPTH_OK;
    STEPTH_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    TRY_RC(ensure(ctx, ctx->depth, n));
    TRY_RC(ensure(ctx, ctx->mask, n));
    TRY_RC(ensure(ctx, ctx->small, 512));
    uint8_t ft[2] = {(uint8_t)(from < 0 ? 0 : from), (uint8_t)(to < 0 ? 255 : to)};  // unwrap_or(0) / unwrap_or(u8::MAX)
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->small.p, ft, 2, cudaMemcpyHostToDevice, st));
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->depth.p, depth, n, cudaMemcpyHostToDevice, st));
    TRY_RC(stepth_cuda_dev_mask_apply(ctx, st, 1, n, (const uint8_t *)ctx->depth.p, (const uint8_t *)ctx->small.p, 2,
                                      nullptr, nullptr, (uint8_t *)ctx->mask.p, nullptr));
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(mask_out, ctx->mask.p, n, cudaMemcpyDeviceToHost, st));
    STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(st));
    return STEPTH_OK;
}

extern "C" int32_t stepth_cuda_select_foreground(stepth_ctx *ctx, const uint8_t *depth, size_t n, uint8_t *mask_out,
                                                 uint8_t *from_to_out) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!depth || !mask_out) return set_err(ctx, STEPTH_ERR_INVALID_ARG, "select_foreground: null argument");
    TRY_RC(split_on_device(ctx, depth, n, 2));
    cudaStream_t st = ctx->stream;
    Small sm(1);
    if (n) {
        TRY_RC(ensure(ctx, ctx->mask, n));
        TRY_RC(stepth_cuda_dev_mask_apply(ctx, st, 1, n, (const uint8_t *)ctx->depth.p, sm.ranges(ctx->small.p), 512,
                                          sm.status(ctx->small.p), nullptr, (uint8_t *)ctx->mask.p, nullptr));
        STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(mask_out, ctx->mask.p, n, cudaMemcpyDeviceToHost, st));
    }
    STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(st));
    int32_t s = sm.status(ctx->h_small)[0];
    if (s != STEPTH_OK) return set_err(ctx, s, "depth_split(2) failed on the depth range (depth_image.rs:212-216)");
    if (from_to_out) { from_to_out[0] = sm.ranges(ctx->h_small)[0]; from_to_out[1] = sm.ranges(ctx->h_small)[1]; }
    return STEPTH_OK;
}

extern "C" int32_t stepth_cuda_apply_mask(stepth_ctx *ctx, uint8_t *rgba_inout, const uint8_t *mask, size_t n) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!rgba_inout || !mask) return set_err(ctx, STEPTH_ERR_INVALID_ARG, "apply_mask: null buffer");
    if (n == 0) return STEPTH_OK;
    STEPTH_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    TRY_RC(ensure(ctx, ctx->main_rgba, n * 4));
    TRY_RC(ensure(ctx, ctx->mask, n));
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->main_rgba.p, rgba_inout, n * 4, cudaMemcpyHostToDevice, st));
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->mask.p, mask, n, cudaMemcpyHostToDevice, st));
    uint32_t grid = (uint32_t)std::min<size_t>((n + 255) / 256, (size_t)ctx->sm_count * 16);
    apply_mask_kernel<<<grid, 256, 0, st>>>((uint32_t *)ctx->main_rgba.p, (const uint8_t *)ctx->mask.p, n);
    LAUNCH_CHECK(ctx);
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(rgba_inout, ctx->main_rgba.p, n * 4, cudaMemcpyDeviceToHost, st));
    STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(st));
    return STEPTH_OK;
}

// ---- editor operations -----------------------------------------------------------------------------
static int32_t highlight(stepth_ctx *ctx, const uint8_t *rgba_in, const uint8_t *plane, size_t n, uint8_t *rgba_out, int mode) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!rgba_in || !plane || !rgba_out) return set_err(ctx, STEPTH_ERR_INVALID_ARG, "highlight: null buffer");
    if (n == 0) return STEPTH_OK;
    STEPTH_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    TRY_RC(ensure(ctx, ctx->main_rgba, n * 4));
    TRY_RC(ensure(ctx, ctx->rgba_out, n * 4));
    TRY_RC(ensure(ctx, ctx->depth, n));
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->main_rgba.p, rgba_in, n * 4, cudaMemcpyHostToDevice, st));
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->depth.p, plane, n, cudaMemcpyHostToDevice, st));
    const uint32_t grid = (uint32_t)std::min<size_t>((n + 255) / 256, (size_t)ctx->sm_count * 16);
    highlight_kernel<<<grid, 256, 0, st>>>((const uint32_t *)ctx->main_rgba.p, (const uint8_t *)ctx->depth.p, n,
                                           (uint32_t *)ctx->rgba_out.p, mode);
    LAUNCH_CHECK(ctx);
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(rgba_out, ctx->rgba_out.p, n * 4, cudaMemcpyDeviceToHost, st));
    STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(st));
    return STEPTH_OK;
}

extern "C" int32_t stepth_cuda_highlight_depth(stepth_ctx *ctx, const uint8_t *rgba_in, const uint8_t *depth, size_t n,
                                               uint8_t *rgba_out) {
    return highlight(ctx, rgba_in, depth, n, rgba_out, 0);
}

extern "C" int32_t stepth_cuda_highlight_mask(stepth_ctx *ctx, const uint8_t *rgba_in, const uint8_t *mask, size_t n,
                                              uint8_t *rgba_out) {
    return highlight(ctx, rgba_in, mask, n, rgba_out, 1);
}

static int32_t mask_logic(stepth_ctx *ctx, uint8_t *mask_inout, const uint8_t *other, size_t n, int op) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!mask_inout || (op != 2 && !other)) return set_err(ctx, STEPTH_ERR_INVALID_ARG, "mask op: null buffer");
    if (n == 0) return STEPTH_OK;
    STEPTH_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    TRY_RC(ensure(ctx, ctx->mask, n));
    TRY_RC(ensure(ctx, ctx->depth, n));
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->mask.p, mask_inout, n, cudaMemcpyHostToDevice, st));
    if (op != 2) STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->depth.p, other, n, cudaMemcpyHostToDevice, st));
    const uint32_t grid = (uint32_t)std::min<size_t>((n + 255) / 256, (size_t)ctx->sm_count * 16);
    mask_logic_kernel<<<grid, 256, 0, st>>>((uint8_t *)ctx->mask.p, (const uint8_t *)ctx->depth.p, n, op);
    LAUNCH_CHECK(ctx);
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(mask_inout, ctx->mask.p, n, cudaMemcpyDeviceToHost, st));
    STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(st));
    return STEPTH_OK;
}

extern "C" int32_t stepth_cuda_mask_and(stepth_ctx *ctx, uint8_t *mask_inout, const uint8_t *other, size_t n) {
    return mask_logic(ctx, mask_inout, other, n, 0);
}
extern "C" int32_t stepth_cuda_mask_or(stepth_ctx *ctx, uint8_t *mask_inout, const uint8_t *other, size_t n) {
    return mask_logic(ctx, mask_inout, other, n, 1);
}
extern "C" int32_t stepth_cuda_mask_not(stepth_ctx *ctx, uint8_t *mask_inout, size_t n) {
    return mask_logic(ctx, mask_inout, nullptr, n, 2);
}

extern "C" int32_t stepth_cuda_image_replace(stepth_ctx *ctx, uint8_t *rgba_inout, const uint8_t *mask, uint32_t w, uint32_t h,
                                             const uint8_t *other_rgba, uint32_t ow, uint32_t oh, uint32_t start_x,
                                             uint32_t start_y) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!rgba_inout || !mask || !other_rgba || w == 0 || h == 0 || ow == 0 || oh == 0 || (uint64_t)w * h > 0xffffffffull)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "image_replace: null or empty buffer");
    // for y in start_y..(start_y + oh).min(h), for x in start_x..(start_x + ow).min(w)      mask_image.rs:101-102
    const uint32_t x1 = (uint32_t)std::min<uint64_t>((uint64_t)start_x + ow, w), y1 = (uint32_t)std::min<uint64_t>((uint64_t)start_y + oh, h);
    if (start_x >= x1 || start_y >= y1) return STEPTH_OK;  // empty ranges: nothing happens
    STEPTH_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const size_t n = (size_t)w * h, no = (size_t)ow * oh;
    TRY_RC(ensure(ctx, ctx->main_rgba, n * 4));
    TRY_RC(ensure(ctx, ctx->rgba_out, no * 4));
    TRY_RC(ensure(ctx, ctx->mask, n));
    TRY_RC(ensure(ctx, ctx->small, 512));
    TRY_RC(ensure_host_small(ctx, 512));
    STEPTH_CUDA_TRY(ctx, cudaMemsetAsync(ctx->small.p, 0, 4, st));
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->main_rgba.p, rgba_inout, n * 4, cudaMemcpyHostToDevice, st));
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->rgba_out.p, other_rgba, no * 4, cudaMemcpyHostToDevice, st));
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->mask.p, mask, n, cudaMemcpyHostToDevice, st));
    const size_t items = (size_t)(x1 - start_x) * (y1 - start_y);
    const uint32_t grid = (uint32_t)std::min<size_t>((items + 255) / 256, (size_t)ctx->sm_count * 16);
    image_replace_kernel<<<grid, 256, 0, st>>>((uint32_t *)ctx->main_rgba.p, (const uint8_t *)ctx->mask.p, w, h,
                                               (const uint32_t *)ctx->rgba_out.p, ow, oh, start_x, start_y, x1, y1,
                                               (int32_t *)ctx->small.p);
    LAUNCH_CHECK(ctx);
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->h_small, ctx->small.p, 4, cudaMemcpyDeviceToHost, st));
    STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(st));
    const int32_t status = *(const int32_t *)ctx->h_small;
    if (status != STEPTH_OK)
        return set_err(ctx, status, "image_replace: a masked pixel of the window lies outside `other` (mask_image.rs:106)");
    STEPTH_CUDA_TRY(ctx, cudaMemcpy(rgba_inout, ctx->main_rgba.p, n * 4, cudaMemcpyDeviceToHost));
    return STEPTH_OK;
}

// ---- operations.rs ------------------------------------------------------------------------------------------
static int32_t brightness(stepth_ctx *ctx, const uint16_t *img1, size_t n1, const uint16_t *img2, size_t n2,
                          double percent, uint16_t *out, uint32_t nch) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!img1 || !img2 || !out || n1 == 0 || n2 == 0)  // len == 0: the reference divides by zero
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "normalize_brightness: null or empty image");
    STEPTH_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const size_t s1 = n1 * nch, s2 = n2 * nch;
    TRY_RC(ensure(ctx, ctx->u16a, s1 * 2));
    TRY_RC(ensure(ctx, ctx->u16b, s2 * 2));
    TRY_RC(ensure(ctx, ctx->u16o, s1 * 2));
    TRY_RC(ensure(ctx, ctx->work, 64));
    unsigned long long *sums = (unsigned long long *)ctx->work.p;
    STEPTH_CUDA_TRY(ctx, cudaMemsetAsync(sums, 0, 48, st));
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->u16a.p, img1, s1 * 2, cudaMemcpyHostToDevice, st));
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->u16b.p, img2, s2 * 2, cudaMemcpyHostToDevice, st));
    auto grid_for = [&](size_t n_sub) {
        return (uint32_t)std::max<size_t>(1, std::min<size_t>((n_sub / 3 + 255) / 256, (size_t)ctx->sm_count * 8));
    };
    u16_channel_sums_kernel<<<grid_for(s1), 256, 0, st>>>((const uint16_t *)ctx->u16a.p, s1, nch, sums);
    LAUNCH_CHECK(ctx);
    u16_channel_sums_kernel<<<grid_for(s2), 256, 0, st>>>((const uint16_t *)ctx->u16b.p, s2, nch, sums + 3);
    LAUNCH_CHECK(ctx);
    u16_scale_kernel<<<grid_for(s1), 256, 0, st>>>((const uint16_t *)ctx->u16a.p, (uint16_t *)ctx->u16o.p, s1, s2, nch,
                                                   percent, sums);
    LAUNCH_CHECK(ctx);
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(out, ctx->u16o.p, s1 * 2, cudaMemcpyDeviceToHost, st));
    STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(st));
    return STEPTH_OK;
}

extern "C" int32_t stepth_cuda_normalize_brightness_luma16(stepth_ctx *ctx, const uint16_t *img1, size_t n1,
                                                           const uint16_t *img2, size_t n2, double percent,
                                                           uint16_t *out) {
    return brightness(ctx, img1, n1, img2, n2, percent, out, 1);
}

extern "C" int32_t stepth_cuda_normalize_brightness_rgb16(stepth_ctx *ctx, const uint16_t *img1, size_t n1,
                                                          const uint16_t *img2, size_t n2, double percent,
                                                          uint16_t *out) {
    return brightness(ctx, img1, n1, img2, n2, percent, out, 3);
}

// measured integer-ALU peak (LOP3/IADD3 class) of the context's device: bench.py's int-roofline denominator
extern "C" int32_t stepth_cuda_int_peak(stepth_ctx *ctx, int32_t logic_only, double *ops_per_second_out) {
    if (!ctx || !ops_per_second_out) return set_err(ctx, STEPTH_ERR_INVALID_ARG, "int_peak: null argument");
    STEPTH_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    return measure_int_peak(ctx, logic_only != 0, ops_per_second_out);
}

// ============================================================================================================
// 4. host-side stand-in for disage::open::rgb_discrete (depth_image.rs:104-109)
// ============================================================================================================
// "disc_v0" is NOT disage (its source is absent): a kd-split of the longer side (ties ->
// width) at floor(len / 2); always split while depth < min_splits; stop at depth == max_splits, at 1x1, or
// when every pixel of the region is within `precision` of the region's per-channel mean (floor).  The block
// value is that mean.  GPU parity never depends on it: blocks are an INPUT of the boundary.
namespace {
struct Disc {
    const uint8_t *px;
    uint32_t stride, w, h;
    uint32_t prec[3];
    uint32_t min_s, max_s;
    stepth_block *out;
    size_t cap, n;

    void emit(uint32_t x, uint32_t y, uint32_t bw, uint32_t bh, const uint32_t mean[3]) {
        if (n < cap) {
            stepth_block b;
            b.x = x; b.y = y; b.w = bw; b.h = bh;
            b.value[0] = (uint8_t)mean[0]; b.value[1] = (uint8_t)mean[1]; b.value[2] = (uint8_t)mean[2];
            b.pad = 0;
            out[n] = b;
        }
        ++n;
    }
    void mean_of(uint32_t x, uint32_t y, uint32_t bw, uint32_t bh, uint32_t mean[3]) const {
        uint64_t s[3] = {0, 0, 0};
        for (uint32_t yy = y; yy < y + bh; ++yy) {
            const uint8_t *row = px + ((size_t)yy * w + x) * stride;
            for (uint32_t xx = 0; xx < bw; ++xx, row += stride) { s[0] += row[0]; s[1] += row[1]; s[2] += row[2]; }
        }
        uint64_t cnt = (uint64_t)bw * bh;
        for (int c = 0; c < 3; ++c) mean[c] = (uint32_t)(s[c] / cnt);
    }
    bool uniform(uint32_t x, uint32_t y, uint32_t bw, uint32_t bh, const uint32_t mean[3]) const {
        for (uint32_t yy = y; yy < y + bh; ++yy) {
            const uint8_t *row = px + ((size_t)yy * w + x) * stride;
            for (uint32_t xx = 0; xx < bw; ++xx, row += stride)
                for (int c = 0; c < 3; ++c) {
                    int d = (int)row[c] - (int)mean[c];
                    if (d < 0) d = -d;
                    if (!((uint32_t)d < prec[c])) return false;
                }
        }
        return true;
    }
    void rec(uint32_t x, uint32_t y, uint32_t bw, uint32_t bh, uint32_t depth) {
        const bool can_split = (uint64_t)bw * bh > 1 && depth < max_s;
        bool split = can_split && depth < min_s;
        uint32_t mean[3] = {0, 0, 0};
        if (!split) {
            mean_of(x, y, bw, bh, mean);
            if (can_split && !uniform(x, y, bw, bh, mean)) split = true;
        }
        if (!split) { emit(x, y, bw, bh, mean); return; }
        if (bw >= bh) {
            uint32_t l = bw / 2;
            rec(x, y, l, bh, depth + 1);
            rec(x + l, y, bw - l, bh, depth + 1);
        } else {
            uint32_t t = bh / 2;
            rec(x, y, bw, t, depth + 1);
            rec(x, y + t, bw, bh - t, depth + 1);
        }
    }
};
}  // namespace

extern "C" size_t stepth_host_discretise_v0(const uint8_t *main_px, uint32_t stride, uint32_t w, uint32_t h,
                                            const uint8_t precision[3], uint32_t min_splits, uint32_t max_splits,
                                            stepth_block *blocks_out, size_t cap) {
    if (!main_px || !precision || (stride != 3 && stride != 4) || w == 0 || h == 0) return 0;
    Disc d;
    d.px = main_px; d.stride = stride; d.w = w; d.h = h;
    d.prec[0] = precision[0]; d.prec[1] = precision[1]; d.prec[2] = precision[2];
    d.min_s = min_splits; d.max_s = max_splits;
    d.out = blocks_out; d.cap = blocks_out ? cap : 0; d.n = 0;
    d.rec(0, 0, w, h, 0);
    return d.n;
}

// ---- disc_v0 on the device (disc.cuh) ----------------------------------------------------------------------------------------
namespace {
// count pass + scan: afterwards a.counts holds every top-level region's offset and, last, the number of blocks
int32_t disc_count(stepth_ctx *ctx, cudaStream_t st, const uint8_t *px_dev, uint32_t stride, uint32_t w, uint32_t h,
                   const uint8_t precision[3], uint32_t min_splits, uint32_t max_splits, stepth::DiscArgs &a) {
    using namespace stepth;
    memset(&a, 0, sizeof a);
    a.px = px_dev; a.stride = stride; a.w = w; a.h = h;
    a.pr = precision[0]; a.pg = precision[1]; a.pb = precision[2];
    a.min_s = min_splits; a.max_s = max_splits;
    a.top = std::min(std::min(min_splits, max_splits), DISC_MAX_TOP);
    const uint32_t n_top = 1u << a.top;
    int32_t rc = ensure(ctx, ctx->disc_counts, sizeof(uint32_t) * ((size_t)n_top + 1));
    if (rc != STEPTH_OK) return rc;
    a.counts = (uint32_t *)ctx->disc_counts.p;
    disc_kernel<false><<<(n_top + 127) / 128, 128, 0, st>>>(a);
    ctx->launches++;
    STEPTH_CUDA_TRY(ctx, cudaGetLastError());
    disc_scan_kernel<<<1, DISC_SCAN_THREADS, 0, st>>>(a.counts, n_top);
    ctx->launches++;
    STEPTH_CUDA_TRY(ctx, cudaGetLastError());
    return STEPTH_OK;
}
int32_t disc_write(stepth_ctx *ctx, cudaStream_t st, stepth::DiscArgs &a, stepth_block *out_dev, size_t cap) {
    using namespace stepth;
    a.out = out_dev; a.cap = cap;
    const uint32_t n_top = 1u << a.top;
    disc_kernel<true><<<(n_top + 127) / 128, 128, 0, st>>>(a);
    ctx->launches++;
    STEPTH_CUDA_TRY(ctx, cudaGetLastError());
    return STEPTH_OK;
}
bool disc_args_ok(const void *px, uint32_t stride, uint32_t w, uint32_t h, const uint8_t *precision) {
    return px && precision && (stride == 3 || stride == 4) && w != 0 && h != 0 && (uint64_t)w * h <= 0xfffffff0ull;
}
// uploads the image and runs the count pass; the number of blocks comes back through the pinned small buffer
int32_t disc_host_count(stepth_ctx *ctx, const uint8_t *main_px, uint32_t stride, uint32_t w, uint32_t h, const uint8_t precision[3],
                        uint32_t min_splits, uint32_t max_splits, stepth::DiscArgs &a, size_t *n_out) {
    using namespace stepth;
    STEPTH_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    const size_t bytes = (size_t)w * h * stride;
    int32_t rc = ensure(ctx, ctx->main_rgba, (size_t)w * h * 4);
    if (rc == STEPTH_OK) rc = ensure_host_small(ctx, 64);
    if (rc != STEPTH_OK) return rc;
    cudaStream_t st = ctx->stream;
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->main_rgba.p, main_px, bytes, cudaMemcpyHostToDevice, st));
    rc = disc_count(ctx, st, (const uint8_t *)ctx->main_rgba.p, stride, w, h, precision, min_splits, max_splits, a);
    if (rc != STEPTH_OK) return rc;
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->h_small, a.counts + (1u << a.top), sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(st));
    *n_out = *(const uint32_t *)ctx->h_small;
    return STEPTH_OK;
}
}  // namespace

extern "C" int32_t stepth_cuda_dev_discretise_v0(stepth_ctx *ctx, void *stream, const uint8_t *main_px, uint32_t stride, uint32_t w,
                                                 uint32_t h, const uint8_t precision[3], uint32_t min_splits, uint32_t max_splits,
                                                 stepth_block *blocks_out, size_t cap, uint32_t *n_out) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!disc_args_ok(main_px, stride, w, h, precision) || ((uintptr_t)blocks_out & 3))
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "dev_discretise_v0: null or invalid argument");
    cudaStream_t st = (cudaStream_t)stream;
    stepth::DiscArgs a;
    TRY_RC(disc_count(ctx, st, main_px, stride, w, h, precision, min_splits, max_splits, a));
    if (blocks_out && cap) TRY_RC(disc_write(ctx, st, a, blocks_out, cap));
    if (n_out) STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(n_out, a.counts + (1u << a.top), sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
    return STEPTH_OK;
}

// ---- the device discretiser without any allocation ---------------------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc (CUDA
// graph capture, tight serving loops).  The per-region counts -- the only scratch of disc_kernel / disc_scan_kernel, which
// stepth_cuda_dev_discretise_v0 keeps in a context buffer that grows with cudaMalloc -- live in a file-scope device array of
// a fixed size instead: every GPU that loads this library holds its STEPTH_NOALLOC_DISC_SCRATCH_BYTES until the process exits.
// Every call on a device shares the array: calls that overlap on that GPU (two streams, two threads) race on it.  Nothing in
// it has to start at zero: pass 1 writes counts[0 .. 2^top), the scan writes counts[2^top], and only those words are read.
static_assert(STEPTH_NOALLOC_DISC_MAX_TOP <= stepth::DISC_MAX_TOP, "the kernels' own limit");
static_assert(STEPTH_NOALLOC_DISC_SCRATCH_BYTES >= sizeof(uint32_t) * ((1u << STEPTH_NOALLOC_DISC_MAX_TOP) + 1), "counts + the total");
static __device__ __align__(256) uint32_t g_noalloc_disc_counts[STEPTH_NOALLOC_DISC_SCRATCH_BYTES / sizeof(uint32_t)];

namespace {
constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mu;
uint32_t *g_noalloc_counts_dev[NOALLOC_MAX_DEVICES] = {};   // the array's DEVICE address, per device (the name above is a host-side stand-in)

// once per device, with that device current: where the array lives in that GPU's memory
int32_t noalloc_counts(stepth_ctx *ctx, uint32_t **out) {
    if (ctx->device < 0 || ctx->device >= NOALLOC_MAX_DEVICES)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "dev_discretise_v0_noalloc: device %d beyond the %d this entry point keeps addresses for",
                       ctx->device, NOALLOC_MAX_DEVICES);
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_counts_dev[ctx->device]) {
        int cur = -1;
        STEPTH_CUDA_TRY(ctx, cudaGetDevice(&cur));
        if (cur != ctx->device) STEPTH_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
        uint32_t *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void **)&p, g_noalloc_disc_counts);
        if (cur != ctx->device && cur >= 0) cudaSetDevice(cur);
        if (e != cudaSuccess || !p)
            return set_err(ctx, STEPTH_ERR_CUDA, "cudaGetSymbolAddress(g_noalloc_disc_counts) failed on device %d: %s", ctx->device,
                           cudaGetErrorString(e));
        g_noalloc_counts_dev[ctx->device] = p;
    }
    *out = g_noalloc_counts_dev[ctx->device];
    return STEPTH_OK;
}
}  // namespace

extern "C" int32_t stepth_cuda_dev_discretise_v0_noalloc(stepth_ctx *ctx, void *stream, const uint8_t *main_px, uint32_t stride,
                                                         uint32_t w, uint32_t h, const uint8_t precision[3], uint32_t min_splits,
                                                         uint32_t max_splits, stepth_block *blocks_out, size_t cap, uint32_t *n_out) {
    using namespace stepth;
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!disc_args_ok(main_px, stride, w, h, precision) || ((uintptr_t)blocks_out & 3))
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "dev_discretise_v0_noalloc: null or invalid argument");
    const uint32_t top = std::min(min_splits, max_splits);
    if (top > STEPTH_NOALLOC_DISC_MAX_TOP)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG,
                       "dev_discretise_v0_noalloc: min(min_splits, max_splits) = %u asks for 2^%u forced regions; the static scratch holds "
                       "2^%d (%d bytes): use stepth_cuda_dev_discretise_v0",
                       top, top, STEPTH_NOALLOC_DISC_MAX_TOP, STEPTH_NOALLOC_DISC_SCRATCH_BYTES);
    if ((uint64_t)w * h > STEPTH_NOALLOC_DISC_MAX_PIXELS)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG,
                       "dev_discretise_v0_noalloc: %u x %u pixels; the limit of this entry point is %llu: use stepth_cuda_dev_discretise_v0", w, h,
                       (unsigned long long)STEPTH_NOALLOC_DISC_MAX_PIXELS);
    uint32_t *counts = nullptr;
    TRY_RC(noalloc_counts(ctx, &counts));
    cudaStream_t st = (cudaStream_t)stream;
    DiscArgs a;
    memset(&a, 0, sizeof a);
    a.px = main_px; a.stride = stride; a.w = w; a.h = h;
    a.pr = precision[0]; a.pg = precision[1]; a.pb = precision[2];
    a.min_s = min_splits; a.max_s = max_splits;
    a.top = top;
    a.counts = counts;
    const uint32_t n_top = 1u << a.top;
    disc_kernel<false><<<(n_top + 127) / 128, 128, 0, st>>>(a);
    ctx->launches++;
    STEPTH_CUDA_TRY(ctx, cudaGetLastError());
    disc_scan_kernel<<<1, DISC_SCAN_THREADS, 0, st>>>(a.counts, n_top);
    ctx->launches++;
    STEPTH_CUDA_TRY(ctx, cudaGetLastError());
    if (blocks_out && cap) {
        a.out = blocks_out; a.cap = cap;
        disc_kernel<true><<<(n_top + 127) / 128, 128, 0, st>>>(a);
        ctx->launches++;
        STEPTH_CUDA_TRY(ctx, cudaGetLastError());
    }
    if (n_out) STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(n_out, a.counts + n_top, sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
    return STEPTH_OK;
}

extern "C" int32_t stepth_cuda_discretise_v0(stepth_ctx *ctx, const uint8_t *main_px, uint32_t stride, uint32_t w, uint32_t h,
                                             const uint8_t precision[3], uint32_t min_splits, uint32_t max_splits,
                                             stepth_block *blocks_out, size_t cap, size_t *n_out) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!disc_args_ok(main_px, stride, w, h, precision) || !n_out)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "discretise_v0: null or invalid argument");
    stepth::DiscArgs a;
    size_t n = 0;
    TRY_RC(disc_host_count(ctx, main_px, stride, w, h, precision, min_splits, max_splits, a, &n));
    *n_out = n;
    const size_t n_write = std::min(n, blocks_out ? cap : (size_t)0);
    if (n_write) {
        TRY_RC(ensure(ctx, ctx->blocks, sizeof(stepth_block) * n_write));
        TRY_RC(disc_write(ctx, ctx->stream, a, (stepth_block *)ctx->blocks.p, n_write));
        STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(blocks_out, ctx->blocks.p, sizeof(stepth_block) * n_write, cudaMemcpyDeviceToHost, ctx->stream));
        STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    }
    return STEPTH_OK;
}

// load_depth_from_additional whole (depth_image.rs:91-136) with the stand-in discretiser: the main image goes up once, is cut
// into blocks on the device, and the block list never visits the host
extern "C" int32_t stepth_cuda_depth_from_additional(stepth_ctx *ctx, const uint8_t *main_rgba, uint32_t w, uint32_t h,
                                                     const uint8_t *add_rgb, uint32_t wa, uint32_t ha, const stepth_params *params,
                                                     uint32_t min_splits, uint32_t max_splits, uint8_t *depth_out, uint8_t *rgba_out,
                                                     uint8_t *mask_out, uint8_t *from_to_out, size_t *n_blocks_out) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!main_rgba || !add_rgb || !depth_out || !params_ok(params) || w == 0 || h == 0 || wa == 0 || ha == 0 ||
        (uint64_t)w * h > 0xfffffff0ull)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "depth_from_additional: null or invalid argument");
    stepth::DiscArgs a;
    size_t n = 0;
    TRY_RC(disc_host_count(ctx, main_rgba, 4, w, h, params->precision, min_splits, max_splits, a, &n));
    if (n_blocks_out) *n_blocks_out = n;
    if (n < 8) return set_err(ctx, STEPTH_ERR_TOO_FEW_BLOCKS, "fewer than 8 blocks: par_chunks_mut(0) panics (depth_image.rs:111-112)");
    TRY_RC(ensure(ctx, ctx->blocks, sizeof(stepth_block) * n));
    TRY_RC(disc_write(ctx, ctx->stream, a, (stepth_block *)ctx->blocks.p, n));
    ctx->blocks_on_device = 1;
    const int32_t rc = run_pairs(ctx, 1, nullptr, w, h, &add_rgb, wa, ha, (const stepth_block *)ctx->blocks.p, n, params, &depth_out,
                                 nullptr, &main_rgba, &rgba_out, mask_out ? &mask_out : nullptr, from_to_out, nullptr);
    ctx->blocks_on_device = 0;
    return rc;
}

// ---- decode / encode on the device (codec.cuh; nvJPEG) -----------------------------------------------------------------------
extern "C" int32_t stepth_cuda_decode_jpeg(stepth_ctx *ctx, const uint8_t *jpeg, size_t n_bytes, uint32_t *w_out, uint32_t *h_out,
                                           uint8_t *rgb_out, size_t cap) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!jpeg || n_bytes < 4 || !w_out || !h_out) return set_err(ctx, STEPTH_ERR_INVALID_ARG, "decode_jpeg: null or invalid argument");
    STEPTH_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    TRY_RC(codec_ensure(ctx));
    NvjpegApi &nj = nvjpeg_api();
    int ncomp = 0, ws[NVJPEG_MAX_COMPONENT] = {0, 0, 0, 0}, hs[NVJPEG_MAX_COMPONENT] = {0, 0, 0, 0};
    nvjpegChromaSubsampling_t css;
    if (nj.GetImageInfo((nvjpegHandle_t)ctx->nvj_handle, jpeg, n_bytes, &ncomp, &css, ws, hs) != NVJPEG_STATUS_SUCCESS || ws[0] <= 0 || hs[0] <= 0)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "Failed to open image");   // the reference's message (src/depth_image.rs:16)
    *w_out = (uint32_t)ws[0];
    *h_out = (uint32_t)hs[0];
    if (!rgb_out) return STEPTH_OK;
    const size_t bytes = (size_t)ws[0] * hs[0] * 3;
    if (cap < bytes) return set_err(ctx, STEPTH_ERR_SIZE_MISMATCH, "decode_jpeg: the output buffer holds %zu of %zu bytes", cap, bytes);
    TRY_RC(ensure(ctx, ctx->codec_a, bytes));
    nvjpegImage_t img;
    memset(&img, 0, sizeof img);
    img.channel[0] = (unsigned char *)ctx->codec_a.p;
    img.pitch[0] = (size_t)ws[0] * 3;
    STEPTH_NVJ_TRY(ctx, nj.Decode((nvjpegHandle_t)ctx->nvj_handle, (nvjpegJpegState_t)ctx->nvj_state, jpeg, n_bytes, NVJPEG_OUTPUT_RGBI, &img, ctx->stream));
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(rgb_out, ctx->codec_a.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return STEPTH_OK;
}

extern "C" int32_t stepth_cuda_encode_jpeg(stepth_ctx *ctx, const uint8_t *pixels, uint32_t w, uint32_t h, uint32_t channels,
                                           int32_t quality, uint8_t *jpeg_out, size_t cap, size_t *n_out) {
    if (!ctx) return STEPTH_ERR_INVALID_ARG;
    if (!pixels || !n_out || w == 0 || h == 0 || (channels != 1 && channels != 3 && channels != 4) || quality < 1 || quality > 100)
        return set_err(ctx, STEPTH_ERR_INVALID_ARG, "encode_jpeg: null or invalid argument");
    STEPTH_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    TRY_RC(codec_ensure(ctx));
    NvjpegApi &nj = nvjpeg_api();
    cudaStream_t st = ctx->stream;
    const size_t px = (size_t)w * h;
    TRY_RC(ensure(ctx, ctx->codec_a, px * channels));
    STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->codec_a.p, pixels, px * channels, cudaMemcpyHostToDevice, st));
    nvjpegHandle_t hd = (nvjpegHandle_t)ctx->nvj_handle;
    nvjpegEncoderState_t es = (nvjpegEncoderState_t)ctx->nvj_enc_state;
    nvjpegEncoderParams_t ep = (nvjpegEncoderParams_t)ctx->nvj_enc_params;
    STEPTH_NVJ_TRY(ctx, nj.EncoderParamsSetQuality(ep, quality, st));
    nvjpegImage_t img;
    memset(&img, 0, sizeof img);
    if (channels == 1) {   // Luma8 (DepthImage::depth().save): one luminance component
        STEPTH_NVJ_TRY(ctx, nj.EncoderParamsSetSamplingFactors(ep, NVJPEG_CSS_GRAY, st));
        img.channel[0] = (unsigned char *)ctx->codec_a.p;
        img.pitch[0] = w;
        STEPTH_NVJ_TRY(ctx, nj.EncodeYUV(hd, es, ep, &img, NVJPEG_CSS_GRAY, (int)w, (int)h, st));
    } else {
        STEPTH_NVJ_TRY(ctx, nj.EncoderParamsSetSamplingFactors(ep, NVJPEG_CSS_444, st));
        const uint8_t *rgb = (const uint8_t *)ctx->codec_a.p;
        if (channels == 4) {
            TRY_RC(ensure(ctx, ctx->codec_b, px * 3));
            rgba_to_rgb_kernel<<<(unsigned)std::min<size_t>((px + 255) / 256, (size_t)ctx->sm_count * 8), 256, 0, st>>>(
                (const uint32_t *)ctx->codec_a.p, (uint8_t *)ctx->codec_b.p, px);
            LAUNCH_CHECK(ctx);
            rgb = (const uint8_t *)ctx->codec_b.p;
        }
        img.channel[0] = (unsigned char *)rgb;
        img.pitch[0] = (size_t)w * 3;
        STEPTH_NVJ_TRY(ctx, nj.EncodeImage(hd, es, ep, &img, NVJPEG_INPUT_RGBI, (int)w, (int)h, st));
    }
    size_t len = 0;
    STEPTH_NVJ_TRY(ctx, nj.EncodeRetrieveBitstream(hd, es, nullptr, &len, st));
    *n_out = len;
    if (!jpeg_out) {   // sizing call: the stream stays in the encoder state until the next encode
        STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(st));
        return STEPTH_OK;
    }
    if (cap < len) return set_err(ctx, STEPTH_ERR_SIZE_MISMATCH, "encode_jpeg: the output buffer holds %zu of %zu bytes", cap, len);
    STEPTH_NVJ_TRY(ctx, nj.EncodeRetrieveBitstream(hd, es, jpeg_out, &len, st));
    STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(st));
    *n_out = len;
    return STEPTH_OK;
}
#include "multi_api.cuh"
