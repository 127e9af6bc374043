// ctx.cuh -- the library context: one device, one stream, scratch buffers that grow on demand, the cached
// Gaussian tap tables and the last error text.  Nothing here is visible through the C ABI except as an
// opaque `stepth_ctx *`.
#pragma once
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "finish.cuh"
#include "staging.cuh"

struct stepth_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;    // compute (and the single-pair copies)
    cudaStream_t copy_in = nullptr;   // batch entry point: host->device copies of the next chunk of pairs
    cudaStream_t copy_out = nullptr;  // batch entry point: device->host copies of the previous chunk
    cudaStream_t stream2 = nullptr;   // batch entry point: second compute stream (chunks alternate, so one chunk's first
                                      // kernels fill the SMs while the previous chunk's last, poorly filled ones drain)
    int slot = 0;                     // which scratch set (planes / worklist) the next search launch uses
    std::vector<cudaEvent_t> events;  // pool for the chunk hand-offs between the three streams
    std::string err;
    uint64_t launches = 0;
    int sm_count = 148;

    struct Buf {
        void *p = nullptr;
        size_t cap = 0;
    };
    // scratch for the host-buffer entry points
    Buf main_rgba, add_rgb, d8, depth, mask, rgba_out, blocks, plane, small, u16a, u16b, u16o, work;
    Buf worklist;  // two [count, cursor, pad..][uint2 entries] lists: queries handed from the near to the far kernel(s)
    int near_rings = 6;  // coarse rings the near kernel walks (STEPTH_NEAR_RINGS)
    int batch_chunks = 0;  // the batch entry point pipelines this many chunks of pairs over its three streams (STEPTH_BATCH_CHUNKS; 0 = 8, or 16 from 64 pairs on)
    cudaEvent_t main_ready = nullptr;  // set by the batch entry point: the dense search waits for it between K0 and K1 (the
                                       // main images are still on their way up while the planes of the additional ones are built)
    int trace = 0;  // STEPTH_TRACE=1: the batch entry point prints each chunk's copy-in / compute / copy-out end times to stderr
    int upload_full_rows = 0;  // host-buffer entry points upload whole rows of the additional image (STEPTH_UPLOAD_FULL_ROWS)
    int near_tile = 3;     // dense mode's near stage (STEPTH_NEAR_TILE): 2 = one warp per origin tile (search_otile.cuh), 0 = the walking near kernel,
                           // 3 = by launch size (origin tiles except for mid-sized launches, search_launch.cuh), 1 = ring 0 + coarse rings 0..1 from
                           // TMA-staged tile records (search_tile.cuh; measured slower than the walk: 2.59 vs 2.26 ms per 16 pairs 1080p)
    int near_rings_otile = 12;  // coarse rings the origin-tile near stage walks, 2..16 (STEPTH_OT_RINGS; 16 pairs 1080p: 6 -> 2.17 ms, 12 -> 2.13, 16 -> 2.13)
    int otile_nit = 0;         // origin-tile stage: origin rows per warp, 8 (a whole tile) / 4 / 2 / 1; 0 = by launch size (STEPTH_OT_NIT)
    int otile_warps_per_sm = 128;  // ... so that the launch has at least this many warps per SM (STEPTH_OT_WARPS_PER_SM)
    int near_compact = 1;  // the near kernel packs the queries that survive ring 0 into full warps (STEPTH_NEAR_COMPACT)
    int mid_rings = 0;   // > near_rings: a second resumption point (FAR-1 walks rings [near_rings, mid_rings)); 0 = the far stage runs once (STEPTH_MID_RINGS)
    Buf planes2, worklist2, work2;  // scratch set of slot 1 (work / work2: the plane between the two streaming finish passes)
    int finish_streaming = 1;       // STEPTH_FINISH_STREAMING: 0 = always the tiled finish kernel
    int two_streams = 1;     // STEPTH_TWO_STREAMS
    stepth::Stager stager;   // pinned rings + copy threads for pageable caller buffers (batch path; created on first use)
    int stage_threads = 8;   // STEPTH_STAGE_THREADS: copy threads (the caller included; default min(16, usable cpus / 2)); 0 = hand pageable pointers to the driver
    Buf main_rgb;            // RGB8 main images as uploaded by the *_rgb entry point, expanded to main_rgba on the device
    uint8_t *finish_mirror[8] = {};  // row bands: where the next dev_finish call also stores its outputs (peer.cuh); cleared by the caller
    uint32_t finish_n_mirror = 0;
    int finish_mirrored = 0;         // set by dev_finish when it did
    int tail_kernel = 1;             // STEPTH_TAIL=0: the last search stage always walks its list with one lane per query
    int k0_tensor_map = 1;           // STEPTH_K0_TMAP=0: K0's strips by per-row bulk copies / plain loads only
    int peer_enabled = 1;            // STEPTH_PEER=0: the row-band exchanges go through NCCL instead of peer windows
    void *nvj_handle = nullptr, *nvj_state = nullptr, *nvj_enc_state = nullptr, *nvj_enc_params = nullptr;   // codec.cuh (nvJPEG, created on first use)
    Buf codec_a, codec_b;            // codec.cuh: pixels and their RGB8 form on the device
    Buf dbg;                         // -DSTEPTH_DEBUG_CHECKS builds: the mask of failed index checks of a search
    Buf disc_counts;         // device discretiser: leaves per top-level region, then their offsets (disc.cuh)
    int blocks_on_device = 0;  // run_pairs: the block list it is handed already lies in HBM (stepth_cuda_depth_from_additional)
    Buf planes;  // two-level match bit-planes of the additional image(s) (tiles.cuh), rebuilt by every search
    // small per-pair state: [max u32][hist 256 u32][status i32][n_out u32][ranges 512 B] per pair
    void *h_small = nullptr;  // pinned mirror
    size_t h_small_cap = 0;

    struct Taps {
        std::vector<stepth::TapRow> host;
        stepth::TapRow *dev = nullptr;
    };
    std::map<uint32_t, Taps> taps;  // keyed by axis length
};

namespace stepth {

inline int32_t set_err(stepth_ctx *ctx, int32_t code, const char *fmt, ...) {
    if (ctx) {
        char buf[512];
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, sizeof buf, fmt, ap);
        va_end(ap);
        ctx->err = buf;
    }
    return code;
}

#define STEPTH_CUDA_TRY(ctx, expr)                                                                          \
    do {                                                                                                    \
        cudaError_t e__ = (expr);                                                                           \
        if (e__ != cudaSuccess)                                                                             \
            return stepth::set_err((ctx), STEPTH_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,                 \
                                   cudaGetErrorString(e__), __FILE__, __LINE__);                            \
    } while (0)

inline int32_t ensure(stepth_ctx *ctx, stepth_ctx::Buf &b, size_t bytes) {
    if (bytes <= b.cap && b.p) return STEPTH_OK;
    if (b.p) STEPTH_CUDA_TRY(ctx, cudaFree(b.p));
    b.p = nullptr;
    b.cap = 0;
    size_t want = bytes < 256 ? 256 : bytes;
    STEPTH_CUDA_TRY(ctx, cudaMalloc(&b.p, want));
    b.cap = want;
    return STEPTH_OK;
}

// image 0.23.14 imageops/sample.rs: gaussian(x, r = 0.5) = 1/(sqrt(2 pi) r) * exp(-x^2 / (2 r^2)), support 3.0.
// Tap tables for a unit-scale resample along an axis of n samples (see finish.cuh).  Plain f32 arithmetic in
// the crate's order; this translation unit is compiled with -fmad=false / no fast-math on the host side.
inline float gaussian_r05(float x) {
    const float r = 0.5f;
    float norm = 1.0f / (sqrtf(2.0f * 3.14159265358979323846f) * r);
    return norm * expf(-(x * x) / (2.0f * (r * r)));
}

inline void build_taps(uint32_t n, std::vector<TapRow> &out) {
    out.resize(n);
    const float ratio = 1.0f, sratio = 1.0f, support = 3.0f * sratio;
    for (uint32_t o = 0; o < n; ++o) {
        float in = ((float)o + 0.5f) * ratio;
        long long left = (long long)floorf(in - support);
        if (left < 0) left = 0;
        if (left > (long long)n - 1) left = (long long)n - 1;
        long long right = (long long)ceilf(in + support);
        if (right < left + 1) right = left + 1;
        if (right > (long long)n) right = (long long)n;
        in = in - 0.5f;
        TapRow t;
        memset(&t, 0, sizeof t);
        t.left = (int32_t)left;
        float sum = 0.0f;
        int cnt = 0;
        for (long long i = left; i < right && cnt < 7; ++i) {
            volatile float wgt = gaussian_r05(((float)i - in) / sratio);
            t.w[cnt++] = wgt;
            sum = sum + wgt;
        }
        for (int k = 0; k < cnt; ++k) t.w[k] = t.w[k] / sum;
        t.cnt = cnt;
        out[o] = t;
    }
}

inline int32_t get_taps(stepth_ctx *ctx, cudaStream_t st, uint32_t n, const TapRow **dev_out, float *interior_out) {
    auto it = ctx->taps.find(n);
    if (it == ctx->taps.end()) {
        stepth_ctx::Taps t;
        build_taps(n, t.host);
        STEPTH_CUDA_TRY(ctx, cudaMalloc((void **)&t.dev, sizeof(TapRow) * (n ? n : 1)));
        // the table is immutable afterwards and shared by every later launch on any stream: the copy is ordered on the
        // caller's (non-blocking) stream and waited for, so the DMA itself -- not just its staging -- is over when this returns
        STEPTH_CUDA_TRY(ctx, cudaMemcpyAsync(t.dev, t.host.data(), sizeof(TapRow) * n, cudaMemcpyHostToDevice, st));
        STEPTH_CUDA_TRY(ctx, cudaStreamSynchronize(st));
        it = ctx->taps.emplace(n, std::move(t)).first;
    }
    for (int k = 0; k < 7; ++k) interior_out[k] = n >= 7 ? it->second.host[3].w[k] : 0.0f;
    *dev_out = it->second.dev;
    return STEPTH_OK;
}

// FinishArgs::wx_edge for a row of n >= 8 samples (the table is already cached by get_taps)
inline void edge_taps(stepth_ctx *ctx, uint32_t n, float (*edge)[7]) {
    const std::vector<TapRow> &t = ctx->taps.find(n)->second.host;
    for (int e = 0; e < 6; ++e) {
        const long long i = e < 3 ? e : (long long)n - 6 + e;   // the output column
        const TapRow &r = t[(size_t)i];
        for (int k = 0; k < 7; ++k) {
            const long long col = i - 3 + k, idx = col - r.left;
            edge[e][k] = (col >= 0 && col < (long long)n && idx >= 0 && idx < r.cnt) ? r.w[idx] : 0.0f;
        }
    }
}

}  // namespace stepth
