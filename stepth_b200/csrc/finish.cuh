// finish.cuh -- K2: normalise by the global max, unit-scale Gaussian resample, optional invert, histogram.
//
// Replaces depth_image.rs:124-134 (+ :225-227 when STEPTH_FLAG_INVERT is set) in ONE pass over the 1 B/px
// distance plane: each CTA stages a (TILE_H+6) x (TILE_W+6) tile, normalised through a 256-entry table
// (`(v * 255 / max) as u8`, depth_image.rs:127), runs the vertical then the horizontal 7-tap pass of
// image 0.23.14's `imageops::resize(.., Gaussian)` out of shared memory, and accumulates the 256-bin
// histogram + that depth_split (depth_image.rs:162-218) needs, so the plane is read once and written once.
//
// Bit-exactness of the f32 part: the per-index tap tables (left index, tap count, weights divided by their
// sequential sum) are computed on the HOST with the same libm `expf` the reference's f32::exp lowers to and
// handed in as arrays; the device only does `t = t + px * w` in tap order with __fmul_rn/__fadd_rn (no FMA
// contraction -- Rust never contracts), clamps to [0,255] and rounds half away from zero like f32::round.
#pragma once
#include "common.cuh"

namespace stepth {

// taps of output index o along one axis (image 0.23.14 imageops/sample.rs, ratio = 1)
struct TapRow {
    int32_t left;  // first contributing input index
    int32_t cnt;   // number of taps (7 in the interior, fewer at the borders / on tiny images)
    float w[7];    // normalised weights, tap order
};

constexpr int FIN_TILE_W = 128;
constexpr int FIN_TILE_H = 32;
constexpr int FIN_HALO = 3;
constexpr int FIN_SW = FIN_TILE_W + 2 * FIN_HALO;  // 134 staged columns
constexpr int FIN_SH = FIN_TILE_H + 2 * FIN_HALO;  // 38 staged rows
constexpr int FIN_SWP = FIN_SW + 2;                // padded row pitch

struct FinishArgs {
    const uint8_t *src;      // [n, src_rows, w] raw `d as u8` plane, global rows [src_y0, src_y0 + src_rows)
    uint8_t *dst;            // [n, out_rows, w] global rows [out_y0, out_y0 + out_rows)
    const uint32_t *max_in;  // [n]
    const TapRow *taps_y;    // [img_h]
    const TapRow *taps_x;    // [w]
    uint32_t *hist;          // [n, 256] accumulated, may be null
    int32_t *status;         // [n], may be null
    uint32_t w, img_h, src_y0, src_rows, out_y0, out_rows;
    uint32_t mode;           // STEPTH_RESAMPLE_*
    uint32_t invert;
};

// clamp(t, 0.0, 255.0) then f32::round (half away from zero), as NumCast::from(FloatNearest(t)) does
__device__ __forceinline__ uint8_t to_u8_nearest(float t) {
    t = fminf(fmaxf(t, 0.0f), 255.0f);  // clamp(t, 0.0, max)
    return (uint8_t)roundf(t);         // NumCast::from(FloatNearest(t)) -> t.round()
}

__global__ void __launch_bounds__(256) finish_kernel(FinishArgs a) {
    __shared__ uint8_t lut[256];
    __shared__ uint8_t s_src[FIN_SH][FIN_SWP];
    __shared__ float s_v[FIN_TILE_H][FIN_SWP];
    __shared__ uint32_t s_hist[256];

    const uint32_t pair = blockIdx.z;
    const uint32_t mx = a.max_in[pair];
    const int tid = threadIdx.x;
    if (mx == 0) {  // depth_image.rs:127 divides by zero -> panic; reported per pair, nothing written
        if (a.status && blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) a.status[pair] = STEPTH_ERR_DIV_BY_ZERO_MAX;
        return;
    }
    lut[tid] = (uint8_t)(((uint32_t)tid * 255u) / mx);  // (v as u64 * 255 / max as u64) as u8; v <= max
    s_hist[tid] = 0;

    const int tx0 = blockIdx.x * FIN_TILE_W;                    // first output column of the tile
    const int ty0 = (int)a.out_y0 + blockIdx.y * FIN_TILE_H;    // first output row (global)
    const int w = (int)a.w, img_h = (int)a.img_h;
    const int y_end = min((int)(a.out_y0 + a.out_rows), ty0 + FIN_TILE_H);
    const uint8_t *src = a.src + (size_t)pair * a.src_rows * a.w;
    uint8_t *dst = a.dst + (size_t)pair * a.out_rows * a.w;
    __syncthreads();

    // stage + normalise; out-of-image cells are never referenced by a tap table, fill with 0
    for (int i = tid; i < FIN_SH * FIN_SW; i += 256) {
        int r = i / FIN_SW, c = i - r * FIN_SW;
        int gy = ty0 - FIN_HALO + r, gx = tx0 - FIN_HALO + c;
        uint8_t v = 0;
        int ly = gy - (int)a.src_y0;
        if (gx >= 0 && gx < w && gy >= 0 && gy < img_h && ly >= 0 && ly < (int)a.src_rows)
            v = lut[src[(size_t)ly * w + gx]];
        s_src[r][c] = v;
    }
    __syncthreads();

    // a thread walks down one column (i += 256 advances two rows): neighbouring outputs are mostly equal, so
    // the shared-memory histogram is fed run-length encoded to keep same-address atomics rare
    uint32_t run_val = 0, run_len = 0;
    auto count = [&](uint8_t o) {
        if (run_len && o != run_val) { atomicAdd(&s_hist[run_val], run_len); run_len = 0; }
        run_val = o;
        ++run_len;
    };
    if (a.mode == STEPTH_RESAMPLE_OFF) {
        for (int i = tid; i < FIN_TILE_H * FIN_TILE_W; i += 256) {
            int r = i / FIN_TILE_W, c = i - r * FIN_TILE_W;
            int gy = ty0 + r, gx = tx0 + c;
            if (gy < y_end && gx < w) {
                uint8_t o = s_src[r + FIN_HALO][c + FIN_HALO];
                if (a.invert) o = (uint8_t)(255 - o);
                dst[(size_t)(gy - (int)a.out_y0) * w + gx] = o;
                if (a.hist) count(o);
            }
        }
    } else {
        // vertical_sample: t = sum_k px(x, left + k) * w_k
        for (int i = tid; i < FIN_TILE_H * FIN_SW; i += 256) {
            int r = i / FIN_SW, c = i - r * FIN_SW;
            int gy = ty0 + r;
            float t = 0.0f;
            if (gy < y_end) {
                const TapRow &tp = a.taps_y[gy];
                int base = tp.left - (ty0 - FIN_HALO);
                for (int k = 0; k < tp.cnt; ++k)
                    t = __fadd_rn(t, __fmul_rn((float)s_src[base + k][c], tp.w[k]));
                if (a.mode == STEPTH_RESAMPLE_U8) t = (float)to_u8_nearest(t);
            }
            s_v[r][c] = t;
        }
        __syncthreads();
        // horizontal_sample
        for (int i = tid; i < FIN_TILE_H * FIN_TILE_W; i += 256) {
            int r = i / FIN_TILE_W, c = i - r * FIN_TILE_W;
            int gy = ty0 + r, gx = tx0 + c;
            if (gy < y_end && gx < w) {
                const TapRow &tp = a.taps_x[gx];
                int base = tp.left - (tx0 - FIN_HALO);
                float t = 0.0f;
                for (int k = 0; k < tp.cnt; ++k) t = __fadd_rn(t, __fmul_rn(s_v[r][base + k], tp.w[k]));
                uint8_t o = to_u8_nearest(t);
                if (a.invert) o = (uint8_t)(255 - o);
                dst[(size_t)(gy - (int)a.out_y0) * w + gx] = o;
                if (a.hist) count(o);
            }
        }
    }
    if (a.hist) {
        if (run_len) atomicAdd(&s_hist[run_val], run_len);
        __syncthreads();
        uint32_t c = s_hist[tid];
        if (c) atomicAdd(&a.hist[(size_t)pair * 256 + tid], c);
    }
}

}  // namespace stepth
