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
    float wy_int[7], wx_int[7];  // the taps of an interior index of each axis (all interior indices share them)
};

// clamp(t, 0.0, 255.0) then f32::round (half away from zero), as NumCast::from(FloatNearest(t)) does
__device__ __forceinline__ uint8_t to_u8_nearest(float t) {
    t = fminf(fmaxf(t, 0.0f), 255.0f);  // clamp(t, 0.0, max)
    return (uint8_t)roundf(t);         // NumCast::from(FloatNearest(t)) -> t.round()
}

// one output sample along an axis: the 7-tap interior case uses the constant weights (all interior indices share them:
// at unit scale the tap offsets i - c are the integers -3 .. 3), the border case walks the host-built table
__device__ __forceinline__ float taps7(const float *v, const float (&w)[7]) {
    float t = 0.0f;
#pragma unroll
    for (int k = 0; k < 7; ++k) t = __fadd_rn(t, __fmul_rn(v[k], w[k]));
    return t;
}

constexpr int FIN_VSEG = 8;    // outputs per thread in the vertical pass (one column, 8 rows)
constexpr int FIN_HSEG = 16;   // outputs per thread in the horizontal pass (one row, 16 columns)
constexpr int FIN_PITCH = 137; // odd row pitch (floats): column walks and row walks are both bank-conflict free
static_assert(FIN_SW <= FIN_PITCH && FIN_TILE_H % FIN_VSEG == 0 && FIN_TILE_W % FIN_HSEG == 0, "tile geometry");
static_assert((FIN_TILE_W / FIN_HSEG) * FIN_TILE_H == 256, "horizontal pass: one item per thread");

__global__ void __launch_bounds__(256) finish_kernel(FinishArgs a) {
    __shared__ float lut[256];
    __shared__ float s_src[FIN_SH][FIN_PITCH];       // normalised input tile + halo, as f32
    __shared__ float s_v[FIN_TILE_H][FIN_PITCH];     // after the vertical pass
    __shared__ __align__(16) uint8_t s_out[FIN_TILE_H][FIN_TILE_W];
    __shared__ uint32_t s_hist[256];

    const uint32_t pair = blockIdx.z;
    const uint32_t mx = a.max_in[pair];
    const int tid = threadIdx.x;
    if (mx == 0) {  // depth_image.rs:127 divides by zero -> panic; reported per pair, nothing written
        if (a.status && blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) a.status[pair] = STEPTH_ERR_DIV_BY_ZERO_MAX;
        return;
    }
    lut[tid] = (float)(uint8_t)(((uint32_t)tid * 255u) / mx);  // (v as u64 * 255 / max as u64) as u8; v <= max
    s_hist[tid] = 0;

    const int tx0 = blockIdx.x * FIN_TILE_W;                    // first output column of the tile
    const int ty0 = (int)a.out_y0 + blockIdx.y * FIN_TILE_H;    // first output row (global)
    const int w = (int)a.w, img_h = (int)a.img_h;
    const int y_end = min((int)(a.out_y0 + a.out_rows), ty0 + FIN_TILE_H);
    const uint8_t *src = a.src + (size_t)pair * a.src_rows * a.w;
    uint8_t *dst = a.dst + (size_t)pair * a.out_rows * a.w;
    __syncthreads();

    // stage + normalise; out-of-image cells are never referenced by a tap table, fill with 0
    if ((w & 3) == 0 && ((uintptr_t)src & 3) == 0) {
        // four pixels per load: words cover columns [tx0 - 4, tx0 + 132); staged column = global column - (tx0 - 3)
        constexpr int WORDS = (FIN_TILE_W + 8) / 4;
        for (int i = tid; i < FIN_SH * WORDS; i += 256) {
            const int r = i / WORDS, wd = i - r * WORDS;
            const int gy = ty0 - FIN_HALO + r, gx0 = tx0 - 4 + 4 * wd;
            const int ly = gy - (int)a.src_y0;
            uint32_t px = 0;
            const bool ok = gx0 >= 0 && gx0 < w && gy >= 0 && gy < img_h && ly >= 0 && ly < (int)a.src_rows;
            if (ok) px = *reinterpret_cast<const uint32_t *>(src + (size_t)ly * w + gx0);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int c = 4 * wd + k - 1;
                if (c >= 0 && c < FIN_SW) s_src[r][c] = ok ? lut[(px >> (8 * k)) & 0xffu] : 0.0f;
            }
        }
    } else {
        for (int i = tid; i < FIN_SH * FIN_SW; i += 256) {
            const int r = i / FIN_SW, c = i - r * FIN_SW;
            const int gy = ty0 - FIN_HALO + r, gx = tx0 - FIN_HALO + c;
            float v = 0.0f;
            const int ly = gy - (int)a.src_y0;
            if (gx >= 0 && gx < w && gy >= 0 && gy < img_h && ly >= 0 && ly < (int)a.src_rows)
                v = lut[src[(size_t)ly * w + gx]];
            s_src[r][c] = v;
        }
    }
    __syncthreads();

    if (a.mode == STEPTH_RESAMPLE_OFF) {
        for (int i = tid; i < FIN_TILE_H * FIN_TILE_W; i += 256) {
            const int r = i / FIN_TILE_W, c = i - r * FIN_TILE_W;
            s_out[r][c] = (uint8_t)s_src[r + FIN_HALO][c + FIN_HALO];
        }
    } else {
        // ---- vertical_sample: t = sum_k px(x, left + k) * w_k; item = (column, 8-row segment), columns fastest ----------
        float wy[7], wx[7];  // tiny images have no interior index and never take the interior path
#pragma unroll
        for (int k = 0; k < 7; ++k) { wy[k] = a.wy_int[k]; wx[k] = a.wx_int[k]; }
        for (int it = tid; it < FIN_SW * (FIN_TILE_H / FIN_VSEG); it += 256) {
            const int seg = it / FIN_SW, c = it - seg * FIN_SW;
            float v[FIN_VSEG + 6];
#pragma unroll
            for (int k = 0; k < FIN_VSEG + 6; ++k) v[k] = s_src[seg * FIN_VSEG + k][c];
#pragma unroll
            for (int j = 0; j < FIN_VSEG; ++j) {
                const int r = seg * FIN_VSEG + j, gy = ty0 + r;
                float t = 0.0f;
                if (gy < y_end) {
                    if (gy >= 3 && gy + 4 <= img_h) {
                        t = taps7(v + j, wy);
                    } else {
                        const TapRow &tp = a.taps_y[gy];
                        const int base = tp.left - (ty0 - FIN_HALO);
                        for (int k = 0; k < tp.cnt; ++k) t = __fadd_rn(t, __fmul_rn(s_src[base + k][c], tp.w[k]));
                    }
                    if (a.mode == STEPTH_RESAMPLE_U8) t = (float)to_u8_nearest(t);
                }
                s_v[r][c] = t;
            }
        }
        __syncthreads();
        // ---- horizontal_sample: item = (row, 16-column segment), rows fastest -------------------------------------------
        {
            const int r = tid & (FIN_TILE_H - 1), seg = tid / FIN_TILE_H;
            const int gy = ty0 + r;
            float v[FIN_HSEG + 6];
#pragma unroll
            for (int k = 0; k < FIN_HSEG + 6; ++k) v[k] = s_v[r][seg * FIN_HSEG + k];
#pragma unroll
            for (int j = 0; j < FIN_HSEG; ++j) {
                const int c = seg * FIN_HSEG + j, gx = tx0 + c;
                uint8_t o = 0;
                if (gy < y_end && gx < w) {
                    float t = 0.0f;
                    if (gx >= 3 && gx + 4 <= w) {
                        t = taps7(v + j, wx);
                    } else {
                        const TapRow &tp = a.taps_x[gx];
                        const int base = tp.left - (tx0 - FIN_HALO);
                        for (int k = 0; k < tp.cnt; ++k) t = __fadd_rn(t, __fmul_rn(s_v[r][base + k], tp.w[k]));
                    }
                    o = to_u8_nearest(t);
                }
                s_out[r][c] = o;
            }
        }
    }
    __syncthreads();
    // ---- invert, store (coalesced: a warp writes 4 rows x 128 bytes), histogram ------------------------------------------
    {
        const int r = tid >> 3, c0 = (tid & 7) * 16;
        const int gy = ty0 + r;
        if (gy < y_end && tx0 + c0 < w) {
            uint4 q = *reinterpret_cast<const uint4 *>(&s_out[r][c0]);
            if (a.invert) { q.x = ~q.x; q.y = ~q.y; q.z = ~q.z; q.w = ~q.w; }
            uint8_t *drow = dst + (size_t)(gy - (int)a.out_y0) * w + tx0 + c0;
            const int n = min(16, w - (tx0 + c0));
            const uint8_t inv = a.invert ? 0xffu : 0u;
            const uint8_t *orow = &s_out[r][c0];
            if (n == 16 && ((uintptr_t)drow & 15) == 0) {
                *reinterpret_cast<uint4 *>(drow) = q;
            } else {
                for (int k = 0; k < n; ++k) drow[k] = orow[k] ^ inv;
            }
            if (a.hist) {
                // neighbouring outputs are mostly equal: feed the shared-memory histogram run-length encoded
                uint32_t run_val = orow[0] ^ inv, run_len = 0;
                for (int k = 0; k < n; ++k) {
                    const uint32_t o = orow[k] ^ inv;
                    if (o != run_val) { atomicAdd(&s_hist[run_val], run_len); run_val = o; run_len = 0; }
                    ++run_len;
                }
                atomicAdd(&s_hist[run_val], run_len);
            }
        }
    }
    if (a.hist) {
        __syncthreads();
        const uint32_t c = s_hist[tid];
        if (c) atomicAdd(&a.hist[(size_t)pair * 256 + tid], c);
    }
}

}  // namespace stepth
