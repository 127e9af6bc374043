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
    // the taps of the three first and three last columns laid on the interior's 7-column footprint (columns i - 3 ..
    // i + 3), 0 where the footprint leaves the row: 0.0 + 0 * 0 and t + 0 * 0 change nothing, so summing all seven in
    // order gives the bits of the clipped sum (rows of 8 or more columns; filled by stepth_cuda_dev_finish)
    float wx_edge[6][7];
    uint32_t seg_rows;       // output rows per thread of the vertical pass: FV_ROWS, or 7 for a launch that would leave SMs idle
    // row bands over several GPUs (peer.cuh): the streaming horizontal pass also stores every output into these planes -- the band's
    // rows inside each rank's assembled map, reached over NVLink -- so the all-gather of the depth bands rides on the pass
    uint8_t *mirror[8];      // [n_mirror] same geometry as dst (n == 1)
    uint32_t n_mirror;
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


// ---- streaming form of K2 for the default mode (STEPTH_RESAMPLE_U8, rows that are whole words) ---------------------------
// The u8 intermediate of the two-pass resample (image 0.23.14 rounds `vertical_sample`'s output to the pixel type before
// `horizontal_sample` reads it) makes the passes two plain streaming kernels with a 1 B/px plane between them: no shared
// memory tile, no barriers, every thread works on four (vertical) or eight (horizontal) neighbouring pixels out of
// registers.  The extra plane costs 2 B/px of HBM traffic -- a few microseconds -- and removes the staging that bounded
// the tiled kernel.  Arithmetic and order are the tiled kernel's (and the reference's): t = t + px * w in tap order.

// exact float of an integer below 2^23 without the (quarter-rate) conversion unit: 2^23 + n has n in its low mantissa bits
__device__ __forceinline__ float small_to_float(uint32_t n) { return __fsub_rn(__uint_as_float(0x4b000000u | n), 8388608.0f); }

// (v * 255 / max) as u8 (depth_image.rs:127) without a table: floor(n / d) = umulhi(n, floor(2^32 / d) + 1) holds for
// n <= 65025, d <= 255 (error term n * d < 2^32); d == 1 would overflow the constant and needs no division
struct NormDiv {
    uint32_t magic, one;
    __device__ __forceinline__ explicit NormDiv(uint32_t mx) : magic(mx == 1u ? 0u : 0xffffffffu / mx + 1u), one(mx == 1u) {}
    __device__ __forceinline__ float operator()(uint32_t v) const {
        const uint32_t n = v * 255u;
        return small_to_float(one ? n : __umulhi(n, magic));
    }
};

// clamp(t, 0, 255) then f32::round (half away from zero), for 0 <= t < 255.5 -- a normalised sum of values in [0, 255]
// never leaves that range, so the clamp is the identity on the rounded result.  t + 2^23 rounds to nearest-even into the
// low mantissa bits; the two differ only on an exact tie that went down, which the exact remainder t - r shows.
__device__ __forceinline__ uint32_t round_u8(float t) {
    const float u = __fadd_rn(t, 8388608.0f);
    const float d = __fsub_rn(t, __fsub_rn(u, 8388608.0f));
    return (__float_as_uint(u) & 0x1ffu) + (d == 0.5f ? 1u : 0u);
}

#ifndef STEPTH_FV_ROWS
#define STEPTH_FV_ROWS 21
#endif
constexpr int FV_ROWS = STEPTH_FV_ROWS;  // output rows per thread of the vertical pass (27 row loads for 21 rows); a multiple of 7

__global__ void __launch_bounds__(128) finish_vert_kernel(FinishArgs a, uint8_t *tmp) {
    const uint32_t pair = blockIdx.z;
    const uint32_t mx = a.max_in[pair];
    if (mx == 0) {  // depth_image.rs:127 divides by zero -> panic; reported per pair, nothing written
        if (a.status && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) a.status[pair] = STEPTH_ERR_DIV_BY_ZERO_MAX;
        return;
    }
    const NormDiv norm(mx);
    const int wq = (int)(a.w >> 2), xq = blockIdx.x * 128 + threadIdx.x;
    if (xq >= wq) return;
    const int seg = (int)a.seg_rows, r0 = blockIdx.y * seg, rows = min(seg, (int)a.out_rows - r0);
    const int gy0 = (int)a.out_y0 + r0, img_h = (int)a.img_h;
    const uint8_t *srcp = a.src + (size_t)pair * a.src_rows * a.w;
    uint32_t *dst = reinterpret_cast<uint32_t *>(tmp + (size_t)pair * a.out_rows * a.w) + (size_t)r0 * wq + xq;
    if (img_h >= 8) {
        // a seven-row window per column slides down in registers.  Rows off the image enter as 0 and the three first / last
        // output rows take their own weights laid on the seven-row footprint (0 where it leaves the image): 0.0 + 0 * 0 and
        // t + 0 * 0 change nothing, so the seven-term sum in order has the bits of the clipped sum
        const uint32_t *col = reinterpret_cast<const uint32_t *>(srcp) + ((ptrdiff_t)(gy0 - 3 - (int)a.src_y0) * wq + xq);
        float win[7][4];
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            const int gr = gy0 - 3 + k;
            const uint32_t px = (gr >= 0 && gr < img_h) ? __ldg(col + (ptrdiff_t)k * wq) : 0u;
#pragma unroll
            for (int c = 0; c < 4; ++c) win[k][c] = norm((px >> (8 * c)) & 0xffu);
        }
        // the window index cycles with period 7: seven rows are unrolled (static register indices), the groups loop
#pragma unroll 1
        for (int j0 = 0; j0 < seg; j0 += 7) {
#pragma unroll
            for (int jj = 0; jj < 7; ++jj) {
                const int j = j0 + jj;
                if (j < rows) {
                    const int gy = gy0 + j;
                    const uint32_t px = gy + 3 < img_h ? __ldg(col + (ptrdiff_t)(j + 6) * wq) : 0u;
#pragma unroll
                    for (int c = 0; c < 4; ++c) win[(jj + 6) % 7][c] = norm((px >> (8 * c)) & 0xffu);
                    uint32_t out = 0;
                    if (gy >= 3 && gy + 4 <= img_h) {
#pragma unroll
                        for (int c = 0; c < 4; ++c) {
                            float t = 0.0f;
#pragma unroll
                            for (int k = 0; k < 7; ++k) t = __fadd_rn(t, __fmul_rn(win[(jj + k) % 7][c], a.wy_int[k]));
                            out |= round_u8(t) << (8 * c);
                        }
                    } else {   // a border row (the same for the whole warp)
                        const TapRow &tp = a.taps_y[gy];
                        float wk[7];
#pragma unroll
                        for (int k = 0; k < 7; ++k) {
                            const int idx = gy - 3 + k - tp.left;
                            wk[k] = (idx >= 0 && idx < tp.cnt) ? tp.w[idx] : 0.0f;
                        }
#pragma unroll
                        for (int c = 0; c < 4; ++c) {
                            float t = 0.0f;
#pragma unroll
                            for (int k = 0; k < 7; ++k) t = __fadd_rn(t, __fmul_rn(win[(jj + k) % 7][c], wk[k]));
                            out |= round_u8(t) << (8 * c);
                        }
                    }
                    dst[(size_t)j * wq] = out;
                }
            }
        }
    } else {
        // images of fewer than 8 rows: per output through the host-built tap rows
        for (int j = 0; j < rows; ++j) {
            const TapRow &tp = a.taps_y[gy0 + j];
            uint32_t out = 0;
            for (int c = 0; c < 4; ++c) {
                float t = 0.0f;
                for (int k = 0; k < tp.cnt; ++k) {
                    const uint32_t v = srcp[(size_t)(tp.left + k - (int)a.src_y0) * a.w + 4 * xq + c];
                    t = __fadd_rn(t, __fmul_rn(norm(v), tp.w[k]));
                }
                out |= round_u8(t) << (8 * c);
            }
            dst[(size_t)j * wq] = out;
        }
    }
}

// horizontal pass + invert + histogram: a thread owns 8 neighbouring outputs of a row, a CTA 256 columns x 32 rows
__global__ void __launch_bounds__(256) finish_horz_kernel(FinishArgs a, const uint8_t *tmp) {
    __shared__ uint32_t s_hist[256];
    const uint32_t pair = blockIdx.z;
    if (a.max_in[pair] == 0) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    s_hist[tid] = 0;
    __syncthreads();
    const int w = (int)a.w, x0 = blockIdx.x * 256 + lane * 8;
    const uint32_t inv = a.invert ? 0xffffffffu : 0u;
    const uint8_t *tplane = tmp + (size_t)pair * a.out_rows * a.w;
    uint8_t *dplane = a.dst + (size_t)pair * a.out_rows * a.w;
    for (int it = 0; it < 4; ++it) {
        const int row = blockIdx.y * 32 + it * 8 + warp;
        if (row >= (int)a.out_rows || x0 >= w) continue;
        const uint8_t *trow = tplane + (size_t)row * w;
        uint32_t o[2] = {0u, 0u};
        {
            // columns x0 - 4 .. x0 + 11 as four words (0 for a word off the row); output j reads columns x0 + j - 3 .. x0 + j + 3
            const uint32_t *tw = reinterpret_cast<const uint32_t *>(trow + x0 - 4);
            float b[16];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int c0 = x0 - 4 + 4 * k;
                const uint32_t px = (c0 >= 0 && c0 + 4 <= w) ? __ldg(tw + k) : 0u;
#pragma unroll
                for (int c = 0; c < 4; ++c)   // one byte permute builds the bits of 2^23 + byte
                    b[4 * k + c] = __fsub_rn(__uint_as_float(__byte_perm(px, 0x4b000000u, 0x7540u | c)), 8388608.0f);
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float t = 0.0f;
#pragma unroll
                for (int k = 0; k < 7; ++k) t = __fadd_rn(t, __fmul_rn(b[j + 1 + k], a.wx_int[k]));
                o[j >> 2] |= round_u8(t) << (8 * (j & 3));
            }
            // the first and the last thread of a row redo their three border outputs with the border's own weights
#define STEPTH_EDGE_OUT(J, E)                                                                          \
    {                                                                                                  \
        float t = 0.0f;                                                                                \
        _Pragma("unroll") for (int k = 0; k < 7; ++k) t = __fadd_rn(t, __fmul_rn(b[(J) + 1 + k], a.wx_edge[E][k])); \
        o[(J) >> 2] = (o[(J) >> 2] & ~(0xffu << (8 * ((J)&3)))) | (round_u8(t) << (8 * ((J)&3)));     \
    }
            if (x0 == 0) {
                STEPTH_EDGE_OUT(0, 0) STEPTH_EDGE_OUT(1, 1) STEPTH_EDGE_OUT(2, 2)
            }
            if (x0 + 8 == w) {
                STEPTH_EDGE_OUT(5, 3) STEPTH_EDGE_OUT(6, 4) STEPTH_EDGE_OUT(7, 5)
            } else if (x0 + 4 == w) {
                STEPTH_EDGE_OUT(1, 3) STEPTH_EDGE_OUT(2, 4) STEPTH_EDGE_OUT(3, 5)
            }
#undef STEPTH_EDGE_OUT
        }
        o[0] ^= inv;
        o[1] ^= inv;
        uint32_t *drow = reinterpret_cast<uint32_t *>(dplane + (size_t)row * w + x0);   // w % 4 == 0: word aligned
        const int n = min(8, w - x0);                                                      // 4 or 8 (w % 4 == 0)
        drow[0] = o[0];
        if (n == 8) drow[1] = o[1];
        if (a.n_mirror) {
            const size_t off = (size_t)row * w + x0;
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                if (q < (int)a.n_mirror) {
                    if (n == 8 && (w & 7) == 0) {
                        *reinterpret_cast<uint2 *>(a.mirror[q] + off) = make_uint2(o[0], o[1]);   // a warp stores 256 contiguous bytes
                    } else {
                        uint32_t *m = reinterpret_cast<uint32_t *>(a.mirror[q] + off);
                        m[0] = o[0];
                        if (n == 8) m[1] = o[1];
                    }
                }
            }
        }
        if (a.hist) {
            // neighbouring outputs are mostly equal: feed the shared-memory histogram run-length encoded
            uint32_t run_val = o[0] & 0xffu, run_len = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (j < n) {
                    const uint32_t v = (o[j >> 2] >> (8 * (j & 3))) & 0xffu;
                    if (v != run_val) { atomicAdd(&s_hist[run_val], run_len); run_val = v; run_len = 0; }
                    ++run_len;
                }
            }
            atomicAdd(&s_hist[run_val], run_len);
        }
    }
    if (a.hist) {
        __syncthreads();
        const uint32_t c = s_hist[tid];
        if (c) atomicAdd(&a.hist[(size_t)pair * 256 + tid], c);
    }
}

}  // namespace stepth
