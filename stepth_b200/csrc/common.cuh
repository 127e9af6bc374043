// common.cuh -- shared device helpers for the stepth sm_100a kernels.
#pragma once
#include <cmath>
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/stepth_cuda.h"

namespace stepth {

// Functions marked STEPTH_HD also compile for the host so that tests/host_emul can unit-test the index math of the
// kernels without a GPU (test infrastructure; the product library never runs them on the CPU).
#define STEPTH_HD __host__ __device__ __forceinline__

STEPTH_HD int ffs32(uint32_t x) {
#ifdef __CUDA_ARCH__
    return __ffs((int)x);
#else
    return __builtin_ffs((int)x);
#endif
}
STEPTH_HD int ffs64(unsigned long long x) {
#ifdef __CUDA_ARCH__
    return __ffsll((long long)x);
#else
    return __builtin_ffsll((long long)x);
#endif
}
STEPTH_HD int clz32(uint32_t x) {  // x != 0
#ifdef __CUDA_ARCH__
    return __clz((int)x);
#else
    return __builtin_clz(x);
#endif
}
STEPTH_HD int clz64(unsigned long long x) {  // x != 0
#ifdef __CUDA_ARCH__
    return __clzll((long long)x);
#else
    return __builtin_clzll(x);
#endif
}
template <typename T>
STEPTH_HD T ldg(const T *p) {
#ifdef __CUDA_ARCH__
    return __ldg(p);
#else
    return *p;
#endif
}

// Per-query colour box.  The reference's predicate `v.substract(what).lt(precision)` (helpers.rs:35) is a
// per-channel interval test for both supported PixelOpps readings:
//   abs/all : |v_c - what_c| < p_c   <=>  what_c - p_c + 1 <= v_c <= what_c + p_c - 1
//   sat/all : sat_sub(v_c, what_c) < p_c  <=>  v_c <= what_c + p_c - 1
// lo/hi are packed one channel per byte (r | g << 8 | b << 16); byte 3 is lo = 0, hi = 255 so a padding
// byte always passes.  p_c == 0 never matches; the host short-circuits that case before any launch.
struct ColourBox {
    uint32_t lo, hi;
};

__host__ __device__ __forceinline__ ColourBox make_box(uint32_t r, uint32_t g, uint32_t b, uint32_t pr,
                                                       uint32_t pg, uint32_t pb, bool sat) {
    uint32_t w[3] = {r, g, b}, p[3] = {pr, pg, pb};
    uint32_t lo = 0, hi = 0xff000000u;
    for (int c = 0; c < 3; ++c) {
        int l = sat ? 0 : (int)w[c] - (int)p[c] + 1;
        int h = (int)w[c] + (int)p[c] - 1;
        if (l < 0) l = 0;
        if (h > 255) h = 255;
        lo |= (uint32_t)l << (8 * c);
        hi |= (uint32_t)h << (8 * c);
    }
    return ColourBox{lo, hi};
}

STEPTH_HD bool box_contains(const ColourBox &bx, uint32_t r, uint32_t g, uint32_t b) {
    return r >= (bx.lo & 0xff) && r <= (bx.hi & 0xff) && g >= ((bx.lo >> 8) & 0xff) && g <= ((bx.hi >> 8) & 0xff) &&
           b >= ((bx.lo >> 16) & 0xff) && b <= ((bx.hi >> 16) & 0xff);
}

// the predicate on one pixel straight from the definition (no box): abs/all |v_c - q_c| < p_c, sat/all sat_sub(v_c, q_c) < p_c
STEPTH_HD bool pixel_in_box(uint32_t r, uint32_t g, uint32_t b, uint32_t col, uint32_t pr, uint32_t pg, uint32_t pb, bool sat) {
    int dr = (int)r - (int)(col & 0xffu), dg = (int)g - (int)((col >> 8) & 0xffu), db = (int)b - (int)((col >> 16) & 0xffu);
    if (sat) { dr = dr < 0 ? 0 : dr; dg = dg < 0 ? 0 : dg; db = db < 0 ? 0 : db; }
    else { dr = dr < 0 ? -dr : dr; dg = dg < 0 ? -dg : dg; db = db < 0 ? -db : db; }
    return dr < (int)pr && dg < (int)pg && db < (int)pb;
}

// floor(sqrt(n)) for n < 2^24: what `((dx^2 + dy^2) as f64).sqrt() as u32` yields (helpers.rs:3-7).
STEPTH_HD uint32_t isqrt_u32(uint32_t n) {
#ifdef __CUDA_ARCH__
    uint32_t r = (uint32_t)__fsqrt_rn((float)n);
#else
    uint32_t r = (uint32_t)sqrtf((float)n);
#endif
    while (r * r > n) --r;
    while ((r + 1) * (r + 1) <= n) ++r;
    return r;
}

// ---- ring-order key -------------------------------------------------------------------------------------
// helpers.rs:26-31 probes ring s as: row y+s (x ascending), row y-s, column x+s (y ascending), column x-s.
// Dropping duplicate probes, a pixel at offset (dx, dy) sits at a unique position in that order:
//   s = max(|dx|, |dy|);  pos = dx+s                         if dy == +s        (bottom row)
//                               (2s+1) + dx+s                if dy == -s, s > 0 (top row)
//                               2(2s+1) + (dy+s-1)           if dx == +s, |dy| < s (right column interior)
//                               2(2s+1) + (2s-1) + (dy+s-1)  if dx == -s, |dy| < s (left column interior)
// The first match in reference order is the matching pixel with the smallest key = (s << 12) | pos
// (pos < 8s <= 4088 for s <= 511).  The search is therefore an arg-min and may visit pixels in ANY order.
STEPTH_HD uint32_t ring_key(int dx, int dy) {
    int ax = dx < 0 ? -dx : dx, ay = dy < 0 ? -dy : dy;
    int s = ax > ay ? ax : ay;
    int pos;
    if (dy == s) pos = dx + s;
    else if (dy == -s) pos = (2 * s + 1) + dx + s;
    else if (dx == s) pos = 2 * (2 * s + 1) + (dy + s - 1);
    else pos = 2 * (2 * s + 1) + (2 * s - 1) + (dy + s - 1);
    return ((uint32_t)s << 12) | (uint32_t)pos;
}

STEPTH_HD void key_to_offset(uint32_t key, int &dx, int &dy) {
    int s = (int)(key >> 12), pos = (int)(key & 0xfff);
    if (pos < 2 * s + 1) { dx = pos - s; dy = s; }
    else if (pos < 2 * (2 * s + 1)) { dx = pos - (2 * s + 1) - s; dy = -s; }
    else if (pos < 2 * (2 * s + 1) + (2 * s - 1)) { dx = s; dy = pos - 2 * (2 * s + 1) - s + 1; }
    else { dx = -s; dy = pos - 2 * (2 * s + 1) - (2 * s - 1) - s + 1; }
}

constexpr uint32_t KEY_NONE = 0xffffffffu;

// d as u8 of the winning offset (depth_image.rs:120-121): None -> 0, wraps mod 256
STEPTH_HD uint8_t key_to_d8(uint32_t key) {
    if (key == KEY_NONE) return 0;
    int dx, dy;
    key_to_offset(key, dx, dy);
    return (uint8_t)isqrt_u32((uint32_t)(dx * dx + dy * dy));
}

}  // namespace stepth
