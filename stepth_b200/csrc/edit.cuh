// edit.cuh -- the elementwise editor operations of SURVEY 8(f) row 3, so that a depth map / mask produced on the device
// can be post-processed without a round trip through the reference's CPU loops:
//   highlight_depth   depth_image.rs:51-63     rgb *= depth / 255 * 2  (f32, clamped, truncated)
//   highlight_mask    mask_image.rs:57-73      where mask == MASK_TRUE: r *= 2, g /= 2, b /= 2
//   mask_and/or/not   mask_image.rs:147-195    on equal-sized masks (the resize branch of the reference is out of scope)
//   image_replace     mask_image.rs:99-109     copy other's pixel where self.mask == MASK_TRUE inside the start window
// f32 arithmetic follows Rust: IEEE single ops in source order (no FMA), `as u8` = truncation of the clamped value.
#pragma once
#include "common.cuh"

namespace stepth {

__device__ __forceinline__ uint32_t scale_u8(uint32_t v, float mul) {
    const float t = fminf(fmaxf(__fmul_rn((float)v, mul), 0.0f), 255.0f);  // (v as f32 * m).max(0.0).min(255.0)
    return (uint32_t)t;                                                    // as u8: truncation
}

// RGB8 -> RGBA8 (alpha 255) for n pixels: the *_rgb entry points upload 3 bytes per main pixel (what to_rgb8() holds,
// depth_image.rs:105) and the search reads whole words.  Four pixels per thread (three words in, four out) when the run
// starts on a word; the rest byte-wise.
__global__ void __launch_bounds__(256) rgb_to_rgba_kernel(const uint8_t *rgb, uint32_t *rgba, size_t n) {
    const size_t stride = (size_t)gridDim.x * blockDim.x, t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool words = (reinterpret_cast<uintptr_t>(rgb) & 3) == 0 && (reinterpret_cast<uintptr_t>(rgba) & 15) == 0;
    const size_t quads = words ? n / 4 : 0;
    const uint32_t *src = reinterpret_cast<const uint32_t *>(rgb);
    for (size_t q = t; q < quads; q += stride) {
        const uint32_t w0 = src[3 * q], w1 = src[3 * q + 1], w2 = src[3 * q + 2];
        uint4 o;
        o.x = w0 | 0xff000000u;
        o.y = __funnelshift_r(w0, w1, 24) | 0xff000000u;
        o.z = __funnelshift_r(w1, w2, 16) | 0xff000000u;
        o.w = (w2 >> 8) | 0xff000000u;
        reinterpret_cast<uint4 *>(rgba)[q] = o;
    }
    for (size_t i = quads * 4 + t; i < n; i += stride)
        rgba[i] = (uint32_t)rgb[3 * i] | ((uint32_t)rgb[3 * i + 1] << 8) | ((uint32_t)rgb[3 * i + 2] << 16) | 0xff000000u;
}

// mode 0: highlight_depth (plane = depth), mode 1: highlight_mask (plane = mask)
__global__ void __launch_bounds__(256) highlight_kernel(const uint32_t *rgba_in, const uint8_t *plane, size_t n, uint32_t *rgba_out,
                                                        int mode) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        uint32_t p = rgba_in[i];
        const uint32_t d = plane[i];
        uint32_t r = p & 0xffu, g = (p >> 8) & 0xffu, b = (p >> 16) & 0xffu;
        if (mode == 0) {
            const float m = __fmul_rn(__fdiv_rn((float)d, 255.0f), 2.0f);  // d as f32 / 255.0 * 2.0
            r = scale_u8(r, m); g = scale_u8(g, m); b = scale_u8(b, m);
        } else if (d == 255u) {
            r = scale_u8(r, 2.0f); g = scale_u8(g, 0.5f); b = scale_u8(b, 0.5f);  // 1.0 / 2.0 == 0.5 exactly
        }
        rgba_out[i] = (p & 0xff000000u) | (b << 16) | (g << 8) | r;
    }
}

// op 0: and, 1: or, 2: not (other unused)
__global__ void __launch_bounds__(256) mask_logic_kernel(uint8_t *mask, const uint8_t *other, size_t n, int op) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const uint32_t a = mask[i];
        uint32_t o;
        if (op == 2) o = 255u - a;
        else {
            const uint32_t b = other[i];
            const bool t = op == 0 ? (a == 255u && b == 255u) : (a == 255u || b == 255u);
            o = t ? 255u : 0u;
        }
        mask[i] = (uint8_t)o;
    }
}

// image_replace: window [sx, min(sx + ow, w)) x [sy, min(sy + oh, h)); other is indexed at the SAME (x, y) as self
// (mask_image.rs:106 -- literal), so a masked pixel with x >= ow or y >= oh is the reference's out-of-bounds panic.
__global__ void __launch_bounds__(256) image_replace_kernel(uint32_t *rgba, const uint8_t *mask, uint32_t w, uint32_t h,
                                                            const uint32_t *other, uint32_t ow, uint32_t oh, uint32_t sx, uint32_t sy,
                                                            uint32_t x1, uint32_t y1, int32_t *status) {
    const uint32_t ww = x1 - sx, n = ww * (y1 - sy);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t x = sx + i % ww, y = sy + i / ww;
        if (mask[(size_t)y * w + x] != 255u) continue;
        if (x >= ow || y >= oh) { *status = STEPTH_ERR_OUT_OF_BOUNDS; continue; }
        rgba[(size_t)y * w + x] = other[(size_t)y * ow + x];
    }
}

}  // namespace stepth
