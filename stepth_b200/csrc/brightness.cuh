// brightness.cuh -- operations.rs:3-36 (normalize_brightness_luma16) and :38-80 (normalize_brightness_rgb16).
// SURVEY 8(f) "next" row 1: the pre-step for a stereo pair; never called by the reference's own hot path.
//
// Two launches, no host round trip: (1) exact u64 channel sums of both images (the reference's f64 running sum
// of u16 values is exact below 2^53, so an integer sum converts to the same f64), (2) every thread derives the
// same f64 ratio(s) with IEEE division and either copies img1 or scales it with `(v as f64 * diff) as u16`
// (Rust float->int casts saturate; NaN -> 0).
#pragma once
#include "common.cuh"

namespace stepth {

// sums[0..3) = img1 channel sums, sums[3..6) = img2 channel sums (nch = 1 uses slots 0 and 3)
__global__ void __launch_bounds__(256) u16_channel_sums_kernel(const uint16_t *img, size_t n_sub, uint32_t nch,
                                                                unsigned long long *sums) {
    unsigned long long acc[3] = {0, 0, 0};
    // a thread keeps a fixed channel phase: stride is a multiple of 3 subpixels
    const size_t stride = (size_t)gridDim.x * blockDim.x * 3;
    for (size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 3; i < n_sub; i += stride) {
#pragma unroll
        for (int c = 0; c < 3; ++c)
            if (i + c < n_sub) acc[nch == 3 ? c : 0] += img[i + c];
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        unsigned long long v = acc[c];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0 && v) atomicAdd(&sums[c], v);
    }
}

__device__ __forceinline__ uint16_t f64_as_u16(double v) {
    if (!(v == v)) return 0;
    if (v <= 0.0) return 0;
    if (v >= 65535.0) return 65535;
    return (uint16_t)v;
}

__global__ void __launch_bounds__(256) u16_scale_kernel(const uint16_t *img1, uint16_t *out, size_t n1_sub,
                                                         size_t n2_sub, uint32_t nch, double percent,
                                                         const unsigned long long *sums) {
    double diff[3];
    bool keep = true;
    if (nch == 1) {
        // fbr = sum1 / len1 (u64 division), sbr likewise; diff = sbr as f64 / fbr as f64     operations.rs:25-27
        unsigned long long fbr = sums[0] / (unsigned long long)n1_sub;
        unsigned long long sbr = sums[3] / (unsigned long long)n2_sub;
        diff[0] = diff[1] = diff[2] = __ddiv_rn((double)sbr, (double)fbr);
        keep = fabs(1.0 - diff[0]) < percent;
    } else {
        // means divide by img.len() = the SUBPIXEL count                                      operations.rs:62-67
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            double m1 = __ddiv_rn((double)sums[c], (double)n1_sub);
            double m2 = __ddiv_rn((double)sums[3 + c], (double)n2_sub);
            diff[c] = __ddiv_rn(m2, m1);
            keep = keep && (fabs(1.0 - diff[c]) < percent);
        }
    }
    const size_t stride = (size_t)gridDim.x * blockDim.x * 3;
    for (size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 3; i < n1_sub; i += stride) {
#pragma unroll
        for (int c = 0; c < 3; ++c)
            if (i + c < n1_sub) {
                uint16_t v = img1[i + c];
                out[i + c] = keep ? v : f64_as_u16(__dmul_rn((double)v, diff[nch == 3 ? c : 0]));
            }
    }
}

}  // namespace stepth
