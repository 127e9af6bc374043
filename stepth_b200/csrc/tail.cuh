// tail.cuh -- K3/K4 and the small elementwise kernels of the tail:
//   depth_split      depth_image.rs:162-218   (from a 256-bin histogram; the reference walks every byte per iteration)
//   slice + apply    depth_image.rs:229-245 + mask_image.rs:205-213 fused: depth is read once, RGBA once
//   invert           depth_image.rs:225-227
//   histogram        256-bin count of a byte plane (input of depth_split when the fused finish pass did not run)
//   paint            depth_image.rs:125-132   normalise + DiscreteImage::collect() + luma of (v,v,v)
#pragma once
#include "common.cuh"

namespace stepth {

// ------------------------------------------------------------------------------------------------------------
// depth_split(zones): one CTA of 256 threads per pair; thread v owns depth value v.
// The Lloyd iteration only depends on the multiset of bytes, so it runs on the histogram:
//   clusters are keyed by centre VALUE (HashMap<u8, Vec<u8>>: duplicate centres collapse)          :166-170
//   a byte goes to the first centre (list order) with the smallest |c - v|                          :171-184
//   new centroid = sum / len.max(1), list sorted ascending                                          :185-189
//   converged iff zip(prev, new) pairwise equal (zip truncates to the shorter)                      :190-196
//   answer: per new centroid c, clusters.get(c).unwrap() -> (min, max); both unwraps can panic      :197-207
// ------------------------------------------------------------------------------------------------------------
struct SplitArgs {
    const uint32_t *hist;  // [n, 256]
    uint8_t *ranges;       // [n, 512]  (min, max) per cluster
    uint32_t *n_out;       // [n]
    int32_t *status;       // [n]  only written on failure (so an earlier DIV_BY_ZERO_MAX survives)
    uint32_t zones;
    uint32_t max_iters;
};

// With the centre list ascending (it is built so, :214-216, and sorted after every iteration, :189) "nearest centre, the
// first one on ties" (:171-179) cuts the value axis into consecutive intervals: centre c_k owns (mid(c_k-1, c_k), mid(c_k, c_k+1)]
// with mid = floor((a + b) / 2).  Sums, counts, minima and maxima of an interval come from prefix tables of the histogram,
// so one Lloyd iteration is O(centres) for a single thread and needs no barrier at all.
__global__ void __launch_bounds__(256) depth_split_kernel(SplitArgs a) {
    __shared__ unsigned long long s_S[257];   // s_S[v + 1] = sum_{u <= v} u * h[u]
    __shared__ uint32_t s_C[257];             // s_C[v + 1] = sum_{u <= v} h[u]
    __shared__ uint32_t s_h[256];
    __shared__ uint16_t s_ge[257], s_le[256]; // next / previous non-empty bin (256 / 0xffff = none)
    __shared__ uint8_t s_prev[256], s_new[256];

    const uint32_t pair = blockIdx.x;
    const int v = threadIdx.x;
    if (a.status[pair] != STEPTH_OK) return;  // e.g. max == 0: the reference panicked before reaching here
    s_h[v] = a.hist[(size_t)pair * 256 + v];
    __syncthreads();
    if (v != 0) return;

    unsigned long long S = 0;
    uint32_t C = 0;
    s_S[0] = 0; s_C[0] = 0;
    int last = 0xffff, gmin = -1, gmax = -1;
    for (int u = 0; u < 256; ++u) {
        const uint32_t h = s_h[u];
        S += (unsigned long long)h * (unsigned long long)u;
        C += h;
        s_S[u + 1] = S; s_C[u + 1] = C;
        if (h) { last = u; if (gmin < 0) gmin = u; gmax = u; }
        s_le[u] = (uint16_t)last;
    }
    int next = 256;
    s_ge[256] = 256;
    for (int u = 255; u >= 0; --u) {
        if (s_h[u]) next = u;
        s_ge[u] = (uint16_t)next;
    }
    int state = 0;  // 0 = iterate, 1 = converged, > 1 = error code + 1
    uint32_t np = 0;
    // :212-216  (img_min..img_max).step_by(((img_max - img_min) / (zones - 1)) as usize - 1)
    if (gmin < 0) {
        state = STEPTH_ERR_DEGENERATE_DEPTH + 1;  // empty image: min().unwrap() on None
    } else {
        const uint32_t q = (uint32_t)(gmax - gmin) / (a.zones - 1);
        if (q <= 1) {
            state = STEPTH_ERR_DEGENERATE_DEPTH + 1;  // usize underflow / step_by(0)
        } else {
            const uint32_t step = q - 1;
            for (uint32_t c = (uint32_t)gmin; c < (uint32_t)gmax; c += step) s_prev[np++] = (uint8_t)c;
        }
    }
    for (uint32_t iter = 1; state == 0; ++iter) {
        // one new centroid per DISTINCT centre value (the HashMap keys, :166-170); equal centres are neighbours
        uint32_t nn = 0;
        int lo = 0;  // first value of the current centre's interval
        for (uint32_t k = 0; k < np; ++k) {
            const int key = s_prev[k];
            if (k > 0 && key == (int)s_prev[k - 1]) continue;
            uint32_t kn = k + 1;
            while (kn < np && (int)s_prev[kn] == key) ++kn;
            const int hi = kn < np ? (key + (int)s_prev[kn]) >> 1 : 255;  // ties go to the earlier (smaller) centre
            const unsigned long long sum = s_S[hi + 1] - s_S[lo];
            const uint32_t cnt = s_C[hi + 1] - s_C[lo];
            s_new[nn++] = (uint8_t)(sum / (cnt ? cnt : 1u));            // :185-188  sum / len.max(1)
            lo = hi + 1;
        }
        for (uint32_t i = 1; i < nn; ++i) {  // new_centroids.sort()   :189
            const uint8_t x = s_new[i];
            int j = (int)i - 1;
            while (j >= 0 && s_new[j] > x) { s_new[j + 1] = s_new[j]; --j; }
            s_new[j + 1] = x;
        }
        bool same = true;   // :190-196  zip truncates to the shorter list
        const uint32_t m = np < nn ? np : nn;
        for (uint32_t k = 0; k < m; ++k)
            if (s_prev[k] != s_new[k]) { same = false; break; }
        if (same) {
            // :197-207  clusters.get(c).unwrap(), then min / max of the cluster: both unwraps can panic
            int st = 1;
            for (uint32_t k = 0; k < nn && st == 1; ++k) {
                const int c = s_new[k];
                // the interval of centre value c among the previous centres (c must be one of them)
                int lo2 = 0, hi2 = -1;
                bool found = false;
                for (uint32_t j = 0; j < np; ++j) {
                    const int key = s_prev[j];
                    if (j > 0 && key == (int)s_prev[j - 1]) continue;
                    uint32_t jn = j + 1;
                    while (jn < np && (int)s_prev[jn] == key) ++jn;
                    hi2 = jn < np ? (key + (int)s_prev[jn]) >> 1 : 255;
                    if (key == c) { found = true; break; }
                    lo2 = hi2 + 1;
                }
                if (!found || s_C[hi2 + 1] == s_C[lo2]) { st = STEPTH_ERR_DEGENERATE_DEPTH + 1; break; }
                a.ranges[(size_t)pair * 512 + 2 * k] = (uint8_t)s_ge[lo2];
                a.ranges[(size_t)pair * 512 + 2 * k + 1] = (uint8_t)s_le[hi2];
            }
            if (st == 1 && nn == 0) st = STEPTH_ERR_DEGENERATE_DEPTH + 1;
            a.n_out[pair] = nn;
            state = st;
        } else if (iter >= a.max_iters) {
            state = STEPTH_ERR_NO_CONVERGENCE + 1;
        } else {
            for (uint32_t k = 0; k < nn; ++k) s_prev[k] = s_new[k];
            np = nn;
        }
    }
    if (state > 1) a.status[pair] = state - 1;
}

// ------------------------------------------------------------------------------------------------------------
// slice(from, to) + apply_mask, 16 pixels per thread: one 16-byte depth load, four 16-byte RGBA loads/stores.
//   mask = 255 iff from <= d <= to else 0        depth_image.rs:232-241
//   rgba = [0;4] where mask == 0                 mask_image.rs:208-210
// ------------------------------------------------------------------------------------------------------------
struct MaskArgs {
    const uint8_t *depth;    // [n, px]
    const uint8_t *from_to;  // [n, stride_ft]  bytes 0,1 = (from, to)
    const int32_t *status;   // [n] may be null; pairs with status != OK are skipped
    const uint8_t *rgba_in;  // [n, px, 4] may be null
    uint8_t *mask_out;       // [n, px] may be null
    uint8_t *rgba_out;       // [n, px, 4] may be null
    size_t px;
    uint32_t stride_ft;
    uint32_t vec_ok;         // every plane 16-byte aligned and px % 16 == 0 (decided by the host wrapper)
};

__device__ __forceinline__ uint32_t mask4(uint32_t d4, uint32_t from, uint32_t to) {
    uint32_t m = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        uint32_t d = (d4 >> (8 * k)) & 0xff;
        m |= (d >= from && d <= to) ? (0xffu << (8 * k)) : 0u;
    }
    return m;
}

__global__ void __launch_bounds__(256) mask_apply_kernel(MaskArgs a) {
    const uint32_t pair = blockIdx.y;
    if (a.status && a.status[pair] != STEPTH_OK) return;
    const uint32_t from = a.from_to[(size_t)pair * a.stride_ft], to = a.from_to[(size_t)pair * a.stride_ft + 1];
    const uint8_t *depth = a.depth + (size_t)pair * a.px;
    const size_t n16 = a.px / 16;
    if (a.vec_ok) {
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) {
            uint4 d = reinterpret_cast<const uint4 *>(depth)[i];
            uint4 m = make_uint4(mask4(d.x, from, to), mask4(d.y, from, to), mask4(d.z, from, to), mask4(d.w, from, to));
            if (a.mask_out) reinterpret_cast<uint4 *>(a.mask_out + (size_t)pair * a.px)[i] = m;
            if (a.rgba_out) {
                const uint4 *in = reinterpret_cast<const uint4 *>(a.rgba_in + (size_t)pair * a.px * 4) + 4 * i;
                uint4 *out = reinterpret_cast<uint4 *>(a.rgba_out + (size_t)pair * a.px * 4) + 4 * i;
                const uint32_t mm[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    uint4 p = in[k];
                    // sign-extend each mask byte to a full pixel word
                    p.x = (mm[k] & 0x000000ffu) ? p.x : 0u;
                    p.y = (mm[k] & 0x0000ff00u) ? p.y : 0u;
                    p.z = (mm[k] & 0x00ff0000u) ? p.z : 0u;
                    p.w = (mm[k] & 0xff000000u) ? p.w : 0u;
                    out[k] = p;
                }
            }
        }
    } else {
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.px; i += (size_t)gridDim.x * blockDim.x) {
            uint32_t d = depth[i];
            bool keep = d >= from && d <= to;
            if (a.mask_out) a.mask_out[(size_t)pair * a.px + i] = keep ? 255 : 0;
            if (a.rgba_out) {
                uint32_t p = reinterpret_cast<const uint32_t *>(a.rgba_in)[(size_t)pair * a.px + i];
                reinterpret_cast<uint32_t *>(a.rgba_out)[(size_t)pair * a.px + i] = keep ? p : 0u;
            }
        }
    }
}

// apply_mask with an arbitrary caller-supplied mask (mask_image.rs:205-213): only mask == 0 clears
__global__ void __launch_bounds__(256) apply_mask_kernel(uint32_t *rgba, const uint8_t *mask, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        if (mask[i] == 0) rgba[i] = 0u;
}

// invert_depth: d := 255 - d  == d ^ 0xff
__global__ void __launch_bounds__(256) invert_kernel(uint8_t *d, size_t n) {
    const size_t n4 = n / 4;
    uint32_t *d4 = reinterpret_cast<uint32_t *>(d);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x)
        d4[i] = ~d4[i];
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) d[n4 * 4 + threadIdx.x] ^= 0xff;
}

__global__ void __launch_bounds__(256) histogram_kernel(const uint8_t *data, size_t n, uint32_t *hist) {
    __shared__ uint32_t s_hist[256];
    s_hist[threadIdx.x] = 0;
    __syncthreads();
    uint32_t run_val = 0, run_len = 0;
    // each thread owns a contiguous chunk so that runs of equal bytes collapse into one atomic
    const size_t total_threads = (size_t)gridDim.x * blockDim.x;
    const size_t chunk = (n + total_threads - 1) / total_threads;
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t b = t * chunk, e = b + chunk < n ? b + chunk : n;
    for (size_t i = b; i < e; ++i) {
        uint32_t o = data[i];
        if (run_len && o != run_val) { atomicAdd(&s_hist[run_val], run_len); run_len = 0; }
        run_val = o;
        ++run_len;
    }
    if (run_len) atomicAdd(&s_hist[run_val], run_len);
    __syncthreads();
    uint32_t c = s_hist[threadIdx.x];
    if (c) atomicAdd(&hist[threadIdx.x], c);
}

// paint: one warp per block rectangle, writing the block's raw `d as u8` (DiscreteImage::collect() +
// to_luma8_from_rgb8 of (v,v,v), depth_image.rs:130-132).  The normalisation of :125-129 is a per-byte map, so
// it commutes with painting and is applied by finish_kernel's table.
struct PaintArgs {
    const stepth_block *blocks;
    const uint8_t *d8;
    uint8_t *plane;
    uint32_t n_blocks, w, h;
};

__global__ void __launch_bounds__(256) paint_kernel(PaintArgs a) {
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= a.n_blocks) return;
    const stepth_block bl = a.blocks[warp];
    const uint8_t v = a.d8[warp];
    const uint32_t area = bl.w * bl.h;
    for (uint32_t i = lane; i < area; i += 32) {
        uint32_t yy = bl.y + i / bl.w, xx = bl.x + i % bl.w;
        if (yy < a.h && xx < a.w) a.plane[(size_t)yy * a.w + xx] = v;
    }
}

}  // namespace stepth
