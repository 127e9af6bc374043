// disc.cuh -- the stand-in discretiser "disc_v0" on the device (depth_image.rs:101-109 is where the
// reference calls disage::open::rgb_discrete, whose source is absent: disc_v0 is NOT disage, see stepth_host_discretise_v0).
//
// disc_v0 is a kd-split: a region splits its longer side (ties -> width) at floor(len / 2); it always splits while
// depth < min_splits, never at depth == max_splits or at 1x1, and otherwise splits unless every pixel lies within `precision`
// of the region's per-channel mean (floor).  Leaves are emitted in depth-first order, left / top child first.
//
// The geometry of a split does not depend on the pixels, so the first `top` = min(min_splits, max_splits, 16) levels are a
// fixed tree: thread t walks the path spelled by the bits of t (most significant first) to its top-level region and runs the
// rest of the recursion there with an explicit stack.  Depth-first order over the whole tree is then ascending t, each
// thread's leaves in its own depth-first order: pass 1 counts the leaves per thread, one CTA scans the counts, pass 2 writes
// each thread's leaves at its offset -- the list is bit-identical to the host stand-in's, order included.
// "all pixels within p of the mean" is evaluated as  max - mean < p  and  mean - min < p  per channel (one pass over the
// region for sum, min and max instead of two).
#pragma once
#include "common.cuh"

namespace stepth {

#if defined(__CUDACC__)

constexpr uint32_t DISC_MAX_TOP = 16;   // at most 65536 top-level regions
constexpr int DISC_STACK = 72;          // a region of w x h <= 2^32 pixels splits at most 64 times

struct DiscArgs {
    const uint8_t *px;     // interleaved, `stride` bytes per pixel (3 or 4), rows of w pixels
    uint32_t stride, w, h;
    uint32_t pr, pg, pb;
    uint32_t min_s, max_s, top;
    uint32_t *counts;      // [2^top + 1]: leaves per top-level region; after the scan exclusive offsets, the total last
    stepth_block *out;
    unsigned long long cap;
};

struct DiscRegion {
    uint32_t x, y, w, h;
};

// halves region r like the host recursion: `second` picks the right / bottom child
__device__ __forceinline__ void disc_child(DiscRegion &r, bool second) {
    if (r.w >= r.h) {
        const uint32_t l = r.w / 2;
        if (second) { r.x += l; r.w -= l; } else r.w = l;
    } else {
        const uint32_t t = r.h / 2;
        if (second) { r.y += t; r.h -= t; } else r.h = t;
    }
}

// the top-level region of thread t and its depth; false when the path ends early in a 1x1 leaf that another thread owns
__device__ __forceinline__ bool disc_top_region(const DiscArgs &a, uint32_t t, DiscRegion &r, uint32_t &depth) {
    r = DiscRegion{0, 0, a.w, a.h};
    depth = 0;
    for (uint32_t d = 0; d < a.top; ++d) {
        // d < top <= min(min_splits, max_splits): a region splits here unless it is a single pixel
        if ((unsigned long long)r.w * r.h <= 1) return (t & ((1u << (a.top - d)) - 1u)) == 0;   // the path of zeros owns the leaf
        disc_child(r, (t >> (a.top - 1 - d)) & 1u);
        depth = d + 1;
    }
    return true;
}

// per-channel floor mean of a region and whether every pixel lies within the precision of it
__device__ __forceinline__ bool disc_mean_uniform(const DiscArgs &a, const DiscRegion &r, uint32_t mean[3]) {
    unsigned long long s0 = 0, s1 = 0, s2 = 0;
    uint32_t lo0 = 255, lo1 = 255, lo2 = 255, hi0 = 0, hi1 = 0, hi2 = 0;
    for (uint32_t yy = r.y; yy < r.y + r.h; ++yy) {
        const uint8_t *row = a.px + ((size_t)yy * a.w + r.x) * a.stride;
        uint32_t t0 = 0, t1 = 0, t2 = 0;   // a row of at most 65535 pixels sums to < 2^24
        for (uint32_t xx = 0; xx < r.w; ++xx, row += a.stride) {
            const uint32_t v0 = row[0], v1 = row[1], v2 = row[2];
            t0 += v0; t1 += v1; t2 += v2;
            lo0 = min(lo0, v0); lo1 = min(lo1, v1); lo2 = min(lo2, v2);
            hi0 = max(hi0, v0); hi1 = max(hi1, v1); hi2 = max(hi2, v2);
        }
        s0 += t0; s1 += t1; s2 += t2;
    }
    const unsigned long long cnt = (unsigned long long)r.w * r.h;
    mean[0] = (uint32_t)(s0 / cnt); mean[1] = (uint32_t)(s1 / cnt); mean[2] = (uint32_t)(s2 / cnt);
    return hi0 - mean[0] < a.pr && mean[0] - lo0 < a.pr && hi1 - mean[1] < a.pg && mean[1] - lo1 < a.pg && hi2 - mean[2] < a.pb &&
           mean[2] - lo2 < a.pb;
}

// the recursion below a top-level region; WRITE: leaves go to out[base ..) (those below cap); returns the number of leaves
template <bool WRITE>
__device__ __forceinline__ uint32_t disc_dfs(const DiscArgs &a, DiscRegion r0, uint32_t depth0, unsigned long long base) {
    DiscRegion stk[DISC_STACK];
    uint8_t dstk[DISC_STACK];
    int sp = 0;
    stk[0] = r0;
    dstk[0] = (uint8_t)depth0;
    sp = 1;
    uint32_t n = 0;
    while (sp > 0) {
        --sp;
        const DiscRegion r = stk[sp];
        const uint32_t depth = dstk[sp];
        const bool can_split = (unsigned long long)r.w * r.h > 1 && depth < a.max_s;
        bool split = can_split && depth < a.min_s;
        uint32_t mean[3] = {0, 0, 0};
        if (!split) {
            const bool uni = disc_mean_uniform(a, r, mean);
            split = can_split && !uni;
        }
        if (!split) {
            if (WRITE && base + n < a.cap) {
                stepth_block b;
                b.x = r.x; b.y = r.y; b.w = r.w; b.h = r.h;
                b.value[0] = (uint8_t)mean[0]; b.value[1] = (uint8_t)mean[1]; b.value[2] = (uint8_t)mean[2];
                b.pad = 0;
                a.out[base + n] = b;
            }
            ++n;
            continue;
        }
        if (sp + 2 > DISC_STACK) continue;   // unreachable for images below 2^32 pixels
        DiscRegion c1 = r, c0 = r;
        disc_child(c1, true);
        disc_child(c0, false);
        stk[sp] = c1; dstk[sp] = (uint8_t)(depth + 1); ++sp;   // the second child waits; the first one is next
        stk[sp] = c0; dstk[sp] = (uint8_t)(depth + 1); ++sp;
    }
    return n;
}

template <bool WRITE>
__global__ void __launch_bounds__(128) disc_kernel(DiscArgs a) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (1u << a.top)) return;
    DiscRegion r;
    uint32_t depth;
    uint32_t n = 0;
    if (disc_top_region(a, t, r, depth)) n = disc_dfs<WRITE>(a, r, depth, WRITE ? (unsigned long long)a.counts[t] : 0ull);
    if (!WRITE) a.counts[t] = n;
}

// exclusive scan of counts[0 .. n) in place, the total into counts[n]; ONE CTA of 512 threads.  A thread owns 64 consecutive counts
// per round (read as sixteen 16-byte words, kept in registers); the 512 per-thread sums are scanned with warp shuffles; rounds
// of 32768 counts carry their total forward.
constexpr uint32_t DISC_SCAN_THREADS = 512;
__global__ void __launch_bounds__(DISC_SCAN_THREADS) disc_scan_kernel(uint32_t *counts, uint32_t n) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    constexpr uint32_t PER = 64, NW = DISC_SCAN_THREADS / 32;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (uint32_t r0 = 0; r0 < n; r0 += DISC_SCAN_THREADS * PER) {
        const uint32_t i0 = r0 + tid * PER;
        const bool vec = i0 + PER <= n && (n & 3u) == 0;   // counts comes from cudaMalloc: 16-byte words are aligned
        uint32_t v[PER];
        if (vec) {
            const uint4 *p = reinterpret_cast<const uint4 *>(counts + i0);
#pragma unroll
            for (uint32_t k = 0; k < PER / 4; ++k) {
                const uint4 q = p[k];
                v[4 * k] = q.x; v[4 * k + 1] = q.y; v[4 * k + 2] = q.z; v[4 * k + 3] = q.w;
            }
        } else {
#pragma unroll
            for (uint32_t k = 0; k < PER; ++k) v[k] = i0 + k < n ? counts[i0 + k] : 0u;
        }
        uint32_t sum = 0;
#pragma unroll
        for (uint32_t k = 0; k < PER; ++k) sum += v[k];
        uint32_t inc = sum;   // inclusive scan of the per-thread sums: inside the warp, then over the warp totals
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
            if ((int)lane >= d) inc += t;
        }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = lane < NW ? s_warp[lane] : 0u;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, w, d);
                if ((int)lane >= d) w += t;
            }
            s_warp[lane] = w;   // inclusive totals of warps 0 .. lane
        }
        __syncthreads();
        uint32_t run = s_carry + inc - sum + (warp ? s_warp[warp - 1] : 0u);   // everything before this thread's first count
#pragma unroll
        for (uint32_t k = 0; k < PER; ++k) {
            const uint32_t c = v[k];
            v[k] = run;
            run += c;
        }
        if (vec) {
            uint4 *p = reinterpret_cast<uint4 *>(counts + i0);
#pragma unroll
            for (uint32_t k = 0; k < PER / 4; ++k) p[k] = make_uint4(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
        } else {
#pragma unroll
            for (uint32_t k = 0; k < PER; ++k)
                if (i0 + k < n) counts[i0 + k] = v[k];
        }
        __syncthreads();   // everybody has read s_carry and s_warp
        if (tid == 0) s_carry += s_warp[NW - 1];
        __syncthreads();
    }
    if (tid == 0) counts[n] = s_carry;
}

#endif  // __CUDACC__

}  // namespace stepth
