// search_literal.cuh -- K1/v0: one thread per query walking the reference's probe order.
//
// Debug / cross-check path (STEPTH_FLAG_LITERAL_SEARCH).  It follows helpers.rs:26-52 ring by ring with the
// duplicate probes removed (they cannot change the first match) and early exit on the first hit, reading the
// caller's interleaved RGB bytes directly.  The production search is search_tiles.cuh; this kernel exists so
// that large inputs, where the CPU oracle needs hours, still have an independent in-order implementation.
#pragma once
#include "common.cuh"

namespace stepth {

struct SearchArgs {
    // queries: dense (one per main pixel) or a block list
    const uint8_t *main_rgba;    // [n_pairs, h, w, 4]
    const stepth_block *blocks;  // [n_queries] (block mode, n_pairs == 1)
    uint32_t n_queries;          // per pair
    uint32_t w, h, y0;           // dense geometry: band width / rows / global row of band row 0
    // additional image
    const uint8_t *add_rgb;      // [n_pairs, ha, wa, 3]
    uint32_t wa, ha, add_y0, add_img_h;
    uint32_t max_rings;
    uint32_t pr, pg, pb;
    uint32_t sat;
    uint8_t *d8;                 // [n_pairs, n_queries]
    uint32_t *max_out;           // [n_pairs]
    unsigned long long *probes_out;  // [n_pairs] or null (literal kernel only): in-bounds pixels the reference order visits, each once
};

// origin + colour box of query q (depth_image.rs:114-118)
__device__ __forceinline__ void load_query(const SearchArgs &a, uint32_t pair, uint32_t q, int &ox, int &oy,
                                           ColourBox &box) {
    uint32_t r, g, b;
    if (a.blocks) {
        stepth_block bl = a.blocks[q];
        ox = (int)((bl.x + bl.w) / 2);  // literal (pos + size) / 2, not the centre
        oy = (int)((bl.y + bl.h) / 2);
        r = bl.value[0]; g = bl.value[1]; b = bl.value[2];
    } else {
        uint32_t x = q % a.w, y = q / a.w;
        ox = (int)((x + 1) / 2);
        oy = (int)((y + a.y0 + 1) / 2);
        uint32_t px = reinterpret_cast<const uint32_t *>(a.main_rgba)[(size_t)pair * a.n_queries + q];
        r = px & 0xff; g = (px >> 8) & 0xff; b = (px >> 16) & 0xff;
    }
    box = make_box(r, g, b, a.pr, a.pg, a.pb, a.sat != 0);
}

__device__ __forceinline__ void publish_max(uint32_t d, uint32_t *max_out) {
    uint32_t m = __reduce_max_sync(__activemask(), d);
    if (m > 0) {
        // one atomic per warp, skipped when it cannot raise the maximum
        uint32_t leader = __ffs(__activemask()) - 1;
        if ((threadIdx.x & 31) == leader) atomicMax(max_out, m);
    }
}

__global__ void __launch_bounds__(256) search_literal_kernel(SearchArgs a) {
    const uint32_t pair = blockIdx.y;
    const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t d = 0;
    unsigned long long n_probes = 0;
    if (q < a.n_queries) {
        int ox, oy;
        ColourBox box;
        load_query(a, pair, q, ox, oy, box);
        const uint8_t *add = a.add_rgb + (size_t)pair * a.ha * a.wa * 3;
        const int wa = (int)a.wa, ha_g = (int)a.add_img_h;
        const int oyl = oy - (int)a.add_y0;  // origin row inside the sub-image held on this device
        uint32_t key = KEY_NONE;
        // ring 0 out of bounds => `still_in_bounds` stays false => break => None   (helpers.rs:49-51)
        if (ox < wa && oy < ha_g) {
            auto probe = [&](int dx, int dy) -> bool {
                int x = ox + dx, yl = oyl + dy;
                if (x < 0 || x >= wa || yl < 0 || yl >= (int)a.ha) return false;
                ++n_probes;
                const uint8_t *v = add + ((size_t)yl * wa + x) * 3;
                return box_contains(box, v[0], v[1], v[2]);
            };
            for (int s = 0; s < (int)a.max_rings && key == KEY_NONE; ++s) {
                if (ox - s < 0 && ox + s >= wa && oy - s < 0 && oy + s >= ha_g) break;
                for (int dx = -s; dx <= s && key == KEY_NONE; ++dx)
                    if (probe(dx, s)) key = ring_key(dx, s);
                if (s == 0) continue;
                for (int dx = -s; dx <= s && key == KEY_NONE; ++dx)
                    if (probe(dx, -s)) key = ring_key(dx, -s);
                for (int dy = -s + 1; dy <= s - 1 && key == KEY_NONE; ++dy)
                    if (probe(s, dy)) key = ring_key(s, dy);
                for (int dy = -s + 1; dy <= s - 1 && key == KEY_NONE; ++dy)
                    if (probe(-s, dy)) key = ring_key(-s, dy);
            }
        }
        d = key_to_d8(key);
        a.d8[(size_t)pair * a.n_queries + q] = (uint8_t)d;
    }
    publish_max(d, a.max_out + pair);
    if (a.probes_out) {   // measurement aid (bench.py's algorithmic-work figure): one atomic per warp
        for (int o = 16; o; o >>= 1) n_probes += __shfl_xor_sync(0xffffffffu, n_probes, o);
        if ((threadIdx.x & 31) == 0 && n_probes) atomicAdd(a.probes_out + pair, n_probes);
    }
}

}  // namespace stepth
