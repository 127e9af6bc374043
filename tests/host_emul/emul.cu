// emul.cu -- TEST INFRASTRUCTURE: the index math of the production search (stepth_b200/csrc/search_hier.cuh) compiled
// for the HOST, so that tests/test_host_emul.py can check it against the oracle in a container without a GPU.
//
// Only the `__host__ __device__` pieces are exercised (ring keys, the diagonal cut of a tile, the coarse-ring walk and
// its stopping rule, the plane membership bits); the planes themselves are built here by a plain scalar loop, NOT by the
// CUDA builder.  This file is never linked into libstepth_cuda.so and nothing in stepth_b200/ loads it: the product has
// no CPU path.
#include <cstdint>
#include <cstring>
#include <vector>

#include "../../stepth_b200/csrc/search_hier.cuh"

using namespace stepth;

namespace {
void build_planes_host(const uint8_t *add, const TileGeom &g, int pr, int pg, int pb, bool sat, std::vector<uint8_t> &buf) {
    buf.assign(g.pair_bytes, 0);
    unsigned long long *fine = reinterpret_cast<unsigned long long *>(buf.data());
    uint32_t *crp = reinterpret_cast<uint32_t *>(buf.data() + g.off_crp);
    uint32_t *ccp = reinterpret_cast<uint32_t *>(buf.data() + g.off_ccp);
    const int prec[3] = {pr, pg, pb};
    for (uint32_t y = 0; y < g.ha; ++y)
        for (uint32_t x = 0; x < g.wa; ++x) {
            const uint32_t TX = x >> 3, TY = y >> 3, bit = (y & 7) * 8 + (x & 7);
            for (int c = 0; c < 3; ++c) {
                uint32_t lo, hi;
                plane_bits(add[((size_t)y * g.src_wa + x) * 3 + c], prec[c], sat, true, lo, hi);
                unsigned long long *rec = fine + (((size_t)TY * g.ntx + TX) * 3 + c) * T_SLOTS;
                for (int j = 0; j < 32; ++j)
                    if (lo & (1u << j)) rec[W_U0 + j] |= 1ull << bit;
                // coarse level: one plane per exact query value q with v inside T(q)
                const int v = add[((size_t)y * g.src_wa + x) * 3 + c], p = prec[c];
                for (int q = 0; q < 256; ++q) {
                    const bool in = sat ? (v <= q + p - 1) : (v >= q - p + 1 && v <= q + p - 1);
                    if (in) {  // window w holds tiles 32w .. 32w+63: a tile sits in the low half of its own window and in
                               // the high half of the previous one
                        crp[(((size_t)TY * g.nxw + (TX >> 5)) * C_CELL + c * C_Q + q) * 2] |= 1u << (TX & 31);
                        if (TX >= 32) crp[(((size_t)TY * g.nxw + (TX >> 5) - 1) * C_CELL + c * C_Q + q) * 2 + 1] |= 1u << (TX & 31);
                        ccp[(((size_t)TX * g.nyw + (TY >> 5)) * C_CELL + c * C_Q + q) * 2] |= 1u << (TY & 31);
                        if (TY >= 32) ccp[(((size_t)TX * g.nyw + (TY >> 5) - 1) * C_CELL + c * C_Q + q) * 2 + 1] |= 1u << (TY & 31);
                    }
                }
                for (int j = 0; j < W_USED - 32; ++j)
                    if (hi & (1u << j)) rec[j ? 32 + j : W_UM1] |= 1ull << bit;
            }
        }
}
}  // namespace

// dense search of a row band, staged exactly like the kernels: ring 0, coarse rings [0, near), [near, mid), [mid, end)
extern "C" int emul_search_dense(const uint8_t *main_rgba, uint32_t w, uint32_t h, uint32_t y0, const uint8_t *add_rgb,
                                 uint32_t wa, uint32_t ha, uint32_t add_y0, uint32_t add_img_h, const uint8_t *precision,
                                 int sat, uint32_t max_rings, int near_rings, int mid_rings, uint8_t *d8_out) {
    uint32_t cover_w = (uint32_t)std::min<uint64_t>(wa, (uint64_t)w / 2 + max_rings);
    const int64_t last_row = (int64_t)(y0 + h) / 2 + max_rings - 1 - (int64_t)add_y0;
    uint32_t cover_h = (uint32_t)std::max<int64_t>(1, std::min<int64_t>(ha, last_row + 1));
    TileGeom g = make_tile_geom(cover_w, cover_h, wa);
    std::vector<uint8_t> planes;
    build_planes_host(add_rgb, g, precision[0], precision[1], precision[2], sat != 0, planes);
    unsigned long long t_le[DIAG_N], t_ge[DIAG_N];
    for (int t = 0; t < DIAG_N; ++t) { t_le[t] = diag_le(t - 9); t_ge[t] = diag_ge(t); }
    const HierView v = make_view(planes.data(), g, 0);
    for (uint32_t y = 0; y < h; ++y)
        for (uint32_t x = 0; x < w; ++x) {
            const uint8_t *px = main_rgba + ((size_t)y * w + x) * 4;
            const uint32_t col = px[0] | (px[1] << 8) | (px[2] << 16);
            const int ox = (int)((x + 1) >> 1), oyg = (int)((y + y0 + 1) >> 1);
            uint32_t key = KEY_NONE;
            if (ox < (int)wa && oyg < (int)add_img_h) {
                const int oyl = oyg - (int)add_y0;
                bool hit0 = false;
                if (oyl >= 0 && oyl < (int)ha) {
                    const uint8_t *a = add_rgb + ((size_t)oyl * wa + ox) * 3;
                    const ColourBox box = make_box(px[0], px[1], px[2], precision[0], precision[1], precision[2], sat != 0);
                    hit0 = box_contains(box, a[0], a[1], a[2]);
                }
                if (hit0) key = 0;
                else if (max_rings > 1) {
                    const HierQuery q = make_query(ox, oyl, col);
                    bool done = hier_rings(v, q, (int)max_rings, 0, near_rings, t_le, t_ge, key);
                    int s = near_rings;
                    if (!done && mid_rings > s) { done = hier_rings(v, q, (int)max_rings, s, mid_rings, t_le, t_ge, key); s = mid_rings; }
                    if (!done) done = hier_rings(v, q, (int)max_rings, s, 1 << 20, t_le, t_ge, key);
                    if (!done) return 1;
                }
            }
            d8_out[(size_t)y * w + x] = key == KEY_NONE ? 0 : key_to_d8(key);
        }
    return 0;
}

// tile_min_key against a brute-force arg-min of ring_key over the set bits
extern "C" int emul_check_tile_min_key(unsigned long long m, int bx, int by, uint32_t *got, uint32_t *want) {
    unsigned long long t_le[DIAG_N], t_ge[DIAG_N];
    for (int t = 0; t < DIAG_N; ++t) { t_le[t] = diag_le(t - 9); t_ge[t] = diag_ge(t); }
    *got = tile_min_key(m, bx, by, t_le, t_ge);
    uint32_t best = KEY_NONE;
    for (int b = 0; b < 64; ++b)
        if (m & (1ull << b)) {
            const uint32_t k = ring_key(bx + (b & 7), by + (b >> 3));
            if (k < best) best = k;
        }
    *want = best;
    return *got == best;
}
