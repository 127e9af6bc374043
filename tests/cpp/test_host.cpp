// test_host.cpp -- the C++ host mirror (include/stepth.hpp) against the CPU oracle, written the way the reference's
// README uses the crate (Readme.md:13-24): open -> depth from additional -> invert -> select_foreground -> apply_mask.
// Built and run by tests/test_gpu_parity.py::test_cpp_host_mirror (needs a B200; links libstepth_cuda.so and, as the
// checker only, oracle/libstepth_oracle.so).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/stepth.hpp"
#include "../../oracle/stepth_oracle.h"

static uint32_t rng_state = 12345;
static uint32_t rnd() { rng_state = rng_state * 1664525u + 1013904223u; return rng_state >> 8; }

#define CHECK(cond)                                                        \
    do {                                                                   \
        if (!(cond)) { std::printf("FAIL %s:%d %s\n", __FILE__, __LINE__, #cond); return 1; } \
    } while (0)

int main() {
    const uint32_t w = 150, h = 90, rings = 255;
    const uint8_t prec[3] = {36, 36, 36};
    // smooth-ish synthetic pair with a few levels so that matches exist at many distances
    std::vector<uint8_t> rgba((size_t)w * h * 4), add((size_t)w * h * 3);
    for (uint32_t y = 0; y < h; ++y)
        for (uint32_t x = 0; x < w; ++x) {
            for (int c = 0; c < 3; ++c) {
                rgba[((size_t)y * w + x) * 4 + c] = (uint8_t)(((x / 9 + y / 7 + c) % 6) * 45 + rnd() % 8);
                add[((size_t)y * w + x) * 3 + c] = (uint8_t)((((x + 20) / 11 + y / 5 + c) % 6) * 45 + rnd() % 8);
            }
            rgba[((size_t)y * w + x) * 4 + 3] = 255;
        }

    // ---- block mode, as load_depth_from_additional does (depth_image.rs:91-136) --------------------------------------
    stepth::DepthImage img = stepth::DepthImage::from_image(rgba, w, h);
    img.load_depth_from_additional(add.data(), w, h, prec);
    const uint32_t max_splits = (uint32_t)std::ceil(std::log2((float)(w * h)));
    size_t nb = stepth_host_discretise_v0(rgba.data(), 4, w, h, prec, 16, max_splits, nullptr, 0);
    std::vector<stepth_block> blocks(nb);
    stepth_host_discretise_v0(rgba.data(), 4, w, h, prec, 16, max_splits, blocks.data(), nb);
    static_assert(sizeof(stepth_block) == sizeof(orc_block), "block layouts must agree");
    std::vector<uint8_t> ref((size_t)w * h);
    CHECK(orc_depth_from_blocks(add.data(), w, h, (const orc_block *)blocks.data(), nb, prec, rings, w, h, ORC_PIXELOPPS_ABS_ALL,
                                ORC_RESAMPLE_U8, 4, ref.data(), nullptr) == ORC_OK);
    CHECK(img.depth() == ref);

    // ---- dense mode + the README tail --------------------------------------------------------------------------------
    img.load_depth_from_additional_dense(add.data(), w, h, prec);
    CHECK(orc_depth_dense(rgba.data(), 4, w, h, add.data(), w, h, prec, rings, ORC_PIXELOPPS_ABS_ALL, ORC_RESAMPLE_U8, 4, ref.data(),
                          nullptr) == ORC_OK);
    CHECK(img.depth() == ref);
    img.invert_depth();
    orc_invert_depth(ref.data(), ref.size());
    CHECK(img.depth() == ref);
    stepth::MaskImage fg = img.select_foreground();
    std::vector<uint8_t> mask_ref((size_t)w * h);
    uint8_t ft[2];
    CHECK(orc_select_foreground(ref.data(), ref.size(), mask_ref.data(), ft) == ORC_OK);
    CHECK(fg.mask == mask_ref);
    fg.apply_mask();
    std::vector<uint8_t> rgba_ref = rgba;
    orc_apply_mask(rgba_ref.data(), mask_ref.data(), mask_ref.size());
    CHECK(fg.image == rgba_ref);
    auto cl = img.depth_split(2);
    CHECK(cl.size() == 2 && cl[0].first && *cl[0].first == ft[0] && *cl[0].second == ft[1]);
    CHECK(img.depth_split(1).size() == 1 && !img.depth_split(1)[0].first);
    stepth::MaskImage sl = img.slice(10, std::nullopt);
    orc_slice(ref.data(), ref.size(), 10, -1, mask_ref.data());
    CHECK(sl.mask == mask_ref);

    // ---- error behaviour (depth_image.rs:41-48, :127) ---------------------------------------------------------------------
    bool threw = false;
    try { img.load_depth(std::vector<uint8_t>(10), 5, 2); } catch (const stepth::IoError &e) { threw = std::strcmp(e.what(), "Sizes don't match") == 0; }
    CHECK(threw);
    threw = false;
    std::vector<uint8_t> far_add((size_t)w * h * 3, 200), dark((size_t)w * h * 4, 10);
    stepth::DepthImage none = stepth::DepthImage::from_image(dark, w, h);
    try { none.load_depth_from_additional_dense(far_add.data(), w, h, prec); } catch (const stepth::Panic &e) { threw = e.code == STEPTH_ERR_DIV_BY_ZERO_MAX; }
    CHECK(threw);
    std::printf("cpp host mirror ok: %zu blocks, foreground range [%u, %u]\n", nb, ft[0], ft[1]);
    return 0;
}
