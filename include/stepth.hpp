// stepth.hpp -- header-only C++17 host side over the C ABI (include/stepth_cuda.h).
//
// The reference is a Rust crate and no Rust toolchain exists in the build image, so the host layer above the C ABI is
// mirrored here in C++ (and in rust/ as source).  Names, argument meaning and error behaviour follow
// the reference's public API (src/lib.rs:3-12):
//
//   DepthImage::from_image / load_depth / depth / width / height        src/depth_image.rs:23-49,138-144
//   DepthImage::load_depth_from_additional(add, precision)              src/depth_image.rs:91-136
//   DepthImage::invert_depth / depth_split / slice / select_foreground  src/depth_image.rs:162-245
//   MaskImage::apply_mask, MASK_TRUE / MASK_FALSE                       src/mask_image.rs:3-4,205-213
//
// Decoding (`DepthImage::open`, `image::open`) is the host image library's business, exactly as in the reference; pixels
// enter here as tightly packed row-major bytes (`ImageBuffer<_, Vec<u8>>`).
//   io::Error(InvalidInput, msg)   -> stepth::IoError      ("Sizes don't match", "Failed to open image")
//   arithmetic panics              -> stepth::Panic        (divide by zero :127, par_chunks_mut(0) :111-112, :212-216)
// There is no CPU fallback: constructing a Context without an sm_100 device throws.
#pragma once
#include <cmath>
#include <cstdint>
#include <memory>
#include <optional>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "stepth_cuda.h"

namespace stepth {

constexpr uint8_t MASK_TRUE = 255;  // src/mask_image.rs:3
constexpr uint8_t MASK_FALSE = 0;   // src/mask_image.rs:4

struct IoError : std::runtime_error {  // io::Error(ErrorKind::InvalidInput, ..)
    using std::runtime_error::runtime_error;
};
struct Panic : std::runtime_error {    // a Rust panic of the reference
    int32_t code;
    Panic(int32_t c, const std::string &m) : std::runtime_error(m), code(c) {}
};
struct CudaError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

class Context {
public:
    explicit Context(int device = 0) {
        if (stepth_cuda_ctx_create(device, &ctx_) != STEPTH_OK)
            throw CudaError(std::string(stepth_cuda_last_error(nullptr)) + " -- stepth has no CPU fallback");
    }
    ~Context() { stepth_cuda_ctx_destroy(ctx_); }
    Context(const Context &) = delete;
    Context &operator=(const Context &) = delete;
    stepth_ctx *get() const { return ctx_; }

    void check(int32_t rc) const {
        switch (rc) {
            case STEPTH_OK: return;
            case STEPTH_ERR_SIZE_MISMATCH: throw IoError("Sizes don't match");
            case STEPTH_ERR_INVALID_ARG: throw IoError(stepth_cuda_last_error(ctx_));
            case STEPTH_ERR_DIV_BY_ZERO_MAX: throw Panic(rc, "attempt to divide by zero");
            case STEPTH_ERR_TOO_FEW_BLOCKS: throw Panic(rc, "chunk_size must not be zero");
            case STEPTH_ERR_DEGENERATE_DEPTH: throw Panic(rc, "depth_split on a degenerate depth range");
            case STEPTH_ERR_NO_CONVERGENCE: throw Panic(rc, "depth_split did not converge");
            case STEPTH_ERR_OUT_OF_BOUNDS: throw Panic(rc, "Image index out of bounds");
            default: throw CudaError(stepth_cuda_last_error(ctx_));
        }
    }

private:
    stepth_ctx *ctx_ = nullptr;
};

inline std::shared_ptr<Context> default_context() {
    static std::shared_ptr<Context> c = std::make_shared<Context>(0);
    return c;
}

// src/mask_image.rs:6-10  { image: RgbaImage, mask: GrayImage }
struct MaskImage {
    uint32_t w = 0, h = 0;
    std::vector<uint8_t> image;  // RGBA8
    std::vector<uint8_t> mask;   // Luma8
    std::shared_ptr<Context> ctx;

    uint32_t width() const { return w; }
    uint32_t height() const { return h; }
    // src/mask_image.rs:57-73 -> RGBA8
    std::vector<uint8_t> highlight_mask() const {
        auto c = ctx ? ctx : default_context();
        std::vector<uint8_t> out(image.size());
        c->check(stepth_cuda_highlight_mask(c->get(), image.data(), mask.data(), (size_t)w * h, out.data()));
        return out;
    }
    // src/mask_image.rs:99-109
    void image_replace(const MaskImage &other, uint32_t start_x, uint32_t start_y) {
        auto c = ctx ? ctx : default_context();
        c->check(stepth_cuda_image_replace(c->get(), image.data(), mask.data(), w, h, other.image.data(), other.w, other.h,
                                           start_x, start_y));
    }
    // src/mask_image.rs:147-197 (equal-sized masks)
    void mask_and(const MaskImage &other) {
        if (other.w != w || other.h != h) throw IoError("Sizes don't match");
        auto c = ctx ? ctx : default_context();
        c->check(stepth_cuda_mask_and(c->get(), mask.data(), other.mask.data(), mask.size()));
    }
    void mask_or(const MaskImage &other) {
        if (other.w != w || other.h != h) throw IoError("Sizes don't match");
        auto c = ctx ? ctx : default_context();
        c->check(stepth_cuda_mask_or(c->get(), mask.data(), other.mask.data(), mask.size()));
    }
    void mask_not() {
        auto c = ctx ? ctx : default_context();
        c->check(stepth_cuda_mask_not(c->get(), mask.data(), mask.size()));
    }
    // src/mask_image.rs:205-213
    void apply_mask() {
        auto c = ctx ? ctx : default_context();
        c->check(stepth_cuda_apply_mask(c->get(), image.data(), mask.data(), (size_t)w * h));
    }
};

// src/depth_image.rs:6-10  { image: RgbaImage, depth: GrayImage }
class DepthImage {
public:
    static constexpr uint32_t MAX_RINGS = 255;   // hard-coded `max`, src/depth_image.rs:119
    static constexpr uint32_t MIN_SPLITS = 16;   // src/depth_image.rs:102

    // src/depth_image.rs:23-27
    static DepthImage from_image(std::vector<uint8_t> rgba, uint32_t w, uint32_t h, std::shared_ptr<Context> ctx = nullptr) {
        if (rgba.size() != (size_t)w * h * 4) throw IoError("Sizes don't match");
        DepthImage d;
        d.w_ = w; d.h_ = h;
        d.image_ = std::move(rgba);
        d.depth_.assign((size_t)w * h, 0);
        d.ctx_ = ctx ? ctx : default_context();
        return d;
    }
    const std::vector<uint8_t> &image() const { return image_; }       // :29-31
    std::vector<uint8_t> depth() const { return depth_; }               // :33-35 (a clone)
    uint32_t width() const { return w_; }                               // :138-140
    uint32_t height() const { return h_; }                              // :142-144

    // src/depth_image.rs:37-49
    void load_depth(std::vector<uint8_t> depth, uint32_t w, uint32_t h) {
        if (w != w_ || h != h_ || depth.size() != (size_t)w * h) throw IoError("Sizes don't match");
        depth_ = std::move(depth);
    }

    // src/depth_image.rs:91-136.  `blocks` = the DiscretePixel list of disage::open::rgb_discrete (:104-110); when absent
    // the documented stand-in discretiser is used (NOT disage, whose source is unavailable).
    void load_depth_from_additional(const uint8_t *add_rgb, uint32_t wa, uint32_t ha, const uint8_t (&precision)[3],
                                    const std::vector<stepth_block> *blocks = nullptr, uint32_t flags = STEPTH_RESAMPLE_U8,
                                    uint32_t max_rings = MAX_RINGS) {
        stepth_params p{};
        p.precision[0] = precision[0]; p.precision[1] = precision[1]; p.precision[2] = precision[2];
        p.max_rings = max_rings;
        p.flags = flags & ~(uint32_t)(STEPTH_FLAG_FOREGROUND | STEPTH_FLAG_APPLY);
        std::vector<uint8_t> depth((size_t)w_ * h_);
        if (!blocks) {   // the stand-in discretiser runs on the device in the same call; its block list never visits the host
            const uint32_t max_splits = (uint32_t)std::ceil(std::log2((float)((size_t)w_ * h_)));  // :101-103
            ctx_->check(stepth_cuda_depth_from_additional(ctx_->get(), image_.data(), w_, h_, add_rgb, wa, ha, &p, MIN_SPLITS, max_splits,
                                                          depth.data(), nullptr, nullptr, nullptr, nullptr));
            load_depth(std::move(depth), w_, h_);  // :135
            return;
        }
        ctx_->check(stepth_cuda_depth_from_blocks(ctx_->get(), add_rgb, wa, ha, blocks->data(), blocks->size(), w_, h_, &p,
                                                  depth.data(), nullptr, nullptr, nullptr, nullptr, nullptr));
        load_depth(std::move(depth), w_, h_);  // :135
    }

    // the dense limit case (every pixel its own block): BASELINE.json's synthetic configs
    void load_depth_from_additional_dense(const uint8_t *add_rgb, uint32_t wa, uint32_t ha, const uint8_t (&precision)[3],
                                          uint32_t flags = STEPTH_RESAMPLE_U8, uint32_t max_rings = MAX_RINGS) {
        stepth_params p{};
        p.precision[0] = precision[0]; p.precision[1] = precision[1]; p.precision[2] = precision[2];
        p.max_rings = max_rings;
        p.flags = flags & ~(uint32_t)(STEPTH_FLAG_FOREGROUND | STEPTH_FLAG_APPLY);
        std::vector<uint8_t> depth((size_t)w_ * h_);
        ctx_->check(stepth_cuda_depth_dense(ctx_->get(), image_.data(), w_, h_, add_rgb, wa, ha, &p, depth.data(), nullptr,
                                            nullptr, nullptr, nullptr));
        load_depth(std::move(depth), w_, h_);
    }

    // src/depth_image.rs:51-63 -> RGBA8
    std::vector<uint8_t> highlight_depth() const {
        std::vector<uint8_t> out(image_.size());
        ctx_->check(stepth_cuda_highlight_depth(ctx_->get(), image_.data(), depth_.data(), depth_.size(), out.data()));
        return out;
    }

    // src/depth_image.rs:225-227
    void invert_depth() { ctx_->check(stepth_cuda_invert_depth(ctx_->get(), depth_.data(), depth_.size())); }

    // src/depth_image.rs:162-218: (Some(min), Some(max)) per cluster; zones < 2 -> [(None, None)]
    using Range = std::pair<std::optional<uint8_t>, std::optional<uint8_t>>;
    std::vector<Range> depth_split(uint8_t zones) const {
        uint8_t ranges[512];
        uint32_t n = 0;
        int32_t has = 0;
        ctx_->check(stepth_cuda_depth_split(ctx_->get(), depth_.data(), depth_.size(), zones, ranges, &n, &has));
        std::vector<Range> out;
        if (!has) { out.emplace_back(std::nullopt, std::nullopt); return out; }
        for (uint32_t i = 0; i < n; ++i) out.emplace_back(ranges[2 * i], ranges[2 * i + 1]);
        return out;
    }

    // src/depth_image.rs:229-245
    MaskImage slice(std::optional<uint8_t> from, std::optional<uint8_t> to) const {
        MaskImage m{w_, h_, image_, std::vector<uint8_t>((size_t)w_ * h_), ctx_};
        ctx_->check(stepth_cuda_slice(ctx_->get(), depth_.data(), depth_.size(), from ? (int32_t)*from : -1,
                                      to ? (int32_t)*to : -1, m.mask.data()));
        return m;
    }

    // src/depth_image.rs:220-223
    MaskImage select_foreground() const {
        MaskImage m{w_, h_, image_, std::vector<uint8_t>((size_t)w_ * h_), ctx_};
        uint8_t ft[2];
        ctx_->check(stepth_cuda_select_foreground(ctx_->get(), depth_.data(), depth_.size(), m.mask.data(), ft));
        return m;
    }

private:
    uint32_t w_ = 0, h_ = 0;
    std::vector<uint8_t> image_, depth_;
    std::shared_ptr<Context> ctx_;
};

}  // namespace stepth
