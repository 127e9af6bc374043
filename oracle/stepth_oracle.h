/*
 * stepth_oracle.h -- CPU restatement of nikilark/stepth's depth-from-additional hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may call it,
 * and only as the checker or as the timed CPU baseline.
 *
 * PARITY UNPINNED: the reference ships no tests, golden vectors or known-answer fixtures
 * (SURVEY.md section 4), cannot be built here (no Rust toolchain) and its arithmetic
 * dependency `disage` @ b2920765b5398ab8259482dc942c2bb7bb94fcd5 is an empty submodule.
 * The restatement follows the reference source line by line (citations on every function);
 * the unpinned third-party pieces (disage PixelOpps, image 0.23.14 Gaussian resample) are
 * behind switches.  Soft pins: assets/depth.jpg and assets/foreground.jpg (tests/golden/).
 */
#ifndef STEPTH_ORACLE_H
#define STEPTH_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* status codes: identical numbering to include/stepth_cuda.h */
enum {
    ORC_OK = 0,
    ORC_INVALID_ARG = 1,
    ORC_SIZE_MISMATCH = 2,     /* "Sizes don't match"            depth_image.rs:41-48 */
    ORC_DIV_BY_ZERO_MAX = 3,   /* max == 0 -> integer division panic   depth_image.rs:127 */
    ORC_TOO_FEW_BLOCKS = 4,    /* len/8 == 0 -> par_chunks_mut(0) panic depth_image.rs:111-112 */
    ORC_DEGENERATE_DEPTH = 5,  /* step_by(0) / usize underflow / unwrap(None) depth_image.rs:212-216 */
    ORC_NO_CONVERGENCE = 6,    /* k-means recursion did not terminate (reference: stack overflow) */
    ORC_OUT_OF_BOUNDS = 7      /* ImageBuffer::get_pixel out of bounds -> panic   mask_image.rs:106 */
};

/* disage::pixels::PixelOpps<[u8;3]> semantics (source absent -- SURVEY.md F6). */
enum {
    ORC_PIXELOPPS_ABS_ALL = 0, /* for all c: |v_c - what_c| <  precision_c      (default, best evidence) */
    ORC_PIXELOPPS_SAT_ALL = 1, /* for all c: sat_sub(v_c, what_c) < precision_c                         */
    ORC_PIXELOPPS_ABS_SUM = 2  /* sum_c |v_c - what_c| < sum_c precision_c       (oracle only)           */
};

/* image 0.23.14 imageops::resize(.., Gaussian) at unit scale (source absent -- SURVEY.md F7). */
enum {
    ORC_RESAMPLE_OFF = 0,
    ORC_RESAMPLE_U8 = 1,  /* vertical pass rounds to u8, then horizontal (0.23.x generic sampler; default) */
    ORC_RESAMPLE_F32 = 2  /* vertical pass keeps f32 (0.24-style Rgba32FImage intermediate)               */
};

typedef struct {
    uint32_t x, y, w, h; /* DiscretePixel.position / .size       depth_image.rs:114-116 */
    uint8_t value[3];    /* DiscretePixel.value ([u8;3])          depth_image.rs:118,121 */
    uint8_t pad;
} orc_block;

/* helpers.rs:3-7 */
uint32_t orc_distance_dot_dot(uint32_t x1, uint32_t y1, uint32_t x2, uint32_t y2);

/* helpers.rs:9-54.  returns 1 = Some, 0 = None.  *probes counts get2d calls that returned Some. */
int orc_distance_dot_array(const uint8_t what[3], const uint8_t *add_rgb, uint32_t wa, uint32_t ha,
                           uint32_t from_x, uint32_t from_y, uint32_t max, const uint8_t precision[3],
                           int pixelopps, uint32_t *dist, uint32_t *pos_x, uint32_t *pos_y,
                           uint64_t *probes);

/* depth_image.rs:110-122: per-block search; writes (d as u8) per block.  threads: 0 = reference-faithful
 * rayon chunking (len/8 -> 8 or 9 tasks), n>0 = n equal chunks.  probes_total may be NULL. */
int32_t orc_search_blocks(const uint8_t *add_rgb, uint32_t wa, uint32_t ha, const orc_block *blocks,
                          size_t n_blocks, const uint8_t precision[3], uint32_t max_rings, int pixelopps,
                          int threads, uint8_t *d8_out, uint64_t *probes_total);

/* dense limit case: every main pixel is a 1x1 block whose value is the pixel (main_stride = 3 or 4 bytes/px). */
int32_t orc_search_dense(const uint8_t *main_px, uint32_t main_stride, uint32_t w, uint32_t h,
                         const uint8_t *add_rgb, uint32_t wa, uint32_t ha, const uint8_t precision[3],
                         uint32_t max_rings, int pixelopps, int threads, uint8_t *d8_out,
                         uint64_t *probes_total);

/* in-bounds probes of the last orc_search_* call with every pixel of a ring counted once (the reference probes the
 * corners of a ring twice and ring 0 four times, helpers.rs:28-31): the algorithmic-work figure of SURVEY.md 8(d).
 * Not thread safe across concurrent searches; bench bookkeeping only. */
uint64_t orc_last_unique_probes(void);

/* as orc_search_dense but only for the listed pixel indices (y*w+x); used for bounded CPU-baseline samples */
int32_t orc_search_dense_subset(const uint8_t *main_px, uint32_t main_stride, uint32_t w, uint32_t h,
                                const uint8_t *add_rgb, uint32_t wa, uint32_t ha, const uint8_t precision[3],
                                uint32_t max_rings, int pixelopps, int threads, const uint32_t *indices,
                                size_t n_idx, uint8_t *d8_out, uint64_t *probes_total);

/* depth_image.rs:124-129: global max + (v*255/max) as u8 in place; ORC_DIV_BY_ZERO_MAX when max == 0 */
int32_t orc_normalise(uint8_t *d8, size_t n, uint8_t *max_out);

/* depth_image.rs:130-132: collect() paints each block rectangle, to_luma8_from_rgb8 of (v,v,v) = v */
int32_t orc_paint_blocks(const orc_block *blocks, const uint8_t *values, size_t n_blocks, uint32_t w,
                         uint32_t h, uint8_t *luma_out);

/* depth_image.rs:133-134 -> image 0.23.14 imageops::sample (vertical_sample then horizontal_sample) */
int32_t orc_gaussian_resample_unit(const uint8_t *src, uint32_t w, uint32_t h, int mode, uint8_t *dst);

/* whole of load_depth_from_additional (depth_image.rs:91-136) from a block list */
int32_t orc_depth_from_blocks(const uint8_t *add_rgb, uint32_t wa, uint32_t ha, const orc_block *blocks,
                              size_t n_blocks, const uint8_t precision[3], uint32_t max_rings, uint32_t w,
                              uint32_t h, int pixelopps, int resample, int threads, uint8_t *depth_out,
                              uint8_t *raw_d8_out /* n_blocks, may be NULL */);

int32_t orc_depth_dense(const uint8_t *main_px, uint32_t main_stride, uint32_t w, uint32_t h,
                        const uint8_t *add_rgb, uint32_t wa, uint32_t ha, const uint8_t precision[3],
                        uint32_t max_rings, int pixelopps, int resample, int threads, uint8_t *depth_out,
                        uint8_t *raw_d8_out /* w*h, may be NULL */);

/* depth_image.rs:225-227 */
void orc_invert_depth(uint8_t *depth, size_t n);

/* depth_image.rs:162-218.  ranges_out[2*i] = min, [2*i+1] = max of cluster i; *n_out clusters;
 * has_values_out = 0 for the zones<2 "[(None,None)]" answer.  centroids_out / iters_out may be NULL. */
int32_t orc_depth_split(const uint8_t *depth, size_t n, uint8_t zones, uint8_t *ranges_out /* 2*256 */,
                        uint32_t *n_out, int *has_values_out, uint8_t *centroids_out /* 256 */,
                        uint32_t *iters_out);

/* depth_image.rs:229-245 (from/to < 0 mean None) */
void orc_slice(const uint8_t *depth, size_t n, int from, int to, uint8_t *mask_out);

/* depth_image.rs:220-223 */
int32_t orc_select_foreground(const uint8_t *depth, size_t n, uint8_t *mask_out, uint8_t from_to_out[2]);

/* mask_image.rs:205-213 */
void orc_apply_mask(uint8_t *rgba, const uint8_t *mask, size_t n);

/* editor operations (SURVEY 8(f) "next" row 3) */
void orc_highlight_depth(const uint8_t *rgba_in, const uint8_t *depth, size_t n, uint8_t *rgba_out); /* depth_image.rs:51-63 */
void orc_highlight_mask(const uint8_t *rgba_in, const uint8_t *mask, size_t n, uint8_t *rgba_out);   /* mask_image.rs:57-73  */
void orc_mask_and(uint8_t *mask, const uint8_t *other, size_t n);                                    /* mask_image.rs:147-169 (equal sizes) */
void orc_mask_or(uint8_t *mask, const uint8_t *other, size_t n);                                     /* mask_image.rs:171-193 (equal sizes) */
void orc_mask_not(uint8_t *mask, size_t n);                                                          /* mask_image.rs:195-197 */
/* mask_image.rs:99-109; returns 7 (out of bounds) where ImageBuffer::get_pixel would panic */
int32_t orc_image_replace(uint8_t *rgba, const uint8_t *mask, uint32_t w, uint32_t h, const uint8_t *other_rgba,
                          uint32_t ow, uint32_t oh, uint32_t start_x, uint32_t start_y);

/* operations.rs:3-36 / 38-80 (SURVEY 8(f) "next" row 1) */
void orc_normalize_brightness_luma16(const uint16_t *img1, size_t n1, const uint16_t *img2, size_t n2,
                                     double percent, uint16_t *out);
void orc_normalize_brightness_rgb16(const uint16_t *img1, size_t n1_px, const uint16_t *img2, size_t n2_px,
                                    double percent, uint16_t *out);

#ifdef __cplusplus
}
#endif
#endif
