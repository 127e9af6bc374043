/*
 * stepth_cuda.h -- C ABI of libstepth_cuda.so, the B200 (sm_100a) implementation of nikilark/stepth's
 * depth-from-additional hot path.  This is the drop-in boundary: the patched `stepth` crate (rust/) binds
 * exactly these symbols through its `stepth-cuda-sys` FFI crate; stepth_b200/ (Python) and include/stepth.hpp
 * (C++) bind the same symbols.  Plain pointers and sizes only; no exceptions cross the boundary.
 *
 * Reference interfaces replaced (file:line into nikilark/stepth):
 *   src/depth_image.rs:110-134  per-block ring search, `as u8`, global max, normalise, paint, luma, Gaussian resample
 *   src/helpers.rs:3-54         distance_dot_dot / distance_dot_array (the search itself)
 *   src/depth_image.rs:225-227  invert_depth
 *   src/depth_image.rs:162-218  depth_split
 *   src/depth_image.rs:220-245  select_foreground / slice
 *   src/mask_image.rs:205-213   MaskImage::apply_mask
 *   src/operations.rs:3-80      normalize_brightness_{luma16,rgb16}   
 *   src/depth_image.rs:51-63, src/mask_image.rs:57-73,99-109,147-197   highlight_*, image_replace, mask_and/or/not (8(f) row 3)
 *
 * Ownership: the caller owns every buffer; the library borrows pointers for the duration of a call and
 * never frees or retains them.  Device buffers, pinned staging and streams live in the context.
 * Threading: one call at a time per context; different contexts may be used from different threads.
 * All `stepth_cuda_*` host-buffer calls are synchronous (they return after the device->host copies finish),
 * like the reference.  `stepth_cuda_dev_*` calls take DEVICE pointers and a CUDA stream and only enqueue work.
 * Host buffers may be ordinary pageable memory (a Rust Vec<u8>): the batch path stages them through a page-locked ring of
 * its own with a few copy threads (STEPTH_STAGE_THREADS, default min(8, cores / 2)); page-locked buffers are used in place.
 *
 * There is NO CPU fallback: every entry point fails with STEPTH_ERR_CUDA when no sm_100 device is usable.
 */
#ifndef STEPTH_CUDA_H
#define STEPTH_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- status codes -------------------------------------------------------------------------------------- */
#define STEPTH_OK 0
#define STEPTH_ERR_INVALID_ARG 1      /* -> io::Error(InvalidInput)                                           */
#define STEPTH_ERR_SIZE_MISMATCH 2    /* -> io::Error(InvalidInput, "Sizes don't match")  depth_image.rs:41-48 */
#define STEPTH_ERR_DIV_BY_ZERO_MAX 3  /* -> panic "attempt to divide by zero"             depth_image.rs:127   */
#define STEPTH_ERR_TOO_FEW_BLOCKS 4   /* -> panic (par_chunks_mut(0))                     depth_image.rs:111-112 */
#define STEPTH_ERR_DEGENERATE_DEPTH 5 /* -> panic (step_by(0) / underflow / unwrap(None)) depth_image.rs:212-216 */
#define STEPTH_ERR_NO_CONVERGENCE 6   /* k-means did not converge (reference: unbounded recursion)             */
#define STEPTH_ERR_OUT_OF_BOUNDS 7    /* -> panic (ImageBuffer::get_pixel out of bounds)  mask_image.rs:106        */
#define STEPTH_ERR_CUDA 100           /* CUDA runtime failure; text in stepth_cuda_last_error()                */
#define STEPTH_ERR_CODEC 102          /* nvJPEG missing or failing (decode / encode entry points only)          */

/* ---- flags (stepth_params.flags) ------------------------------------------------------------------------ */
#define STEPTH_RESAMPLE_OFF 0u        /* skip depth_image.rs:133 (debug / raw distance maps)                   */
#define STEPTH_RESAMPLE_U8 1u         /* image 0.23.x sampler: u8 intermediate between the two passes (default) */
#define STEPTH_RESAMPLE_F32 2u        /* 0.24-style f32 intermediate                                           */
#define STEPTH_RESAMPLE_MASK 3u
#define STEPTH_FLAG_INVERT (1u << 2)      /* fuse invert_depth                depth_image.rs:225-227 */
#define STEPTH_FLAG_FOREGROUND (1u << 3)  /* fuse select_foreground -> mask   depth_image.rs:220-245 */
#define STEPTH_FLAG_APPLY (1u << 4)       /* fuse MaskImage::apply_mask       mask_image.rs:205-213  */
#define STEPTH_FLAG_PIXELOPPS_SAT (1u << 5) /* PixelOpps = saturating-sub / all-lt instead of abs-diff / all-lt */
#define STEPTH_FLAG_LITERAL_SEARCH (1u << 6) /* debug: one thread per query walking the reference's probe order */

typedef struct stepth_ctx stepth_ctx;

/* disage::DiscretePixel<[u8;3]> as it crosses the boundary (depth_image.rs:110-121): 20 bytes */
typedef struct {
    uint32_t x, y, w, h; /* position.x, position.y, size.width, size.height */
    uint8_t value[3];    /* block value ([u8;3]) BEFORE it is overwritten by the distance */
    uint8_t pad;
} stepth_block;

typedef struct {
    uint8_t precision[3]; /* depth_precision: [u8;3]             depth_image.rs:79 */
    uint8_t pad;
    uint32_t max_rings;   /* `max` of distance_dot_array; the reference hard-codes 255   depth_image.rs:119 */
    uint32_t flags;
} stepth_params;

/* ---- context --------------------------------------------------------------------------------------------- */
int32_t stepth_cuda_device_count(int32_t *count_out);
int32_t stepth_cuda_ctx_create(int32_t device, stepth_ctx **ctx_out);
void stepth_cuda_ctx_destroy(stepth_ctx *ctx);
const char *stepth_cuda_last_error(const stepth_ctx *ctx); /* ctx may be NULL: last create() error */
const char *stepth_cuda_version(void);

/* ---- the hot path, host buffers (what load_depth_from_additional calls) ---------------------------------- */

/* depth_image.rs:110-134 from a block list (the disage discretiser stays on the host).
 *   add_rgb   : additional image, RGB8 row-major, wa*ha*3
 *   blocks    : n_blocks DiscretePixels that tile the w*h main image
 *   depth_out : w*h Luma8 (the value handed to load_depth, depth_image.rs:135)
 *   raw_d8_out: optional n_blocks bytes, `d as u8` per block before normalisation
 * fused tail (flags): rgba_in (w*h*4, required for APPLY), rgba_out (w*h*4), mask_out (w*h), from_to_out[2]. */
int32_t stepth_cuda_depth_from_blocks(stepth_ctx *ctx, const uint8_t *add_rgb, uint32_t wa, uint32_t ha,
                                      const stepth_block *blocks, size_t n_blocks, uint32_t w, uint32_t h,
                                      const stepth_params *params, uint8_t *depth_out, uint8_t *raw_d8_out,
                                      const uint8_t *rgba_in, uint8_t *rgba_out, uint8_t *mask_out,
                                      uint8_t *from_to_out);

/* limit case: every main pixel is its own 1x1 block whose value is the pixel (alpha ignored). */
int32_t stepth_cuda_depth_dense(stepth_ctx *ctx, const uint8_t *main_rgba, uint32_t w, uint32_t h,
                                const uint8_t *add_rgb, uint32_t wa, uint32_t ha, const stepth_params *params,
                                uint8_t *depth_out, uint8_t *raw_d8_out, uint8_t *rgba_out, uint8_t *mask_out,
                                uint8_t *from_to_out);

/* a batch of independent pairs of identical shape, one launch sequence for all of them.
 * status_out[i] receives the per-pair status (a pair with max == 0 does not fail the others). */
int32_t stepth_cuda_depth_dense_batch(stepth_ctx *ctx, uint32_t n_pairs, const uint8_t *const *main_rgba,
                                      uint32_t w, uint32_t h, const uint8_t *const *add_rgb, uint32_t wa,
                                      uint32_t ha, const stepth_params *params, uint8_t *const *depth_out,
                                      uint8_t *const *rgba_out, uint8_t *const *mask_out, uint8_t *from_to_out,
                                      int32_t *status_out);

/* the same for main images held as RGB8 -- what `self.image.clone()...to_rgb8()` holds at depth_image.rs:105; the search never
 * reads alpha, so depth-only calls (no STEPTH_FLAG_APPLY: that one returns RGBA pixels) upload 3 instead of 4 bytes per pixel. */
int32_t stepth_cuda_depth_dense_batch_rgb(stepth_ctx *ctx, uint32_t n_pairs, const uint8_t *const *main_rgb,
                                          uint32_t w, uint32_t h, const uint8_t *const *add_rgb, uint32_t wa,
                                          uint32_t ha, const stepth_params *params, uint8_t *const *depth_out,
                                          uint8_t *const *mask_out, uint8_t *from_to_out, int32_t *status_out);

/* ---- tail operations, host buffers ----------------------------------------------------------------------- */
int32_t stepth_cuda_invert_depth(stepth_ctx *ctx, uint8_t *depth_inout, size_t n);             /* :225-227 */
/* ranges_out: 2*256 bytes, (min,max) per cluster; has_values_out = 0 for the zones<2 answer [(None,None)] */
int32_t stepth_cuda_depth_split(stepth_ctx *ctx, const uint8_t *depth, size_t n, uint8_t zones,
                                uint8_t *ranges_out, uint32_t *n_out, int32_t *has_values_out); /* :162-218 */
/* from / to < 0 mean None */
int32_t stepth_cuda_slice(stepth_ctx *ctx, const uint8_t *depth, size_t n, int32_t from, int32_t to,
                          uint8_t *mask_out);                                                   /* :229-245 */
int32_t stepth_cuda_select_foreground(stepth_ctx *ctx, const uint8_t *depth, size_t n, uint8_t *mask_out,
                                      uint8_t *from_to_out);                                    /* :220-223 */
int32_t stepth_cuda_apply_mask(stepth_ctx *ctx, uint8_t *rgba_inout, const uint8_t *mask, size_t n); /* mask_image.rs:205-213 */

/* ---- editor operations on host buffers (n = pixel counts) -------------------------------- */
int32_t stepth_cuda_highlight_depth(stepth_ctx *ctx, const uint8_t *rgba_in, const uint8_t *depth, size_t n,
                                    uint8_t *rgba_out);                                  /* depth_image.rs:51-63 */
int32_t stepth_cuda_highlight_mask(stepth_ctx *ctx, const uint8_t *rgba_in, const uint8_t *mask, size_t n,
                                   uint8_t *rgba_out);                                   /* mask_image.rs:57-73  */
/* equal-sized masks only (the reference resizes `other` otherwise: out of scope -> callers resize first) */
int32_t stepth_cuda_mask_and(stepth_ctx *ctx, uint8_t *mask_inout, const uint8_t *other, size_t n); /* :147-169 */
int32_t stepth_cuda_mask_or(stepth_ctx *ctx, uint8_t *mask_inout, const uint8_t *other, size_t n);  /* :171-193 */
int32_t stepth_cuda_mask_not(stepth_ctx *ctx, uint8_t *mask_inout, size_t n);                        /* :195-197 */
/* mask_image.rs:99-109; STEPTH_ERR_OUT_OF_BOUNDS when a masked pixel of the window lies outside `other` */
int32_t stepth_cuda_image_replace(stepth_ctx *ctx, uint8_t *rgba_inout, const uint8_t *mask, uint32_t w, uint32_t h,
                                  const uint8_t *other_rgba, uint32_t ow, uint32_t oh, uint32_t start_x,
                                  uint32_t start_y);

/* operations.rs:3-36 / 38-80 (n = pixel counts) */
int32_t stepth_cuda_normalize_brightness_luma16(stepth_ctx *ctx, const uint16_t *img1, size_t n1,
                                                const uint16_t *img2, size_t n2, double percent, uint16_t *out);
int32_t stepth_cuda_normalize_brightness_rgb16(stepth_ctx *ctx, const uint16_t *img1, size_t n1,
                                               const uint16_t *img2, size_t n2, double percent, uint16_t *out);

/* ---- host-side stand-in for disage::open::rgb_discrete (depth_image.rs:104-109; disage source absent) ----- */
/* kd-splits the main image into blocks ("disc_v0", documented in DESIGN.md as NOT disage).  Returns the number
 * of blocks, writing at most `cap` of them; call with cap = 0 to size the buffer. */
size_t stepth_host_discretise_v0(const uint8_t *main_px, uint32_t stride, uint32_t w, uint32_t h,
                                 const uint8_t precision[3], uint32_t min_splits, uint32_t max_splits,
                                 stepth_block *blocks_out, size_t cap);

/* the same stand-in on the device (csrc/disc.cuh): the block list is bit-identical to stepth_host_discretise_v0's, order included.
 * Host buffers; *n_out = the number of blocks, at most `cap` of them written (blocks_out may be NULL to size the buffer). */
int32_t stepth_cuda_discretise_v0(stepth_ctx *ctx, const uint8_t *main_px, uint32_t stride, uint32_t w, uint32_t h,
                                  const uint8_t precision[3], uint32_t min_splits, uint32_t max_splits,
                                  stepth_block *blocks_out, size_t cap, size_t *n_out);

/* load_depth_from_additional whole (depth_image.rs:91-136) in ONE call: to_rgb8 + the stand-in for rgb_discrete (:104-109, on the
 * device) + the block search + normalise + paint + resample (+ the fused tail by flags).  main_rgba: w*h*4; precision doubles as
 * the discretiser's (:103); min_splits / max_splits as at :101-102 (16, ceil(log2(w*h))).  n_blocks_out: optional. */
int32_t stepth_cuda_depth_from_additional(stepth_ctx *ctx, const uint8_t *main_rgba, uint32_t w, uint32_t h,
                                          const uint8_t *add_rgb, uint32_t wa, uint32_t ha, const stepth_params *params,
                                          uint32_t min_splits, uint32_t max_splits, uint8_t *depth_out, uint8_t *rgba_out,
                                          uint8_t *mask_out, uint8_t *from_to_out, size_t *n_blocks_out);

/* ---- decode / encode on the device (csrc/codec.cuh) -------------------------------------------------- */
/* image::open (src/depth_image.rs:14,66,81) / save (src/mask_image.rs:197-199) for JPEG by nvJPEG, resolved at run time.  NOT
 * on the parity path: decoded pixels differ from jpeg-decoder 0.1.22's by a level or two, as between any two JPEG decoders.
 * decode: *w_out, *h_out always; rgb_out (w*h*3, interleaved RGB8) when not NULL.  A stream nvJPEG cannot parse answers
 * STEPTH_ERR_INVALID_ARG with the reference's "Failed to open image".  encode: channels 1 (Luma8), 3 (RGB8) or 4 (RGBA8, alpha
 * dropped); jpeg_out NULL sizes the buffer.  Status 102 = nvJPEG absent or failing. */
int32_t stepth_cuda_decode_jpeg(stepth_ctx *ctx, const uint8_t *jpeg, size_t n_bytes, uint32_t *w_out, uint32_t *h_out,
                                uint8_t *rgb_out, size_t cap);
int32_t stepth_cuda_encode_jpeg(stepth_ctx *ctx, const uint8_t *pixels, uint32_t w, uint32_t h, uint32_t channels,
                                int32_t quality, uint8_t *jpeg_out, size_t cap, size_t *n_out);

/* ---- device-resident stage API (device pointers; work is enqueued on `stream`, a cudaStream_t) ------------ */
/* Used by the batch / row-band orchestration (one process per GPU) and by the device-resident benchmark.
 * ONE stream in flight per context: the stages keep scratch in the context (match planes, work lists, the plane between the
 * two finish passes), so calls on one context must be ordered -- the same stream, or streams the caller orders with events.
 * Use one context per concurrent stream. */

typedef struct {
    uint32_t n_pairs;  /* images in the batch (contiguous, identical shape)                          */
    uint32_t w, h;     /* main band: width and rows held on this device                              */
    uint32_t img_h;    /* rows of the whole main image (== h unless row-banded)                      */
    uint32_t y0;       /* global row index of band row 0                                             */
    uint32_t wa, ha;   /* additional sub-image held on this device                                   */
    uint32_t add_y0;   /* global row index of its row 0                                              */
    uint32_t add_img_h;/* rows of the whole additional image (for the origin-out-of-bounds rule)     */
} stepth_dev_shape;

/* K0+K1: ring search for every main pixel of the band.  main_rgba [n,h,w,4], add_rgb [n,ha,wa,3],
 * d8_out [n,h,w]; max_out [n] u32 receives max(d8) per pair (zeroed by the call). */
int32_t stepth_cuda_dev_search_dense(stepth_ctx *ctx, void *stream, const stepth_dev_shape *shape,
                                     const stepth_params *params, const uint8_t *main_rgba,
                                     const uint8_t *add_rgb, uint8_t *d8_out, uint32_t *max_out);

/* block-list search: blocks [n_blocks] (device), add_rgb [ha,wa,3]; d8_out [n_blocks]; max_out [1] */
int32_t stepth_cuda_dev_search_blocks(stepth_ctx *ctx, void *stream, const stepth_params *params,
                                      const uint8_t *add_rgb, uint32_t wa, uint32_t ha,
                                      const stepth_block *blocks, size_t n_blocks, uint8_t *d8_out,
                                      uint32_t *max_out);

/* the stand-in discretiser on the device (csrc/disc.cuh): main_px [h,w,stride] (device), blocks_out [cap] (device, may be NULL),
 * n_out [1] u32 (device, may be NULL) = the number of blocks; the list equals stepth_host_discretise_v0's, order included */
int32_t stepth_cuda_dev_discretise_v0(stepth_ctx *ctx, void *stream, const uint8_t *main_px, uint32_t stride, uint32_t w,
                                      uint32_t h, const uint8_t precision[3], uint32_t min_splits, uint32_t max_splits,
                                      stepth_block *blocks_out, size_t cap, uint32_t *n_out);

/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  stepth_cuda_dev_discretise_v0 with the same
 * arguments and the same result (block list, order, *n_out), except that this call allocates nothing -- no cudaMalloc, cudaFree
 * or allocating library routine, on the first call or any later one: the kernels' scratch (the per-region counts) is a
 * file-scope device array of STEPTH_NOALLOC_DISC_SCRATCH_BYTES that every GPU which loads this library holds until exit.
 * Limits, rejected with STEPTH_ERR_INVALID_ARG and a text in stepth_cuda_last_error(ctx) where the allocating call would clamp
 * or carry on:  min(min_splits, max_splits) <= STEPTH_NOALLOC_DISC_MAX_TOP (2^16 forced regions fill the scratch; the
 * reference's own min_splits is 16, depth_image.rs:101) and w * h <= STEPTH_NOALLOC_DISC_MAX_PIXELS (at most 2^16 threads
 * walk the image: 1024 pixels apiece at the limit).
 * EVERY call on a device shares the one array: calls that overlap on that GPU (two streams, two host threads, two contexts)
 * race on it -- serialise them.  The array keeps the last call's contents; nothing in it has to be cleared (a call writes every
 * word it reads).  The first call on a device looks the array's address up (cudaGetSymbolAddress, no allocation): make one
 * call outside a capture before capturing. */
#define STEPTH_NOALLOC_DISC_MAX_TOP 16
#define STEPTH_NOALLOC_DISC_SCRATCH_BYTES 262400 /* (2^16 + 1) u32 counts, rounded up to 256 */
#define STEPTH_NOALLOC_DISC_MAX_PIXELS 67108864ull /* 2^26 */
int32_t stepth_cuda_dev_discretise_v0_noalloc(stepth_ctx *ctx, void *stream, const uint8_t *main_px, uint32_t stride, uint32_t w,
                                              uint32_t h, const uint8_t precision[3], uint32_t min_splits, uint32_t max_splits,
                                              stepth_block *blocks_out, size_t cap, uint32_t *n_out);

/* paint: plane_out[h,w] = d8[b] over each block rectangle (DiscreteImage::collect + luma, depth_image.rs:130-132).
 * The plane holds RAW distances: the normalisation of :125-129 is a per-byte map and is applied by dev_finish. */
int32_t stepth_cuda_dev_paint_blocks(stepth_ctx *ctx, void *stream, const stepth_block *blocks,
                                     const uint8_t *d8, size_t n_blocks, uint32_t w, uint32_t h,
                                     uint8_t *plane_out);

/* K2: normalise by max_in[n] (depth_image.rs:125-129), unit-scale Gaussian (:133), optional invert, histogram.
 * src [n, src_rows, w] holds global rows [src_y0, src_y0+src_rows); depth_out [n, out_rows, w] receives global
 * rows [out_y0, out_y0+out_rows); hist_out [n,256] u32 is ACCUMULATED into (zero it first); status_out [n]. */
int32_t stepth_cuda_dev_finish(stepth_ctx *ctx, void *stream, uint32_t n_pairs, uint32_t w, uint32_t img_h,
                               const uint8_t *src, uint32_t src_y0, uint32_t src_rows, const uint32_t *max_in,
                               uint32_t flags, uint8_t *depth_out, uint32_t out_y0, uint32_t out_rows,
                               uint32_t *hist_out, int32_t *status_out);

/* K3: depth_split(zones) from 256-bin histograms [n,256]; ranges_out [n, 512] bytes, n_out [n], status_out [n] */
int32_t stepth_cuda_dev_depth_split(stepth_ctx *ctx, void *stream, uint32_t n_pairs, const uint32_t *hist,
                                    uint8_t zones, uint8_t *ranges_out, uint32_t *n_out, int32_t *status_out);

/* K4: slice + apply_mask.  from_to [n, stride_ft] bytes (device): cluster 0's (min,max).
 * mask_out [n,px] and/or rgba_out [n,px,4] (rgba_in may equal rgba_out) */
int32_t stepth_cuda_dev_mask_apply(stepth_ctx *ctx, void *stream, uint32_t n_pairs, size_t px,
                                   const uint8_t *depth, const uint8_t *from_to, uint32_t stride_ft,
                                   const int32_t *status_in, const uint8_t *rgba_in, uint8_t *mask_out,
                                   uint8_t *rgba_out);

/* 256-bin histogram of a device buffer, accumulated into hist_out[256] */
int32_t stepth_cuda_dev_histogram(stepth_ctx *ctx, void *stream, const uint8_t *data, size_t n,
                                  uint32_t *hist_out);

/* number of kernel launches issued by this context so far (bench.py's gpu_launches) */
uint64_t stepth_cuda_launch_count(const stepth_ctx *ctx);

/* Measurement aid (no reference counterpart): the rate at which this device retires 32-bit logic/add instructions
 * (thread-level ops per second), measured with a register-only kernel: logic_only = 0 mixes LOP3 with adds (both
 * integer pipes, the highest rate any integer mix reaches), 1 = LOP3 + SHF (the alu pipe alone, the search's class).  bench.py divides the search's algorithmic
 * integer work (2 ops per probe of the reference order, src/helpers.rs:26-47) by it. */
int32_t stepth_cuda_int_peak(stepth_ctx *ctx, int32_t logic_only, double *ops_per_second_out);

/* ---- multi-GPU: pair batches dealt to the devices, or ONE giant pair in row bands ------------------------ */
#define STEPTH_ERR_NCCL 101 /* NCCL missing or a collective failed; text in the owner's last_error()                 */

/* main rows [out[0], out[1]) of rank `rank` of `n_ranks` for an image of h rows, and the additional rows [out[2], out[3]) those
 * rows can reach: origins are half-coordinates (depth_image.rs:114-117), widened by max_rings - 1 (helpers.rs:26). */
void stepth_host_band_plan(uint32_t h, uint32_t add_h, uint32_t max_rings, uint32_t n_ranks, uint32_t rank, uint32_t out[4]);

/* One rank of a row-band group: a context + a bootstrap (NCCL) communicator (one process per GPU; rank 0 makes the 128-byte id and
 * hands it to the others by whatever channel the host has).  n_ranks == 1 needs no id and no NCCL.  Destroy a group only after
 * every rank's last band call has completed (stepth_cuda_comm_band_result returned): ranks map each other's exchange windows. */
typedef struct stepth_comm stepth_comm;
int32_t stepth_cuda_comm_unique_id(uint8_t id_out[128]);
int32_t stepth_cuda_comm_create(stepth_ctx *ctx, int32_t n_ranks, int32_t rank, const uint8_t id[128], stepth_comm **comm_out);
void stepth_cuda_comm_destroy(stepth_comm *comm);

/* This rank's share of depth_image.rs:110-134 (+ fused tail) for ONE pair whose main rows are banded over the group (device
 * pointers, enqueued on `stream`; every rank of the group must make the same call).  w, h: the whole main image; wa: width of
 * the additional rows held here; add_h: rows of the whole additional image.  main_band [rows, w, 4] and add_band [arows, wa, 3]
 * are the rows stepth_host_band_plan gives this rank.  Exchanges: the maximum over all bands for the normaliser (:124), 3-row
 * halos of the raw distance plane for the Gaussian (:133), and -- when depth_full_out [h, w] is given, by EVERY rank or by none --
 * the gather of the depth bands: all three as stores into the other ranks' HBM over NVLink by this library's kernels (csrc/peer.cuh;
 * the gather rides on the finish pass), or as NCCL all-reduce / send-recv / all-gather when the GPUs cannot map each other's
 * memory; all-reduce SUM of the histogram for depth_split (:162-218, NCCL).  depth_band_out [rows, w]; mask_band_out /
 * rgba_band_out optional (fused tail).  Status 101 from stepth_cuda_comm_band_result: a peer did not arrive within 20 s. */
int32_t stepth_cuda_dev_depth_dense_band(stepth_comm *comm, void *stream, uint32_t w, uint32_t h, uint32_t wa, uint32_t add_h,
                                         const stepth_params *params, const uint8_t *main_band, const uint8_t *add_band,
                                         uint8_t *depth_band_out, uint8_t *depth_full_out, uint8_t *mask_band_out,
                                         uint8_t *rgba_band_out);
/* waits for `stream` and reports the last band call's status (0, 3 = max == 0, 5 / 6 = depth_split) and cluster 0's (from, to) */
int32_t stepth_cuda_comm_band_result(stepth_comm *comm, void *stream, int32_t *status_out, uint8_t from_to_out[2]);

/* 1 when this group's band exchanges are stores into the peers' HBM over NVLink (csrc/peer.cuh), 0 when they go through NCCL
 * (no peer access between the GPUs, or STEPTH_PEER=0), -1 before the first multi-rank band call has decided. */
int32_t stepth_cuda_comm_peer_mode(const stepth_comm *comm);
/* The assembled depth map inside this rank's window (device pointer, w*h bytes): handing it to stepth_cuda_dev_depth_dense_band
 * as depth_full_out saves the copy out of the window.  COLLECTIVE.  *ptr_out = NULL when the group runs over NCCL. */
int32_t stepth_cuda_comm_full_map(stepth_comm *comm, void *stream, uint32_t w, uint32_t h, uint32_t add_h, uint32_t max_rings,
                                  uint8_t **ptr_out);
/* ms_out[4]: device time the last band call spent in search | MAX all-reduce + halo exchange | finish | tail + gather */
int32_t stepth_cuda_comm_band_stage_ms(stepth_comm *comm, float ms_out[4]);

/* One process driving several GPUs (what a Rust host links): a context per device, a host thread per device per call. */
typedef struct stepth_multi stepth_multi;
int32_t stepth_cuda_multi_create(const int32_t *devices, uint32_t n_devices, stepth_multi **multi_out);
void stepth_cuda_multi_destroy(stepth_multi *multi);
const char *stepth_cuda_multi_last_error(const stepth_multi *multi); /* multi may be NULL: last create() error */
uint64_t stepth_cuda_multi_launch_count(const stepth_multi *multi);
/* stepth_cuda_depth_dense_batch with pair i on device i mod n_devices; no collective on the data path */
int32_t stepth_cuda_multi_depth_dense_batch(stepth_multi *multi, uint32_t n_pairs, const uint8_t *const *main_rgba, uint32_t w,
                                            uint32_t h, const uint8_t *const *add_rgb, uint32_t wa, uint32_t ha,
                                            const stepth_params *params, uint8_t *const *depth_out, uint8_t *const *rgba_out,
                                            uint8_t *const *mask_out, uint8_t *from_to_out, int32_t *status_out);
/* stepth_cuda_depth_dense for ONE pair too large (or too slow) for one GPU: row bands over the devices, host buffers */
int32_t stepth_cuda_multi_depth_dense_banded(stepth_multi *multi, const uint8_t *main_rgba, uint32_t w, uint32_t h,
                                             const uint8_t *add_rgb, uint32_t wa, uint32_t ha, const stepth_params *params,
                                             uint8_t *depth_out, uint8_t *rgba_out, uint8_t *mask_out, uint8_t *from_to_out);

/* Measurement aid (no reference counterpart): runs the in-order kernel (one thread per query, the probe order of
 * src/helpers.rs:26-47 with the duplicate probes dropped) and counts, per pair, the in-bounds pixels that order visits up to
 * and including the first match -- the "probes the reference order needs" of SURVEY.md 8(d).  d8_out / max_out as
 * stepth_cuda_dev_search_dense; probes_out [n_pairs] u64 (device). */
int32_t stepth_cuda_dev_count_probes(stepth_ctx *ctx, void *stream, const stepth_dev_shape *shape,
                                     const stepth_params *params, const uint8_t *main_rgba, const uint8_t *add_rgb,
                                     uint8_t *d8_out, uint32_t *max_out, uint64_t *probes_out);

#ifdef __cplusplus
}
#endif
#endif /* STEPTH_CUDA_H */
