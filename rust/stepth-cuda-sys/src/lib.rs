//! Raw bindings to `libstepth_cuda.so` -- one declaration per symbol of `include/stepth_cuda.h`, in the header's order:
//! the host-buffer entry points the patched `stepth` crate calls, the batch and multi-GPU entry points a batch tool built
//! on the crate would use, and the device-pointer stage API (`stepth_cuda_dev_*`, `stepth_comm`) for in-process GPU
//! pipelines and one-process-per-GPU launchers.  `tests/test_abi.py::test_rust_sys_crate_matches_the_header` keeps names,
//! arities and parameter types in lock-step with the header (there is no Rust toolchain in the build image).
#![allow(non_camel_case_types)]
use std::os::raw::{c_char, c_double, c_float, c_void};

pub const STEPTH_OK: i32 = 0;
pub const STEPTH_ERR_INVALID_ARG: i32 = 1;
pub const STEPTH_ERR_SIZE_MISMATCH: i32 = 2;
pub const STEPTH_ERR_DIV_BY_ZERO_MAX: i32 = 3;
pub const STEPTH_ERR_TOO_FEW_BLOCKS: i32 = 4;
pub const STEPTH_ERR_DEGENERATE_DEPTH: i32 = 5;
pub const STEPTH_ERR_NO_CONVERGENCE: i32 = 6;
pub const STEPTH_ERR_OUT_OF_BOUNDS: i32 = 7;
pub const STEPTH_ERR_CUDA: i32 = 100;
pub const STEPTH_ERR_NCCL: i32 = 101;
pub const STEPTH_ERR_CODEC: i32 = 102;

pub const STEPTH_RESAMPLE_OFF: u32 = 0;
pub const STEPTH_RESAMPLE_U8: u32 = 1;
pub const STEPTH_RESAMPLE_F32: u32 = 2;
pub const STEPTH_RESAMPLE_MASK: u32 = 3;
pub const STEPTH_FLAG_INVERT: u32 = 1 << 2;
pub const STEPTH_FLAG_FOREGROUND: u32 = 1 << 3;
pub const STEPTH_FLAG_APPLY: u32 = 1 << 4;
pub const STEPTH_FLAG_PIXELOPPS_SAT: u32 = 1 << 5;
pub const STEPTH_FLAG_LITERAL_SEARCH: u32 = 1 << 6;

/// limits of `stepth_cuda_dev_discretise_v0_noalloc`
pub const STEPTH_NOALLOC_DISC_MAX_TOP: u32 = 16;
pub const STEPTH_NOALLOC_DISC_SCRATCH_BYTES: usize = 262400;
pub const STEPTH_NOALLOC_DISC_MAX_PIXELS: u64 = 67108864;

#[repr(C)]
pub struct stepth_ctx {
    _private: [u8; 0],
}

/// one process driving several GPUs: a context and a host thread per device
#[repr(C)]
pub struct stepth_multi {
    _private: [u8; 0],
}

/// one rank of a row-band group (one process per GPU): a context + the group's bootstrap communicator
#[repr(C)]
pub struct stepth_comm {
    _private: [u8; 0],
}

/// `disage::DiscretePixel<[u8;3]>` as it crosses the boundary (20 bytes)
#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct stepth_block {
    pub x: u32,
    pub y: u32,
    pub w: u32,
    pub h: u32,
    pub value: [u8; 3],
    pub pad: u8,
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct stepth_params {
    pub precision: [u8; 3],
    pub pad: u8,
    pub max_rings: u32,
    pub flags: u32,
}

/// shape of a batch / row band held on one device (`stepth_cuda_dev_search_dense`)
#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct stepth_dev_shape {
    pub n_pairs: u32,
    pub w: u32,
    pub h: u32,
    pub img_h: u32,
    pub y0: u32,
    pub wa: u32,
    pub ha: u32,
    pub add_y0: u32,
    pub add_img_h: u32,
}

extern "C" {
    pub fn stepth_cuda_device_count(count_out: *mut i32) -> i32;
    pub fn stepth_cuda_ctx_create(device: i32, ctx_out: *mut *mut stepth_ctx) -> i32;
    pub fn stepth_cuda_ctx_destroy(ctx: *mut stepth_ctx);
    pub fn stepth_cuda_last_error(ctx: *const stepth_ctx) -> *const c_char;
    pub fn stepth_cuda_version() -> *const c_char;

    pub fn stepth_cuda_depth_from_blocks(
        ctx: *mut stepth_ctx, add_rgb: *const u8, wa: u32, ha: u32, blocks: *const stepth_block, n_blocks: usize,
        w: u32, h: u32, params: *const stepth_params, depth_out: *mut u8, raw_d8_out: *mut u8, rgba_in: *const u8,
        rgba_out: *mut u8, mask_out: *mut u8, from_to_out: *mut u8,
    ) -> i32;
    pub fn stepth_cuda_depth_dense(
        ctx: *mut stepth_ctx, main_rgba: *const u8, w: u32, h: u32, add_rgb: *const u8, wa: u32, ha: u32,
        params: *const stepth_params, depth_out: *mut u8, raw_d8_out: *mut u8, rgba_out: *mut u8, mask_out: *mut u8,
        from_to_out: *mut u8,
    ) -> i32;
    pub fn stepth_cuda_depth_dense_batch(
        ctx: *mut stepth_ctx, n_pairs: u32, main_rgba: *const *const u8, w: u32, h: u32, add_rgb: *const *const u8,
        wa: u32, ha: u32, params: *const stepth_params, depth_out: *const *mut u8, rgba_out: *const *mut u8,
        mask_out: *const *mut u8, from_to_out: *mut u8, status_out: *mut i32,
    ) -> i32;

    /// `load_depth_from_additional` whole (src/depth_image.rs:91-136) with the stand-in discretiser on the device
    pub fn stepth_cuda_depth_from_additional(
        ctx: *mut stepth_ctx, main_rgba: *const u8, w: u32, h: u32, add_rgb: *const u8, wa: u32, ha: u32,
        params: *const stepth_params, min_splits: u32, max_splits: u32, depth_out: *mut u8, rgba_out: *mut u8,
        mask_out: *mut u8, from_to_out: *mut u8, n_blocks_out: *mut usize,
    ) -> i32;
    pub fn stepth_cuda_discretise_v0(
        ctx: *mut stepth_ctx, main_px: *const u8, stride: u32, w: u32, h: u32, precision: *const u8, min_splits: u32,
        max_splits: u32, blocks_out: *mut stepth_block, cap: usize, n_out: *mut usize,
    ) -> i32;
    pub fn stepth_host_discretise_v0(
        main_px: *const u8, stride: u32, w: u32, h: u32, precision: *const u8, min_splits: u32, max_splits: u32,
        blocks_out: *mut stepth_block, cap: usize,
    ) -> usize;

    pub fn stepth_cuda_depth_dense_batch_rgb(
        ctx: *mut stepth_ctx, n_pairs: u32, main_rgb: *const *const u8, w: u32, h: u32, add_rgb: *const *const u8,
        wa: u32, ha: u32, params: *const stepth_params, depth_out: *const *mut u8, mask_out: *const *mut u8,
        from_to_out: *mut u8, status_out: *mut i32,
    ) -> i32;

    // several GPUs from one process: pair batches dealt to the devices; one giant pair in row bands
    pub fn stepth_cuda_multi_create(devices: *const i32, n_devices: u32, multi_out: *mut *mut stepth_multi) -> i32;
    pub fn stepth_cuda_multi_destroy(multi: *mut stepth_multi);
    pub fn stepth_cuda_multi_last_error(multi: *const stepth_multi) -> *const c_char;
    pub fn stepth_cuda_multi_depth_dense_batch(
        multi: *mut stepth_multi, n_pairs: u32, main_rgba: *const *const u8, w: u32, h: u32,
        add_rgb: *const *const u8, wa: u32, ha: u32, params: *const stepth_params, depth_out: *const *mut u8,
        rgba_out: *const *mut u8, mask_out: *const *mut u8, from_to_out: *mut u8, status_out: *mut i32,
    ) -> i32;
    pub fn stepth_cuda_multi_depth_dense_banded(
        multi: *mut stepth_multi, main_rgba: *const u8, w: u32, h: u32, add_rgb: *const u8, wa: u32, ha: u32,
        params: *const stepth_params, depth_out: *mut u8, rgba_out: *mut u8, mask_out: *mut u8, from_to_out: *mut u8,
    ) -> i32;

    pub fn stepth_cuda_invert_depth(ctx: *mut stepth_ctx, depth_inout: *mut u8, n: usize) -> i32;
    pub fn stepth_cuda_depth_split(
        ctx: *mut stepth_ctx, depth: *const u8, n: usize, zones: u8, ranges_out: *mut u8, n_out: *mut u32,
        has_values_out: *mut i32,
    ) -> i32;
    pub fn stepth_cuda_slice(ctx: *mut stepth_ctx, depth: *const u8, n: usize, from: i32, to: i32, mask_out: *mut u8) -> i32;
    pub fn stepth_cuda_select_foreground(
        ctx: *mut stepth_ctx, depth: *const u8, n: usize, mask_out: *mut u8, from_to_out: *mut u8,
    ) -> i32;
    pub fn stepth_cuda_apply_mask(ctx: *mut stepth_ctx, rgba_inout: *mut u8, mask: *const u8, n: usize) -> i32;

    pub fn stepth_cuda_highlight_depth(ctx: *mut stepth_ctx, rgba_in: *const u8, depth: *const u8, n: usize, rgba_out: *mut u8) -> i32;
    pub fn stepth_cuda_highlight_mask(ctx: *mut stepth_ctx, rgba_in: *const u8, mask: *const u8, n: usize, rgba_out: *mut u8) -> i32;
    pub fn stepth_cuda_mask_and(ctx: *mut stepth_ctx, mask_inout: *mut u8, other: *const u8, n: usize) -> i32;
    pub fn stepth_cuda_mask_or(ctx: *mut stepth_ctx, mask_inout: *mut u8, other: *const u8, n: usize) -> i32;
    pub fn stepth_cuda_mask_not(ctx: *mut stepth_ctx, mask_inout: *mut u8, n: usize) -> i32;
    pub fn stepth_cuda_image_replace(
        ctx: *mut stepth_ctx, rgba_inout: *mut u8, mask: *const u8, w: u32, h: u32, other_rgba: *const u8, ow: u32,
        oh: u32, start_x: u32, start_y: u32,
    ) -> i32;

    pub fn stepth_cuda_normalize_brightness_luma16(
        ctx: *mut stepth_ctx, img1: *const u16, n1: usize, img2: *const u16, n2: usize, percent: c_double, out: *mut u16,
    ) -> i32;
    pub fn stepth_cuda_normalize_brightness_rgb16(
        ctx: *mut stepth_ctx, img1: *const u16, n1: usize, img2: *const u16, n2: usize, percent: c_double, out: *mut u16,
    ) -> i32;

    // decode / encode on the device through nvJPEG (off the parity path; include/stepth_cuda.h)
    pub fn stepth_cuda_decode_jpeg(
        ctx: *mut stepth_ctx, jpeg: *const u8, n_bytes: usize, w_out: *mut u32, h_out: *mut u32, rgb_out: *mut u8, cap: usize,
    ) -> i32;
    pub fn stepth_cuda_encode_jpeg(
        ctx: *mut stepth_ctx, pixels: *const u8, w: u32, h: u32, channels: u32, quality: i32, jpeg_out: *mut u8, cap: usize,
        n_out: *mut usize,
    ) -> i32;

    // device-resident stage API: DEVICE pointers, work enqueued on `stream` (a cudaStream_t); one stream in flight per context
    pub fn stepth_cuda_dev_search_dense(
        ctx: *mut stepth_ctx, stream: *mut c_void, shape: *const stepth_dev_shape, params: *const stepth_params,
        main_rgba: *const u8, add_rgb: *const u8, d8_out: *mut u8, max_out: *mut u32,
    ) -> i32;
    pub fn stepth_cuda_dev_search_blocks(
        ctx: *mut stepth_ctx, stream: *mut c_void, params: *const stepth_params, add_rgb: *const u8, wa: u32, ha: u32,
        blocks: *const stepth_block, n_blocks: usize, d8_out: *mut u8, max_out: *mut u32,
    ) -> i32;
    pub fn stepth_cuda_dev_discretise_v0(
        ctx: *mut stepth_ctx, stream: *mut c_void, main_px: *const u8, stride: u32, w: u32, h: u32, precision: *const u8,
        min_splits: u32, max_splits: u32, blocks_out: *mut stepth_block, cap: usize, n_out: *mut u32,
    ) -> i32;
    /// no counterpart in the reference: the same call with its scratch in a fixed device array (allocates nothing; calls that
    /// overlap on one GPU race on it; `min(min_splits, max_splits) <= 16`, `w * h <= 2^26`)
    pub fn stepth_cuda_dev_discretise_v0_noalloc(
        ctx: *mut stepth_ctx, stream: *mut c_void, main_px: *const u8, stride: u32, w: u32, h: u32, precision: *const u8,
        min_splits: u32, max_splits: u32, blocks_out: *mut stepth_block, cap: usize, n_out: *mut u32,
    ) -> i32;
    pub fn stepth_cuda_dev_paint_blocks(
        ctx: *mut stepth_ctx, stream: *mut c_void, blocks: *const stepth_block, d8: *const u8, n_blocks: usize, w: u32, h: u32,
        plane_out: *mut u8,
    ) -> i32;
    pub fn stepth_cuda_dev_finish(
        ctx: *mut stepth_ctx, stream: *mut c_void, n_pairs: u32, w: u32, img_h: u32, src: *const u8, src_y0: u32,
        src_rows: u32, max_in: *const u32, flags: u32, depth_out: *mut u8, out_y0: u32, out_rows: u32, hist_out: *mut u32,
        status_out: *mut i32,
    ) -> i32;
    pub fn stepth_cuda_dev_depth_split(
        ctx: *mut stepth_ctx, stream: *mut c_void, n_pairs: u32, hist: *const u32, zones: u8, ranges_out: *mut u8,
        n_out: *mut u32, status_out: *mut i32,
    ) -> i32;
    pub fn stepth_cuda_dev_mask_apply(
        ctx: *mut stepth_ctx, stream: *mut c_void, n_pairs: u32, px: usize, depth: *const u8, from_to: *const u8,
        stride_ft: u32, status_in: *const i32, rgba_in: *const u8, mask_out: *mut u8, rgba_out: *mut u8,
    ) -> i32;
    pub fn stepth_cuda_dev_histogram(ctx: *mut stepth_ctx, stream: *mut c_void, data: *const u8, n: usize, hist_out: *mut u32) -> i32;
    pub fn stepth_cuda_launch_count(ctx: *const stepth_ctx) -> u64;
    pub fn stepth_cuda_multi_launch_count(multi: *const stepth_multi) -> u64;

    // measurement aids (no reference counterpart)
    pub fn stepth_cuda_int_peak(ctx: *mut stepth_ctx, logic_only: i32, ops_per_second_out: *mut c_double) -> i32;
    pub fn stepth_cuda_dev_count_probes(
        ctx: *mut stepth_ctx, stream: *mut c_void, shape: *const stepth_dev_shape, params: *const stepth_params,
        main_rgba: *const u8, add_rgb: *const u8, d8_out: *mut u8, max_out: *mut u32, probes_out: *mut u64,
    ) -> i32;

    // one rank of a row-band group (one process per GPU)
    pub fn stepth_host_band_plan(h: u32, add_h: u32, max_rings: u32, n_ranks: u32, rank: u32, out: *mut u32);
    pub fn stepth_cuda_comm_unique_id(id_out: *mut u8) -> i32;
    pub fn stepth_cuda_comm_create(
        ctx: *mut stepth_ctx, n_ranks: i32, rank: i32, id: *const u8, comm_out: *mut *mut stepth_comm,
    ) -> i32;
    pub fn stepth_cuda_comm_destroy(comm: *mut stepth_comm);
    pub fn stepth_cuda_dev_depth_dense_band(
        comm: *mut stepth_comm, stream: *mut c_void, w: u32, h: u32, wa: u32, add_h: u32, params: *const stepth_params,
        main_band: *const u8, add_band: *const u8, depth_band_out: *mut u8, depth_full_out: *mut u8, mask_band_out: *mut u8,
        rgba_band_out: *mut u8,
    ) -> i32;
    pub fn stepth_cuda_comm_band_result(
        comm: *mut stepth_comm, stream: *mut c_void, status_out: *mut i32, from_to_out: *mut u8,
    ) -> i32;
    pub fn stepth_cuda_comm_peer_mode(comm: *const stepth_comm) -> i32;
    pub fn stepth_cuda_comm_full_map(
        comm: *mut stepth_comm, stream: *mut c_void, w: u32, h: u32, add_h: u32, max_rings: u32, ptr_out: *mut *mut u8,
    ) -> i32;
    pub fn stepth_cuda_comm_band_stage_ms(comm: *mut stepth_comm, ms_out: *mut c_float) -> i32;
}
