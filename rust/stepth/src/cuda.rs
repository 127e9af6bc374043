//! `stepth::cuda` -- the module the patched crate adds.  It owns one `stepth_ctx` per thread and turns the C-ABI status
//! codes back into the reference's error behaviour: `io::Error(InvalidInput, ..)` for 1-2, the original panics for 3-5.
//! Source only: the build image has no Rust toolchain.  See INTEGRATION.md for the call sites.
use std::cell::RefCell;
use std::ffi::CStr;
use std::io::{Error, ErrorKind};

use stepth_cuda_sys as sys;

pub struct Ctx(*mut sys::stepth_ctx);

impl Ctx {
    fn new() -> Ctx {
        let device = std::env::var("STEPTH_DEVICE").ok().and_then(|v| v.parse().ok()).unwrap_or(0);
        let mut raw = std::ptr::null_mut();
        let rc = unsafe { sys::stepth_cuda_ctx_create(device, &mut raw) };
        if rc != sys::STEPTH_OK {
            let why = unsafe { CStr::from_ptr(sys::stepth_cuda_last_error(std::ptr::null())) }.to_string_lossy().into_owned();
            panic!("stepth: no usable B200 ({why}); this build has no CPU fallback");
        }
        Ctx(raw)
    }
    pub fn raw(&self) -> *mut sys::stepth_ctx {
        self.0
    }
    /// status -> the reference's behaviour
    pub fn check(&self, rc: i32) -> Result<(), Error> {
        match rc {
            sys::STEPTH_OK => Ok(()),
            sys::STEPTH_ERR_SIZE_MISMATCH => Err(Error::new(ErrorKind::InvalidInput, "Sizes don't match")),
            sys::STEPTH_ERR_INVALID_ARG => Err(Error::new(ErrorKind::InvalidInput, self.last_error())),
            sys::STEPTH_ERR_DIV_BY_ZERO_MAX => panic!("attempt to divide by zero"), // src/depth_image.rs:127
            sys::STEPTH_ERR_TOO_FEW_BLOCKS => panic!("chunk_size must not be zero"), // src/depth_image.rs:111-112
            sys::STEPTH_ERR_DEGENERATE_DEPTH => panic!("depth_split: degenerate depth range"), // src/depth_image.rs:212-216
            sys::STEPTH_ERR_NO_CONVERGENCE => panic!("depth_split: no convergence"),
            sys::STEPTH_ERR_OUT_OF_BOUNDS => panic!("Image index out of bounds"), // src/mask_image.rs:106
            _ => Err(Error::new(ErrorKind::Other, self.last_error())),
        }
    }
    fn last_error(&self) -> String {
        unsafe { CStr::from_ptr(sys::stepth_cuda_last_error(self.0)) }.to_string_lossy().into_owned()
    }
}

impl Drop for Ctx {
    fn drop(&mut self) {
        unsafe { sys::stepth_cuda_ctx_destroy(self.0) }
    }
}

thread_local! {
    static CTX: RefCell<Option<Ctx>> = RefCell::new(None);
}

/// Runs `f` with this thread's context (created on first use).  `&mut self` methods of `DepthImage` / `MaskImage`
/// already serialise calls per image; different threads get different contexts.
pub fn with_ctx<R>(f: impl FnOnce(&Ctx) -> R) -> R {
    CTX.with(|c| {
        let mut slot = c.borrow_mut();
        let ctx = slot.get_or_insert_with(Ctx::new);
        f(ctx)
    })
}

/// `DiscretePixel<&mut [u8;3]>` list -> the 20-byte records of the boundary (src/depth_image.rs:110-117)
pub fn blocks_of<'a>(
    pixels: impl Iterator<Item = (u32, u32, u32, u32, [u8; 3])> + 'a,
) -> Vec<sys::stepth_block> {
    pixels.map(|(x, y, w, h, value)| sys::stepth_block { x, y, w, h, value, pad: 0 }).collect()
}
