"""stepth_b200 -- B200 (sm_100a) implementation of nikilark/stepth's depth-from-additional hot path.

The product is libstepth_cuda.so (C ABI in include/stepth_cuda.h, CUDA in stepth_b200/csrc/); this package is the
Python host side that mirrors the reference's public API over that ABI.  There is no CPU fallback.
"""
from . import _ffi
from ._ffi import (FLAG_APPLY, FLAG_FOREGROUND, FLAG_INVERT, FLAG_LITERAL_SEARCH, FLAG_PIXELOPPS_SAT, RESAMPLE_F32,
                   RESAMPLE_OFF, RESAMPLE_U8, BLOCK_DTYPE, StepthError)
from .context import Context, Multi, StepthIOError, StepthPanic, default_context, discretise_v0
from .depth_image import MASK_FALSE, MASK_TRUE, DepthImage, MaskImage

__version__ = "0.1.0"
__all__ = [
    "Context", "Multi", "DepthImage", "MaskImage", "MASK_TRUE", "MASK_FALSE", "StepthError", "StepthIOError", "StepthPanic",
    "default_context", "discretise_v0", "BLOCK_DTYPE", "RESAMPLE_OFF", "RESAMPLE_U8", "RESAMPLE_F32", "FLAG_INVERT",
    "FLAG_FOREGROUND", "FLAG_APPLY", "FLAG_PIXELOPPS_SAT", "FLAG_LITERAL_SEARCH",
]
