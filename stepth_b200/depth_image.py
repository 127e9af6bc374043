"""`DepthImage` / `MaskImage`: the reference's public API (src/lib.rs:3-12) with the hot-path bodies swapped
for calls into libstepth_cuda.so.  Names, argument meaning and error behaviour follow the reference:

    img = DepthImage.open("assets/main.jpg")                                   # depth_image.rs:13-21
    img.open_depth_from_additional("assets/additional.jpg", [255 // 7] * 3)    # :76-89
    img.depth().save("depth.jpg")                                              # :33-35
    img.invert_depth()                                                         # :225-227
    fg = img.select_foreground()                                               # :220-223
    fg.apply_mask()                                                            # mask_image.rs:205-213

Decoding / encoding stays on the host (PIL here, the `image` crate in the reference) exactly as in the
reference; pixels cross the C ABI as tightly packed row-major bytes (`ImageBuffer<_, Vec<u8>>`).
"""
from __future__ import annotations

import numpy as np

from . import _ffi
from .context import Context, StepthIOError, default_context, discretise_v0

MASK_TRUE = 255   # mask_image.rs:3
MASK_FALSE = 0    # mask_image.rs:4


def _open_image(path: str, ctx: Context | None = None):
    """image::open (depth_image.rs:14).  STEPTH_GPU_DECODE=1: JPEG files are decoded on the GPU (csrc/codec.cuh, nvJPEG) -- faster, but
    a level or two away from the host decoder's pixels, like any two JPEG decoders: off by default, the parity path decodes on the host."""
    import os

    if os.environ.get("STEPTH_GPU_DECODE") == "1" and path.lower().endswith((".jpg", ".jpeg")):
        try:
            with open(path, "rb") as f:
                data = f.read()
        except OSError:
            raise StepthIOError("Failed to open image") from None
        return (ctx or default_context()).decode_jpeg(data)
    from PIL import Image

    try:
        im = Image.open(path)
        im.load()
        return im
    except Exception:
        raise StepthIOError("Failed to open image") from None  # depth_image.rs:14-17


def _to_rgba8(img) -> np.ndarray:
    if isinstance(img, np.ndarray):
        a = np.ascontiguousarray(img, dtype=np.uint8)
        if a.ndim == 2:
            a = np.stack([a, a, a, np.full_like(a, 255)], axis=-1)
        elif a.shape[-1] == 3:
            a = np.concatenate([a, np.full(a.shape[:2] + (1,), 255, np.uint8)], axis=-1)
        return np.ascontiguousarray(a)
    return np.ascontiguousarray(np.asarray(img.convert("RGBA"), dtype=np.uint8))


def _to_rgb8(img) -> np.ndarray:
    if isinstance(img, np.ndarray):
        a = np.ascontiguousarray(img, dtype=np.uint8)
        if a.ndim == 2:
            a = np.stack([a, a, a], axis=-1)
        return np.ascontiguousarray(a[..., :3])
    return np.ascontiguousarray(np.asarray(img.convert("RGB"), dtype=np.uint8))


class MaskImage:
    """mask_image.rs:6-10  { image: RgbaImage, mask: GrayImage }"""

    def __init__(self, image: np.ndarray, mask: np.ndarray, ctx: Context | None = None):
        self.image = image
        self.mask = mask
        self._ctx = ctx

    @classmethod
    def from_image(cls, img, ctx: Context | None = None) -> "MaskImage":  # mask_image.rs:21-25
        image = _to_rgba8(img)
        return cls(image, np.full(image.shape[:2], MASK_TRUE, np.uint8), ctx)

    @classmethod
    def open(cls, image_path: str, ctx: Context | None = None) -> "MaskImage":  # mask_image.rs:13-19
        return cls.from_image(_open_image(image_path, ctx), ctx)

    def width(self) -> int:
        return int(self.image.shape[1])

    def height(self) -> int:
        return int(self.image.shape[0])

    def load_mask(self, mask: np.ndarray):  # mask_image.rs:31-43
        mask = np.ascontiguousarray(mask, dtype=np.uint8)
        if mask.shape != self.image.shape[:2]:
            raise StepthIOError("Sizes don't match")
        self.mask = mask

    def mask_reset(self):  # mask_image.rs:201-203
        self.mask = np.full(self.image.shape[:2], MASK_TRUE, np.uint8)

    def _context(self) -> Context:
        return self._ctx or default_context()

    def highlight_mask(self) -> np.ndarray:  # mask_image.rs:57-73 -> RGBA8
        return self._context().highlight_mask(self.image, self.mask)

    def image_replace(self, other: "MaskImage", start_point) -> None:  # mask_image.rs:99-109
        self.image = self._context().image_replace(self.image, self.mask, other.image, start_point[0], start_point[1])

    def mask_and(self, other: "MaskImage") -> None:  # mask_image.rs:147-169 (equal sizes)
        self.mask = self._context().mask_logic("and", self.mask, other.mask)

    def mask_or(self, other: "MaskImage") -> None:  # mask_image.rs:171-193 (equal sizes)
        self.mask = self._context().mask_logic("or", self.mask, other.mask)

    def mask_not(self) -> None:  # mask_image.rs:195-197
        self.mask = self._context().mask_logic("not", self.mask)

    def apply_mask(self) -> None:  # mask_image.rs:205-213
        ctx = self._ctx or default_context()
        self.image = ctx.apply_mask(self.image, self.mask)

    def to_pil(self):
        from PIL import Image

        return Image.fromarray(self.image, "RGBA")

    def save(self, path: str):  # mask_image.rs:197-199
        im = self.to_pil()
        if path.lower().endswith((".jpg", ".jpeg")):
            im = im.convert("RGB")
        im.save(path)


class DepthImage:
    """depth_image.rs:6-10  { image: RgbaImage, depth: GrayImage }"""

    #: hard-coded in the reference (depth_image.rs:119, :102); exposed for the synthetic configs of BASELINE.json
    MAX_RINGS = 255
    MIN_SPLITS = 16

    def __init__(self, image: np.ndarray, depth: np.ndarray, ctx: Context | None = None):
        self.image = image
        self.depth_plane = depth
        self._ctx = ctx

    # ---- constructors / accessors (host only) -------------------------------------------------------------
    @classmethod
    def from_image(cls, img, ctx: Context | None = None) -> "DepthImage":  # :23-27
        image = _to_rgba8(img)
        return cls(image, np.zeros(image.shape[:2], np.uint8), ctx)

    @classmethod
    def open(cls, image_path: str, ctx: Context | None = None) -> "DepthImage":  # :13-21
        return cls.from_image(_open_image(image_path, ctx), ctx)

    def _context(self) -> Context:
        if self._ctx is None:
            self._ctx = default_context()
        return self._ctx

    def width(self) -> int:  # :138-140
        return int(self.image.shape[1])

    def height(self) -> int:  # :142-144
        return int(self.image.shape[0])

    def dimensions(self):  # :155-160
        return {"width": self.width(), "height": self.height()}

    def image_pil(self):  # :29-31
        from PIL import Image

        return Image.fromarray(self.image, "RGBA")

    def depth(self):  # :33-35 -- a clone, as DynamicImage::ImageLuma8
        from PIL import Image

        return Image.fromarray(self.depth_plane.copy(), "L")

    def highlight_depth(self) -> np.ndarray:  # :51-63 -> RGBA8
        return self._context().highlight_depth(self.image, self.depth_plane)

    def load_depth(self, depth: np.ndarray) -> None:  # :37-49
        depth = np.ascontiguousarray(depth, dtype=np.uint8)
        if depth.ndim == 2 and depth.shape[1] == self.width() and depth.shape[0] == self.height():
            self.depth_plane = depth
            return
        raise StepthIOError("Sizes don't match")

    def open_depth(self, depth_path: str) -> None:  # :65-74
        im = _open_image(depth_path, self._ctx)
        self.load_depth(np.ascontiguousarray(im[..., 0]) if isinstance(im, np.ndarray) else np.asarray(im.convert("L"), dtype=np.uint8))

    # ---- the hot path ----------------------------------------------------------------------------------------
    def open_depth_from_additional(self, add_path: str, precision, **kw) -> None:  # :76-89
        self.load_depth_from_additional(_open_image(add_path, self._ctx), precision, **kw)

    def load_depth_from_additional(self, add_image, precision, *, mode: str = "blocks", blocks=None,
                                   max_rings: int | None = None, flags: int = _ffi.RESAMPLE_U8) -> None:
        """depth_image.rs:91-136.

        mode="blocks": `blocks` may carry the DiscretePixel list produced by the real disage on the host (steps 3-5); by default
        the documented stand-in `discretise_v0` runs on the GPU in the same call as steps 6-10.  mode="dense": every pixel is its own block (the limit case the
        benchmark configs use).
        """
        add_rgb = _to_rgb8(add_image)  # :96
        ctx = self._context()
        max_rings = self.MAX_RINGS if max_rings is None else max_rings
        flags = flags & ~(_ffi.FLAG_FOREGROUND | _ffi.FLAG_APPLY)
        if mode == "dense":
            out = ctx.depth_dense(self.image, add_rgb, precision, max_rings, flags)
        elif mode == "blocks":
            if blocks is None:   # :101-109 on the GPU too (the stand-in discretiser; its block list never visits the host)
                out = ctx.depth_from_additional(self.image, add_rgb, precision, max_rings, flags, self.MIN_SPLITS)
            else:
                out = ctx.depth_from_blocks(add_rgb, blocks, self.width(), self.height(), precision, max_rings, flags)
        else:
            raise ValueError("mode must be 'blocks' or 'dense'")
        self.load_depth(out["depth"])  # :135

    # ---- tail --------------------------------------------------------------------------------------------------
    def invert_depth(self) -> None:  # :225-227
        self.depth_plane = self._context().invert_depth(self.depth_plane)

    def depth_split(self, zones: int):  # :162-218
        return self._context().depth_split(self.depth_plane, zones)

    def slice(self, frm, to) -> MaskImage:  # :229-245
        mask = self._context().slice(self.depth_plane, frm, to)
        return MaskImage(self.image.copy(), mask, self._ctx)

    def select_foreground(self) -> MaskImage:  # :220-223
        mask, _ = self._context().select_foreground(self.depth_plane)
        return MaskImage(self.image.copy(), mask, self._ctx)
