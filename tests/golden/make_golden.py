"""Generates tests/golden/assets.npz from the reference's shipped images (run in the build container only).

The reference has no tests or golden vectors; its four JPEG assets are the only result
artefacts it ships, so they are the soft pins of the oracle.  They are decoded ONCE here with PIL (libjpeg) and the
decoded pixels are committed: parity is defined on decoded buffers (jpeg-decoder 0.1.22 != libjpeg bit for bit).

    python tests/golden/make_golden.py        # needs /root/reference (absent on the GPU box)
"""
import os

import numpy as np
from PIL import Image

REF = "/root/reference/assets"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "assets.npz")


def main():
    arrs = {}
    for name, mode in [("main", "RGB"), ("additional", "RGB"), ("depth", "L"), ("foreground", "RGB")]:
        im = Image.open(os.path.join(REF, name + ".jpg"))
        assert im.mode == mode, (name, im.mode)
        arrs[name] = np.asarray(im, dtype=np.uint8)
    np.savez_compressed(OUT, **arrs)
    print("wrote", OUT, {k: v.shape for k, v in arrs.items()}, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
