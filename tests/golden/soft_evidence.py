"""What the reference's shipped assets say about each UNPINNED choice of the oracle.

    python tests/golden/soft_evidence.py > tests/golden/soft_evidence.txt        # CPU only: oracle + tests/golden/assets.npz

assets/depth.jpg is the reference's own output for assets/main.jpg + assets/additional.jpg at precision 255/7 (Readme.md:13-16),
JPEG-compressed: it cannot pin a byte, but it ranks the alternatives.  Every row runs the C restatement (oracle/stepth_oracle.c)
with ONE choice changed and scores the result against the decoded depth.jpg: Pearson r, mean absolute error in grey levels, the
fraction of exact zeros (asset: see the first line).  TEST INFRASTRUCTURE -- nothing here is on the product path.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import oracle as orc  # noqa: E402


def score(out, ref):
    o, r = out.reshape(-1).astype(np.float64), ref.reshape(-1).astype(np.float64)
    return np.corrcoef(o, r)[0, 1], np.abs(o - r).mean(), float((out == 0).mean())


def main():
    a = np.load(os.path.join(HERE, "assets.npz"))
    main_px, add, ref = a["main"], a["additional"], a["depth"]
    h, w = ref.shape
    p = [255 // 7] * 3
    print(f"asset depth.jpg: {w}x{h}, mean {ref.mean():.2f}, zeros {(ref == 0).mean():.3f}, <=2 {(ref <= 2).mean():.3f}")
    rows = []

    def add_row(what, out):
        r, mae, z = score(out, ref)
        rows.append((what, r, mae, z))
        print(f"{what:<78s} r {r:.4f}  MAE {mae:5.2f}  zeros {z:.3f}", flush=True)

    # PixelOpps (disage, absent): dense mode, u8 resample
    for name, po in (("abs-diff / all < p   (default)", orc.PIXELOPPS_ABS_ALL), ("saturating-sub / all < p", orc.PIXELOPPS_SAT_ALL),
                     ("abs-diff / sum < sum(p)", orc.PIXELOPPS_ABS_SUM)):
        add_row("PixelOpps = " + name + ", dense, u8 resample", orc.depth_dense(main_px, add, p, 255, pixelopps=po, resample=orc.RESAMPLE_U8))
    # the unit-scale Gaussian of image 0.23.14: off / u8 intermediate / f32 intermediate
    for name, rs in (("no resample", orc.RESAMPLE_OFF), ("u8 intermediate (default)", orc.RESAMPLE_U8), ("f32 intermediate", orc.RESAMPLE_F32)):
        add_row("resample = " + name + ", dense, abs/all", orc.depth_dense(main_px, add, p, 255, resample=rs))
    # the block list (disage::open::rgb_discrete, absent): 1x1 blocks vs the stand-in disc_v0 at the reference's parameters
    try:
        import stepth_b200
        blocks = stepth_b200.discretise_v0(main_px, p)          # host stand-in, no GPU involved
        add_row(f"blocks = disc_v0 stand-in ({len(blocks)} blocks, min 16 / max 18 splits), abs/all, u8 resample",
                orc.depth_from_blocks(add, blocks, p, w, h, 255))
        for ms in (4, 8, 12):
            b = stepth_b200.discretise_v0(main_px, p, min_splits=ms)
            add_row(f"blocks = disc_v0 with min_splits {ms} ({len(b)} blocks)", orc.depth_from_blocks(add, b, p, w, h, 255))
    except Exception as e:   # the stand-in lives in the product library (host code); without it the rows above stand alone
        print("block rows skipped:", e)
    # max (the reference hard-codes 255, depth_image.rs:119) and the precision of the README call
    for mr in (128, 200):
        add_row(f"max = {mr} instead of 255, dense, abs/all, u8 resample", orc.depth_dense(main_px, add, p, mr))
    for pp in (30, 42):
        add_row(f"precision = {pp} instead of 36, dense", orc.depth_dense(main_px, add, [pp] * 3, 255))
    best = max(rows, key=lambda t: t[1])
    print("highest r:", best[0])


if __name__ == "__main__":
    main()
