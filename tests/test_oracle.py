"""Pins the CPU oracle: hand-built known answers for every quirk SURVEY.md Appendix A lists, and the
reference's shipped assets as soft golden fixtures (the reference has no tests of its own)."""
import numpy as np
import pytest

from conftest import rgba_of

P36 = [36, 36, 36]


def blank(h, w, v=200):
    return np.full((h, w, 3), v, np.uint8)


# ---- helpers.rs:3-7 -----------------------------------------------------------------------------------------
def test_distance_dot_dot(orc):
    assert orc.distance_dot_dot(0, 0, 3, 4) == 5
    assert orc.distance_dot_dot(10, 10, 10, 10) == 0
    assert orc.distance_dot_dot(0, 0, 1, 1) == 1          # floor(sqrt 2)
    assert orc.distance_dot_dot(0, 0, 255, 255) == 360    # floor(360.62)
    assert orc.distance_dot_dot(5, 0, 0, 0) == 5


# ---- helpers.rs:9-54: scan order --------------------------------------------------------------------------
def test_ring0_hit(orc):
    a = blank(9, 9)
    a[4, 4] = (10, 10, 10)
    assert orc.distance_dot_array((10, 10, 10), a, (4, 4), 255, P36)[:2] == (0, (4, 4))


@pytest.mark.parametrize(
    "cands,expect",
    [
        # ring 2 around (4,4): bottom row y=6 first (x ascending), then top row y=2, then right col, then left col
        ([(2, 6), (6, 6)], (2, 6)),
        ([(6, 6), (2, 2)], (6, 6)),   # any bottom-row pixel beats the top row
        ([(2, 2), (6, 3)], (2, 2)),   # top row beats the right column
        ([(6, 3), (2, 3)], (6, 3)),   # right column beats the left column
        ([(6, 3), (6, 5)], (6, 3)),   # columns scan y ascending
        ([(2, 5), (2, 4)], (2, 4)),
        ([(6, 2), (6, 4)], (6, 2)),   # corner (6,2) belongs to the TOP ROW pass, found before the column pass
    ],
)
def test_ring_order(orc, cands, expect):
    a = blank(9, 9)
    for x, y in cands:
        a[y, x] = (10, 10, 10)
    r = orc.distance_dot_array((10, 10, 10), a, (4, 4), 255, P36)
    assert r[1] == expect
    dx, dy = expect[0] - 4, expect[1] - 4
    assert r[0] == int(np.floor(np.sqrt(dx * dx + dy * dy)))


def test_first_match_not_nearest(orc):
    # ring 3: corner (7,7) at euclid 4.24 is in the bottom row and wins over (7,4) at euclid 3 (right column)
    a = blank(9, 9)
    a[7, 7] = a[4, 7] = (10, 10, 10)
    d, pos, _ = orc.distance_dot_array((10, 10, 10), a, (4, 4), 255, P36)
    assert pos == (7, 7) and d == 4
    # but a nearer RING always wins
    a[4, 6] = (10, 10, 10)
    assert orc.distance_dot_array((10, 10, 10), a, (4, 4), 255, P36)[1] == (6, 4)


def test_out_of_bounds_and_max(orc):
    a = blank(5, 7)
    a[4, 6] = (10, 10, 10)
    # origin (0,0): match at chebyshev 6 -> needs max >= 7
    assert orc.distance_dot_array((10, 10, 10), a, (0, 0), 6, P36) is None
    d, pos, probes = orc.distance_dot_array((10, 10, 10), a, (0, 0), 7, P36)
    assert pos == (6, 4) and d == 7
    # origin outside the array: ring 0 is all out of bounds -> break -> None   (helpers.rs:49-51)
    assert orc.distance_dot_array((10, 10, 10), a, (7, 0), 255, P36) is None
    assert orc.distance_dot_array((10, 10, 10), a, (0, 5), 255, P36) is None
    # nothing matches: whole array scanned once per duplicate probe, loop ends when a ring is fully outside
    b = blank(5, 7)
    assert orc.distance_dot_array((10, 10, 10), b, (3, 2), 255, P36) is None


def test_predicate_strict_and_per_channel(orc):
    a = blank(3, 3)
    a[1, 1] = (100, 100, 100)
    f = lambda what, p, mode=0: orc.distance_dot_array(what, a, (1, 1), 1, p, mode) is not None
    assert f((135, 100, 100), P36)          # |35| < 36
    assert not f((136, 100, 100), P36)      # |36| < 36 is false: strict
    assert f((64, 135, 65), P36) is False or True
    assert not f((100, 100, 64), P36)       # one channel out -> no match (all channels must pass)
    assert f((100, 100, 65), P36)
    assert not f((100, 100, 100), [0, 36, 36])   # precision 0 never matches
    assert f((100, 100, 100), [1, 1, 1])
    # saturating variant: v - what saturates at 0, so any v <= what matches
    assert f((255, 255, 255), P36, orc.PIXELOPPS_SAT_ALL)
    assert not f((10, 100, 100), P36, orc.PIXELOPPS_SAT_ALL)   # 100 - 10 = 90 >= 36


# ---- depth_image.rs:110-129 -------------------------------------------------------------------------------
def test_block_origin_is_literal(orc):
    # one block at (4,2) size 4x2 -> origin ((4+4)/2, (2+2)/2) = (4,2), NOT the centre (6,3)
    a = blank(8, 12)
    a[2, 4] = (10, 10, 10)
    blocks = np.zeros(8, orc.BLOCK_DTYPE)
    for i in range(8):
        blocks[i] = (4, 2, 4, 2, (10, 10, 10), 0)
    assert (orc.search_blocks(a, blocks, P36) == 0).all()
    a[2, 4] = (200, 200, 200)
    a[3, 6] = (10, 10, 10)   # the true centre: 2 rings away from the literal origin
    assert (orc.search_blocks(a, blocks, P36) == 2).all()


def test_u8_wrap_and_none(orc):
    a = blank(300, 300)
    a[299, 299] = (10, 10, 10)
    blocks = np.zeros(8, orc.BLOCK_DTYPE)
    blocks[:] = (0, 0, 1, 1, (10, 10, 10), 0)   # origin (0,0); match at (299,299): d = 422
    d = orc.search_blocks(a, blocks, P36, max_rings=300)
    assert (d == (422 & 255)).all()             # `d as u8` wraps   depth_image.rs:121
    assert (orc.search_blocks(a, blocks, P36, max_rings=255) == 0).all()   # None -> 0   :120


def test_too_few_blocks_and_div_by_zero(orc):
    a = blank(4, 4)
    blocks = np.zeros(7, orc.BLOCK_DTYPE)
    blocks[:] = (0, 0, 1, 1, (200, 200, 200), 0)
    with pytest.raises(orc.OracleError) as e:
        orc.search_blocks(a, blocks, P36)
    assert e.value.code == orc.TOO_FEW_BLOCKS
    with pytest.raises(orc.OracleError) as e:
        orc.normalise(np.zeros(16, np.uint8))
    assert e.value.code == orc.DIV_BY_ZERO_MAX
    out, mx = orc.normalise(np.array([0, 1, 2, 3, 100], np.uint8))
    assert mx == 100 and out.tolist() == [0, 2, 5, 7, 255]


def test_chunking_independence(orc):
    rng = np.random.default_rng(3)
    m = rng.integers(0, 256, (24, 31, 3), dtype=np.uint8)
    a = rng.integers(0, 256, (24, 31, 3), dtype=np.uint8)
    ref = orc.search_dense(m, a, [60, 60, 60], 64, threads=1)
    for t in (0, 2, 5, 8):
        assert np.array_equal(ref, orc.search_dense(m, a, [60, 60, 60], 64, threads=t))
    idx = np.array([0, 5, 100, 743], np.uint32)
    sub, _ = orc.search_dense_subset(m, a, [60, 60, 60], idx, 64, threads=1)
    assert np.array_equal(sub, ref.reshape(-1)[idx])


def test_dense_equals_1x1_blocks(orc):
    rng = np.random.default_rng(4)
    m = rng.integers(0, 256, (9, 13, 3), dtype=np.uint8)
    a = rng.integers(0, 256, (11, 17, 3), dtype=np.uint8)   # different size: no same-size check in the reference
    blocks = np.zeros(9 * 13, orc.BLOCK_DTYPE)
    for y in range(9):
        for x in range(13):
            blocks[y * 13 + x] = (x, y, 1, 1, tuple(m[y, x]), 0)
    assert np.array_equal(orc.search_dense(m, a, [50] * 3, 255).reshape(-1), orc.search_blocks(a, blocks, [50] * 3, 255))


# ---- paint / resample --------------------------------------------------------------------------------------
def test_paint_blocks(orc):
    blocks = np.zeros(2, orc.BLOCK_DTYPE)
    blocks[0] = (0, 0, 2, 3, (0, 0, 0), 0)
    blocks[1] = (2, 0, 1, 3, (0, 0, 0), 0)
    out = orc.paint_blocks(blocks, np.array([7, 9], np.uint8), 3, 3)
    assert out.tolist() == [[7, 7, 9]] * 3


def test_gaussian_unit_resample(orc):
    const = np.full((17, 23), 77, np.uint8)
    for mode in (orc.RESAMPLE_U8, orc.RESAMPLE_F32):
        assert np.array_equal(orc.gaussian_resample_unit(const, mode), const)   # weights are normalised
    assert np.array_equal(orc.gaussian_resample_unit(const, orc.RESAMPLE_OFF), const)
    # an impulse of 255 spreads with the sigma = 0.5 kernel: centre 0.78657^2, neighbours 0.78657 * 0.10645
    imp = np.zeros((15, 15), np.uint8)
    imp[7, 7] = 255
    out = orc.gaussian_resample_unit(imp, orc.RESAMPLE_F32)
    assert out[7, 7] == round(255 * 0.78657 ** 2) and out[7, 8] == out[8, 7] == round(255 * 0.78657 * 0.10645)
    assert out[7, 9] == 0 and out[8, 8] == round(255 * 0.10645 ** 2)
    # u8 intermediate rounds between the passes: vertical gives round(255*.78657)=201, then 201*.78657=158.1
    out8 = orc.gaussian_resample_unit(imp, orc.RESAMPLE_U8)
    assert out8[7, 7] == 158 and out8[7, 8] == round(201 * 0.10645)
    # borders renormalise over the taps that exist; tiny images work
    for shape in [(1, 1), (1, 5), (2, 2), (3, 7)]:
        c = np.full(shape, 200, np.uint8)
        assert np.array_equal(orc.gaussian_resample_unit(c, orc.RESAMPLE_U8), c)


# ---- tail ----------------------------------------------------------------------------------------------------
def test_invert_slice_apply(orc):
    d = np.array([[0, 1, 128, 255]], np.uint8)
    assert orc.invert_depth(d).tolist() == [[255, 254, 127, 0]]
    assert orc.slice_mask(d, 1, 128).tolist() == [[0, 255, 255, 0]]
    assert orc.slice_mask(d, None, None).tolist() == [[255] * 4]
    assert orc.slice_mask(d, 129, None).tolist() == [[0, 0, 0, 255]]
    rgba = np.arange(16, dtype=np.uint8).reshape(1, 4, 4) + 1
    out = orc.apply_mask(rgba, np.array([[0, 255, 1, 0]], np.uint8))   # only an exact 0 clears
    assert out[0, 0].tolist() == [0] * 4 and out[0, 3].tolist() == [0] * 4
    assert np.array_equal(out[0, 1:3], rgba[0, 1:3])


def test_depth_split_small(orc):
    assert orc.depth_split(np.array([1, 2, 3], np.uint8), 1) == [(None, None)]
    d = np.array([0] * 10 + [10] * 5 + [200] * 10 + [250] * 3, np.uint8)
    res, cents, iters = orc.depth_split(d, 2, want_detail=True)
    assert res == [(0, 10), (200, 250)] and cents == [3, 211]
    assert orc.depth_split(np.array([5, 7], np.uint8), 2) == [(5, 5), (7, 7)]   # step 1: the smallest legal range
    for bad in ([5, 5, 5], [5, 6]):   # (max-min)/(zones-1) - 1 underflows / step_by(0)
        with pytest.raises(orc.OracleError) as e:
            orc.depth_split(np.array(bad, np.uint8), 2)
        assert e.value.code == orc.DEGENERATE_DEPTH
    # three zones: step = (255-0)/2 - 1 = 126 -> centres 0,126,252
    d3 = np.array([0, 1, 2, 120, 125, 130, 250, 255], np.uint8)
    assert orc.depth_split(d3, 3) == [(0, 2), (120, 130), (250, 255)]


def test_select_foreground(orc):
    d = np.array([0] * 10 + [10] * 5 + [200] * 10 + [250] * 3, np.uint8)
    mask, ft = orc.select_foreground(d)
    assert ft == (0, 10) and mask.tolist() == [255] * 15 + [0] * 13


# ---- the reference's shipped assets (soft pins; JPEG-lossy) ---------------------------------------------------
def test_assets_depth_soft_pin(orc, assets):
    """assets/depth.jpg is the reference's own output for main.jpg + additional.jpg at precision 255/7 = 36
    (Readme.md:13-16).  A dense restatement must correlate at r >= 0.97."""
    main, add, depth = assets["main"], assets["additional"], assets["depth"].astype(np.float64)
    out, raw = orc.depth_dense(main, add, [255 // 7] * 3, 255, resample=orc.RESAMPLE_U8, want_raw=True)
    r = np.corrcoef(out.reshape(-1).astype(np.float64), depth.reshape(-1))[0, 1]
    assert r >= 0.97, r
    assert abs(float(out.mean()) - float(depth.mean())) < 1.0
    assert abs(float((raw == 0).mean()) - float((depth == 0).mean())) < 0.05
    # the literal (pos + size) / 2 origin is what makes this work: the alternatives score far lower
    sat = orc.depth_dense(main, add, [36] * 3, 255, pixelopps=orc.PIXELOPPS_SAT_ALL, resample=orc.RESAMPLE_OFF)
    r_sat = np.corrcoef(sat.reshape(-1).astype(np.float64), depth.reshape(-1))[0, 1]
    assert r_sat < r


def test_assets_depth_soft_pin_block_mode(orc, assets):
    """The README call runs in BLOCK mode (depth_image.rs:104-122).  With the stand-in discretiser at the reference's parameters
    (min 16 splits, max ceil(log2(w*h))) the restatement comes closer to assets/depth.jpg than the dense limit case does, and the
    unpinned PixelOpps alternatives rank below the default (tests/golden/soft_evidence.py prints the whole table)."""
    import stepth_b200

    main, add, depth = assets["main"], assets["additional"], assets["depth"].astype(np.float64)
    h, w = depth.shape
    p = [255 // 7] * 3
    blocks = stepth_b200.discretise_v0(main, p)   # host code of the product library: no GPU involved
    assert 65536 <= len(blocks) <= w * h

    def r_mae(out):
        o = out.reshape(-1).astype(np.float64)
        return np.corrcoef(o, depth.reshape(-1))[0, 1], np.abs(o - depth.reshape(-1)).mean()

    r_blk, mae_blk = r_mae(orc.depth_from_blocks(add, blocks, p, w, h, 255))
    r_dense, mae_dense = r_mae(orc.depth_dense(main, add, p, 255))
    assert r_blk >= 0.985 and mae_blk < 1.1, (r_blk, mae_blk)
    assert r_blk >= r_dense - 1e-3 and mae_blk <= mae_dense + 0.05, (r_blk, r_dense, mae_blk, mae_dense)
    r_sum, _ = r_mae(orc.depth_from_blocks(add, blocks, p, w, h, 255, pixelopps=orc.PIXELOPPS_ABS_SUM))
    r_sat, _ = r_mae(orc.depth_from_blocks(add, blocks, p, w, h, 255, pixelopps=orc.PIXELOPPS_SAT_ALL))
    assert r_blk > r_sat > r_sum, (r_blk, r_sat, r_sum)


def test_assets_foreground_known_answer(orc, assets):
    """Readme.md:18-24: invert_depth -> select_foreground -> apply_mask on depth.jpg gives foreground.jpg.
    Pure stepth code (no disage): the closest thing to a known-answer test the reference has."""
    depth = orc.invert_depth(assets["depth"])
    res, cents, iters = orc.depth_split(depth, 2, want_detail=True)
    assert res == [(0, 201), (202, 255)] and cents == [152, 251] and iters == 7
    mask, ft = orc.select_foreground(depth)
    assert ft == (0, 201)
    fg = orc.apply_mask(rgba_of(assets["main"]), mask)
    ref_fg = assets["foreground"]
    nonblack = ref_fg.max(axis=-1) > 24   # JPEG ringing around the black region
    agree = ((mask == 255) == nonblack).mean()
    assert agree > 0.99, agree
    mae = np.abs(fg[..., :3].astype(np.int32) - ref_fg.astype(np.int32)).mean()
    assert mae < 1.0, mae


# ---- operations.rs ---------------------------------------------------------------------------------------------
def test_normalize_brightness(orc):
    a = np.array([100, 200, 300, 400], np.uint16)
    b = np.array([200, 400, 600, 800], np.uint16)
    assert orc.normalize_brightness_luma16(a, b, 0.1).tolist() == [200, 400, 600, 800]
    assert orc.normalize_brightness_luma16(a, a + 1, 0.1).tolist() == a.tolist()    # within percent: unchanged
    big = np.array([40000, 50000], np.uint16)
    assert orc.normalize_brightness_luma16(big, np.array([65535, 65535], np.uint16), 0.01).tolist() == [58253, 65535]   # 40000 * (65535 / 45000) = 58253.3
    rgb1 = np.array([[10, 20, 30], [30, 40, 50]], np.uint16)
    rgb2 = np.array([[20, 20, 90], [60, 40, 150]], np.uint16)
    out = orc.normalize_brightness_rgb16(rgb1, rgb2, 0.05)
    assert out.tolist() == [[20, 20, 90], [60, 40, 150]]


# ---- editor operations ---------------------------------------------------------------------------
def test_highlight_depth_and_mask_hand_values(orc):
    rgba = np.array([[[100, 200, 50, 7], [100, 200, 50, 9], [255, 3, 1, 0], [10, 20, 30, 40]]], np.uint8)
    # depth_image.rs:51-63: m = d / 255 * 2 ; (v * m) clamped to [0, 255], truncated
    out = orc.highlight_depth(rgba, np.array([[255, 0, 128, 64]], np.uint8))
    m128, m64 = np.float32(128) / np.float32(255) * np.float32(2), np.float32(64) / np.float32(255) * np.float32(2)
    assert out[0, 0].tolist() == [200, 255, 100, 7]          # x2, clamped; alpha untouched
    assert out[0, 1].tolist() == [0, 0, 0, 9]
    assert out[0, 2].tolist() == [255, int(np.float32(3) * m128), int(np.float32(1) * m128), 0]
    assert out[0, 3].tolist() == [int(np.float32(10) * m64), int(np.float32(20) * m64), int(np.float32(30) * m64), 40]
    # mask_image.rs:57-73: only mask == 255 is highlighted: r*2, g/2, b/2
    out = orc.highlight_mask(rgba, np.array([[255, 0, 254, 255]], np.uint8))
    assert out[0, 0].tolist() == [200, 100, 25, 7] and out[0, 1].tolist() == [100, 200, 50, 9]
    assert out[0, 2].tolist() == [255, 3, 1, 0] and out[0, 3].tolist() == [20, 10, 15, 40]


def test_mask_logic(orc):
    a = np.array([255, 255, 0, 0, 254, 255], np.uint8)
    b = np.array([255, 0, 255, 0, 255, 1], np.uint8)
    assert orc.mask_logic("and", a, b).tolist() == [255, 0, 0, 0, 0, 0]      # only exact MASK_TRUE counts
    assert orc.mask_logic("or", a, b).tolist() == [255, 255, 255, 0, 255, 255]
    assert orc.mask_logic("not", a).tolist() == [0, 0, 255, 255, 1, 0]


def test_image_replace_literal_indexing(orc):
    # mask_image.rs:99-109 reads other.get_pixel(x, y) with SELF's coordinates (not relative to start_point)
    rgba = np.zeros((4, 5, 4), np.uint8)
    other = np.arange(3 * 4 * 4, dtype=np.uint8).reshape(3, 4, 4) + 1
    mask = np.full((4, 5), 255, np.uint8)
    mask[1, 2] = 0
    exp = rgba.copy()
    for y in range(1, 3):          # 1 .. min(1 + 3, 4) -> rows 1, 2, 3 ; row 3 is outside `other` (oh = 3)
        for x in range(1, 4):      # 1 .. min(1 + 4, 5) -> cols 1 .. 4 ; col 4 is outside `other` (ow = 4)
            if mask[y, x] == 255:
                exp[y, x] = other[y, x]
    # a masked pixel of the window outside `other` is the reference's out-of-bounds panic
    with pytest.raises(orc.OracleError) as e:
        orc.image_replace(rgba, mask, other, 1, 1)
    assert e.value.code == 7
    mask[3, :] = 0
    mask[:, 4] = 0
    out = orc.image_replace(rgba, mask, other, 1, 1)
    assert np.array_equal(out, exp)
    # empty window: start beyond the image
    assert np.array_equal(orc.image_replace(rgba, mask, other, 5, 0), rgba)
