"""GPU parity: libstepth_cuda.so (through the C ABI) against the CPU oracle, bit for bit, on the same bytes.

Everything here is integer / byte / index work except the unit-scale Gaussian resample (f32, image 0.23.14);
that one is also required to be bit-exact because both sides use the same tap tables, tap order and rounding.
"""
import numpy as np
import pytest
import torch

from conftest import rgba_of

pytestmark = pytest.mark.gpu

P36 = [36, 36, 36]


def synth_pair(kind, w, h, seed=1):
    from stepth_b200 import synth

    m, a = synth.make_pair(kind, w, h, seed)
    return m.numpy(), a.numpy()


def rand_pair(w, h, wa=None, ha=None, seed=0, levels=256):
    rng = np.random.default_rng(seed)
    wa, ha = wa or w, ha or h
    step = 256 // levels
    m = (rng.integers(0, levels, (h, w, 3)) * step).astype(np.uint8)
    a = (rng.integers(0, levels, (ha, wa, 3)) * step).astype(np.uint8)
    return rgba_of(m), a


def gpu_raw(ctx, m, a, prec, rings, flags=0):
    import stepth_b200 as sb

    try:
        return ctx.depth_dense(m, a, prec, rings, flags | sb.RESAMPLE_OFF, want_raw=True)["raw"]
    except sb.StepthPanic as e:   # max == 0: the raw plane is still returned
        assert e.code == 3
        return e.partial["raw"]


# ---- K1: raw distances --------------------------------------------------------------------------------------
@pytest.mark.parametrize("kind", ["nat", "geo", "worst"])
@pytest.mark.parametrize("w,h,rings", [(96, 64, 128), (161, 77, 255), (64, 200, 40)])
def test_search_dense_vs_oracle(ctx, orc, kind, w, h, rings):
    import stepth_b200 as sb

    m, a = synth_pair(kind, w, h, seed=w + h)
    ref = orc.search_dense(m, a, P36, rings, threads=8)
    assert np.array_equal(gpu_raw(ctx, m, a, P36, rings), ref)
    assert np.array_equal(gpu_raw(ctx, m, a, P36, rings, sb.FLAG_LITERAL_SEARCH), ref)


@pytest.mark.parametrize(
    "w,h,wa,ha",
    [(8, 1, 8, 1), (1, 8, 1, 8), (3, 5, 3, 5), (17, 33, 17, 33), (129, 67, 129, 67), (40, 30, 7, 5), (40, 30, 90, 70),
     (33, 9, 4, 40), (2, 4, 300, 2), (257, 3, 129, 2)],
)
def test_search_dense_ragged_shapes(ctx, orc, w, h, wa, ha):
    m, a = rand_pair(w, h, wa, ha, seed=w * 31 + h, levels=6)
    for prec in ([60, 60, 60], [44, 90, 130]):
        ref = orc.search_dense(m, a, prec, 255, threads=4)
        assert np.array_equal(gpu_raw(ctx, m, a, prec, 255), ref), (w, h, wa, ha, prec)


@pytest.mark.parametrize("seed", range(40))
def test_search_seeded_random_cases(ctx, orc, seed):
    """Shapes, additional-image shapes, colour depth, per-channel precision and ring limit all drawn from a seed."""
    rng = np.random.default_rng(1000 + seed)
    w, h = int(rng.integers(1, 280)), int(rng.integers(1, 200))
    if rng.random() < 0.5:
        wa, ha = w, h
    else:
        wa, ha = int(rng.integers(1, 320)), int(rng.integers(1, 240))
    levels = int(rng.choice([3, 6, 16, 64, 256, 256]))
    prec = [int(rng.choice([1, 2, 4, 9, 20, 36, 70, 128, 255])) for _ in range(3)]
    if seed % 13 == 0:
        prec[seed % 3] = 0   # nothing matches: every query runs out of rings
    rings = int(rng.choice([1, 2, 5, 6, 7, 10, 11, 15, 16, 17, 33, 64, 128, 255, 300]))
    m, a = rand_pair(w, h, wa, ha, seed=seed, levels=levels)
    ref = orc.search_dense(m, a, prec, rings, threads=4)
    assert np.array_equal(gpu_raw(ctx, m, a, prec, rings), ref), (w, h, wa, ha, levels, prec, rings)


@pytest.mark.parametrize("env", [
    {"STEPTH_NEAR_RINGS": "1", "STEPTH_MID_RINGS": "2"},
    {"STEPTH_NEAR_RINGS": "3", "STEPTH_MID_RINGS": "3"},
    {"STEPTH_NEAR_RINGS": "40", "STEPTH_MID_RINGS": "0"},
    {"STEPTH_NEAR_COMPACT": "0"},
    {"STEPTH_NEAR_TILE": "0"},
    {"STEPTH_NEAR_TILE": "0", "STEPTH_NEAR_RINGS": "1", "STEPTH_MID_RINGS": "4"},
    {"STEPTH_NEAR_TILE": "1", "STEPTH_NEAR_RINGS": "2", "STEPTH_MID_RINGS": "0"},
    {"STEPTH_NEAR_TILE": "1", "STEPTH_NEAR_RINGS": "3", "STEPTH_MID_RINGS": "9"},
    {"STEPTH_BATCH_CHUNKS": "3", "STEPTH_TWO_STREAMS": "0", "STEPTH_TRACE": "1"},
    {"STEPTH_BATCH_CHUNKS": "16", "STEPTH_UPLOAD_FULL_ROWS": "1", "STEPTH_FINISH_STREAMING": "0"},
])
def test_tuning_knobs_do_not_change_results(orc, env, monkeypatch):
    """The stage boundaries, compaction, chunking and upload knobs are schedule only: same bytes out."""
    import stepth_b200 as sb

    for k, v in env.items():
        monkeypatch.setenv(k, v)
    c = sb.Context(0)
    n, w, h = 5, 232, 120
    pairs = [synth_pair("nat" if i % 2 else "geo", w, h, seed=70 + i) for i in range(n)]
    got = c.depth_dense_batch([p[0] for p in pairs], [p[1] for p in pairs], P36, 128, sb.RESAMPLE_U8)
    for i, (m, a) in enumerate(pairs):
        want = orc.depth_dense(m, a, P36, 128, resample=orc.RESAMPLE_U8, threads=4)
        assert got["status"][i] == 0
        assert np.array_equal(got["depth"][i], want), (env, i)


@pytest.mark.parametrize("w,h,rings,rho", [(200, 120, 64, 0.0), (272, 96, 128, 0.002), (130, 130, 255, 0.0005)])
def test_search_every_tile_flagged(ctx, orc, w, h, rings, rho):
    """Adversarial for the tile search: every pixel of the additional image matches the query colour in exactly two
    channels (which ones varies per pixel), so the per-channel tile-occupancy level flags EVERY tile while the exact test
    rejects it.  A warp then lists far more tiles than its task array holds (the overflow passes) and, once a planted
    match is found, drops most of them by the range masks.  rho = density of planted true matches."""
    rng = np.random.default_rng(w * 7 + rings)
    base = np.array([60, 130, 200], np.uint8)
    m = np.broadcast_to(base, (h, w, 3)).copy()
    m += rng.integers(0, 5, (h, w, 3), dtype=np.uint8)          # query colours within a few levels of the base
    a = np.broadcast_to(base, (h, w, 3)).copy()
    off = rng.integers(0, 3, (h, w))                             # the channel that is far off (never within precision 36)
    a[np.arange(h)[:, None], np.arange(w)[None, :], off] = np.array([160, 30, 100], np.uint8)[off]
    planted = rng.random((h, w)) < rho
    a[planted] = base
    a[h // 2 + 3, w // 2 + 5] = base                             # at least one match: max != 0
    m = rgba_of(m)
    ref = orc.search_dense(m, a, P36, rings, threads=8)
    assert np.array_equal(gpu_raw(ctx, m, a, P36, rings), ref)


@pytest.mark.parametrize("rings", [1, 2, 3, 17, 255, 256, 300, 512])
def test_search_max_rings_and_u8_wrap(ctx, orc, rings):
    # a far planted match: origins near (0,0), the only matching pixels sit 280..420 away -> d >= 256 wraps
    h, w = 24, 24
    m = rgba_of(np.full((h, w, 3), 10, np.uint8))
    a = np.full((420, 420, 3), 200, np.uint8)
    a[300:, 300:] = 10
    a[5, 290] = 10
    ref = orc.search_dense(m, a, P36, rings, threads=8)
    got = gpu_raw(ctx, m, a, P36, rings)
    assert np.array_equal(got, ref)
    if rings >= 300:
        assert (ref.astype(int) != 0).any()


@pytest.mark.parametrize("prec", [[0, 36, 36], [1, 1, 1], [255, 255, 255], [36, 1, 255], [128, 128, 128]])
def test_search_precision_edges(ctx, orc, prec):
    import stepth_b200 as sb

    m, a = rand_pair(50, 40, seed=7, levels=256)
    m[..., :3][::3, ::5] = a[::3, ::5][:, :, :]   # some exact matches so that precision 1 finds something
    for flags in (0, sb.FLAG_PIXELOPPS_SAT):
        mode = orc.PIXELOPPS_SAT_ALL if flags else orc.PIXELOPPS_ABS_ALL
        ref = orc.search_dense(m, a, prec, 64, pixelopps=mode, threads=4)
        assert np.array_equal(gpu_raw(ctx, m, a, prec, 64, flags), ref), (prec, flags)


def test_probe_counter_equals_the_oracle_count(ctx, orc):
    """bench.py's algorithmic-work figure: the device-side count of the probes the reference order needs == the oracle's"""
    from stepth_b200 import device

    pairs = [synth_pair(k, 150, 90, seed=3 + i) for i, k in enumerate(("nat", "geo", "worst"))]
    m = torch.from_numpy(np.stack([p[0] for p in pairs])).cuda()
    a = torch.from_numpy(np.stack([p[1] for p in pairs])).cuda()
    d8, probes = device.count_probes(ctx, m, a, P36, 64)
    for i, (mi, ai) in enumerate(pairs):
        ref, _ = orc.search_dense(mi, ai, P36, 64, threads=4, want_probes=True)
        assert np.array_equal(d8[i].cpu().numpy(), ref)
        assert int(probes[i]) == orc.last_unique_probes(), i


def test_origin_outside_additional(ctx, orc):
    # additional much smaller than main/2: origins beyond it give None -> 0 (helpers.rs:49-51)
    m, a = rand_pair(64, 48, 10, 6, seed=5, levels=3)
    ref = orc.search_dense(m, a, [100] * 3, 255, threads=4)
    assert np.array_equal(gpu_raw(ctx, m, a, [100] * 3, 255), ref)
    assert (ref[20:, :] == 0).all()


# ---- whole path ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("resample", ["off", "u8", "f32"])
@pytest.mark.parametrize("kind,w,h", [("nat", 150, 90), ("geo", 131, 70)])
def test_depth_dense_vs_oracle(ctx, orc, resample, kind, w, h):
    import stepth_b200 as sb

    m, a = synth_pair(kind, w, h, seed=11)
    g = {"off": sb.RESAMPLE_OFF, "u8": sb.RESAMPLE_U8, "f32": sb.RESAMPLE_F32}[resample]
    o = {"off": orc.RESAMPLE_OFF, "u8": orc.RESAMPLE_U8, "f32": orc.RESAMPLE_F32}[resample]
    ref, raw = orc.depth_dense(m, a, P36, 128, resample=o, threads=8, want_raw=True)
    out = ctx.depth_dense(m, a, P36, 128, g, want_raw=True)
    assert np.array_equal(out["raw"], raw)
    assert np.array_equal(out["depth"], ref)


@pytest.mark.parametrize("shape", [(8, 1), (1, 9), (2, 7), (5, 5), (6, 300), (300, 6), (40, 129),
                                   # whole-word rows take the streaming form of K2: several row segments / column blocks
                                   (129, 40), (50, 1040), (7, 8), (30, 12), (3, 16), (22, 516), (64, 260)])
def test_finish_small_and_border_shapes(ctx, orc, shape):
    """the Gaussian's border renormalisation and tiny images (fewer than 7 taps), through the band-capable K2"""
    from stepth_b200 import device

    h, w = shape
    rng = np.random.default_rng(h * 1000 + w)
    raw = rng.integers(0, 200, (h, w), dtype=np.uint8)
    raw[0, 0] = 199
    norm, mx = orc.normalise(raw)
    src = torch.from_numpy(raw).cuda().unsqueeze(0)
    mxt = torch.tensor([mx], dtype=torch.int32, device="cuda")
    for g, o in ((1, orc.RESAMPLE_U8), (2, orc.RESAMPLE_F32), (0, orc.RESAMPLE_OFF)):
        depth, status = device.finish(ctx, src, mxt, h, g)
        assert int(status[0]) == 0
        assert np.array_equal(depth[0].cpu().numpy(), orc.gaussian_resample_unit(norm, o)), (shape, g)


def test_finish_normalise_every_max(ctx, orc):
    """(v * 255 / max) as u8 for every possible max: the streaming K2 divides by a multiply-high constant, not a table"""
    from stepth_b200 import device

    h, w = 9, 256
    for mx in list(range(1, 256)):
        raw = (np.arange(h * w, dtype=np.int64).reshape(h, w) * 7919 % (mx + 1)).astype(np.uint8)
        raw[4, 100] = mx
        norm, got_mx = orc.normalise(raw)
        assert got_mx == mx
        src = torch.from_numpy(raw).cuda().unsqueeze(0)
        mxt = torch.tensor([mx], dtype=torch.int32, device="cuda")
        depth, status = device.finish(ctx, src, mxt, h, 1)
        assert int(status[0]) == 0
        assert np.array_equal(depth[0].cpu().numpy(), orc.gaussian_resample_unit(norm, orc.RESAMPLE_U8)), mx


def test_fused_tail_vs_oracle(ctx, orc):
    import stepth_b200 as sb

    m, a = synth_pair("nat", 200, 120, seed=3)
    flags = sb.RESAMPLE_U8 | sb.FLAG_INVERT | sb.FLAG_FOREGROUND | sb.FLAG_APPLY
    out = ctx.depth_dense(m, a, P36, 128, flags)
    depth = orc.invert_depth(orc.depth_dense(m, a, P36, 128, resample=orc.RESAMPLE_U8, threads=8))
    mask, ft = orc.select_foreground(depth)
    assert np.array_equal(out["depth"], depth)
    assert out["from_to"] == ft
    assert np.array_equal(out["mask"], mask)
    assert np.array_equal(out["rgba"], orc.apply_mask(m, mask))


def test_blocks_vs_oracle(ctx, orc):
    import stepth_b200 as sb

    m, a = synth_pair("nat", 180, 100, seed=9)
    blocks = sb.discretise_v0(m, P36, min_splits=10)
    assert 1024 <= len(blocks) < 180 * 100
    ref, raw = orc.depth_from_blocks(a, blocks, P36, 180, 100, 255, resample=orc.RESAMPLE_U8, want_raw=True)
    out = ctx.depth_from_blocks(a, blocks, 180, 100, P36, 255, sb.RESAMPLE_U8, want_raw=True)
    assert np.array_equal(out["raw"], raw)
    assert np.array_equal(out["depth"], ref)
    # fused tail in block mode
    flags = sb.RESAMPLE_U8 | sb.FLAG_INVERT | sb.FLAG_FOREGROUND | sb.FLAG_APPLY
    out = ctx.depth_from_blocks(a, blocks, 180, 100, P36, 255, flags, rgba_in=m)
    depth = orc.invert_depth(ref)
    mask, ft = orc.select_foreground(depth)
    assert out["from_to"] == ft and np.array_equal(out["mask"], mask)
    assert np.array_equal(out["rgba"], orc.apply_mask(m, mask))


def test_blocks_handmade_literal_origin(ctx, orc):
    a = np.full((40, 60, 3), 200, np.uint8)
    a[2, 4] = (10, 10, 10)
    a[30, 50] = (90, 10, 10)
    blocks = np.zeros(9, orc.BLOCK_DTYPE)
    blocks[0] = (0, 0, 30, 20, (10, 10, 10), 0)      # origin (15, 10)
    blocks[1] = (30, 0, 30, 20, (90, 12, 8), 0)      # origin (30, 10)
    blocks[2] = (0, 20, 60, 20, (10, 10, 10), 0)     # origin (30, 20)
    for i in range(3, 9):                            # zero-area filler blocks (>= 8 needed): origin (x/2, y/2)
        blocks[i] = (i, i, 0, 0, (200, 200, 200), 0)
    raw = orc.search_blocks(a, blocks, P36, 255)
    out = ctx.depth_from_blocks(a, blocks, 60, 40, P36, 255, 0, want_raw=True)
    assert np.array_equal(out["raw"], raw)
    assert np.array_equal(out["depth"], orc.depth_from_blocks(a, blocks, P36, 60, 40, 255, resample=orc.RESAMPLE_OFF))


# ---- error behaviour -----------------------------------------------------------------------------------------
def test_errors_mirror_the_reference_panics(ctx, orc):
    import stepth_b200 as sb

    same = np.full((16, 16, 3), 50, np.uint8)
    with pytest.raises(sb.StepthPanic) as e:           # every distance 0 -> max 0 -> divide by zero (:127)
        ctx.depth_dense(rgba_of(same), same, P36, 255)
    assert e.value.code == 3 and "divide by zero" in str(e.value)
    with pytest.raises(sb.StepthPanic) as e:           # precision 0: nothing matches -> None -> 0 -> max 0
        ctx.depth_dense(*rand_pair(16, 16), [0, 0, 0], 255)
    assert e.value.code == 3
    with pytest.raises(sb.StepthPanic) as e:           # < 8 blocks: par_chunks_mut(0)
        ctx.depth_dense(rgba_of(same[:1, :7]), same, P36, 255)
    assert e.value.code == 4
    blocks = np.zeros(7, sb.BLOCK_DTYPE)
    blocks[:] = (0, 0, 1, 1, (1, 2, 3), 0)
    with pytest.raises(sb.StepthPanic):
        ctx.depth_from_blocks(same, blocks, 16, 16, P36)
    with pytest.raises(sb.StepthIOError):              # a block outside the image
        bad = np.zeros(8, sb.BLOCK_DTYPE)
        bad[:] = (10, 10, 10, 10, (1, 2, 3), 0)
        ctx.depth_from_blocks(same, bad, 16, 16, P36)
    with pytest.raises(sb.StepthPanic) as e:           # flat depth: depth_split underflows
        ctx.select_foreground(np.full((4, 4), 9, np.uint8))
    assert e.value.code == 5
    with pytest.raises(sb.StepthPanic):
        ctx.depth_split(np.array([5, 6], np.uint8), 2)
    with pytest.raises(sb.StepthIOError):
        ctx.apply_mask(np.zeros((4, 4, 4), np.uint8), np.zeros((3, 4), np.uint8))
    d = sb.DepthImage.from_image(np.zeros((4, 6, 3), np.uint8), ctx)
    with pytest.raises(sb.StepthIOError, match="Sizes don't match"):
        d.load_depth(np.zeros((6, 4), np.uint8))
    with pytest.raises(sb.StepthIOError, match="Failed to open image"):
        sb.DepthImage.open("/nonexistent/file.jpg")


# ---- tail, stand-alone ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("n", [1, 3, 4, 17, 1000, 4099, 65536 + 5])
def test_invert_slice_apply(ctx, orc, n):
    rng = np.random.default_rng(n)
    d = rng.integers(0, 256, n, dtype=np.uint8)
    assert np.array_equal(ctx.invert_depth(d), orc.invert_depth(d))
    for frm, to in [(None, None), (10, 200), (200, 10), (0, 0), (255, None), (None, 0)]:
        assert np.array_equal(ctx.slice(d, frm, to), orc.slice_mask(d, frm, to))
    rgba = rng.integers(0, 256, (n, 4), dtype=np.uint8)
    mask = rng.choice(np.array([0, 1, 255], np.uint8), n)
    assert np.array_equal(ctx.apply_mask(rgba, mask), orc.apply_mask(rgba, mask))


@pytest.mark.parametrize("zones", [1, 2, 3, 4, 7, 16, 100, 255])
def test_depth_split_vs_oracle(ctx, orc, zones, assets):
    rng = np.random.default_rng(zones)
    planes = [assets["depth"], orc.invert_depth(assets["depth"]),
              rng.integers(0, 256, 5000, dtype=np.uint8),
              np.clip(rng.normal(120, 30, 20000), 0, 255).astype(np.uint8),
              np.array([0] * 50 + [255] * 3 + [100] * 20, np.uint8),
              # sparse value sets: empty clusters, centres that collapse, neighbours one apart
              np.array([3] * 7 + [4] * 9 + [250] * 2, np.uint8),
              np.array([10, 12, 14, 200, 202, 255] * 5, np.uint8),
              rng.choice(np.array([0, 1, 2, 128, 254, 255], np.uint8), 3000),
              np.repeat(np.arange(0, 256, 17, dtype=np.uint8), 11)]
    import stepth_b200 as sb

    for d in planes:
        try:
            ref = orc.depth_split(d, zones)
        except orc.OracleError as e:
            with pytest.raises(sb.StepthPanic) as g:
                ctx.depth_split(d, zones)
            assert g.value.code == e.code
            continue
        assert ctx.depth_split(d, zones) == ref


def test_assets_through_the_public_api(ctx, orc, assets, tmp_path):
    """Readme.md:9-24 end to end with the reference's own images (decoded pixels from tests/golden)."""
    import stepth_b200 as sb

    main, add = assets["main"], assets["additional"]
    img = sb.DepthImage.from_image(main, ctx)
    img.load_depth_from_additional(add, [255 // 7] * 3, mode="dense")
    ref = orc.depth_dense(rgba_of(main), add, [36] * 3, 255, resample=orc.RESAMPLE_U8, threads=8)
    assert np.array_equal(img.depth_plane, ref)
    r = np.corrcoef(img.depth_plane.reshape(-1).astype(float), assets["depth"].reshape(-1).astype(float))[0, 1]
    assert r >= 0.97
    # block mode with the stand-in discretiser == oracle on the same block list
    blocks = sb.discretise_v0(img.image, [36] * 3)
    img2 = sb.DepthImage.from_image(main, ctx)
    img2.load_depth_from_additional(add, [36] * 3, blocks=blocks)
    assert np.array_equal(img2.depth_plane, orc.depth_from_blocks(add, blocks, [36] * 3, 600, 400, 255))
    # the tail on the reference's own depth map: the known answer (0, 201)
    img.load_depth(assets["depth"])
    img.invert_depth()
    assert img.depth_split(2) == [(0, 201), (202, 255)]
    fg = img.select_foreground()
    fg.apply_mask()
    mask_ref, ft = orc.select_foreground(orc.invert_depth(assets["depth"]))
    assert ft == (0, 201) and np.array_equal(fg.mask, mask_ref)
    assert np.array_equal(fg.image, orc.apply_mask(rgba_of(main), mask_ref))
    fg.save(str(tmp_path / "fg.png"))
    assert img.depth().size == (600, 400)


def test_batch_and_per_pair_status(ctx, orc):
    import stepth_b200 as sb

    pairs = [synth_pair("nat", 120, 80, seed=s) for s in (5, 2, 3)]  # seed 1 has max == 0 by itself
    same = np.full((80, 120, 3), 77, np.uint8)
    pairs.insert(1, (rgba_of(same), same))   # this pair has max == 0
    flags = sb.RESAMPLE_U8 | sb.FLAG_INVERT | sb.FLAG_FOREGROUND | sb.FLAG_APPLY
    out = ctx.depth_dense_batch([p[0] for p in pairs], [p[1] for p in pairs], P36, 128, flags)
    assert out["status"].tolist() == [0, 3, 0, 0]
    for i in (0, 2, 3):
        m, a = pairs[i]
        depth = orc.invert_depth(orc.depth_dense(m, a, P36, 128, resample=orc.RESAMPLE_U8, threads=8))
        mask, ft = orc.select_foreground(depth)
        assert np.array_equal(out["depth"][i], depth)
        assert tuple(out["from_to"][i]) == ft
        assert np.array_equal(out["mask"][i], mask)
        assert np.array_equal(out["rgba"][i], orc.apply_mask(m, mask))


@pytest.mark.parametrize("threads", ["4", "1", "0"])
def test_pageable_and_pinned_host_buffers_agree(orc, threads, monkeypatch):
    """Ordinary (malloc) host memory goes through the library's pinned ring and copy threads, page-locked memory is used in
    place: same bytes out, for every output, with images larger than a staged piece and more outputs than the ring holds."""
    import stepth_b200 as sb

    monkeypatch.setenv("STEPTH_STAGE_THREADS", threads)
    monkeypatch.setenv("STEPTH_BATCH_CHUNKS", "3")
    c = sb.Context(0)
    n, w, h, rings = 7, 1536, 1100, 40          # main RGBA 6.8 MB (two pieces), additional window pitched
    flags = sb.RESAMPLE_U8 | sb.FLAG_INVERT | sb.FLAG_FOREGROUND | sb.FLAG_APPLY
    pairs = [synth_pair("nat", w, h, seed=40 + i) for i in range(n)]
    pg = c.depth_dense_batch([p[0] for p in pairs], [p[1] for p in pairs], P36, rings, flags)
    pin_m = [torch.from_numpy(p[0]).pin_memory() for p in pairs]
    pin_a = [torch.from_numpy(p[1]).pin_memory() for p in pairs]
    out = {"depth": [torch.empty((h, w), dtype=torch.uint8).pin_memory() for _ in range(n)],
           "mask": [torch.empty((h, w), dtype=torch.uint8).pin_memory() for _ in range(n)],
           "rgba": [torch.empty((h, w, 4), dtype=torch.uint8).pin_memory() for _ in range(n)]}
    pn = c.depth_dense_batch(pin_m, pin_a, P36, rings, flags, out=out)
    assert pg["status"].tolist() == [0] * n and pn["status"].tolist() == [0] * n
    assert np.array_equal(pg["from_to"], pn["from_to"])
    for i in range(n):
        for k in ("depth", "mask", "rgba"):
            assert np.array_equal(pg[k][i], pn[k][i].numpy()), (k, i)
    m, a = pairs[n - 1]
    depth = orc.invert_depth(orc.depth_dense(m, a, P36, rings, resample=orc.RESAMPLE_U8, threads=8))
    mask, ft = orc.select_foreground(depth)
    assert np.array_equal(pg["depth"][n - 1], depth) and np.array_equal(pg["mask"][n - 1], mask)
    assert np.array_equal(pg["rgba"][n - 1], orc.apply_mask(m, mask))
    # the rings are reused by the next call; a shorter additional image (pitched window upload) through the same path
    again = c.depth_dense_batch([pairs[0][0], pairs[1][0]], [pairs[0][1][:-7], pairs[1][1][:-7]], P36, rings, flags)
    want = orc.invert_depth(orc.depth_dense(pairs[1][0], pairs[1][1][:-7], P36, rings, resample=orc.RESAMPLE_U8, threads=8))
    assert np.array_equal(again["depth"][1], want)


@pytest.mark.parametrize("w,h", [(232, 120), (233, 121), (1920, 1080)])
def test_rgb_main_entry_point_equals_rgba(ctx, orc, w, h):
    """depth_image.rs:105 holds the main image as RGB8: the *_rgb entry point uploads 3 bytes per pixel"""
    import stepth_b200 as sb

    n = 3
    pairs = [synth_pair("nat" if i % 2 else "geo", w, h, seed=90 + i) for i in range(n)]
    flags = sb.RESAMPLE_U8 | sb.FLAG_INVERT | sb.FLAG_FOREGROUND
    a4 = ctx.depth_dense_batch([p[0] for p in pairs], [p[1] for p in pairs], P36, 64, flags)
    a3 = ctx.depth_dense_batch([np.ascontiguousarray(p[0][..., :3]) for p in pairs], [p[1] for p in pairs], P36, 64, flags)
    assert a3["status"].tolist() == a4["status"].tolist() == [0] * n
    assert np.array_equal(a3["from_to"], a4["from_to"])
    for i in range(n):
        assert np.array_equal(a3["depth"][i], a4["depth"][i]) and np.array_equal(a3["mask"][i], a4["mask"][i])
    want = orc.invert_depth(orc.depth_dense(pairs[0][0], pairs[0][1], P36, 64, resample=orc.RESAMPLE_U8, threads=8))
    assert np.array_equal(a3["depth"][0], want)
    with pytest.raises(sb.StepthIOError):
        ctx.depth_dense_batch([np.ascontiguousarray(p[0][..., :3]) for p in pairs], [p[1] for p in pairs], P36, 64, flags | sb.FLAG_APPLY)


def test_apply_needs_both_rgba_buffers(ctx):
    """APPLY with a NULL rgba_in (block mode) or a NULL rgba_out is rejected, not silently dropped (C callers)"""
    import ctypes as C

    import stepth_b200 as sb
    from stepth_b200 import _ffi

    m, a = synth_pair("nat", 120, 80, seed=5)
    blocks = sb.discretise_v0(m, P36, min_splits=6)
    p = _ffi.make_params(P36, 32, sb.RESAMPLE_U8 | sb.FLAG_FOREGROUND | sb.FLAG_APPLY)
    depth, rgba, mask, ft = np.empty((80, 120), np.uint8), np.empty((80, 120, 4), np.uint8), np.empty((80, 120), np.uint8), np.zeros(2, np.uint8)
    lib = ctx._lib
    rc = lib.stepth_cuda_depth_from_blocks(ctx.handle, _ffi.ptr(a), 120, 80, blocks.ctypes.data, len(blocks), 120, 80, C.byref(p),
                                           _ffi.ptr(depth), None, None, _ffi.ptr(rgba), _ffi.ptr(mask), _ffi.ptr(ft))
    assert rc == _ffi.ERR_INVALID_ARG
    rc = lib.stepth_cuda_depth_from_blocks(ctx.handle, _ffi.ptr(a), 120, 80, blocks.ctypes.data, len(blocks), 120, 80, C.byref(p),
                                           _ffi.ptr(depth), None, _ffi.ptr(m), None, _ffi.ptr(mask), _ffi.ptr(ft))
    assert rc == _ffi.ERR_INVALID_ARG
    rc = lib.stepth_cuda_depth_dense(ctx.handle, _ffi.ptr(m), 120, 80, _ffi.ptr(a), 120, 80, C.byref(p), _ffi.ptr(depth), None, None,
                                     _ffi.ptr(mask), _ffi.ptr(ft))
    assert rc == _ffi.ERR_INVALID_ARG
    rc = lib.stepth_cuda_depth_dense(ctx.handle, _ffi.ptr(m), 120, 80, _ffi.ptr(a), 120, 80, C.byref(p), _ffi.ptr(depth), None,
                                     _ffi.ptr(rgba), _ffi.ptr(mask), _ffi.ptr(ft))
    assert rc == 0


def test_device_pipeline_matches_host_api(ctx, orc):
    import stepth_b200 as sb
    from stepth_b200 import device

    pairs = [synth_pair("geo", 100, 60, seed=s) for s in (4, 5)]
    flags = sb.RESAMPLE_U8 | sb.FLAG_INVERT | sb.FLAG_FOREGROUND | sb.FLAG_APPLY
    mains = torch.from_numpy(np.stack([p[0] for p in pairs])).cuda()
    adds = torch.from_numpy(np.stack([p[1] for p in pairs])).cuda()
    pipe = device.DensePipeline(ctx, 2, 100, 60, 100, 60, P36, 64, flags)
    pipe.run(mains, adds)
    torch.cuda.synchronize()
    for i, (m, a) in enumerate(pairs):
        host = ctx.depth_dense(m, a, P36, 64, flags)
        assert np.array_equal(pipe.depth[i].cpu().numpy(), host["depth"])
        assert np.array_equal(pipe.mask[i].cpu().numpy(), host["mask"])
        assert np.array_equal(pipe.rgba_out[i].cpu().numpy(), host["rgba"])
        assert tuple(pipe.ranges[i, :2].tolist()) == host["from_to"]


def test_stage_api_under_cuda_graph_capture(ctx, orc):
    """The stage API only enqueues work on the caller's stream and, once it has seen a shape, neither allocates nor synchronises:
    a caller may capture a whole step into a CUDA graph and replay it (bench.py: graph_replay).  Dense mode with the fused tail,
    replayed on fresh inputs of the same shape, and block mode with the device discretiser, against the host API and the oracle."""
    import stepth_b200 as sb
    from stepth_b200 import device

    w, h, rings = 100, 60, 64
    flags = sb.RESAMPLE_U8 | sb.FLAG_INVERT | sb.FLAG_FOREGROUND | sb.FLAG_APPLY
    first, second = synth_pair("geo", w, h, seed=4), synth_pair("geo", w, h, seed=5)
    mains = torch.from_numpy(first[0][None]).cuda().contiguous()
    adds = torch.from_numpy(first[1][None]).cuda().contiguous()
    pipe = device.DensePipeline(ctx, 1, w, h, w, h, P36, rings, flags)
    # block mode: its own (constant) inputs, the list rebuilt on the device inside the graph
    blocks = sb.discretise_v0(first[0], P36, min_splits=8)
    bmain, badd = mains[0].clone(), adds[0].clone()
    bpipe = device.BlockPipeline(ctx, len(blocks), w, h, w, h, P36, rings, sb.RESAMPLE_U8)
    blocks_t = torch.zeros((len(blocks), 20), dtype=torch.uint8, device="cuda")
    want_blocks = orc.depth_from_blocks(first[1], blocks, P36, w, h, rings)

    def step():
        pipe.run(mains, adds)
        bpipe.discretise(bmain, blocks_t, min_splits=8)
        bpipe.run(blocks_t, badd)

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        step()   # the first call of a shape sizes the library's scratch: outside the capture
    side.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side, capture_error_mode="thread_local"):
        step()
    torch.cuda.synchronize()
    for pair in (second, first):
        mains.copy_(torch.from_numpy(pair[0][None]))
        adds.copy_(torch.from_numpy(pair[1][None]))
        for t in (pipe.depth, pipe.mask, pipe.rgba_out, bpipe.depth, blocks_t):
            t.zero_()
        g.replay()
        torch.cuda.synchronize()
        host = ctx.depth_dense(pair[0], pair[1], P36, rings, flags)
        want = orc.invert_depth(orc.depth_dense(pair[0], pair[1], P36, rings, resample=orc.RESAMPLE_U8))
        assert np.array_equal(pipe.depth[0].cpu().numpy(), want)
        assert np.array_equal(pipe.depth[0].cpu().numpy(), host["depth"])
        assert np.array_equal(pipe.mask[0].cpu().numpy(), host["mask"])
        assert np.array_equal(pipe.rgba_out[0].cpu().numpy(), host["rgba"])
        assert int(bpipe.n_blocks[0]) == len(blocks) and blocks_t.cpu().numpy().tobytes() == blocks.tobytes()
        assert int(bpipe.status[0]) == 0
        assert np.array_equal(bpipe.depth[0].cpu().numpy(), want_blocks)


def test_row_bands_equal_whole_image(ctx, orc):
    """config 5's decomposition on one GPU: search + finish per row band (with halos) == the whole image"""
    from stepth_b200 import device

    m, a = synth_pair("nat", 96, 120, seed=21)
    rings = 20
    whole = orc.depth_dense(m, a, P36, rings, resample=orc.RESAMPLE_U8, threads=8)
    mt, at = torch.from_numpy(m).cuda(), torch.from_numpy(a).cuda()
    h = 120
    bands = [(0, 37), (37, 80), (80, 120)]
    d8 = torch.empty((1, h, 96), dtype=torch.uint8, device="cuda")
    mx_all = torch.zeros(1, dtype=torch.int32, device="cuda")
    for y0, y1 in bands:
        ay0 = max(0, (y0 + 1) // 2 - rings)
        ay1 = min(h, y1 // 2 + rings + 1)
        band_d8, mx = device.search_dense(ctx, mt[None, y0:y1].contiguous(), at[None, ay0:ay1].contiguous(), P36, rings,
                                          y0=y0, add_y0=ay0, add_img_h=h, img_h=h)
        d8[0, y0:y1] = band_d8[0]
        mx_all = torch.maximum(mx_all, mx)   # the 1-byte MAX all-reduce of the multi-GPU path
    out = torch.empty((h, 96), dtype=torch.uint8, device="cuda")
    for y0, y1 in bands:
        s0, s1 = max(0, y0 - 3), min(h, y1 + 3)   # 3-row halo for the vertical pass
        depth, status = device.finish(ctx, d8[:, s0:s1].contiguous(), mx_all, h, 1, src_y0=s0, out_y0=y0, out_rows=y1 - y0)
        out[y0:y1] = depth[0]
    assert np.array_equal(out.cpu().numpy(), whole)


def test_row_band_without_reachable_additional_rows(ctx, orc):
    """a band whose origins all lie below the additional image: every query is None -> 0, as in the unbanded call"""
    import stepth_b200 as sb
    from stepth_b200 import bands

    m, a = synth_pair("nat", 64, 80, seed=4)
    a = a[:10].copy()
    whole = gpu_raw(ctx, m, a, P36, 6)
    assert np.array_equal(whole, orc.search_dense(m, a, P36, 6, threads=4))
    plan = bands.band_plan(80, 10, 6, 2)
    assert plan[1].ay1 <= plan[1].ay0
    st = bands.DeviceStages(ctx)
    mt, at = torch.from_numpy(m).cuda(), torch.from_numpy(a).cuda()
    for b in plan:
        d8, mx = st.search(mt[None, b.y0:b.y1].contiguous(), at[None, b.ay0:b.ay1].contiguous(), P36, 6, 0, b.y0, b.ay0, 10, 80)
        assert np.array_equal(d8[0].cpu().numpy(), whole[b.y0:b.y1])
    assert (whole[plan[1].y0:] == 0).all()


def test_normalize_brightness_vs_oracle(ctx, orc):
    rng = np.random.default_rng(0)
    for n1, n2 in [(1000, 1000), (12345, 999), (7, 5000)]:
        a = rng.integers(0, 65536, n1, dtype=np.uint16)
        b = (rng.integers(0, 65536, n2) * 0.6).astype(np.uint16)
        for pct in (0.01, 0.9):
            assert np.array_equal(ctx.normalize_brightness_luma16(a, b, pct), orc.normalize_brightness_luma16(a, b, pct))
        a3 = rng.integers(0, 65536, (n1, 3), dtype=np.uint16)
        b3 = (rng.integers(0, 65536, (n2, 3)) * np.array([0.5, 1.0, 1.7])).clip(0, 65535).astype(np.uint16)
        for pct in (0.01, 0.9):
            assert np.array_equal(ctx.normalize_brightness_rgb16(a3, b3, pct), orc.normalize_brightness_rgb16(a3, b3, pct))


# ---- BASELINE.json's full sizes: size-independent properties + sampled oracle checks -----------------------------
@pytest.mark.parametrize("kind,w,h,rings", [("nat", 1920, 1080, 128), ("geo", 1920, 1080, 128), ("nat", 3840, 2160, 256),
                                            # 256 rings: coarse rings 16..32 take the per-lane window loop (S > 15)
                                            ("geo", 3840, 2160, 256), ("worst", 3840, 2160, 256)])
def test_full_size_production_equals_in_order_kernel(ctx, orc, kind, w, h, rings):
    import stepth_b200 as sb
    from stepth_b200 import device, synth

    m, a = synth.make_pair(kind, w, h, 1, device="cuda")
    d8, mx = device.search_dense(ctx, m[None], a[None], P36, rings)
    d8_lit, mx_lit = device.search_dense(ctx, m[None], a[None], P36, rings, flags=sb.FLAG_LITERAL_SEARCH)
    assert torch.equal(d8, d8_lit) and torch.equal(mx, mx_lit)
    assert int(mx[0]) == int(d8.max())
    # sampled oracle check on the same bytes
    rng = np.random.default_rng(1)
    idx = rng.choice(w * h, 3000 if kind == "nat" else 1600, replace=False).astype(np.uint32)
    if kind == "worst":   # the queries that can see the planted pixel at (w/4, h/4) are the only non-zero ones
        ys, xs = np.mgrid[h // 2 - 120:h // 2 + 120:11, w // 2 - 120:w // 2 + 120:11]
        idx = np.concatenate([idx, (ys * w + xs).reshape(-1).astype(np.uint32)])
    sub, _ = orc.search_dense_subset(m.cpu().numpy(), a.cpu().numpy(), P36, idx, rings, threads=8)
    assert np.array_equal(d8[0].reshape(-1).cpu().numpy()[idx], sub)
    assert (sub != 0).any()
    # the FINISHED plane, whole, against the oracle's normalise + unit-scale Gaussian of the same raw distances
    # (depth_image.rs:124-134): the streaming finish_vert / finish_horz path at 1920 / 3840 columns
    norm, mx_ref = orc.normalise(d8[0].cpu().numpy())
    assert mx_ref == int(mx[0])
    want_depth = orc.gaussian_resample_unit(norm, orc.RESAMPLE_U8)
    for mode_g, mode_o in ((sb.RESAMPLE_F32, orc.RESAMPLE_F32), (sb.RESAMPLE_OFF, orc.RESAMPLE_OFF)):
        got, st = device.finish(ctx, d8, mx, h, mode_g)
        assert int(st[0]) == 0 and np.array_equal(got[0].cpu().numpy(), orc.gaussian_resample_unit(norm, mode_o)), mode_g
    # the fused tail's invariants: invert is an involution, the mask is a pure function of depth and (from,to),
    # apply only ever clears, and clears exactly the masked-out pixels
    flags = sb.RESAMPLE_U8 | sb.FLAG_INVERT | sb.FLAG_FOREGROUND | sb.FLAG_APPLY
    pipe = device.DensePipeline(ctx, 1, w, h, w, h, P36, rings, flags)
    pipe.run(m[None], a[None])
    plain = device.DensePipeline(ctx, 1, w, h, w, h, P36, rings, sb.RESAMPLE_U8)
    plain.run(m[None], a[None])
    torch.cuda.synchronize()
    assert int(pipe.status[0]) == 0
    assert np.array_equal(plain.depth[0].cpu().numpy(), want_depth)
    assert torch.equal(pipe.depth, 255 - plain.depth)
    hist = torch.bincount(pipe.depth.reshape(-1).int(), minlength=256).int()
    assert torch.equal(hist, pipe.hist[0])
    frm, to = int(pipe.ranges[0, 0]), int(pipe.ranges[0, 1])
    assert orc.depth_split(pipe.depth[0].cpu().numpy(), 2)[0] == (frm, to)
    keep = (pipe.depth >= frm) & (pipe.depth <= to)
    assert torch.equal(pipe.mask, torch.where(keep, 255, 0).to(torch.uint8))
    assert torch.equal(pipe.rgba_out, m[None] * keep.unsqueeze(-1))


@pytest.mark.parametrize("n,kind", [(1, "nat"), (2, "geo"), (4, "nat"), (8, "nat"), (8, "geo")])
def test_launch_shapes_of_every_batch_size(ctx, orc, n, kind):
    """The search picks patches per CTA (1, 2, 4) and one or two far stages by the size of the launch: every choice
    against the in-order kernel on the same 1080p pairs, and sampled against the oracle."""
    import stepth_b200 as sb
    from stepth_b200 import device, synth

    w, h, rings = 1920, 1080, 128
    pairs = [synth.make_pair(kind, w, h, 20 + i, device="cuda") for i in range(n)]
    m = torch.stack([p[0] for p in pairs])
    a = torch.stack([p[1] for p in pairs])
    d8, mx = device.search_dense(ctx, m, a, P36, rings)
    d8_lit, mx_lit = device.search_dense(ctx, m, a, P36, rings, flags=sb.FLAG_LITERAL_SEARCH)
    assert torch.equal(d8, d8_lit) and torch.equal(mx, mx_lit)
    rng = np.random.default_rng(n)
    idx = rng.choice(w * h, 1500, replace=False).astype(np.uint32)
    i = n - 1
    sub, _ = orc.search_dense_subset(m[i].cpu().numpy(), a[i].cpu().numpy(), P36, idx, rings, threads=8)
    assert np.array_equal(d8[i].reshape(-1).cpu().numpy()[idx], sub)


def test_full_size_worst_case_sampled(ctx, orc):
    from stepth_b200 import device, synth

    w, h, rings = 1920, 1080, 128
    m, a = synth.make_pair("worst", w, h, 1, device="cuda")
    d8, mx = device.search_dense(ctx, m[None], a[None], P36, rings)
    rng = np.random.default_rng(2)
    idx = rng.choice(w * h, 1500, replace=False).astype(np.uint32)
    # make sure some sampled queries can see the planted pixel at (w/4, h/4): they are the only non-zero ones
    ys, xs = np.mgrid[h // 2 - 60:h // 2 + 60:7, w // 2 - 60:w // 2 + 60:7]
    idx = np.concatenate([idx, (ys * w + xs).reshape(-1).astype(np.uint32)])
    sub, _ = orc.search_dense_subset(m.cpu().numpy(), a.cpu().numpy(), P36, idx, rings, threads=8)
    got = d8[0].reshape(-1).cpu().numpy()[idx]
    assert np.array_equal(got, sub)
    assert (sub != 0).any() and int(mx[0]) > 0


def test_batch_of_4k_pairs_sliced_like_the_bench(ctx, orc):
    """configs[3]'s shape: 4K pairs, 256 rings, fused tail, more pairs than one launch sequence holds (the bench cuts a
    rank's pairs into sequences of 16): pairs of the first and of the second sequence against the oracle."""
    import stepth_b200 as sb
    from stepth_b200 import device, synth

    w, h, rings, n, sub = 3840, 2160, 256, 18, 16
    flags = sb.RESAMPLE_U8 | sb.FLAG_INVERT | sb.FLAG_FOREGROUND | sb.FLAG_APPLY
    pairs = [synth.make_pair("nat", w, h, 100 + i, device="cuda") for i in range(n)]
    m = torch.stack([p[0] for p in pairs])
    a = torch.stack([p[1] for p in pairs])
    del pairs
    pipes = {k: device.DensePipeline(ctx, k, w, h, w, h, P36, rings, flags) for k in (sub, n - sub)}
    rng = np.random.default_rng(18)
    for i0 in range(0, n, sub):
        k = min(sub, n - i0)
        pipe = pipes[k]
        pipe.run(m[i0:i0 + k], a[i0:i0 + k])
        torch.cuda.synchronize()
        assert int(pipe.status.abs().sum()) == 0
        for j in sorted({0, k - 1}):
            mi, ai = m[i0 + j].cpu().numpy(), a[i0 + j].cpu().numpy()
            idx = rng.choice(w * h, 1500, replace=False).astype(np.uint32)
            subset, _ = orc.search_dense_subset(mi, ai, P36, idx, rings, threads=8)
            raw = pipe.d8[j].cpu().numpy()
            assert np.array_equal(raw.reshape(-1)[idx], subset), (i0, j)
            norm, mx = orc.normalise(raw)
            assert mx == int(pipe.max[j])
            depth = orc.invert_depth(orc.gaussian_resample_unit(norm, orc.RESAMPLE_U8))
            assert np.array_equal(pipe.depth[j].cpu().numpy(), depth), (i0, j)
            mask, ft = orc.select_foreground(depth)
            assert tuple(pipe.ranges[j, :2].tolist()) == ft
            assert np.array_equal(pipe.mask[j].cpu().numpy(), mask)
            assert np.array_equal(pipe.rgba_out[j].cpu().numpy(), orc.apply_mask(mi, mask))


def test_giant_pair_in_row_bands_vs_oracle(ctx, orc):
    """configs[4]'s shape on ONE GPU: the 16384x8192 pair cut into the 8 row bands the 8-GPU run uses (add_y0 != 0, halo
    rows of the neighbours), band after band; raw distances sampled against the oracle on the whole image, the finished
    plane compared whole."""
    import stepth_b200 as sb
    from stepth_b200 import bands, device, synth

    w, h, rings = 16384, 8192, 256
    m, a = synth.make_pair("nat", w, h, 1, device="cuda")
    plan = bands.band_plan(h, h, rings, 8)
    d8 = torch.empty((1, h, w), dtype=torch.uint8, device="cuda")
    mx_all = torch.zeros(1, dtype=torch.int32, device="cuda")
    for b in plan:
        band_d8, mx = device.search_dense(ctx, m[None, b.y0:b.y1].contiguous(), a[None, b.ay0:b.ay1].contiguous(), P36, rings,
                                          y0=b.y0, add_y0=b.ay0, add_img_h=h, img_h=h)
        d8[0, b.y0:b.y1] = band_d8[0]
        mx_all = torch.maximum(mx_all, mx.to(torch.int32))
    rng = np.random.default_rng(5)
    idx = rng.choice(w * h, 4000, replace=False).astype(np.uint32)
    seam = np.concatenate([np.arange(b.y0 - 2, b.y0 + 2) for b in plan[1:]])           # rows either side of every seam
    idx = np.concatenate([idx, (seam[:, None].astype(np.int64) * w + rng.integers(0, w, (len(seam), 40))).reshape(-1).astype(np.uint32)])
    m_np, a_np = m.cpu().numpy(), a.cpu().numpy()
    del m, a
    subset, _ = orc.search_dense_subset(m_np, a_np, P36, idx, rings, threads=8)
    raw = d8[0].cpu().numpy()
    assert np.array_equal(raw.reshape(-1)[idx], subset)
    del m_np, a_np
    out = torch.empty((h, w), dtype=torch.uint8, device="cuda")
    for b in plan:
        s0, s1 = max(0, b.y0 - 3), min(h, b.y1 + 3)
        depth, status = device.finish(ctx, d8[:, s0:s1].contiguous(), mx_all, h, sb.RESAMPLE_U8, src_y0=s0, out_y0=b.y0,
                                      out_rows=b.y1 - b.y0)
        assert int(status[0]) == 0
        out[b.y0:b.y1] = depth[0]
    norm, mx_ref = orc.normalise(raw)
    assert mx_ref == int(mx_all[0])
    assert np.array_equal(out.cpu().numpy(), orc.gaussian_resample_unit(norm, orc.RESAMPLE_U8))


# ---- multi-GPU orchestration (stepth_b200/bands.py) on the CUDA stage API ---------------------------------------------
def test_band_runner_one_rank_equals_host_api(ctx, orc):
    import stepth_b200 as sb
    from stepth_b200 import bands

    m, a = synth_pair("nat", 120, 90, seed=33)
    flags = sb.RESAMPLE_U8 | sb.FLAG_INVERT | sb.FLAG_FOREGROUND | sb.FLAG_APPLY
    host = ctx.depth_dense(m, a, P36, 40, flags)
    runner = bands.BandRunner(bands.DeviceStages(ctx), 120, 90, 90, P36, 40, flags)
    out = runner.run(torch.from_numpy(m).cuda()[None], torch.from_numpy(a).cuda()[None])
    assert out["status"] == 0 and out["from_to"] == host["from_to"]
    assert np.array_equal(out["depth_band"][0].cpu().numpy(), host["depth"])
    assert np.array_equal(out["mask_band"][0].cpu().numpy(), host["mask"])
    assert np.array_equal(out["rgba_band"][0].cpu().numpy(), host["rgba"])


def test_multi_handle_equals_one_context(ctx, orc):
    """`stepth_multi` (one process, a context + host thread per device) over every GPU of the box: pair batches dealt to the
    devices and ONE pair in row bands, against the single-context calls.  With one GPU the same code runs on one device."""
    import stepth_b200 as sb

    devs = list(range(torch.cuda.device_count()))
    mu = sb.Multi(devs)
    flags = sb.RESAMPLE_U8 | sb.FLAG_INVERT | sb.FLAG_FOREGROUND | sb.FLAG_APPLY
    pairs = [synth_pair("nat" if i % 3 else "geo", 200, 120, seed=50 + i) for i in range(2 * len(devs) + 1)]
    got = mu.depth_dense_batch([p[0] for p in pairs], [p[1] for p in pairs], P36, 64, flags)
    want = ctx.depth_dense_batch([p[0] for p in pairs], [p[1] for p in pairs], P36, 64, flags)
    assert got["status"].tolist() == want["status"].tolist() and np.array_equal(got["from_to"], want["from_to"])
    for i in range(len(pairs)):
        for k in ("depth", "mask", "rgba"):
            assert np.array_equal(got[k][i], want[k][i]), (k, i)
    # one pair in row bands (each band >= 3 rows); a shorter additional image so that the last bands reach nothing
    for (w, h, ha, rings, seed) in [(320, 250, 250, 48, 9), (120, 96, 14, 8, 5)]:
        m, a = synth_pair("nat", w, h, seed=seed)
        a = a[:ha].copy()
        one = ctx.depth_dense(m, a, P36, rings, flags)
        banded = mu.depth_dense_banded(m, a, P36, rings, flags)
        assert banded["from_to"] == one["from_to"]
        for k in ("depth", "mask", "rgba"):
            assert np.array_equal(banded[k], one[k]), (k, w, h)
    same = np.full((40, 64, 3), 50, np.uint8)
    with pytest.raises(sb.StepthPanic) as e:      # max == 0 is a property of the whole image: every band reports it
        mu.depth_dense_banded(rgba_of(same), same, P36, 32, sb.RESAMPLE_U8)
    assert e.value.code == 3
    assert mu.launch_count() > 0
    mu.close()


def test_comm_band_runner_one_rank_equals_host_api(ctx, orc):
    """the band function below the boundary (stepth_cuda_dev_depth_dense_band) with a group of one"""
    import stepth_b200 as sb
    from stepth_b200 import bands

    m, a = synth_pair("nat", 120, 90, seed=33)
    flags = sb.RESAMPLE_U8 | sb.FLAG_INVERT | sb.FLAG_FOREGROUND | sb.FLAG_APPLY
    host = ctx.depth_dense(m, a, P36, 40, flags)
    runner = bands.CommBandRunner(ctx, 120, 90, 90, P36, 40, flags)
    b = runner.band   # one rank owns every main row; only the additional rows a query can reach belong to the band
    out = runner.run(torch.from_numpy(m).cuda()[None], torch.from_numpy(a[b.ay0:b.ay1]).cuda()[None].contiguous(), gather_depth=True)
    assert out["status"] == 0 and out["from_to"] == host["from_to"]
    for got, want in ((out["depth_band"], host["depth"]), (out["depth"], host["depth"]), (out["mask_band"], host["mask"]),
                      (out["rgba_band"], host["rgba"])):
        assert np.array_equal(got[0].cpu().numpy(), want)
    runner.close()


def _band_worker(rank, world, port, w, h, rings, q):
    import os

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    import torch.distributed as dist

    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        import stepth_b200 as sb
        from stepth_b200 import bands, synth

        c = sb.Context(rank)
        m, a = synth.make_pair("nat", w, h, 9)
        flags = sb.RESAMPLE_U8 | sb.FLAG_INVERT | sb.FLAG_FOREGROUND | sb.FLAG_APPLY
        runner = bands.BandRunner(bands.DeviceStages(c), w, h, h, P36, rings, flags)
        b = runner.band
        mb, ab = m[None, b.y0:b.y1].cuda().contiguous(), a[None, b.ay0:b.ay1].cuda().contiguous()
        out = runner.run(mb, ab, gather_depth=True)
        torch.cuda.synchronize()
        # the same through the C ABI's own communicator (csrc/multi_api.cuh: NCCL called by the library)
        def same_as_torch(res):
            return (res["status"] == out["status"] and res.get("from_to") == out.get("from_to")
                    and all(torch.equal(res[k], out[k]) for k in ("depth", "depth_band", "mask_band", "rgba_band")))

        crun = bands.CommBandRunner(c, w, h, h, P36, rings, flags)
        same = True
        for _ in range(3):   # three calls: the mailboxes and halo zones of the peer windows alternate with the call's parity
            same = same and same_as_torch(crun.run(mb, ab, gather_depth=True))
        peer_mode = crun.peer_mode()
        # a second image through the same group: other bytes in the same windows
        m2, a2 = synth.make_pair("nat", w, h, 10)
        one = sb.Context(rank).depth_dense(m2.numpy(), a2.numpy(), P36, rings, flags) if rank == 0 else None
        res2 = crun.run(m2[None, b.y0:b.y1].cuda().contiguous(), a2[None, b.ay0:b.ay1].cuda().contiguous(), gather_depth=True)
        if rank == 0:
            same = same and res2["status"] == 0 and res2["from_to"] == one["from_to"] and np.array_equal(res2["depth"][0].cpu().numpy(), one["depth"])
        crun.close()
        # the same exchanges through NCCL (what a box without peer access between its GPUs runs)
        os.environ["STEPTH_PEER"] = "0"
        c_nccl = sb.Context(rank)
        del os.environ["STEPTH_PEER"]
        nrun = bands.CommBandRunner(c_nccl, w, h, h, P36, rings, flags)
        same_nccl = same_as_torch(nrun.run(mb, ab, gather_depth=True)) and nrun.peer_mode() != 1
        nrun.close()
        q.put({"rank": rank, "band": (b.y0, b.y1), "status": out["status"], "from_to": out.get("from_to"),
               "depth": out["depth"][0].cpu().numpy(), "mask_band": out["mask_band"][0].cpu().numpy(),
               "rgba_band": out["rgba_band"][0].cpu().numpy(), "comm_runner_same": bool(same), "peer_mode": peer_mode,
               "nccl_runner_same": bool(same_nccl)})
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("w", [320, 324, 321])   # rows of whole 8-byte words, of whole 4-byte words (mirror stores word by word),
                                                  # odd (the tiled finish pass does not mirror: the band goes out by a copy kernel)
def test_row_bands_over_nccl_equal_one_gpu(ctx, orc, w):
    """configs[4]'s decomposition on every GPU of the box: bands + the normaliser's maximum + halo exchange + the gather of the
    depth bands -- by torch.distributed, by the library over peer windows (NVLink stores, csrc/peer.cuh) and by the library over
    NCCL -- against one GPU"""
    world = torch.cuda.device_count()
    if world < 2:
        pytest.skip("needs at least 2 GPUs")
    import socket

    import torch.multiprocessing as mp

    import stepth_b200 as sb

    h, rings = 250, 48
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mpc = mp.get_context("spawn")
    q = mpc.Queue()
    procs = [mpc.Process(target=_band_worker, args=(r, world, port, w, h, rings, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=600) for _ in range(world)]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    m, a = synth_pair("nat", w, h, seed=9)
    host = ctx.depth_dense(m, a, P36, rings, sb.RESAMPLE_U8 | sb.FLAG_INVERT | sb.FLAG_FOREGROUND | sb.FLAG_APPLY)
    for r in results:
        y0, y1 = r["band"]
        assert r["status"] == 0 and r["from_to"] == host["from_to"]
        assert r["comm_runner_same"], "stepth_cuda_dev_depth_dense_band disagrees with the torch.distributed orchestration"
        assert r["nccl_runner_same"], "the NCCL fallback of the band exchanges disagrees"
        assert r["peer_mode"] in (0, 1)
        assert np.array_equal(r["depth"], host["depth"])
        assert np.array_equal(r["mask_band"], host["mask"][y0:y1])
        assert np.array_equal(r["rgba_band"], host["rgba"][y0:y1])


def test_cpp_host_mirror(ctx, orc, tmp_path):
    """include/stepth.hpp (the C++ host side over the C ABI) against the oracle: tests/cpp/test_host.cpp"""
    import os
    import subprocess

    from stepth_b200 import _build

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "test_host")
    lib_dir, orc_dir = os.path.dirname(_build.LIB_PATH), os.path.join(root, "oracle")
    subprocess.run(["g++", "-std=c++17", "-O1", "-o", exe, os.path.join(root, "tests", "cpp", "test_host.cpp"),
                    f"-L{lib_dir}", "-lstepth_cuda", f"-L{orc_dir}", "-lstepth_oracle",
                    f"-Wl,-rpath,{lib_dir}", f"-Wl,-rpath,{orc_dir}", "-L/usr/local/cuda/lib64", "-lcudart"],
                   check=True, capture_output=True)
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "cpp host mirror ok" in r.stdout


# ---- editor operations ---------------------------------------------------------------------------
@pytest.mark.parametrize("n_side", [(1, 1), (7, 3), (64, 33), (401, 257)])
def test_editor_ops_vs_oracle(ctx, orc, n_side):
    import stepth_b200 as sb

    w, h = n_side
    rng = np.random.default_rng(w * 31 + h)
    rgba = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    depth = rng.integers(0, 256, (h, w), dtype=np.uint8)
    depth.reshape(-1)[: min(3, depth.size)] = [255, 0, 128][: min(3, depth.size)]
    mask = rng.choice(np.array([0, 255, 254, 1], np.uint8), (h, w))
    other = rng.choice(np.array([0, 255, 7], np.uint8), (h, w))
    assert np.array_equal(ctx.highlight_depth(rgba, depth), orc.highlight_depth(rgba, depth))
    assert np.array_equal(ctx.highlight_mask(rgba, mask), orc.highlight_mask(rgba, mask))
    for op in ("and", "or"):
        assert np.array_equal(ctx.mask_logic(op, mask, other), orc.mask_logic(op, mask, other))
    assert np.array_equal(ctx.mask_logic("not", mask), orc.mask_logic("not", mask))
    with pytest.raises(sb.StepthIOError):
        ctx.mask_logic("and", mask, np.zeros((h + 1, w), np.uint8))


def test_highlight_depth_every_value_pair(ctx, orc):
    # all 256 x 256 (channel value, depth) pairs: the f32 product / clamp / truncation must agree everywhere
    v, d = np.meshgrid(np.arange(256, dtype=np.uint8), np.arange(256, dtype=np.uint8))
    rgba = np.stack([v, v[::-1], v.T, np.full_like(v, 77)], axis=-1)
    assert np.array_equal(ctx.highlight_depth(rgba, d), orc.highlight_depth(rgba, d))
    assert np.array_equal(ctx.highlight_mask(rgba, d), orc.highlight_mask(rgba, d))


def test_image_replace_vs_oracle(ctx, orc):
    import stepth_b200 as sb

    rng = np.random.default_rng(3)
    for (w, h, ow, oh, sx, sy) in [(50, 40, 50, 40, 0, 0), (50, 40, 60, 45, 10, 5), (33, 17, 40, 30, 32, 16), (20, 20, 5, 5, 25, 0)]:
        rgba = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
        other = rng.integers(0, 256, (oh, ow, 4), dtype=np.uint8)
        mask = rng.choice(np.array([0, 255, 200], np.uint8), (h, w))
        assert np.array_equal(ctx.image_replace(rgba, mask, other, sx, sy), orc.image_replace(rgba, mask, other, sx, sy))
    # the reference's out-of-bounds panic: a masked pixel of the window at x >= other.width
    rgba = np.zeros((10, 10, 4), np.uint8)
    other = np.ones((4, 4, 4), np.uint8)
    mask = np.full((10, 10), 255, np.uint8)
    with pytest.raises(sb.StepthPanic) as e:
        ctx.image_replace(rgba, mask, other, 2, 2)
    assert e.value.code == 7
    with pytest.raises(orc.OracleError):
        orc.image_replace(rgba, mask, other, 2, 2)
    # through the public classes
    a, b = sb.MaskImage.from_image(rgba, ctx), sb.MaskImage.from_image(other, ctx)
    a.mask[:] = 0
    a.mask[:4, :4] = 255
    a.image_replace(b, (0, 0))
    assert a.image[:4, :4].min() == 1 and a.image[4:].max() == 0
    a.mask_not()
    assert a.mask[0, 0] == 0 and a.mask[9, 9] == 255


# ---- the stand-in discretiser on the device (csrc/disc.cuh) ------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize(
    "w,h,min_s,max_s",
    [(600, 400, 16, None), (96, 64, 16, None), (161, 77, 10, None), (33, 9, 4, 11), (1, 17, 16, None), (5, 1, 16, None),
     (2, 2, 16, None), (64, 64, 0, 7), (64, 64, 3, 3), (300, 200, 18, 18), (257, 3, 16, 40), (1920, 1080, 16, None)],
)
def test_device_discretiser_equals_host_stand_in(ctx, assets, w, h, min_s, max_s):
    """identical block list, order included, for the README image, synthetic images, ragged shapes and split bounds that force
    leaves above, at and below the top level"""
    import stepth_b200 as sb

    if (w, h) == (600, 400):
        imgs = [rgba_of(assets["main"]), np.ascontiguousarray(assets["main"])]           # RGBA and RGB strides
    else:
        rng = np.random.default_rng(w * 131 + h)
        imgs = [synth_pair("nat", w, h, seed=w + h)[0], synth_pair("geo", w, h, seed=3)[0],
                np.full((h, w, 3), 77, np.uint8), (rng.integers(0, 4, (h, w, 3)) * 60).astype(np.uint8)]
    for img in imgs:
        for prec in (P36, [1, 200, 17], [0, 36, 36]):
            want = sb.discretise_v0(img, prec, min_s, max_s)
            got = ctx.discretise_v0(img, prec, min_s, max_s)
            assert len(got) == len(want)
            assert got.tobytes() == want.tobytes()


@pytest.mark.gpu
def test_depth_from_additional_equals_host_discretiser_plus_blocks(ctx, orc, assets):
    """load_depth_from_additional in one call (the block list never leaves the device) == host stand-in + depth_from_blocks"""
    import stepth_b200 as sb

    main, add = rgba_of(assets["main"]), assets["additional"]
    blocks = sb.discretise_v0(main, P36)
    flags = sb.RESAMPLE_U8 | sb.FLAG_INVERT | sb.FLAG_FOREGROUND | sb.FLAG_APPLY
    want = ctx.depth_from_blocks(add, blocks, 600, 400, P36, 255, flags, rgba_in=main)
    got = ctx.depth_from_additional(main, add, P36, 255, flags)
    assert got["n_blocks"] == len(blocks) and got["from_to"] == want["from_to"]
    for k in ("depth", "mask", "rgba"):
        assert np.array_equal(got[k], want[k]), k
    plain = ctx.depth_from_additional(main, add, P36, 255, sb.RESAMPLE_U8)
    assert np.array_equal(plain["depth"], orc.depth_from_blocks(add, blocks, P36, 600, 400, 255, resample=orc.RESAMPLE_U8, threads=8))
    # the public API takes the same route when no block list is handed in
    img = sb.DepthImage.from_image(assets["main"], ctx)
    img.load_depth_from_additional(add, P36)
    assert np.array_equal(img.depth_plane, plain["depth"])
    # fewer than 8 blocks: par_chunks_mut(0) panics (depth_image.rs:111-112)
    with pytest.raises(sb.StepthPanic):
        ctx.depth_from_additional(np.full((1, 2, 4), 9, np.uint8), np.zeros((4, 4, 3), np.uint8), P36, 255, sb.RESAMPLE_U8)


# ---- decode / encode on the device (csrc/codec.cuh; SURVEY 8(f) row 4: NOT on the parity path) --------------------------------
@pytest.mark.gpu
def test_jpeg_decode_and_encode_on_the_device(ctx, assets, tmp_path):
    """nvJPEG against the host decoder (PIL / libjpeg): same size, pixels within a few levels (two JPEG decoders never agree bit for
    bit -- which is why the parity path keeps decoding on the host); encode -> host decode round trip close to the source"""
    import io

    from PIL import Image

    import stepth_b200 as sb

    main = np.ascontiguousarray(assets["main"])                       # 400 x 600 x 3, the README image's decoded pixels
    buf = io.BytesIO()
    Image.fromarray(main, "RGB").save(buf, format="JPEG", quality=92, subsampling=0)
    data = buf.getvalue()
    host = np.asarray(Image.open(io.BytesIO(data)).convert("RGB"), dtype=np.uint8)
    got = ctx.decode_jpeg(data)
    assert got.shape == host.shape == main.shape
    diff = np.abs(got.astype(np.int16) - host.astype(np.int16))
    assert diff.max() <= 6 and diff.mean() < 0.6, (int(diff.max()), float(diff.mean()))
    # 4:2:0 streams and grey streams decode too
    buf = io.BytesIO()
    Image.fromarray(main, "RGB").save(buf, format="JPEG", quality=85, subsampling=2)
    g420 = ctx.decode_jpeg(buf.getvalue())
    h420 = np.asarray(Image.open(io.BytesIO(buf.getvalue())).convert("RGB"), dtype=np.uint8)
    assert g420.shape == main.shape and np.abs(g420.astype(np.int16) - h420.astype(np.int16)).mean() < 2.0
    buf = io.BytesIO()
    Image.fromarray(assets["depth"], "L").save(buf, format="JPEG", quality=90)
    grey = ctx.decode_jpeg(buf.getvalue())
    hgrey = np.asarray(Image.open(io.BytesIO(buf.getvalue())).convert("L"), dtype=np.uint8)
    assert grey.shape == hgrey.shape + (3,) and np.abs(grey[..., 0].astype(np.int16) - hgrey.astype(np.int16)).max() <= 3
    with pytest.raises(sb.StepthIOError, match="Failed to open image"):
        ctx.decode_jpeg(b"this is not a jpeg stream at all")
    # encode: RGB, RGBA (alpha dropped), Luma -> a stream the host decoder reads back close to the source
    def psnr(a, b):
        mse = np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)
        return 99.0 if mse == 0 else 10 * np.log10(255.0 ** 2 / mse)

    for px, mode in ((main, "RGB"), (rgba_of(main), "RGB"), (np.ascontiguousarray(assets["depth"]), "L")):
        stream = ctx.encode_jpeg(px, quality=90)
        back = np.asarray(Image.open(io.BytesIO(stream)).convert(mode), dtype=np.uint8)
        want = px[..., :3] if px.ndim == 3 else px
        assert back.shape == want.shape and psnr(back, want) > 32.0, (mode, psnr(back, want))
    # the public API takes the device decoder when asked to
    path = tmp_path / "main.jpg"
    path.write_bytes(data)
    import os

    os.environ["STEPTH_GPU_DECODE"] = "1"
    try:
        img = sb.DepthImage.open(str(path), ctx)
    finally:
        del os.environ["STEPTH_GPU_DECODE"]
    assert img.image.shape == (400, 600, 4) and np.array_equal(img.image[..., :3], got)


# ---- the -DSTEPTH_DEBUG_CHECKS build (compute-sanitizer is closed on this pool) ----------------------------------------------
_DEBUG_CASE = r"""
import sys
sys.path.insert(0, %r)
import numpy as np
import stepth_b200 as sb
from stepth_b200 import synth
from oracle import oracle
ctx = sb.Context(0)
for (w, h, rings, kind, tile) in ((161, 77, 64, "nat", 2), (96, 200, 255, "geo", 2), (257, 33, 300, "nat", 2), (640, 360, 128, "nat", 0), (64, 64, 40, "worst", 2)):
    import os
    m, a = synth.make_pair(kind, w, h, 5)
    m, a = m.numpy(), a.numpy()
    got = ctx.depth_dense(m, a, [36] * 3, rings, sb.RESAMPLE_U8, want_raw=True)
    _, raw = oracle.depth_dense(m, a, [36] * 3, rings, resample=oracle.RESAMPLE_U8, threads=4, want_raw=True)
    assert np.array_equal(got["raw"], raw), (kind, w, h)
m, a = synth.make_pair("nat", 200, 120, 3)
m, a = m.numpy(), a.numpy()
blocks = sb.discretise_v0(m, [36] * 3, 10)
gb = ctx.depth_from_blocks(a, blocks, 200, 120, [36] * 3, 255, sb.RESAMPLE_U8, want_raw=True)
_, rb = oracle.depth_from_blocks(a, blocks, [36] * 3, 200, 120, 255, resample=oracle.RESAMPLE_U8, threads=4, want_raw=True)
assert np.array_equal(gb["raw"], rb)
print("debug-checks build ok")
"""


@pytest.mark.gpu
def test_search_kernels_index_checks_build(tmp_path):
    """the same library compiled with -DSTEPTH_DEBUG_CHECKS: every index the search kernels form (shared-memory lists, tile grid,
    outputs) is checked on the device and a failed check fails the call; a set of dense / block cases runs clean and bit-exact"""
    import os
    import subprocess
    import sys

    from stepth_b200 import _build

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lib = _build.build_debug()
    env = dict(os.environ, STEPTH_LIB=lib)
    for near in ("2", "0"):   # the origin-tile stage and the walking kernel
        env["STEPTH_NEAR_TILE"] = near
        r = subprocess.run([sys.executable, "-c", _DEBUG_CASE % root], env=env, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0 and "debug-checks build ok" in r.stdout, r.stdout + r.stderr
