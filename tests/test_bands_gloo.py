"""The N>1 host logic (stepth_b200/bands.py: band plan, MAX all-reduce, 3-row halo exchange, histogram all-reduce,
band all-gather, pair sharding) over `gloo` on the CPU.  The compute stages are a stand-in built on the oracle (the
product passes the CUDA stage API instead); the exchange code under test is the product's own."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT  # noqa: F401  (puts the repo root on sys.path)

P36 = [36, 36, 36]


class OracleStages:
    """Same four methods as bands.DeviceStages, computed by the oracle from the WHOLE pair (test double)."""

    def __init__(self, main, add, precision, max_rings):
        from oracle import oracle

        self.o = oracle
        self.d8_full = oracle.search_dense(main, add, precision, max_rings, threads=2)

    def search(self, main_band, add_band, precision, max_rings, flags, y0, add_y0, add_img_h, img_h):
        rows = main_band.shape[1]
        band = self.d8_full[y0:y0 + rows]
        return torch.from_numpy(band.copy())[None], torch.tensor([int(band.max()) if band.size else 0], dtype=torch.int32)

    def finish(self, src, mx, img_h, flags, src_y0, out_y0, out_rows, hist):
        import stepth_b200 as sb

        o = self.o
        n, rows, w = src.shape
        m = int(mx[0])
        if m == 0:
            return torch.zeros((1, out_rows, w), dtype=torch.uint8), torch.tensor([3], dtype=torch.int32)
        plane = np.zeros((img_h, w), np.uint8)   # rows outside [src_y0, src_y0+rows) are poison-free zeros: the halo must suffice
        plane[src_y0:src_y0 + rows] = (src[0].numpy().astype(np.uint32) * 255 // m).astype(np.uint8)
        mode = flags & 3
        full = o.gaussian_resample_unit(plane, mode) if mode else plane
        out = full[out_y0:out_y0 + out_rows]
        if flags & sb.FLAG_INVERT:
            out = 255 - out
        if hist is not None:
            hist += torch.from_numpy(np.bincount(out.reshape(-1), minlength=256).astype(np.int32))[None]
        return torch.from_numpy(np.ascontiguousarray(out))[None], torch.zeros(1, dtype=torch.int32)

    def depth_split(self, hist):
        vals = np.repeat(np.arange(256, dtype=np.uint8), hist[0].numpy())
        ranges = torch.zeros((1, 512), dtype=torch.uint8)
        try:
            cl = self.o.depth_split(vals, 2)
        except self.o.OracleError as e:
            return ranges, torch.zeros(1, dtype=torch.int32), torch.tensor([e.code], dtype=torch.int32)
        for i, (a, b) in enumerate(cl):
            ranges[0, 2 * i], ranges[0, 2 * i + 1] = a, b
        return ranges, torch.tensor([len(cl)], dtype=torch.int32), torch.zeros(1, dtype=torch.int32)

    def mask_apply(self, depth, from_to, rgba):
        d = depth[0].numpy()
        mask = self.o.slice_mask(d, int(from_to[0, 0]), int(from_to[0, 1]))
        out = None
        if rgba is not None:
            out = torch.from_numpy(self.o.apply_mask(rgba[0].numpy(), mask))[None]
        return torch.from_numpy(mask)[None], out


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, w, h, rings, seed, q, add_h=None):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import stepth_b200 as sb
        from stepth_b200 import bands, synth

        m, a = synth.make_pair("nat", w, h, seed)
        m, a = m.numpy(), a.numpy()[: add_h or h].copy()
        flags = sb.RESAMPLE_U8 | sb.FLAG_INVERT | sb.FLAG_FOREGROUND | sb.FLAG_APPLY
        runner = bands.BandRunner(OracleStages(m, a, P36, rings), w, h, a.shape[0], P36, rings, flags)
        b = runner.band
        mb = torch.from_numpy(m[b.y0:b.y1].copy())[None]
        ab = torch.from_numpy(a[b.ay0:b.ay1].copy())[None]
        out = runner.run(mb, ab, gather_depth=True)
        res = {"rank": rank, "band": (b.y0, b.y1), "status": out["status"], "depth": out["depth"][0].numpy(),
               "from_to": out.get("from_to")}
        if "mask_band" in out and out["mask_band"] is not None:
            res["mask_band"] = out["mask_band"][0].numpy()
            res["rgba_band"] = out["rgba_band"][0].numpy()
        q.put(res)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,w,h,rings,add_h", [(2, 64, 46, 12, None), (3, 40, 10, 6, None),
                                                   (2, 64, 48, 12, None),    # equal bands (direct gather)
                                                   (2, 64, 80, 6, 10)])      # rank 1 can reach no additional row at all
def test_row_bands_over_gloo_equal_the_whole_image(orc, world, w, h, rings, add_h):
    import stepth_b200 as sb  # noqa: F401
    from stepth_b200 import synth

    seed = 5
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, w, h, rings, seed, q, add_h)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    m, a = synth.make_pair("nat", w, h, seed)
    m, a = m.numpy(), a.numpy()[: add_h or h].copy()
    if add_h:
        from stepth_b200.bands import band_plan

        assert any(b.ay1 <= b.ay0 and b.y1 > b.y0 for b in band_plan(h, add_h, rings, world))
    whole = orc.invert_depth(orc.depth_dense(m, a, P36, rings, resample=orc.RESAMPLE_U8, threads=2))
    mask_ref, ft_ref = orc.select_foreground(whole)
    rgba_ref = orc.apply_mask(m, mask_ref)
    for r in results:
        assert r["status"] == 0
        assert np.array_equal(r["depth"], whole), f"rank {r['rank']}: assembled depth differs"
        assert r["from_to"] == tuple(ft_ref)
        y0, y1 = r["band"]
        if y1 > y0:
            assert np.array_equal(r["mask_band"], mask_ref[y0:y1])
            assert np.array_equal(r["rgba_band"], rgba_ref[y0:y1])


def test_band_plan_covers_every_reachable_row():
    from stepth_b200.bands import band_plan, shard_pairs

    for h, add_h, rings, world in [(8192, 8192, 256, 8), (1080, 1080, 128, 4), (46, 46, 12, 2), (10, 10, 6, 3), (5, 40, 3, 8),
                                   (2160, 1000, 256, 8)]:
        plan = band_plan(h, add_h, rings, world)
        assert [b.rank for b in plan] == list(range(world))
        assert plan[0].y0 == 0 and plan[-1].y1 == h
        for p, nxt in zip(plan, plan[1:]):
            assert p.y1 == nxt.y0
        for b in plan:
            for y in range(b.y0, b.y1):
                oy = (y + 1) // 2
                lo, hi = max(0, oy - (rings - 1)), min(add_h - 1, oy + rings - 1)
                if lo <= hi:
                    assert b.ay0 <= lo and hi < b.ay1, (h, add_h, rings, world, b, y)
    assert sorted(sum((shard_pairs(64, r, 8) for r in range(8)), [])) == list(range(64))
    assert [len(shard_pairs(10, r, 4)) for r in range(4)] == [3, 3, 2, 2]
