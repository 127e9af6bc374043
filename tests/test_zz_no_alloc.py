"""stepth_cuda_dev_discretise_v0_noalloc: the device discretiser whose scratch is a file-scope device array (include/stepth_cuda.h;
no counterpart in the reference, added for callers that cannot call cudaMalloc).

Collected last on purpose: the file makes its own context and does not lean on the session
fixtures' state, so it passes alone in a fresh process and after every other test alike.  Through the C ABI, bit-exact (block
lists are integer work) against stepth_cuda_dev_discretise_v0 on the same inputs, two calls in a row on the shared array."""
import ctypes as C
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
P36 = [36, 36, 36]
ERR_INVALID_ARG = 1
MAX_TOP = 16                 # STEPTH_NOALLOC_DISC_MAX_TOP
MAX_PIXELS = 1 << 26         # STEPTH_NOALLOC_DISC_MAX_PIXELS
BLOCK_BYTES = 20             # sizeof(stepth_block)


def _images():
    from stepth_b200 import synth

    g = np.load(os.path.join(ROOT, "tests", "golden", "assets.npz"))
    readme = np.ascontiguousarray(g["main"])                                                  # 400 x 600 x 3
    rgba = np.ascontiguousarray(np.concatenate([readme, np.full(readme.shape[:2] + (1,), 255, np.uint8)], -1))
    ragged = np.ascontiguousarray(synth.make_pair("nat", 321, 203, 5)[0].numpy())                    # ragged shape, RGBA
    flat = np.full((64, 48, 3), 77, np.uint8)                                                 # 3072 pixels < 2^16 regions: early 1x1 leaves
    return [("readme_rgb", readme, 16, None), ("readme_rgba", rgba, 16, None), ("ragged", ragged, 16, None), ("flat_small", flat, 16, None),
            ("shallow", ragged, 3, 9)]


def _call(lib, name, ctx, img_dev, w, h, prec, min_s, max_s, blocks_dev, cap, n_dev):
    import torch

    from stepth_b200 import _ffi

    fn = getattr(lib, name)
    return fn(ctx.handle, C.c_void_p(torch.cuda.current_stream().cuda_stream), _ffi.ptr(img_dev), 4 if img_dev is None else int(img_dev.shape[2]), w, h, _ffi.ptr(prec),
              min_s, max_s, _ffi.ptr(blocks_dev), cap, _ffi.ptr(n_dev))


@pytest.mark.gpu
def test_noalloc_discretiser_equals_the_allocating_entry_point_twice_in_a_row():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import stepth_b200 as sb
    from stepth_b200 import _ffi
    from stepth_b200.context import _max_splits

    lib = _ffi.lib()
    assert hasattr(lib, "stepth_cuda_dev_discretise_v0_noalloc"), "libstepth_cuda.so lacks the entry point (stale build?)"
    ctx = sb.Context(0)
    dev = torch.device("cuda", 0)
    prec = np.asarray(P36, np.uint8)
    cases = []
    for name, img, min_s, max_s in _images():
        h, w = img.shape[:2]
        max_s = _max_splits(w, h) if max_s is None else max_s
        img_dev = torch.from_numpy(img).to(dev)
        n_dev = torch.zeros(1, dtype=torch.int32, device=dev)
        # the allocating entry point: count, then write
        ctx.check(_call(lib, "stepth_cuda_dev_discretise_v0", ctx, img_dev, w, h, prec, min_s, max_s, None, 0, n_dev))
        n = int(n_dev.item())
        assert n > 0
        want = torch.zeros((n, BLOCK_BYTES), dtype=torch.uint8, device=dev)
        ctx.check(_call(lib, "stepth_cuda_dev_discretise_v0", ctx, img_dev, w, h, prec, min_s, max_s, want, n, n_dev))
        assert int(n_dev.item()) == n
        host = sb.discretise_v0(img, P36, min_s, max_s)                                       # and the host stand-in behind both
        assert want.cpu().numpy().tobytes() == host.tobytes(), name
        got = [torch.full((n, BLOCK_BYTES), 0xEE, dtype=torch.uint8, device=dev) for _ in range(2)]
        n_got = [torch.full((1,), -1, dtype=torch.int32, device=dev) for _ in range(2)]
        cases.append((name, img_dev, w, h, min_s, max_s, n, want, got, n_got))
    torch.cuda.synchronize()

    # one call outside the measurement looks up the array's address and loads the kernels (no allocation either, but the driver's
    # lazy module loading is not this library's memory)
    name, img_dev, w, h, min_s, max_s, n, want, got, n_got = cases[0]
    ctx.check(_call(lib, "stepth_cuda_dev_discretise_v0_noalloc", ctx, img_dev, w, h, prec, min_s, max_s, got[0], n, n_got[0]))
    torch.cuda.synchronize()
    got[0].fill_(0xEE)
    n_got[0].fill_(-1)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(0)[0]

    # two calls in a row per image, images interleaved: every call finds the previous call's (another image's) counts in the array
    for rep in range(2):
        for name, img_dev, w, h, min_s, max_s, n, want, got, n_got in cases:
            ctx.check(_call(lib, "stepth_cuda_dev_discretise_v0_noalloc", ctx, img_dev, w, h, prec, min_s, max_s, got[rep], n, n_got[rep]))
    # and back to back on one image, no other work between them
    name, img_dev, w, h, min_s, max_s, n, want, got, n_got = cases[1]
    for rep in range(2):
        ctx.check(_call(lib, "stepth_cuda_dev_discretise_v0_noalloc", ctx, img_dev, w, h, prec, min_s, max_s, got[rep], n, n_got[rep]))
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info(0)[0]
    assert free_after >= free_before, f"device memory shrank by {free_before - free_after} bytes across the no-alloc calls"

    for name, img_dev, w, h, min_s, max_s, n, want, got, n_got in cases:
        for rep in range(2):
            assert int(n_got[rep].item()) == n, (name, rep)
            assert torch.equal(got[rep], want), (name, rep)                                   # bit-exact, order included

    # count only (no output buffer), then a short output buffer: at most `cap` blocks written, the count still whole
    name, img_dev, w, h, min_s, max_s, n, want, got, n_got = cases[2]
    n_only = torch.full((1,), -1, dtype=torch.int32, device=dev)
    ctx.check(_call(lib, "stepth_cuda_dev_discretise_v0_noalloc", ctx, img_dev, w, h, prec, min_s, max_s, None, 0, n_only))
    assert int(n_only.item()) == n
    short = torch.full((n, BLOCK_BYTES), 0xEE, dtype=torch.uint8, device=dev)
    ctx.check(_call(lib, "stepth_cuda_dev_discretise_v0_noalloc", ctx, img_dev, w, h, prec, min_s, max_s, short, n // 2, n_only))
    assert int(n_only.item()) == n
    assert torch.equal(short[: n // 2], want[: n // 2]) and bool((short[n // 2:] == 0xEE).all())

    # beyond the static scratch: a clear error, nothing launched
    launches = ctx.launch_count()
    rc = _call(lib, "stepth_cuda_dev_discretise_v0_noalloc", ctx, img_dev, w, h, prec, MAX_TOP + 1, MAX_TOP + 4, short, n, n_only)
    assert rc == ERR_INVALID_ARG and "static scratch" in ctx.last_error() and "2^16" in ctx.last_error()
    big_w, big_h = 8193, 8192                                                                 # > 2^26 pixels; rejected before the pixels are touched
    assert big_w * big_h > MAX_PIXELS
    rc = _call(lib, "stepth_cuda_dev_discretise_v0_noalloc", ctx, img_dev, big_w, big_h, prec, 16, 27, short, n, n_only)
    assert rc == ERR_INVALID_ARG and "limit of this entry point" in ctx.last_error()
    rc = _call(lib, "stepth_cuda_dev_discretise_v0_noalloc", ctx, None, w, h, prec, 16, 27, short, n, n_only)
    assert rc == ERR_INVALID_ARG
    assert ctx.launch_count() == launches
    torch.cuda.synchronize()
