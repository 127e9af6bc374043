"""The index math of the production search (search_hier.cuh: ring keys, the diagonal cut of a tile, the coarse-ring walk
and its stopping rule, the plane membership bits) compiled for the HOST and checked against the oracle.

TEST INFRASTRUCTURE: this is how the kernel logic is unit-tested in a container without a GPU; tests/host_emul/emul.cu is
never linked into the product library (the product has no CPU path).  The GPU parity tests are in test_gpu_parity.py.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from conftest import rgba_of

HERE = os.path.dirname(os.path.abspath(__file__))
EMUL_DIR = os.path.join(HERE, "host_emul")
LIB = os.path.join(EMUL_DIR, "libemul.so")
CSRC = os.path.join(os.path.dirname(HERE), "stepth_b200", "csrc")


@pytest.fixture(scope="module")
def emul():
    srcs = [os.path.join(EMUL_DIR, "emul.cu")] + [os.path.join(CSRC, f) for f in ("search_hier.cuh", "tiles.cuh", "common.cuh")]
    if not os.path.exists(LIB) or any(os.path.getmtime(s) > os.path.getmtime(LIB) for s in srcs):
        subprocess.run(["nvcc", "-O2", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC",
                        "-shared", "-o", LIB, srcs[0]], check=True, capture_output=True)
    L = C.CDLL(LIB)
    L.emul_check_tile_min_key.argtypes = [C.c_ulonglong, C.c_int, C.c_int, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    L.emul_search_dense.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint32, C.c_uint32,
                                    C.c_uint32, C.c_uint32, C.c_void_p, C.c_int, C.c_uint32, C.c_int, C.c_int, C.c_void_p]
    return L


def emul_dense(L, m, a, prec, rings, sat=0, near=2, mid=6, y0=0, add_y0=0, add_img_h=None):
    m = np.ascontiguousarray(m)
    a = np.ascontiguousarray(a)
    h, w = m.shape[:2]
    ha, wa = a.shape[:2]
    out = np.zeros((h, w), np.uint8)
    p = np.asarray(prec, np.uint8)
    rc = L.emul_search_dense(m.ctypes.data, w, h, y0, a.ctypes.data, wa, ha, add_y0, add_img_h or ha, p.ctypes.data, sat,
                             rings, near, mid, out.ctypes.data)
    assert rc == 0
    return out


def test_tile_min_key_equals_brute_force(emul):
    rng = np.random.default_rng(0)
    got, want = C.c_uint32(), C.c_uint32()
    for bx in range(-20, 14):
        for by in range(-20, 14):
            for t in range(8):
                m = 0
                for _ in range((1, 2, 4, 8, 32, 64, 3, 5)[t]):
                    m |= 1 << int(rng.integers(0, 64))
                if t == 7:
                    m = (1 << 64) - 1
                if m:
                    assert emul.emul_check_tile_min_key(m, bx, by, C.byref(got), C.byref(want)), (hex(m), bx, by)


@pytest.mark.parametrize("kind", ["nat", "geo", "worst"])
@pytest.mark.parametrize("w,h,rings", [(96, 64, 128), (161, 77, 255), (64, 200, 40)])
def test_walk_vs_oracle_synth(emul, orc, kind, w, h, rings):
    from stepth_b200 import synth

    m, a = synth.make_pair(kind, w, h, w + h)
    m, a = m.numpy(), a.numpy()
    ref = orc.search_dense(m, a, [36] * 3, rings, threads=8)
    assert np.array_equal(emul_dense(emul, m, a, [36] * 3, rings), ref)
    assert np.array_equal(emul_dense(emul, m, a, [36] * 3, rings, near=1, mid=0), ref)
    assert np.array_equal(emul_dense(emul, m, a, [36] * 3, rings, near=3, mid=4), ref)


@pytest.mark.parametrize(
    "w,h,wa,ha",
    [(8, 1, 8, 1), (1, 8, 1, 8), (3, 5, 3, 5), (17, 33, 17, 33), (129, 67, 129, 67), (40, 30, 7, 5), (40, 30, 90, 70),
     (33, 9, 4, 40), (2, 4, 300, 2), (257, 3, 129, 2)],
)
def test_walk_ragged_shapes(emul, orc, w, h, wa, ha):
    rng = np.random.default_rng(w * 1000 + h)
    for levels, prec, rings in [(4, [36, 36, 36], 255), (256, [90, 70, 50], 64), (16, [1, 200, 17], 300), (8, [255, 255, 255], 7)]:
        step = 256 // levels
        m = rgba_of((rng.integers(0, levels, (h, w, 3)) * step).astype(np.uint8))
        a = (rng.integers(0, levels, (ha, wa, 3)) * step).astype(np.uint8)
        ref = orc.search_dense(m, a, prec, rings, threads=4)
        assert np.array_equal(emul_dense(emul, m, a, prec, rings), ref), (levels, prec, rings)
        ref_sat = orc.search_dense(m, a, prec, rings, pixelopps=orc.PIXELOPPS_SAT_ALL, threads=4)
        assert np.array_equal(emul_dense(emul, m, a, prec, rings, sat=1), ref_sat), ("sat", levels, prec, rings)


def test_walk_u8_wrap_far_match(emul, orc):
    # one matching pixel far away: d = floor(sqrt(dx^2 + dy^2)) wraps mod 256 (depth_image.rs:121)
    w = h = 80
    m = rgba_of(np.full((h, w, 3), 10, np.uint8))
    a = np.full((400, 400, 3), 200, np.uint8)
    a[260, 262] = 12
    for rings in (255, 300, 511):
        ref = orc.search_dense(m, a, [36] * 3, rings, threads=4)
        assert np.array_equal(emul_dense(emul, m, a, [36] * 3, rings), ref)
    assert ref.max() > 0


def test_walk_row_bands(emul, orc):
    from stepth_b200 import synth

    m, a = synth.make_pair("nat", 96, 120, 21)
    m, a = m.numpy(), a.numpy()
    rings, h = 20, 120
    ref = orc.search_dense(m, a, [36] * 3, rings, threads=8)
    for y0, y1 in [(0, 37), (37, 80), (80, 120)]:
        ay0 = max(0, (y0 + 1) // 2 - rings)
        ay1 = min(h, y1 // 2 + rings + 1)
        got = emul_dense(emul, m[y0:y1], a[ay0:ay1], [36] * 3, rings, y0=y0, add_y0=ay0, add_img_h=h)
        assert np.array_equal(got, ref[y0:y1])
