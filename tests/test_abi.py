"""The C-ABI boundary without a GPU: the library builds for sm_100a, loads, exports every symbol the header
declares, fails loudly (no CPU fallback) when no device is present, and the host-side pieces work."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

from conftest import ROOT

HEADER = os.path.join(ROOT, "include", "stepth_cuda.h")
no_gpu = not torch.cuda.is_available()


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(stepth_(?:cuda|host)_\w+)\s*\(", src)))


def test_library_builds_and_exports_every_declared_symbol():
    from stepth_b200 import _build, _ffi

    if no_gpu:
        _build.build()          # stale-aware: nvcc cross-compiles for sm_100a here
    lib = _ffi.lib()
    syms = declared_symbols()
    assert len(syms) >= 24
    for name in syms:
        assert hasattr(lib, name), f"{name} declared in include/stepth_cuda.h but not exported"
    # and the Python binding covers exactly the header
    assert sorted(_ffi.SIGNATURES) == syms
    assert b"sm_100a" in lib.stepth_cuda_version()


def test_library_carries_sm100a_sass_only():
    import shutil
    import subprocess

    from stepth_b200 import _build, _ffi

    _ffi.lib()
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    out = subprocess.run([cuobjdump, "-lelf", _build.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


@pytest.mark.skipif(not no_gpu, reason="needs a box without a GPU")
def test_no_cpu_fallback_without_a_device():
    import stepth_b200
    from stepth_b200 import _ffi

    with pytest.raises(stepth_b200.StepthError) as e:
        stepth_b200.Context(0)
    assert e.value.code == _ffi.ERR_CUDA
    assert "no CPU fallback" in str(e.value)
    lib = _ffi.lib()
    # every compute entry point rejects a null context instead of computing anything on the host
    assert lib.stepth_cuda_invert_depth(None, None, 0) == _ffi.ERR_INVALID_ARG
    assert lib.stepth_cuda_launch_count(None) == 0


def test_product_package_never_imports_the_oracle():
    import stepth_b200  # noqa: F401

    pkg = os.path.join(ROOT, "stepth_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text and "stepth_oracle" not in text, f


def test_params_struct_layout():
    from stepth_b200 import _ffi

    assert C.sizeof(_ffi.Params) == 12
    assert C.sizeof(_ffi.DevShape) == 36
    assert _ffi.BLOCK_DTYPE.itemsize == 20
    p = _ffi.make_params([1, 2, 3], 128, _ffi.RESAMPLE_U8 | _ffi.FLAG_INVERT)
    assert bytes(p)[:4] == b"\x01\x02\x03\x00" and p.max_rings == 128
    with pytest.raises(ValueError):
        _ffi.make_params([1, 2], 5)


def test_discretise_v0_tiles_the_image():
    import stepth_b200

    rng = np.random.default_rng(0)
    h, w = 37, 53
    img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    img[:20, :30] = (90, 91, 92)   # a uniform region
    for stride_img in (img, np.concatenate([img, np.full((h, w, 1), 255, np.uint8)], -1)):
        blocks = stepth_b200.discretise_v0(stride_img, [36, 36, 36], min_splits=4, max_splits=11)
        cover = np.zeros((h, w), np.int32)
        for b in blocks:
            cover[b["y"]:b["y"] + b["h"], b["x"]:b["x"] + b["w"]] += 1
            region = img[b["y"]:b["y"] + b["h"], b["x"]:b["x"] + b["w"]].reshape(-1, 3).astype(np.int64)
            assert b["value"].tolist() == (region.sum(0) // len(region)).tolist()   # floor mean
        assert (cover == 1).all()
        assert len(blocks) >= 16   # forced splits down to depth 4
    # the reference's parameters: min 16 splits -> at least 2^16 blocks once the image has that many pixels
    big = rng.integers(0, 256, (300, 400, 3), dtype=np.uint8)
    blocks = stepth_b200.discretise_v0(big, [36, 36, 36])
    assert 65536 <= len(blocks) <= 300 * 400
    assert int((blocks["w"].astype(np.int64) * blocks["h"]).sum()) == 300 * 400
    # identical pixels everywhere: stops right at the forced depth
    flat = np.full((256, 256, 3), 7, np.uint8)
    assert len(stepth_b200.discretise_v0(flat, [36, 36, 36])) == 65536


# ---- the Rust FFI crate (source only: no Rust toolchain in the build image) against the header --------------------------------
_C2RUST = {"int32_t": "i32", "uint32_t": "u32", "uint64_t": "u64", "size_t": "usize", "double": "c_double", "float": "c_float",
           "uint8_t": "u8", "uint16_t": "u16", "char": "c_char", "void": "c_void"}


def _c_param_to_rust(p):
    """`const uint8_t *const *main_rgba` -> `*const *const u8`; `uint8_t id_out[128]` -> `*mut u8`; `uint32_t w` -> `u32`"""
    p = p.strip()
    arr = re.search(r"\[\d*\]$", p)
    if arr:
        p = p[:arr.start()]
    m = re.match(r"^(const\s+)?(\w+)\s*((?:\*\s*(?:const\s*)?)*)\s*(\w+)?$", p)
    assert m, p
    base_const, base, stars, _name = m.groups()
    ty = _C2RUST.get(base, base)   # struct names (stepth_ctx ..) are the same on both sides
    # pointers, innermost first: the first `*` points at the (const?) base; a `const` after a `*` qualifies THAT pointer,
    # i.e. the pointee of the next `*`
    levels = re.findall(r"\*\s*(const)?", stars)
    pointee_const = bool(base_const)
    for lv in levels:
        ty = ("*const " if pointee_const else "*mut ") + ty
        pointee_const = lv == "const"
    if arr:
        ty = ("*const " if base_const else "*mut ") + ty
    return ty


def _header_prototypes():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = "\n".join(l for l in src.splitlines() if not l.lstrip().startswith("#"))
    protos = {}
    for ret, name, params in re.findall(r"([\w\s\*]+?)\b(stepth_(?:cuda|host)_\w+)\s*\(([^)]*)\)\s*;", src):
        ret = ret.replace("extern \"C\" {", "").strip()
        ps = [] if params.strip() == "void" else [_c_param_to_rust(p) for p in params.split(",")]
        rret = None if ret == "void" else _c_param_to_rust(ret + " _")
        assert name not in protos, name
        protos[name] = (ps, rret)
    return protos


def _rust_prototypes():
    src = open(os.path.join(ROOT, "rust", "stepth-cuda-sys", "src", "lib.rs")).read()
    src = re.sub(r"//[^\n]*", "", src)
    block = src[src.index('extern "C" {'):]
    out = []
    for name, params, ret in re.findall(r"pub fn (\w+)\s*\(([^)]*)\)\s*(?:->\s*([^;]+))?;", block):
        ps = [re.sub(r"\s+", " ", p.split(":", 1)[1].strip()) for p in params.split(",") if p.strip()]
        out.append((name, ps, ret.strip() if ret else None))
    return out


def test_c_to_rust_type_mapping():
    assert _c_param_to_rust("const uint8_t *const *main_rgba") == "*const *const u8"
    assert _c_param_to_rust("uint8_t *const *depth_out") == "*const *mut u8"
    assert _c_param_to_rust("stepth_ctx **ctx_out") == "*mut *mut stepth_ctx"
    assert _c_param_to_rust("const uint8_t precision[3]") == "*const u8"
    assert _c_param_to_rust("float ms_out[4]") == "*mut c_float"
    assert _c_param_to_rust("void *stream") == "*mut c_void"
    assert _c_param_to_rust("uint32_t w") == "u32"


def test_rust_sys_crate_matches_the_header():
    """rust/stepth-cuda-sys/src/lib.rs declares every function of include/stepth_cuda.h exactly once, with the same number of
    parameters, the same parameter types and the same return type (the cgo / bindgen stand-in while no Rust toolchain exists)."""
    header = _header_prototypes()
    rust = _rust_prototypes()
    names = [n for n, _, _ in rust]
    assert len(names) == len(set(names)), sorted(n for n in set(names) if names.count(n) > 1)   # a duplicate does not compile
    assert sorted(names) == sorted(header), (sorted(set(header) - set(names)), sorted(set(names) - set(header)))
    for name, ps, ret in rust:
        hps, hret = header[name]
        assert ps == hps, (name, ps, hps)
        assert ret == hret, (name, ret, hret)
    # constants: every numeric #define of the header exists on the Rust side with the same value
    hsrc = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    rsrc = open(os.path.join(ROOT, "rust", "stepth-cuda-sys", "src", "lib.rs")).read()
    for name, val in re.findall(r"#define\s+(STEPTH_[A-Z0-9_]+)\s+(.+)", hsrc):
        if name == "STEPTH_CUDA_H":
            continue
        v = val.strip().replace("ull", "").replace("u", "")
        want = eval(v)   # `(1 << 2)`, `100`, `262400`
        m = re.search(rf"pub const {name}: \w+ = ([^;]+);", rsrc)
        assert m, f"{name} missing in lib.rs"
        assert eval(m.group(1)) == want, (name, m.group(1), want)
    # struct sizes on the Rust side follow the same field lists (repr(C))
    for struct, fields in (("stepth_dev_shape", 9), ("stepth_params", 4), ("stepth_block", 6)):
        body = re.search(rf"pub struct {struct} \{{(.*?)\}}", rsrc, flags=re.S).group(1)
        assert len(re.findall(r"pub \w+:", body)) == fields, struct


def test_headers_compile_as_plain_c_and_cpp(tmp_path):
    """include/stepth_cuda.h is a C header (C99, pedantic, no warnings) with the record sizes the bindings assume; include/stepth.hpp
    (the C++ host mirror) and its test driver parse as C++17.  The GPU suite links and runs the driver (test_cpp_host_mirror)."""
    import shutil
    import subprocess

    if not shutil.which("gcc") or not shutil.which("g++"):
        pytest.skip("no host compiler")
    inc = os.path.join(ROOT, "include")
    src = tmp_path / "abi.c"
    src.write_text('#include "stepth_cuda.h"\n'
                   "int main(void) { return sizeof(stepth_block) == 20 && sizeof(stepth_params) == 12 && "
                   "sizeof(stepth_dev_shape) == 36 ? 0 : 1; }\n")
    exe = tmp_path / "abi"
    subprocess.run(["gcc", "-std=c99", "-pedantic", "-Wall", "-Wextra", "-Werror", f"-I{inc}", str(src), "-o", str(exe)],
                   check=True, capture_output=True)
    assert subprocess.run([str(exe)]).returncode == 0
    r = subprocess.run(["g++", "-std=c++17", "-fsyntax-only", f"-I{inc}", f"-I{os.path.join(ROOT, 'oracle')}",
                        os.path.join(ROOT, "tests", "cpp", "test_host.cpp")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
