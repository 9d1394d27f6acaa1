"""The drop-in boundary, checked without a GPU: the C-ABI library loads, exports every symbol
include/bitnet_b200.h declares, validates arguments with the reference's messages, and FAILS
LOUDLY (no CPU fallback) when no sm_100 device is present."""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _no_gpu():
    import torch
    return not torch.cuda.is_available()


def test_library_loads_and_exports_every_declared_symbol(b200):
    hdr = open(os.path.join(ROOT, "include", "bitnet_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(bitnet_(?:cpp|b200)_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    lib = b200.lib()
    for sym in sorted(declared):
        assert hasattr(lib, sym), "libbitnet_b200.so does not export " + sym
    from bitnet_rs_b200 import ffi
    assert declared == set(ffi.SYMBOLS)
    assert b"sm_100a" in lib.bitnet_b200_version()


def test_reference_bridge_symbols_match_bridge_rs(b200):
    # crates/bitnet-kernels/src/ffi/bridge.rs:17-39 binds exactly these six
    for sym in ("bitnet_cpp_init", "bitnet_cpp_cleanup", "bitnet_cpp_is_available",
                "bitnet_cpp_matmul_i2s", "bitnet_cpp_quantize", "bitnet_cpp_get_last_error"):
        assert hasattr(b200.lib(), sym)


def test_gemv_qk256_dimension_errors_match_reference(b200):
    # i2s_qk256.rs:690-742 -- validated before any device work, so this runs on CPU too
    with pytest.raises(ValueError, match="x length"):
        b200.gemv_qk256(np.zeros(64, np.uint8), np.ones(246, np.float32), np.zeros(1, np.float32), 1, 256, 64)
    with pytest.raises(ValueError, match="too short"):
        b200.gemv_qk256(np.zeros(64, np.uint8), np.ones(256, np.float32), np.zeros(2, np.float32), 2, 256, 64)
    with pytest.raises(ValueError, match="y_out length"):
        b200.gemv_qk256(np.zeros(128, np.uint8), np.ones(256, np.float32), np.zeros(1, np.float32), 2, 256, 64)
    with pytest.raises(ValueError, match="row bytes mismatch"):
        b200.gemv_qk256(np.zeros(128, np.uint8), np.ones(256, np.float32), np.zeros(1, np.float32), 1, 256, 128)


def test_quantize_argument_errors_match_reference(b200):
    lib = b200.lib()
    x = np.ones(32, np.float32)
    out = np.zeros(1, np.uint8)
    sc = np.zeros(1, np.float32)
    assert lib.bitnet_cpp_quantize(x.ctypes.data, 32, out.ctypes.data, 1, sc.ctypes.data, 1, 0) == -1
    assert b"Output buffer too small for I2_S" in lib.bitnet_cpp_get_last_error()
    assert lib.bitnet_cpp_quantize(x.ctypes.data, 32, out.ctypes.data, 1, sc.ctypes.data, 1, 7) == -1
    assert b"Invalid quantization type" in lib.bitnet_cpp_get_last_error()
    assert lib.bitnet_cpp_quantize(None, 32, out.ctypes.data, 1, sc.ctypes.data, 1, 0) == -1
    assert b"Null pointer" in lib.bitnet_cpp_get_last_error()
    assert lib.bitnet_cpp_matmul_i2s(x.ctypes.data, out.ctypes.data, sc.ctypes.data, 0, 1, 1) == -1
    assert b"Invalid matrix dimensions" in lib.bitnet_cpp_get_last_error()


def test_i2s_qk256_container(b200):
    # I2SQk256NoScale::new, i2s_qk256.rs:85-107, test :498-541
    exact = 512 * 256
    t = b200.I2SQk256NoScale(512, 1024, np.zeros(exact, np.uint8))
    assert t.row_stride_bytes == 256 and t.row_bytes(3).size == 256
    b200.I2SQk256NoScale(512, 1024, np.zeros(exact + 128, np.uint8))
    with pytest.raises(ValueError, match="data size mismatch"):
        b200.I2SQk256NoScale(512, 1024, np.zeros(exact + 129, np.uint8))


def test_no_cpu_fallback_without_a_device(b200):
    if not _no_gpu():
        pytest.skip("a GPU is present")
    lib = b200.lib()
    assert lib.bitnet_cpp_is_available() == 0
    assert lib.bitnet_cpp_init() != 0
    assert b"not available" in lib.bitnet_cpp_get_last_error()
    with pytest.raises(b200.KernelError):
        b200.B200Kernel()
    mgr = b200.KernelManager()
    assert mgr.list_available_providers() == []
    with pytest.raises(b200.KernelError, match="No available kernel provider"):
        mgr.select_best()
    y = np.zeros(1, np.float32)
    with pytest.raises(ValueError, match="no CUDA device"):
        b200.gemv_qk256(np.zeros(64, np.uint8), np.ones(256, np.float32), y, 1, 256, 64)
    a = np.ones(4, np.int8)
    bb = np.ones(4, np.uint8)
    c = np.zeros(4, np.float32)
    assert lib.bitnet_cpp_matmul_i2s(a.ctypes.data, bb.ctypes.data, c.ctypes.data, 2, 2, 2) == -1
    assert b"no CUDA device" in lib.bitnet_cpp_get_last_error()
    with pytest.raises(b200.B200Error, match="no CUDA device"):
        b200.Context(0, stream=None)


def test_product_never_imports_the_oracle():
    # the product path must not route through oracle/ (tests, smoke and bench's cpu_baseline only)
    pkg = os.path.join(ROOT, "bitnet-rs_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.lower() or f == "__init__.py" and "oracle" not in src, (dirpath, f)


def _build_c_caller(tmp_path):
    import subprocess
    exe = str(tmp_path / "abi_caller")
    libdir = os.path.join(ROOT, "bitnet-rs_b200")
    cmd = ["gcc", "-std=c11", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "c", "abi_caller.c"), "-o", exe, "-L", libdir, "-lbitnet_b200",
           "-Wl,-rpath," + libdir]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_header_compiles_as_c11_and_a_c_program_calls_through_it(b200, tmp_path):
    # the boundary is a C ABI: the header must be valid C (not only C++), and a C caller must link and run
    import subprocess
    exe = _build_c_caller(tmp_path)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, "abi_caller exit %d: %s %s" % (r.returncode, r.stdout, r.stderr)
    assert ("loud failure ok" in r.stdout) if _no_gpu() else ("matmul_i2s ok" in r.stdout)


def test_rust_bindings_name_every_header_symbol(b200):
    # bitnet-b200-sys ships as source (no Rust toolchain in the image): keep its extern block in step with the header
    hdr = open(os.path.join(ROOT, "include", "bitnet_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(bitnet_(?:cpp|b200)_[a-z0-9_]+)\s*\(", hdr))
    rs = open(os.path.join(ROOT, "bitnet-b200-sys", "src", "lib.rs")).read()
    bound = set(re.findall(r"pub fn (bitnet_(?:cpp|b200)_[a-z0-9_]+)\s*\(", rs))
    assert bound == declared, "missing in lib.rs: %s; not in the header: %s" % (sorted(declared - bound), sorted(bound - declared))
    for f in ("Cargo.toml", "build.rs", os.path.join("src", "b200.rs")):
        assert os.path.exists(os.path.join(ROOT, "bitnet-b200-sys", f))
    # the config struct is bound field for field
    fields = re.search(r"typedef struct \{(.*?)\} bitnet_b200_model_config", hdr, flags=re.S).group(1)
    names = re.findall(r"\b([a-z_]+)\s*[,;]", fields)
    for n in names:
        assert re.search(r"pub %s: c_" % n, rs), "model_config field %s not bound" % n
