#!/usr/bin/env python
"""Writes tests/golden/linear_golden.npz: inputs and expected outputs of the hot path (per-token absmax int8 quantiser ->
packed ternary linear -> scaled f32) for a few small cases of both I2_S layouts, computed by the CPU oracle
(oracle/oracle.cpp: quantize_symmetric_i8 after crates/bitnet-kernels/src/cpu/quantize.rs:32-49, gemv_qk256 after
crates/bitnet-quantization/src/i2s_qk256.rs:346-372, the BitNet32-F16 block decode after
crates/bitnet-models/src/quant/i2s.rs:66-140) once it had passed tests/test_oracle_golden.py and tests/test_oracle_ref.py
in the container that holds the reference tree.

The reference is a Rust workspace and cannot be built or imported here (no rustc), so these are not outputs of the
reference binary: they freeze the pinned oracle, so that neither the oracle nor the CUDA path can drift unnoticed
(tests/test_golden_fixtures.py checks both sides against the file; nothing here reads /root/reference).

    python tests/golden/make_golden.py        # rewrites the fixture; commit the result"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

# (name, layout, rows, cols, M): decode, batched-decode and prefill regimes of every kernel behind linear_i8, a ragged
# row count, and a QK256 column count that is not a multiple of the 256-element block (the reference's tail case, :419-437)
CASES = [("qk256_m%d" % m, "qk256", 192, 256, m) for m in (1, 4, 12, 40, 72, 130)] + [
    ("qk256_ragged", "qk256", 33, 768, 9), ("qk256_tail", "qk256", 64, 300, 3)] + [
    ("bn32_m%d" % m, "bitnet32", 192, 256, m) for m in (1, 4, 12, 40, 72, 130)] + [("bn32_ragged", "bitnet32", 33, 320, 9)]


def compute(oracle, name, layout, rows, cols, M, seed):
    x = oracle.fill_normal(M * cols, 1000 + seed, 1.5).reshape(M, cols)
    x[0, 0] = 0.5 * np.abs(x[0]).max()     # a value that lands on a rounding boundary region
    if M > 2:
        x[2] = 0.0                          # an all-zero token: scale 0, outputs 0
    q, s = oracle.quantize_rows_i8(x)
    out = {"x": x, "q": q, "s": s}
    if layout == "qk256":
        stride = -(-cols // 256) * 64
        w = oracle.fill_bytes(rows * stride, 0x601D0000 + seed)
        acc, y = oracle.linear_qk256_i8(w, rows, cols, q, s)
        # the reference's own call shape: f32 activations straight into gemv_qk256 (i2s_qk256.rs:346-372), row by row
        y_f32 = np.stack([oracle.gemv_qk256(w, x[i], rows, cols, stride, "scalar") for i in range(M)])
        out.update(w=w, acc=acc, y=y, y_f32=y_f32)
    else:
        w = oracle.make_bitnet32(rows, cols, 0x601D0000 + seed)
        y = oracle.linear_bitnet32_i8(w, rows, cols, q, s)
        out.update(w=w, y=y, y_dequant=oracle.linear_bitnet32_dequant(w, rows, cols, q, s))
    return out


def main():
    from oracle import pyoracle
    pyoracle.build()
    arrays = {}
    for seed, (name, layout, rows, cols, M) in enumerate(CASES):
        for k, v in compute(pyoracle, name, layout, rows, cols, M, seed).items():
            arrays["%s.%s" % (name, k)] = np.ascontiguousarray(v)
    path = os.path.join(HERE, "linear_golden.npz")
    np.savez_compressed(path, **arrays)
    print("wrote", path, os.path.getsize(path), "bytes,", len(arrays), "arrays")


if __name__ == "__main__":
    main()
