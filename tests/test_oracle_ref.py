"""Validates the oracle's restatement against the reference's OWN C code, compiled from
/root/reference into oracle/_ref/ by oracle/Makefile (ggml-quants.c: LUT map A, fp16 decode,
dequantize_row_iq2_s). Skipped when oracle/_ref was not built. CPU only."""
import ctypes as C

import numpy as np
import pytest


@pytest.fixture(scope="module")
def ggml(oracle):
    lib = oracle.ref_lib()
    if lib is None:
        pytest.skip("oracle/_ref/libggml_quants_ref.so not built (reference tree absent)")
    return lib


def _iq2s_blocks(d_bits, qs):
    # block_iq2_s {fp16 d; u8 qs[64]; u8 qh[8]; u8 scales[8]} = 82 bytes (ggml-quants.h:17-22)
    nb = len(d_bits)
    blk = np.zeros((nb, 82), np.uint8)
    blk[:, 0:2] = np.asarray(d_bits, np.uint16).view(np.uint8).reshape(nb, 2)
    blk[:, 2:66] = qs.reshape(nb, 64)
    return blk


def test_lut_and_layout_match_ggml(oracle, ggml):
    # crates/bitnet-ggml-ffi/csrc/ggml/src/ggml-quants.c:58-73 (qmap = {-2,-1,1,2}, LSB-first)
    nb = 6
    qs = oracle.fill_bytes(nb * 64, 2024)
    one = int(np.array([1.0], np.float16).view(np.uint16)[0])
    blk = _iq2s_blocks([one] * nb, qs)
    y = np.zeros(nb * 256, np.float32)
    ggml.bitnet_dequantize_row_iq2_s(blk.ctypes.data, y.ctypes.data, nb * 256)
    codes = np.concatenate([oracle.unpack_qk256_block(qs[i * 64:(i + 1) * 64]) for i in range(nb)])
    assert np.array_equal(y, oracle.MAP_A[codes].astype(np.float32))
    # a QK256 row dotted with x == ggml-dequantised row dotted with x (integer x -> exact)
    x = np.round(oracle.fill_normal(nb * 256, 5) * 30).clip(-127, 127).astype(np.float32)
    got = oracle.gemv_qk256(qs, x, 1, nb * 256, nb * 64, "scalar")[0]
    assert got == np.float32(np.dot(y.astype(np.float64), x.astype(np.float64)))


def test_f16_decode_matches_ggml(oracle, ggml):
    # ggml-quants.c:31-56 fp16_to_fp32 over every finite half, through dequantize_row_iq2_s
    bits = np.array([b for b in range(65536) if (b >> 10) & 0x1F != 31], np.uint16)
    qs = np.full((bits.size, 64), 0xAA, np.uint8)          # code 2 -> +1.0 -> y = d
    blk = _iq2s_blocks(bits, qs)
    y = np.zeros(bits.size * 256, np.float32)
    ggml.bitnet_dequantize_row_iq2_s(blk.ctypes.data, y.ctypes.data, bits.size * 256)
    d = y.reshape(-1, 256)[:, 0]
    mine = np.array([oracle.f16_to_f32(b) for b in bits], np.float32)
    assert np.array_equal(d.view(np.uint32), mine.view(np.uint32))


def test_bitnet32_block_matches_ggml_scaling(oracle, ggml):
    # a 256-element inline-f16 block (66 B) decoded by the oracle == ggml's d * qmap[q]
    # for scales inside the reference's clamp envelope [1e-3, 1e3] (quant/i2s.rs:15-16)
    qs = oracle.fill_bytes(64, 9)
    for val in (1.0, 0.02, 3.5, 0.001953125):
        hb = int(np.array([val], np.float16).view(np.uint16)[0])
        blk = _iq2s_blocks([hb], qs)
        y = np.zeros(256, np.float32)
        ggml.bitnet_dequantize_row_iq2_s(blk.ctypes.data, y.ctypes.data, 256)
        mine = oracle.i2s_dequant_block(qs, 256, hb)
        assert np.array_equal(mine, y)
