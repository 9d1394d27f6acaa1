"""Pins the CPU oracle against every known-answer vector the reference's own tests hold
for the I2_S hot path. Each test cites the reference test it restates;
paths are relative to the BitNet-rs tree. CPU only."""
import numpy as np
import pytest


def _lcg_vals(fn, n):
    return np.array([fn(i) for i in range(n)])


# ---------------------------------------------------------------- KernelProvider::matmul_i2s
def test_matmul_identity_2x2(oracle):
    # crates/bitnet-kernels/src/cpu/fallback.rs:305-318; x86.rs:811-827, 922-939
    a = np.array([1, 2, 3, 4], np.int8)
    b = np.array([1, 0, 0, 1], np.uint8)
    for v in ("fallback", "avx2", "avx512"):
        if v == "avx2" and not oracle.lib().oracle_has_avx2():
            continue
        if v == "avx512" and not oracle.lib().oracle_has_avx512():
            continue
        c = oracle.matmul_i2s(a, b, 2, 2, 2, variant=v)
        assert c.ravel().tolist() == [1.0, 2.0, 3.0, 4.0]


def test_matmul_dimension_validation(oracle):
    # fallback.rs:320-331
    with pytest.raises(oracle.OracleError, match="Matrix A dimension mismatch"):
        oracle.matmul_i2s(np.array([1, 2], np.int8), np.array([1, 0], np.uint8), 2, 2, 2)
    with pytest.raises(oracle.OracleError, match="Matrix C dimension mismatch"):
        oracle.matmul_i2s(np.zeros(4, np.int8), np.zeros(4, np.uint8), 2, 2, 2, c_len=3)


def test_avx512_equals_avx2_8x8(oracle):
    # x86.rs:831-856
    if not (oracle.lib().oracle_has_avx2() and oracle.lib().oracle_has_avx512()):
        pytest.skip("needs AVX2 + AVX-512")
    a = _lcg_vals(lambda i: i % 127, 64).astype(np.int8)
    b = _lcg_vals(lambda i: i % 255, 64).astype(np.uint8)
    c512 = oracle.matmul_i2s(a, b, 8, 8, 8, "avx512")
    c2 = oracle.matmul_i2s(a, b, 8, 8, 8, "avx2")
    assert np.array_equal(c512, c2)


@pytest.mark.parametrize("m,n,k", [(3, 5, 7), (9, 13, 15), (17, 19, 33), (31, 27, 65)])
def test_avx512_matches_fallback_non_aligned(oracle, m, n, k):
    # x86.rs:860-895
    if not oracle.lib().oracle_has_avx512():
        pytest.skip("needs AVX-512")
    a = _lcg_vals(lambda i: (i % 11) - 5, m * k).astype(np.int8)
    b = _lcg_vals(lambda i: (i * 3 + 7) % 251, k * n).astype(np.uint8)
    assert np.array_equal(oracle.matmul_i2s(a, b, m, n, k, "avx512"),
                          oracle.matmul_i2s(a, b, m, n, k, "fallback"))


@pytest.mark.parametrize("m,n,k", [(2, 2, 2), (8, 8, 8), (16, 16, 16), (7, 9, 11), (32, 32, 32),
                                   (33, 33, 33)])
def test_avx2_matches_fallback(oracle, m, n, k):
    # x86.rs:943-997 (tolerance 1e-6 there; integer data makes it exact)
    if not oracle.lib().oracle_has_avx2():
        pytest.skip("needs AVX2")
    a = _lcg_vals(lambda i: (i % 5) - 2, m * k).astype(np.int8)
    b = _lcg_vals(lambda i: i % 3, k * n).astype(np.uint8)
    c2 = oracle.matmul_i2s(a, b, m, n, k, "avx2")
    cf = oracle.matmul_i2s(a, b, m, n, k, "fallback")
    assert np.array_equal(c2, cf)
    assert np.array_equal(cf, (a.reshape(m, k).astype(np.int64) @ b.reshape(k, n).astype(np.int64)))


def test_gpu_validation_inputs(oracle):
    # gpu/validation.rs:215-260: a=(i%3)-1, b=i%2
    m, n, k = 64, 64, 64
    a = _lcg_vals(lambda i: (i % 3) - 1, m * k).astype(np.int8)
    b = _lcg_vals(lambda i: i % 2, k * n).astype(np.uint8)
    c = oracle.matmul_i2s(a, b, m, n, k, "fallback")
    assert np.array_equal(c, a.reshape(m, k).astype(np.int32) @ b.reshape(k, n).astype(np.int32))


# ---------------------------------------------------------------- KernelProvider::quantize
def test_quantize_i2s_basic(oracle):
    # fallback.rs:334-348
    x = [1.5, -1.0, 0.5, -0.5, 0.0, 2.0, -2.0, 0.1]
    out, scales = oracle.quantize(x, 0, out_len=2, scales_len=1)
    assert scales[0] > 0 and out.any()
    # hand-derived: absmax 2 -> scale 4/3; v/scale = 1.125,-.75,.375,-.375,0,1.5,-1.5,.075
    assert np.float32(scales[0]) == np.float32(2.0) / np.float32(1.5)
    codes = [(out[i // 4] >> (2 * (i % 4))) & 3 for i in range(8)]
    assert codes == [1, 3, 0, 0, 0, 1, 3, 0]


def test_quantize_buffer_validation(oracle):
    # fallback.rs:350-360
    with pytest.raises(oracle.OracleError, match="Output buffer too small for I2_S"):
        oracle.quantize(np.ones(32, np.float32), 0, out_len=1, scales_len=1)
    with pytest.raises(oracle.OracleError, match="Scales buffer too small"):
        oracle.quantize(np.ones(64, np.float32), 0, out_len=16, scales_len=1)


def test_quantize_i2s_edge_list(oracle):
    # tests/gpu_quantization_parity.rs:55-80 edge list; thresholds +-0.5 on v/scale, scale=absmax/1.5
    edge = np.array([0.0, 0.5, -0.5, 0.499, -0.499, 0.501, -0.501, 1.0, -1.0, 1.5, -1.5, 0.75,
                     -0.75, 0.1, -0.1, 1e-8], np.float32)
    out, scales = oracle.quantize(edge, 0)
    s = np.float32(1.5) / np.float32(1.5)
    assert scales[0] == s
    codes = np.array([(out[i // 4] >> (2 * (i % 4))) & 3 for i in range(edge.size)])
    nrm = edge / s
    want = np.where(nrm > 0.5, 1, np.where(nrm < -0.5, 3, 0))
    assert np.array_equal(codes, want)
    # OR semantics: caller zeroes; a pre-set output keeps its bits (fallback.rs:153)
    pre = np.full(4, 0xFF, np.uint8)
    out2, _ = oracle.quantize(edge, 0, out=pre)
    assert np.all(out2 == 0xFF)


def test_quantize_all_zero_block(oracle):
    out, scales = oracle.quantize(np.zeros(32, np.float32), 0)
    assert scales[0] == 1.0 and not out.any()


def test_quantize_tl1_tl2(oracle):
    x = np.linspace(-1.5, 1.5, 128).astype(np.float32)
    for qt, blk, lut in ((1, 64, [-1.0, -0.33, 0.33, 1.0]), (2, 128, [-1.2, -0.4, 0.4, 1.2])):
        out, scales = oracle.quantize(x, qt)
        assert scales.size == 128 // blk
        lut = np.array(lut, np.float32)
        for i in range(x.size):
            s = scales[i // blk]
            want = int(np.argmin(np.abs(np.float32(x[i] / s) - lut)))
            assert ((out[i // 4] >> (2 * (i % 4))) & 3) == want


# ---------------------------------------------------------------- int8 activation quantiser
def test_symmetric_i8_known_answers(oracle):
    # cpu/quantize.rs:184-257
    q, s = oracle.quantize_symmetric_i8([3.0])
    assert q.tolist() == [127] and s == np.float32(3.0) / np.float32(127.0)
    _, s = oracle.quantize_symmetric_i8([-10.0, 5.0, 10.0])
    assert s == np.float32(10.0) / np.float32(127.0)
    q, s = oracle.quantize_symmetric_i8(np.zeros(8))
    assert s == 0.0 and not q.any()
    x = np.array([-1.0, -0.5, 0.0, 0.5, 1.0], np.float32)
    for bits, tol in ((8, 0.01), (4, 0.15)):
        q, s = oracle.quantize_symmetric_i8(x, bits)
        assert np.max(np.abs(q.astype(np.float32) * s - x)) < tol
    q, s = oracle.quantize_symmetric_i8([-1.0, 0.0, 1.0], 2)
    assert (q.astype(np.float32) * s).tolist() == [-1.0, 0.0, 1.0]
    for bad in (1, 9):
        with pytest.raises(oracle.OracleError, match=r"bits must be in \[2, 8\]"):
            oracle.quantize_symmetric_i8([1.0], bad)


def test_symmetric_i8_is_multiply_by_reciprocal(oracle):
    # quantize.rs:41-46: q = round(v * (1/scale)), NOT round(v/scale); half away from zero
    rng = np.random.default_rng(0)
    x = rng.standard_normal(4096).astype(np.float32)
    q, s = oracle.quantize_symmetric_i8(x)
    inv = np.float32(1.0) / s
    t = (x * inv).astype(np.float32)
    want = np.clip(np.sign(t) * np.floor(np.abs(t) + np.float32(0.5)), -127, 127).astype(np.int8)
    assert np.array_equal(q, want)
    assert np.abs(q).max() == 127


# ---------------------------------------------------------------- QK256
def test_qk256_unpack_and_lut(oracle):
    # i2s_qk256.rs:378-397 (unpack smoke), :468-474 / :557-563 (LUT), :574-621 (golden block)
    qs = np.array([(0b11100100 + (i & 3)) & 0xFF for i in range(64)], np.uint8)
    codes = oracle.unpack_qk256_block(qs)
    assert codes.max() < 4 and codes[:4].tolist() == [0, 1, 2, 3]
    assert [float(oracle.lib().ref_code_to_f32(c)) for c in range(4)] == [-2.0, -1.0, 1.0, 2.0]
    qs = np.zeros(64, np.uint8)
    for i in range(64):
        b = i * 4
        qs[i] = (b % 4) | (((b + 1) % 4) << 2) | (((b + 2) % 4) << 4) | (((b + 3) % 4) << 6)
    codes = oracle.unpack_qk256_block(qs)
    assert np.array_equal(codes, np.arange(256) % 4)


def test_qk256_row_sums(oracle):
    # i2s_qk256.rs:399-416 (0xAA -> sum x), :419-437 (tail cols=300)
    row = np.full(128, 0xAA, np.uint8)
    x = (np.arange(512) * np.float32(0.01)).astype(np.float32)
    assert abs(oracle.gemv_qk256_row(row, x, 512) - x.sum(dtype=np.float32)) < 1e-3
    x = (np.arange(300) % 7).astype(np.float32)
    assert abs(oracle.gemv_qk256_row(row, x, 300) - x.sum()) < 1e-3


@pytest.mark.parametrize("variant", ["dispatch", "scalar", "avx2"])
def test_qk256_multi_row_and_tiny(oracle, variant):
    # i2s_qk256.rs:440-465 (0x55 -> -sum x), :633-683 (1x256 ones -> 256.0)
    if variant == "avx2" and not oracle.lib().oracle_has_avx2():
        pytest.skip("needs AVX2")
    qs = np.full(3 * 64, 0x55, np.uint8)
    x = np.arange(256, dtype=np.float32)
    y = oracle.gemv_qk256(qs, x, 3, 256, 64, variant)
    assert np.all(np.abs(y + x.sum()) < 1e-3)
    y = oracle.gemv_qk256(np.full(64, 0xAA, np.uint8), np.ones(256, np.float32), 1, 256, 64, variant)
    assert y[0] == 256.0


@pytest.mark.parametrize("code", [0, 1, 2, 3])
def test_qk256_integration_code_times_256(oracle, code):
    # crates/bitnet-models/tests/qk256_integration.rs:117-148: code c everywhere, x=1 -> LUT[c]*256
    byte = code | (code << 2) | (code << 4) | (code << 6)
    y = oracle.gemv_qk256(np.full(4 * 64, byte, np.uint8), np.ones(256, np.float32), 4, 256, 64)
    assert np.all(y == [-2.0, -1.0, 1.0, 2.0][code] * 256)


def test_qk256_dimension_errors(oracle):
    # i2s_qk256.rs:476-484, 690-742: messages contain "x length", "too short", "y_out length"
    with pytest.raises(oracle.OracleError, match="x length"):
        oracle.gemv_qk256(np.zeros(64, np.uint8), np.ones(246, np.float32), 1, 256, 64)
    with pytest.raises(oracle.OracleError, match="too short"):
        oracle.gemv_qk256(np.zeros(64, np.uint8), np.ones(256, np.float32), 2, 256, 64)
    with pytest.raises(oracle.OracleError, match="y_out length"):
        oracle.gemv_qk256(np.zeros(128, np.uint8), np.ones(256, np.float32), 2, 256, 64, y_len=1)
    with pytest.raises(oracle.OracleError, match="row bytes mismatch"):
        oracle.gemv_qk256(np.zeros(128, np.uint8), np.ones(256, np.float32), 1, 256, 128)


def test_qk256_size_tolerance(oracle):
    # i2s_qk256.rs:498-541: +-128 B accepted, +129 rejected
    exact = 512 * 256
    assert oracle.qk256_new_check(512, 1024, exact) == 256
    oracle.qk256_new_check(512, 1024, exact + 32)
    oracle.qk256_new_check(512, 1024, exact + 128)
    with pytest.raises(oracle.OracleError):
        oracle.qk256_new_check(512, 1024, exact + 129)
    with pytest.raises(oracle.OracleError):
        oracle.qk256_new_check(512, 1024, exact // 2)


def test_qk256_legacy_dispatch(oracle):
    # qk256_dispatch.rs:106-119: 256x256, 0x55, x=0.5 -> every y == -128.0.
    # The wrapper qk256_gemv (:20-58) ignores scales and calls gemv_qk256 (map A: code 1 -> -1).
    y = oracle.gemv_qk256(np.full(256 * 64, 0x55, np.uint8), np.full(256, 0.5, np.float32), 256, 256, 64)
    assert np.all(y == -128.0)
    # the bench-only map-D kernel (:60-99): code 1 -> 0
    y = oracle.qk256_gemv_scalar_legacy(2, 256, np.full(128, 0x55, np.uint8), np.ones(2, np.float32),
                                        np.full(256, 0.5, np.float32))
    assert np.all(y == 0.0)
    y = oracle.qk256_gemv_scalar_legacy(1, 256, np.full(64, 0xAA, np.uint8), np.full(1, 2.0, np.float32),
                                        np.ones(256, np.float32))
    assert y[0] == 512.0


@pytest.mark.parametrize("rows,cols,seed", [(1, 256, 1), (3, 512, 2), (5, 1024, 3), (2, 4096, 4),
                                            (4, 300, 5), (3, 777, 6), (2, 2560, 7)])
def test_qk256_avx2_vs_scalar_property(oracle, rows, cols, seed):
    # qk256_avx2_correctness.rs / i2s_qk256_avx2.rs:370-432: |d| <= 1e-4 abs-or-rel
    if not oracle.lib().oracle_has_avx2():
        pytest.skip("needs AVX2")
    stride = -(-cols // 256) * 64
    qs = oracle.fill_bytes(rows * stride, 0xB2000000 + seed)
    x = oracle.fill_normal(cols, 42 + seed)
    ys = oracle.gemv_qk256(qs, x, rows, cols, stride, "scalar")
    ya = oracle.gemv_qk256(qs, x, rows, cols, stride, "avx2")
    assert np.all((np.abs(ys - ya) <= 1e-4) | (np.abs(ys - ya) <= 1e-4 * np.abs(ys)))
    # integer-valued activations: both are exact and identical
    xi = np.round(x * 40).clip(-127, 127).astype(np.float32)
    assert np.array_equal(oracle.gemv_qk256(qs, xi, rows, cols, stride, "scalar"),
                          oracle.gemv_qk256(qs, xi, rows, cols, stride, "avx2"))


def test_qk256_row_equals_dequant_dot(oracle):
    # qk256_property_tests.rs: row == dequant . x
    cols = 768
    qs = oracle.fill_bytes(192, 99)
    x = oracle.fill_normal(cols, 7)
    codes = np.concatenate([oracle.unpack_qk256_block(qs[i * 64:(i + 1) * 64]) for i in range(3)])
    w = oracle.MAP_A[codes].astype(np.float64)
    assert abs(oracle.gemv_qk256_row(qs, x, cols) - float(w @ x.astype(np.float64))) < 1e-3


# ---------------------------------------------------------------- BitNet32-F16 decode
def _pack(codes):
    codes = np.asarray(codes, np.uint8)
    pad = (-codes.size) % 4
    c = np.concatenate([codes, np.zeros(pad, np.uint8)]).reshape(-1, 4)
    return (c[:, 0] | (c[:, 1] << 2) | (c[:, 2] << 4) | (c[:, 3] << 6)).astype(np.uint8)


def _f16(v):
    return int(np.array([v], np.float16).view(np.uint16)[0])


def test_i2s_lut_sym_k1(oracle):
    # quant/i2s.rs:922-945
    dst = oracle.i2s_dequant_block(_pack([0, 1, 2, 3]), 4, _f16(1.0))
    assert dst.tolist() == [-2.0, -1.0, 1.0, 2.0]


def test_i2s_code_boundary_values(oracle):
    # quant/i2s.rs:1162-1192
    for code in range(4):
        dst = oracle.i2s_dequant_block(_pack([code] * 4), 4, _f16(1.0))
        assert np.all(dst == [-2.0, -1.0, 1.0, 2.0][code])


def test_i2s_extreme_scales(oracle):
    # quant/i2s.rs:1007-1057: zero -> clamp 1e-3; inf -> clamp 1e3; tiny, 65504 finite
    q = _pack([0, 1, 2, 3])
    dst = oracle.i2s_dequant_block(q, 4, _f16(0.0))
    assert np.array_equal(dst, np.float32(1e-3) * np.array([-2, -1, 1, 2], np.float32))
    dst = oracle.i2s_dequant_block(q, 4, 0x7C00)
    assert np.array_equal(dst, np.float32(1e3) * np.array([-2, -1, 1, 2], np.float32))
    for bits in (0x0001, _f16(65504.0), 0x8400):
        assert np.all(np.isfinite(oracle.i2s_dequant_block(q, 4, bits)))
    # negative scale -> abs (quant/i2s.rs:78)
    assert np.array_equal(oracle.i2s_dequant_block(q, 4, _f16(-0.5)),
                          oracle.i2s_dequant_block(q, 4, _f16(0.5)))
    # inv=true: tiny -> 1.0, else 1/s (quant/i2s.rs:80-87)
    assert oracle.i2s_dequant_block(q, 4, _f16(4.0), inv=True).tolist() == [-0.5, -0.25, 0.25, 0.5]
    assert oracle.i2s_dequant_block(q, 4, _f16(0.0), inv=True).tolist() == [-2.0, -1.0, 1.0, 2.0]


def test_i2s_block_sizes_and_infer(oracle):
    # quant/i2s.rs:1085-1107, 1136-1160
    assert oracle.i2s_infer_block_size(66, 1, 256) == 256
    assert oracle.i2s_infer_block_size(34, 1, 128) == 128
    assert oracle.i2s_infer_block_size(1, 1, 256) is None
    assert oracle.i2s_expected_bytes(1, 1024 * 1024, 256) == 4096 * 66
    for (rows, cols), block in (((1, 256), 256), ((1, 128), 128), ((1, 64), 64), ((2, 32), 32)):
        nb = rows * (cols // block)
        data = np.concatenate([np.concatenate([_pack(np.arange(block) % 4),
                                               np.array([_f16(2.0)], np.uint16).view(np.uint8)])
                               for _ in range(nb)])
        w = oracle.dequantize_i2s(data, rows, cols, block=block)
        assert np.array_equal(w.ravel(), np.tile(2.0 * oracle.MAP_A[np.arange(block) % 4], nb))
    with pytest.raises(oracle.OracleError, match="internal size mismatch"):
        oracle.dequantize_i2s(np.zeros(11, np.uint8), 1, 32, block=32)


def test_i2s_small_block8_with_tail(oracle):
    # quant/i2s.rs:948-1004 geometry: rows=3, cols=10, block=8 (tail block of 2 is read compactly)
    rows, cols, block = 3, 10, 8
    per_q = 2
    sb = np.array([_f16(2.0)], np.uint16).view(np.uint8)
    blk = np.concatenate([_pack([0, 1, 2, 3, 0, 1, 2, 3]), sb])
    data = np.tile(blk, rows * 2)
    assert data.size == rows * 2 * (per_q + 2)
    w = oracle.dequantize_i2s(data, rows, cols, block=block)
    assert w.shape == (3, 10) and np.all(np.isfinite(w))
    assert w[0, :8].tolist() == [-4.0, -2.0, 2.0, 4.0] * 2


def test_bitnet32_production_layout(oracle):
    # production value: block 32 -> 10 B/block, blocks along K within each row
    rows, cols = 4, 128
    data = oracle.make_bitnet32(rows, cols, 123)
    assert data.size == rows * (cols // 32) * 10
    assert oracle.i2s_infer_block_size(data.size, rows, cols) == 32
    w = oracle.dequantize_i2s(data, rows, cols)
    blk = data.reshape(-1, 10)
    for b in range(blk.shape[0]):
        s = np.float32(abs(np.float32(blk[b, 8:10].copy().view(np.float16)[0])))
        s = np.clip(s, np.float32(1e-3), np.float32(1e3))
        codes = np.array([(blk[b, e // 4] >> (2 * (e % 4))) & 3 for e in range(32)])
        assert np.array_equal(w.ravel()[b * 32:(b + 1) * 32], s * oracle.MAP_A[codes].astype(np.float32))


def test_f16_conversion_exhaustive(oracle):
    bits = np.arange(65536, dtype=np.uint16)
    want = bits.view(np.float16).astype(np.float32)
    got = np.array([oracle.f16_to_f32(b) for b in bits], np.float32)
    ok = (got == want) | (np.isnan(got) & np.isnan(want))
    assert ok.all()
    rng = np.random.default_rng(1)
    xs = np.concatenate([rng.standard_normal(2000).astype(np.float32) * 10,
                         np.array([0.0, -0.0, 65504.0, 65520.0, 1e-8, 6e-8, 5.96e-8, 1e10], np.float32)])
    with np.errstate(over="ignore"):
        for v in xs:
            assert oracle.f32_to_f16(v) == int(np.array([v], np.float32).astype(np.float16).view(np.uint16)[0])


# ---------------------------------------------------------------- layout detection
def test_tolerance_doc_tests(oracle):
    # crates/bitnet-quantization/src/lib.rs:78-84
    assert oracle.qk256_tolerance_bytes(1_000_000) == 1000
    assert oracle.qk256_tolerance_bytes(131_072) == 132
    assert oracle.qk256_tolerance_bytes(100_000) == 100
    assert oracle.qk256_tolerance_bytes(1_000) == 8
    assert oracle.qk256_tolerance_bytes(20) == 8


def test_detect_flavor_reference_vectors(oracle):
    # crates/bitnet-models/tests/i2s_flavor_detection.rs (nelems, bytes, sibling -> flavour)
    cases = [
        (64, 20, False, "BitNet32F16"),            # :27-34
        (64, 16, True, "Split32WithSibling"),      # :42-49
        (160, 40, False, "Split32WithSibling"),    # :63-69 (no sibling, warn)
        (512, 128, False, "GgmlQk256NoScale"),     # :105-112, :198-204
        (64, 24, False, "BitNet32F16"),            # :120-125 (20 + 4 padding)
        (2048, 640, False, "BitNet32F16"),         # :133-138
        (2080, 520, True, "Split32WithSibling"),   # :150-155
        (100, 40, False, "BitNet32F16"),           # :162-167
        (96, 50, False, None),                     # :177-187
        (32, 10, False, "BitNet32F16"),            # :232-237
        (800, 256, False, "GgmlQk256NoScale"),     # :255-260
        (64, 28, False, "BitNet32F16"),            # :268-274 (tolerance boundary)
        (96, 39, False, None),                     # :282-287 (beyond tolerance)
    ]
    for nelems, avail, sib, want in cases:
        assert oracle.detect_i2s_flavor(avail, sib, nelems) == want, (nelems, avail, sib)


def test_detect_flavor_production_shapes(oracle):
    # formats/gguf/types.rs:868-1066 on the 2B-4T gate/up tensor
    n = 2560 * 6912
    assert oracle.detect_i2s_flavor(n // 256 * 64, False, n) == "GgmlQk256NoScale"
    assert oracle.detect_i2s_flavor(n // 32 * 10, False, n) == "BitNet32F16"
    # split_need == qk256_need when nelems % 256 == 0: exact match prefers QK256 even with a sibling
    assert n // 32 * 8 == n // 256 * 64
    assert oracle.detect_i2s_flavor(n // 32 * 8, True, n) == "GgmlQk256NoScale"
    assert oracle.detect_i2s_flavor(n // 256 * 64 + 32, False, n) == "GgmlQk256NoScale"
    assert oracle.detect_i2s_flavor(n // 32 * 10 + 64, False, n) == "BitNet32F16"
    assert oracle.detect_i2s_flavor(n // 32 * 10 + 64, False, n, strict=True) is None
    assert oracle.detect_i2s_flavor(12345, False, n) is None


# ---------------------------------------------------------------- packed + scaled matmul (map C)
def _pack_weight_matrix(w, k, n):
    # cpu/quantized_matmul.rs:266-292 (column-major packing, code 1 -> 0b01, -1 -> 0b11)
    packed_k = -(-k // 4)
    packed = np.zeros(packed_k * n, np.uint8)
    for col in range(n):
        for row in range(k):
            v = w[row * n + col]
            code = 1 if v == 1 else (3 if v == -1 else 0)
            packed[col * packed_k + row // 4] |= code << ((row % 4) * 2)
    return packed


def _run_all(oracle, act, packed, scales, m, n, k, bs, expected, tol):
    for v in ("f32", "blocked", "dequant"):
        out = oracle.i2s_matmul(act, packed, scales, m, n, k, bs, v)
        assert np.max(np.abs(out.ravel() - np.asarray(expected, np.float32).ravel())) <= tol, v


def test_i2s_matmul_snapshots(oracle):
    # tests/snapshot_kernel_outputs.rs:152-177 -> [10,10] and [6]
    p = oracle.pack_i2s([1, 1, 1, 1])
    out = oracle.i2s_matmul([1, 2, 3, 4], [p, p], [1.0, 1.0], 1, 2, 4, 4)
    assert out.ravel().tolist() == [10.0, 10.0]
    out = oracle.i2s_matmul([1, 2, 3, 4], [oracle.pack_i2s([1, -1, 0, 1])], [2.0], 1, 1, 4, 4)
    assert out.ravel().tolist() == [6.0]
    # tests/snapshot_wave6.rs:269 -> 0x4d
    assert oracle.pack_i2s([1, -1, 0, 1]) == 0x4D
    # tests/snapshot_wave6.rs:230-262 -> 3,3,1,-3
    act = [1.0, 2.0, 3.0, 4.0, 0.5, -1.0, 1.5, -0.5]
    w = [1, 0, -1, 1, 0, -1, 1, 1]
    out = oracle.i2s_matmul(act, _pack_weight_matrix(w, 4, 2), np.ones(2), 2, 2, 4, 32)
    assert out.ravel().tolist() == [3.0, 3.0, 1.0, -3.0]


def test_i2s_matmul_known_answers(oracle):
    # cpu/quantized_matmul.rs:343-631
    act = np.array([3.0, -2.0, 5.0, 7.0], np.float32)
    for bs in (32, 256):
        _run_all(oracle, act, _pack_weight_matrix([1, 0, 0, 1], 2, 2), np.ones(2), 2, 2, 2, bs, act, 1e-6)
    act = np.arange(16, dtype=np.float32)
    _run_all(oracle, act, _pack_weight_matrix([1] * 16, 4, 4), np.ones(4), 4, 4, 4, 32,
             act.reshape(4, 4) @ np.ones((4, 4), np.float32), 1e-5)
    _run_all(oracle, np.ones(12), _pack_weight_matrix([-1] * 12, 4, 3), np.ones(3), 3, 3, 4, 32,
             np.full((3, 3), -4.0), 1e-5)
    w = [1, 0, -1, 1, 0, 1, -1, 0, 1, -1]
    act = np.arange(15, dtype=np.float32) + 0.5
    _run_all(oracle, act, _pack_weight_matrix(w, 5, 2), np.ones(2), 3, 2, 5, 32,
             act.reshape(3, 5) @ np.array(w, np.float32).reshape(5, 2), 1e-5)
    # :555-586 scales [2.0, 0.5] -> [8,2,8,2]
    packed = _pack_weight_matrix([1] * 8, 4, 2)
    out = oracle.i2s_matmul(np.ones(8), packed, [2.0, 0.5], 2, 2, 4, 32)
    assert out.ravel().tolist() == [8.0, 2.0, 8.0, 2.0]
    # :589-613 two k-blocks 1*32 + 3*32 = 128
    out = oracle.i2s_matmul(np.ones(64), np.full(16, 0b01010101, np.uint8), [1.0, 3.0], 1, 1, 64, 32)
    assert out.ravel().tolist() == [128.0]
    # :618-631 integer-exact, tol 0.0
    k, n, m = 8, 4, 4
    w = [[1, -1, 0][i % 3] for i in range(k * n)]
    act = np.array([i % 5 for i in range(m * k)], np.float32)
    _run_all(oracle, act, _pack_weight_matrix(w, k, n), np.ones(n), m, n, k, 32,
             act.reshape(m, k) @ np.array(w, np.float32).reshape(k, n), 0.0)


def test_i2s_matmul_validation(oracle):
    with pytest.raises(oracle.OracleError, match="block_size must be > 0"):
        oracle.i2s_matmul([1.0], [0], [1.0], 1, 1, 1, 0)
    with pytest.raises(oracle.OracleError, match="dimensions must be > 0"):
        oracle.i2s_matmul([1.0], [0], [1.0], 0, 1, 1, 32)
    with pytest.raises(oracle.OracleError, match="activations too small"):
        oracle.i2s_matmul([1.0], [0], [1.0], 1, 1, 4, 32)
    with pytest.raises(oracle.OracleError, match="scales too small"):
        oracle.i2s_matmul(np.ones(64), np.zeros(16, np.uint8), [1.0], 1, 1, 64, 32)


def test_map_b_unpack(oracle):
    # crates/bitnet-quantization/src/utils.rs:57-91, test :267-272 (round trip through [-2,1])
    vals = np.array([-2, -1, 0, 1, 1, 0, -1, -2, 0], np.int8)
    packed = oracle.pack_2bit_values(vals)
    assert np.array_equal(oracle.unpack_2bit_values(packed, vals.size), vals)
    assert oracle.unpack_2bit_values([0b11100100], 4).tolist() == [-2, -1, 0, 1]
    # quantized_linear.rs:1828-1844: clamp[-2,1].round
    x = np.array([-9, -2.2, -1.6, -0.2, 0.49, 1.49, 100], np.float32)
    q = np.sign(np.clip(x, -2, 1)) * np.floor(np.abs(np.clip(x, -2, 1)) + 0.5)
    assert q.tolist() == [-2, -2, -2, 0, 0, 1, 1]


# ---------------------------------------------------------------- RMSNorm
def test_rms_norm(oracle):
    # cpu/layer_norm.rs:118-137,162-169: x * (1/sqrt(mean(x^2)+eps)) * gamma, f64 mean
    x = np.array([1.0, 2.0, 3.0, 4.0], np.float32)
    out = oracle.rms_norm(x, np.ones(4, np.float32), 1e-5)
    inv = np.float32(1.0) / np.sqrt(np.float32(7.5) + np.float32(1e-5), dtype=np.float32)
    assert np.array_equal(out, x * inv)
    x2 = oracle.fill_normal(2 * 2560, 5).reshape(2, 2560)
    g = oracle.fill_uniform(2560, 6, 0.5, 1.5)
    out = oracle.rms_norm(x2, g, 1e-5)
    ref = x2.astype(np.float64) / np.sqrt((x2.astype(np.float64) ** 2).mean(1, keepdims=True) + 1e-5) * g
    assert np.max(np.abs(out - ref)) < 1e-5
    with pytest.raises(oracle.OracleError):
        oracle.rms_norm(np.ones(5, np.float32), np.ones(4, np.float32))
    with pytest.raises(oracle.OracleError):
        oracle.rms_norm(np.ones(4, np.float32), np.ones(4, np.float32), 0.0)


def test_fused_rmsnorm_linear_snapshot(oracle):
    # tests/snapshot_wave5.rs:344-351 -> "0.365148, 1.460593"
    out = oracle.fused_rmsnorm_linear([1, 2, 3, 4], [1, 0, 0, 0, 0, 0, 0, 1], np.ones(4), 1e-5)
    assert ["%.6f" % v for v in out] == ["0.365148", "1.460593"]


# ---------------------------------------------------------------- composed int8 linears
@pytest.mark.parametrize("rows,cols,m", [(8, 256, 1), (16, 2560, 3), (5, 300, 2), (7, 6912, 1)])
def test_composed_qk256_matches_reference_kernel(oracle, rows, cols, m):
    stride = -(-cols // 256) * 64
    qs = oracle.fill_bytes(rows * stride, 0xB2000001)
    x = oracle.fill_normal(m * cols, 42).reshape(m, cols)
    q, s = oracle.quantize_rows_i8(x)
    acc, y = oracle.linear_qk256_i8(qs, rows, cols, q, s)
    # integer truth, independent of the reference kernels
    codes = np.stack([(qs.reshape(rows, stride)[:, c // 4] >> (2 * (c % 4))) & 3 for c in range(cols)], 1)
    want = q.astype(np.int64) @ oracle.MAP_A[codes].astype(np.int64).T
    assert np.array_equal(acc, want)
    for variant in ("scalar", "avx2"):
        if variant == "avx2" and not oracle.lib().oracle_has_avx2():
            continue
        for i in range(m):
            yf = oracle.gemv_qk256(qs, q[i].astype(np.float32), rows, cols, stride, variant)
            assert np.array_equal(yf.astype(np.int32), acc[i])
    assert np.array_equal(y, s[:, None] * acc.astype(np.float32))
    for lut in (oracle.MAP_B, oracle.MAP_C):
        acc2, _ = oracle.linear_qk256_i8(qs, rows, cols, q, s, lut)
        assert np.array_equal(acc2, q.astype(np.int64) @ lut[codes].astype(np.int64).T)


def test_composed_map_c_matches_i2s_matmul(oracle):
    # the packed [n, k/4] orientation of i2s_matmul_f32 == QK256's [rows, cols/4]
    rows, cols, m = 6, 512, 2
    qs = oracle.fill_bytes(rows * cols // 4, 77)
    q = (oracle.fill_normal(m * cols, 3) * 30).clip(-127, 127).round().astype(np.int8).reshape(m, cols)
    acc, _ = oracle.linear_qk256_i8(qs, rows, cols, q, np.ones(m, np.float32), oracle.MAP_C)
    out = oracle.i2s_matmul(q.astype(np.float32), qs, np.ones(rows * 2), m, rows, cols, 256)
    assert np.array_equal(out.astype(np.int32), acc)


@pytest.mark.parametrize("rows,cols,m", [(4, 32, 1), (9, 2560, 2), (3, 6912, 1)])
def test_composed_bitnet32(oracle, rows, cols, m):
    data = oracle.make_bitnet32(rows, cols, 0xB2000002)
    x = oracle.fill_normal(m * cols, 43).reshape(m, cols)
    q, s = oracle.quantize_rows_i8(x)
    acc, y = oracle.linear_bitnet32_i8(data, rows, cols, q, s, want_blocks=True)
    blk = data.reshape(rows, cols // 32, 10)
    codes = np.stack([(blk[:, :, e // 4] >> (2 * (e % 4))) & 3 for e in range(32)], 2)  # rows,nb,32
    wq = oracle.MAP_A[codes].astype(np.int64)
    want = np.einsum("mbe,rbe->mrb", q.reshape(m, cols // 32, 32).astype(np.int64), wq)
    assert np.array_equal(acc, want)
    # within 1e-3 relative of the dequantise-then-multiply order the reference uses (quant/i2s.rs:111)
    yd = oracle.linear_bitnet32_dequant(data, rows, cols, q, s)
    scale = np.abs(yd).max() + 1e-12
    assert np.max(np.abs(y - yd)) / scale < 1e-3


def test_splitmix_generator_is_stable(oracle):
    # fixed bytes so the CPU and GPU harnesses can regenerate identical inputs
    assert oracle.fill_bytes(8, 0).tolist() == list(int(0xE220A8397B1DCDAF).to_bytes(8, "little"))
    assert oracle.fill_bytes(16, 1234567)[8:].tolist() == oracle.fill_bytes(24, 1234567)[8:16].tolist()


def test_round_half_away_trick():
    """The device quantiser rounds with trunc(y + copysign(0.5 - 2^-25, y)) (decode_mega.cu,
    mega_quant_i8): it must equal round-half-away-from-zero (Rust f32::round, cpu/quantize.rs:44)
    for EVERY float with |y| <= 130 -- checked exhaustively, both signs."""
    half_m = np.nextafter(np.float32(0.5), np.float32(0))
    hi = int(np.float32(130.0).view(np.uint32))
    step = 1 << 24
    for lo in range(0, hi + 1, step):
        y = np.arange(lo, min(lo + step, hi + 1), dtype=np.uint32).view(np.float32)
        for ys in (y, -y):
            ref = np.sign(ys) * np.floor(np.abs(ys).astype(np.float64) + 0.5)
            got = np.trunc((ys + np.copysign(half_m, ys)).astype(np.float32))
            assert np.array_equal(ref.astype(np.float32), got)
