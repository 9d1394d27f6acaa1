"""Parity of the CUDA path (through the C ABI) against the CPU oracle. Needs a B200.

Bar (BASELINE.json north_star): int32 accumulators bit-exact against the reference CPU
kernel on identical packed weights and quantised activations; scaled fp32 outputs within
1e-3 relative; int8 activation quantiser bit-exact."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def ctx(b200):
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    c = b200.Context(0)
    yield c
    c.close()


@pytest.fixture(scope="module")
def kernel(b200):
    k = b200.B200Kernel()
    assert k.is_available() and k.name() == "b200"
    yield k
    k.close()


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _f(fn, n):
    return np.array([fn(i) for i in range(n)])


# =========================================================================================
# KernelProvider::matmul_i2s (unpacked bytes) -- reference tests of cpu/fallback.rs, x86.rs
# =========================================================================================
def test_provider_matmul_identity(kernel):
    # fallback.rs:305-318
    c = np.zeros(4, np.float32)
    kernel.matmul_i2s(np.array([1, 2, 3, 4], np.int8), np.array([1, 0, 0, 1], np.uint8), c, 2, 2, 2)
    assert c.tolist() == [1.0, 2.0, 3.0, 4.0]


def test_provider_matmul_dimension_validation(kernel, b200):
    # fallback.rs:320-331
    with pytest.raises(b200.KernelError, match="Matrix A dimension mismatch"):
        kernel.matmul_i2s(np.array([1, 2], np.int8), np.array([1, 0], np.uint8), np.zeros(4, np.float32), 2, 2, 2)


@pytest.mark.parametrize("m,n,k", [(3, 5, 7), (9, 13, 15), (17, 19, 33), (31, 27, 65), (2, 2, 2), (8, 8, 8),
                                   (16, 16, 16), (7, 9, 11), (32, 32, 32), (33, 33, 33), (64, 64, 64)])
def test_provider_matmul_reference_shapes(kernel, oracle, m, n, k):
    # x86.rs:860-895 inputs (a=(idx%11)-5, b=(idx*3+7)%251): arbitrary byte values, not just codes
    a = _f(lambda i: (i % 11) - 5, m * k).astype(np.int8)
    b = _f(lambda i: (i * 3 + 7) % 251, k * n).astype(np.uint8)
    c = np.full(m * n, np.nan, np.float32)
    kernel.matmul_i2s(a, b, c, m, n, k)
    assert np.array_equal(c.reshape(m, n), oracle.matmul_i2s(a, b, m, n, k, "fallback"))


def test_provider_matmul_config1_shape(kernel, oracle):
    # BASELINE.json configs[0]: 2560 -> 6912, batch 1, codes as the caller's `+2` bytes
    m, k, n = 1, 2560, 6912
    a = (oracle.fill_normal(m * k, 42) * 40).round().clip(-127, 127).astype(np.int8)
    b = (oracle.fill_bytes(k * n, 0xB2000000) & 3).astype(np.uint8)
    c = np.zeros(m * n, np.float32)
    kernel.matmul_i2s(a, b, c, m, n, k)
    ref = oracle.matmul_i2s(a, b, m, n, k, "avx2" if oracle.lib().oracle_has_avx2() else "fallback")
    assert np.array_equal(c.reshape(m, n), ref)


# =========================================================================================
# KernelProvider::quantize -- weight quantisers
# =========================================================================================
def _lcg_input(size, seed):
    # tests/gpu_quantization_parity.rs:19-52 style: LCG + diverse ranges
    st = seed
    out = np.zeros(size, np.float32)
    for i in range(size):
        st = (st * 1103515245 + 12345) % (1 << 64)
        nrm = np.float32(st) / np.float32(2 ** 64)
        out[i] = [2 * nrm - 1, 0.5 * (2 * nrm - 1), 1.5 if nrm > 0.5 else -1.5, 0.6 if nrm > 0.5 else -0.6,
                  0.4 if nrm > 0.5 else -0.4, 0.0, 1e-6 * (2 * nrm - 1), 3 * (2 * nrm - 1)][i % 8]
    return out


@pytest.mark.parametrize("qtype", [0, 1, 2])
@pytest.mark.parametrize("size", [32, 128, 1000, 4096, 65536])
def test_provider_quantize_matches_cpu(kernel, oracle, b200, qtype, size):
    x = _lcg_input(size, 12345 + size)
    out = np.zeros((size + 3) // 4, np.uint8)
    block = [32, 64, 128][qtype]
    scales = np.zeros(-(-size // block), np.float32)
    kernel.quantize(x, out, scales, b200.QuantizationType(qtype))
    ro, rs = oracle.quantize(x, qtype)
    assert np.array_equal(scales, rs)
    assert np.array_equal(out, ro)


def test_provider_quantize_edge_list_and_or_semantics(kernel, oracle, b200):
    edge = np.array([0.0, 0.5, -0.5, 0.499, -0.499, 0.501, -0.501, 1.0, -1.0, 1.5, -1.5, 2.0, -2.0,
                     1e-8, -1e-8, 1e8, -1e8, np.inf, -np.inf, 0.1, -0.1, 0.75, -0.75, 1e-30] + [0.0] * 8,
                    np.float32)
    out = np.zeros(edge.size // 4, np.uint8)
    scales = np.zeros(1, np.float32)
    kernel.quantize(edge, out, scales, b200.QuantizationType.I2S)
    ro, rs = oracle.quantize(edge, 0)
    assert np.array_equal(out, ro) and np.array_equal(scales, rs)
    pre = np.full(edge.size // 4, 0xFF, np.uint8)
    kernel.quantize(edge, pre, scales, b200.QuantizationType.I2S)
    assert np.all(pre == 0xFF)      # ORs into the caller's bytes (fallback.rs:153)
    with pytest.raises(b200.KernelError, match="Output buffer too small"):
        kernel.quantize(np.ones(32, np.float32), np.zeros(1, np.uint8), np.zeros(1, np.float32), 0)


# =========================================================================================
# gemv_qk256 on host slices -- the live transformer's call
# =========================================================================================
def test_gemv_qk256_reference_known_answers(b200):
    # i2s_qk256.rs:440-465, :633-683; qk256_integration.rs:117-148; qk256_dispatch.rs:106-119
    x = np.arange(256, dtype=np.float32)
    y = np.zeros(3, np.float32)
    b200.gemv_qk256(np.full(3 * 64, 0x55, np.uint8), x, y, 3, 256, 64)
    assert np.all(np.abs(y + x.sum()) < 1e-3)
    for code in range(4):
        byte = code | (code << 2) | (code << 4) | (code << 6)
        y = np.zeros(4, np.float32)
        b200.gemv_qk256(np.full(4 * 64, byte, np.uint8), np.ones(256, np.float32), y, 4, 256, 64)
        assert np.all(y == [-2.0, -1.0, 1.0, 2.0][code] * 256)
    y = np.zeros(256, np.float32)
    b200.gemv_qk256(np.full(256 * 64, 0x55, np.uint8), np.full(256, 0.5, np.float32), y, 256, 256, 64)
    assert np.all(y == -128.0)
    # tail: cols = 300 (i2s_qk256.rs:419-437)
    x = (np.arange(300) % 7).astype(np.float32)
    y = np.zeros(1, np.float32)
    b200.gemv_qk256(np.full(128, 0xAA, np.uint8), x, y, 1, 300, 128)
    assert abs(y[0] - x.sum()) < 1e-3


@pytest.mark.parametrize("rows,cols", [(1, 256), (16, 512), (33, 300), (7, 777), (640, 2560), (2560, 6912)])
def test_gemv_qk256_vs_oracle(b200, oracle, rows, cols):
    stride = -(-cols // 256) * 64
    qs = oracle.fill_bytes(rows * stride, 0xB2000000 + rows)
    x = oracle.fill_normal(cols, 42)
    y = np.zeros(rows, np.float32)
    b200.gemv_qk256(qs, x, y, rows, cols, stride)
    ref = oracle.gemv_qk256(qs, x, rows, cols, stride, "scalar")
    # the reference's own AVX2-vs-scalar bound (qk256_avx2_correctness.rs:53): 1e-4 abs-or-rel
    d = np.abs(y - ref)
    assert np.all((d <= 1e-4 * np.sqrt(cols / 256)) | (d <= 1e-4 * np.abs(ref)))
    xi = (x * 40).round().clip(-127, 127).astype(np.float32)
    b200.gemv_qk256(qs, xi, y, rows, cols, stride)
    assert np.array_equal(y, oracle.gemv_qk256(qs, xi, rows, cols, stride))   # exact


# =========================================================================================
# per-token absmax int8 activation quantiser -- bit exact
# =========================================================================================
@pytest.mark.parametrize("M,K", [(1, 2560), (3, 6912), (16, 4096), (2, 14336), (5, 300), (4, 7), (64, 2560)])
def test_actquant_bit_exact(ctx, oracle, M, K):
    x = oracle.fill_normal(M * K, 42 + K).reshape(M, K)
    x[0, : min(K, 8)] = [0.5, -0.5, 1.5, -1.5, 2.5, -2.5, 0.49999997, 126.5][: min(K, 8)]
    q, s = ctx.actquant_i8(dev(x))
    rq, rs = oracle.quantize_rows_i8(x)
    assert np.array_equal(s.cpu().numpy(), rs)
    assert np.array_equal(q.cpu().numpy(), rq)


def test_actquant_edge_rows(ctx, oracle):
    K = 2560
    x = np.zeros((6, K), np.float32)
    x[1, 17] = 3.0                                  # single spike -> 127
    x[2] = oracle.fill_normal(K, 1)
    x[2, 5] = x[2, 9] = -np.abs(x[2]).max()         # +-absmax ties
    x[3] = (np.arange(K) % 255 - 127) * 0.5         # .5 rounding boundaries: absmax 63.5 -> inv 2
    x[4] = 1e-30 * oracle.fill_normal(K, 2)         # tiny
    x[5] = 1e30 * oracle.fill_normal(K, 3)          # huge
    q, s = ctx.actquant_i8(dev(x))
    rq, rs = oracle.quantize_rows_i8(x)
    assert np.array_equal(s.cpu().numpy(), rs) and np.array_equal(q.cpu().numpy(), rq)
    assert s[0].item() == 0.0 and not q[0].any().item()     # all-zero row -> scale 0 (quantize.rs:38-40)
    assert q[1, 17].item() == 127


@pytest.mark.parametrize("M,K", [(1, 2560), (4, 4096), (3, 6912), (2, 300)])
def test_rmsnorm_actquant(ctx, oracle, M, K):
    x = oracle.fill_normal(M * K, 7, 2.0).reshape(M, K)     # pre-norm hidden ~ N(0, 4)
    g = oracle.fill_uniform(K, 8, 0.5, 1.5)
    q, s, normed = ctx.rmsnorm_actquant(dev(x), dev(g), 1e-5, want_normed=True)
    normed = normed.cpu().numpy()
    ref = oracle.rms_norm(x, g, 1e-5)
    assert np.max(np.abs(normed - ref) / (np.abs(ref) + 1e-6)) < 1e-6     # f64 mean, order differs
    # the fused epilogue is exactly quantize_symmetric_i8 of the normalised row it produced
    rq, rs = oracle.quantize_rows_i8(normed)
    assert np.array_equal(s.cpu().numpy(), rs) and np.array_equal(q.cpu().numpy(), rq)
    q2, s2 = ctx.rmsnorm_actquant(dev(x), dev(g), 1e-5)
    assert torch.equal(q2, q) and torch.equal(s2, s)


def test_rmsnorm_actquant_rejects_bad_eps(ctx, b200):
    x = torch.zeros(1, 32, device="cuda")
    with pytest.raises(b200.B200Error, match="eps"):
        ctx.rmsnorm_actquant(x, torch.ones(32, device="cuda"), 0.0)


# =========================================================================================
# the ternary linear, QK256 -- int32 accumulators bit-exact
# =========================================================================================
SHAPES_2B = [(2560, 2560), (640, 2560), (3840, 2560), (6912, 2560), (13824, 2560), (2560, 6912)]


def _qk256_case(oracle, rows, cols, M, seed):
    stride = -(-cols // 256) * 64
    qs = oracle.fill_bytes(rows * stride, 0xB2000000 + seed)
    x = oracle.fill_normal(M * cols, 42 + seed).reshape(M, cols)
    q, s = oracle.quantize_rows_i8(x)
    return qs, q, s


@pytest.mark.parametrize("rows,cols", [(8, 256), (9, 512), (33, 300), (7, 777), (100, 1024), (17, 4096)])
@pytest.mark.parametrize("M", [1, 2, 3, 4, 5, 8, 9, 16, 17])
def test_linear_qk256_small(ctx, b200, oracle, rows, cols, M, monkeypatch):
    qs, q, s = _qk256_case(oracle, rows, cols, M, rows + M)
    w = ctx.weight(qs, b200.LAYOUT_QK256, rows, cols)
    y, acc = ctx.linear_i8(w, dev(q), dev(s), want_acc=True)
    racc, ry = oracle.linear_qk256_i8(qs, rows, cols, q, s)
    assert np.array_equal(acc.cpu().numpy(), racc)
    assert np.array_equal(y.cpu().numpy(), ry)          # one exact int->f32 and one multiply
    monkeypatch.setenv("BITNET_B200_SMALLM_MAX", "0")   # the GEMV (M <= 8) / tcgen05 GEMM alone, without gemm_small.cu
    y2, acc2 = ctx.linear_i8(w, dev(q), dev(s), want_acc=True)
    assert torch.equal(acc2, acc) and torch.equal(y2, y)


@pytest.mark.parametrize("rows,cols", SHAPES_2B + [(4096, 4096), (1024, 14336)])
@pytest.mark.parametrize("M", [1, 8, 16])
def test_linear_qk256_model_shapes(ctx, b200, oracle, rows, cols, M, monkeypatch):
    qs, q, s = _qk256_case(oracle, rows, cols, M, rows ^ cols)
    w = ctx.weight(qs, b200.LAYOUT_QK256, rows, cols)
    assert w.nbytes == rows * cols // 4
    y, acc = ctx.linear_i8(w, dev(q), dev(s), want_acc=True)
    racc, ry = oracle.linear_qk256_i8(qs, rows, cols, q, s)   # the reference kernel itself (AVX2 gemv_qk256)
    assert np.array_equal(acc.cpu().numpy(), racc)
    assert np.array_equal(y.cpu().numpy(), ry)
    y2 = ctx.linear_i8(w, dev(q), dev(s))                      # the no-accumulator path
    assert torch.equal(y2, y)
    monkeypatch.setenv("BITNET_B200_SMALLM_MAX", "0")          # the GEMV / tcgen05 GEMM alone, without gemm_small.cu
    y3, acc3 = ctx.linear_i8(w, dev(q), dev(s), want_acc=True)
    assert torch.equal(acc3, acc) and torch.equal(y3, y)


# ---- M > 16: the tcgen05 kind::i8 GEMM (gemm_tcgen05.cu); same bit-exact bar as the GEMV -------
@pytest.mark.parametrize("rows,cols,M", [
    (128, 128, 128), (128, 256, 17), (8, 256, 64), (130, 1024, 129), (256, 512, 300), (100, 1024, 33),
    (640, 2560, 64), (3840, 2560, 128), (2560, 6912, 64), (6912, 2560, 130), (2560, 2560, 512), (1024, 14336, 64),
    (384, 128, 40), (2560, 2560, 257), (6912, 2560, 1000), (3840, 2560, 9), (2560, 6912, 12), (4096, 512, 64),
    (13824, 2560, 2048),
])
def test_linear_qk256_gemm_tcgen05(ctx, b200, oracle, rows, cols, M, monkeypatch):
    monkeypatch.setenv("BITNET_B200_SMALLM_MAX", "0")   # M <= 64 would otherwise take gemm_small.cu (tested below)
    qs, q, s = _qk256_case(oracle, rows, cols, M, rows + cols + M)
    w = ctx.weight(qs, b200.LAYOUT_QK256, rows, cols)
    y, acc = ctx.linear_i8(w, dev(q), dev(s), want_acc=True)
    racc, ry = oracle.linear_qk256_i8(qs, rows, cols, q, s)
    assert np.array_equal(acc.cpu().numpy(), racc)
    assert np.array_equal(y.cpu().numpy(), ry)
    y2 = ctx.linear_i8(w, dev(q), dev(s))
    assert torch.equal(y2, y)


# ---- M = 1..64: the mma.sync batched-decode kernel (gemm_small.cu); same bit-exact bar -------------------
@pytest.mark.parametrize("rows,cols,M", [
    (640, 2560, 1), (4096, 4096, 3), (2560, 6912, 8), (8, 256, 9), (9, 512, 16), (100, 1024, 17), (17, 4096, 24), (33, 768, 25), (65, 1280, 33), (128, 256, 48),
    (640, 2560, 64), (3840, 2560, 9), (3840, 2560, 32), (2560, 2560, 64), (13824, 2560, 16), (13824, 2560, 64),
    (2560, 6912, 12), (2560, 6912, 64), (4096, 4096, 64), (1024, 14336, 40), (4096, 2048, 64), (6144, 4096, 20),
    (4096, 7168, 64), (1000, 3072, 57),
])
def test_linear_qk256_gemm_small(ctx, b200, oracle, rows, cols, M, monkeypatch):
    qs, q, s = _qk256_case(oracle, rows, cols, M, rows + cols + M)
    w = ctx.weight(qs, b200.LAYOUT_QK256, rows, cols)
    racc, ry = oracle.linear_qk256_i8(qs, rows, cols, q, s)
    y, acc = ctx.linear_i8(w, dev(q), dev(s), want_acc=True)   # first use of a shape: times the candidate plans
    assert np.array_equal(acc.cpu().numpy(), racc)
    assert np.array_equal(y.cpu().numpy(), ry)
    l0 = b200.launch_count()
    y2 = ctx.linear_i8(w, dev(q), dev(s))                       # the chosen plan (or the tcgen05 GEMM, if that won)
    assert b200.launch_count() - l0 == 1
    assert torch.equal(y2, y)
    monkeypatch.setenv("BITNET_B200_SMALLM_AUTOTUNE", "0")      # the cost model's pick, on a shape not seen before
    w1 = ctx.weight(qs[:(rows - 1) * (cols // 4)], b200.LAYOUT_QK256, rows - 1, cols) if rows > 1 else None
    if w1 is not None:
        y4, acc4 = ctx.linear_i8(w1, dev(q), dev(s), want_acc=True)
        assert np.array_equal(acc4.cpu().numpy(), racc[:, :rows - 1])
    # and the same integers from the tcgen05 GEMM on the other side of the routing boundary
    monkeypatch.setenv("BITNET_B200_SMALLM_MAX", "0")
    y3, acc3 = ctx.linear_i8(w, dev(q), dev(s), want_acc=True)
    assert b200.last_linear_kernel() in ("gemm_tcgen05", "gemv_fast", "gemv")
    assert torch.equal(acc3, acc) and torch.equal(y3, y)


# both sides of every routing boundary of linear_i8 on one weight: GEMV | gemm_small (M = 8 | 9), gemm_small | tcgen05
# (M = 64 | 65), and the token-tile edges in between -- whichever kernel the first-use timing picks, the same integers
@pytest.mark.parametrize("rows,cols", [(3840, 2560), (2560, 6912), (6144, 4096)])
def test_linear_qk256_every_routing_boundary(ctx, b200, oracle, rows, cols):
    qs = oracle.fill_bytes(rows * cols // 4, 0xB2000000 + rows)
    w = ctx.weight(qs, b200.LAYOUT_QK256, rows, cols)
    seen = set()
    for M in (1, 4, 8, 9, 16, 17, 32, 33, 48, 49, 64, 65, 128):
        x = oracle.fill_normal(M * cols, 1000 + M).reshape(M, cols)
        q, s = oracle.quantize_rows_i8(x)
        racc, ry = oracle.linear_qk256_i8(qs, rows, cols, q, s)
        for _ in range(2):      # first call of a shape may be the timing pass, the second runs the chosen kernel
            y, acc = ctx.linear_i8(w, dev(q), dev(s), want_acc=True)
            assert np.array_equal(acc.cpu().numpy(), racc), (M, b200.last_linear_kernel())
            assert np.array_equal(y.cpu().numpy(), ry), (M, b200.last_linear_kernel())
        seen.add(b200.last_linear_kernel())
    assert "gemm_tcgen05" in seen and ("gemm_small" in seen or "gemm_small_resident" in seen)


# a spread of the planner's search space (row warps, k warps, segments per warp, cluster K split): warp grids, ragged row
# groups, ragged last chunks (zero-filled by the TMA unit), ring wrap-around (more chunks than stages), DSMEM reduction
@pytest.mark.parametrize("rows,cols,M,plans", [
    (200, 2560, 64, ["1,5,2,1", "1,3,2,2", "1,10,1,2", "1,4,2,3", "1,10,1,1", "2,3,2,2", "1,8,1,3", "1,9,2,2", "2,2,1,2",
                     "2,8,1,1", "1,16,1,2", "2,5,4,1"]),
    (2560, 6912, 16, ["1,4,4,2", "1,15,2,2", "2,6,2,5", "2,5,2,6", "2,4,4,2", "1,3,2,3", "1,7,2,2", "1,4,2,7", "1,13,2,3",
                      "2,8,1,7", "1,2,1,1"]),
    (4096, 4096, 33, ["1,4,1,1", "2,3,2,6", "2,8,1,4", "1,2,4,1", "1,6,1,6", "1,3,4,1", "2,2,2,2", "1,16,1,2", "2,2,1,4",
                      "2,8,1,1", "2,8,2,2"]),
    (3840, 2560, 5, ["1,10,2,1", "2,5,4,1", "1,4,1,5", "1,16,4,1", "2,8,1,3", "1,2,1,5"]),
    (77, 14336, 24, ["1,6,1,1", "1,4,1,1", "1,11,2,2", "1,5,4,1", "2,7,2,1", "2,5,2,4", "2,5,4,6", "2,2,4,2", "1,9,4,4",
                     "2,8,1,7", "1,16,1,7"]),
])
def test_linear_qk256_gemm_small_every_plan(ctx, b200, oracle, rows, cols, M, plans, monkeypatch):
    qs, q, s = _qk256_case(oracle, rows, cols, M, rows * 3 + M)
    w = ctx.weight(qs, b200.LAYOUT_QK256, rows, cols)
    racc, ry = oracle.linear_qk256_i8(qs, rows, cols, q, s)
    for plan in plans:
        monkeypatch.setenv("BITNET_B200_SMALLM_PLAN", plan)
        l0 = b200.launch_count()
        y, acc = ctx.linear_i8(w, dev(q), dev(s), want_acc=True)
        assert b200.launch_count() - l0 == 1, plan
        assert b200.last_linear_kernel() == "gemm_small"
        assert np.array_equal(acc.cpu().numpy(), racc), plan
        assert np.array_equal(y.cpu().numpy(), ry), plan


# the activation-resident persistent variant (more row groups than SMs, all tokens' activations kept in shared memory):
# every warp grid, ragged last row group, ring wrap-around (up to 7 groups per CTA), M below / at a token-tile edge
@pytest.mark.parametrize("rows,cols,M", [(13824, 2560, 16), (13001, 2560, 32), (14336, 4096, 5), (9999, 6912, 8),
                                         (28672, 4096, 9), (6000, 256, 24)])
def test_linear_qk256_gemm_small_resident(ctx, b200, oracle, rows, cols, M, monkeypatch):
    qs, q, s = _qk256_case(oracle, rows, cols, M, rows + 7 * M)
    w = ctx.weight(qs, b200.LAYOUT_QK256, rows, cols)
    racc, ry = oracle.linear_qk256_i8(qs, rows, cols, q, s)
    ran = 0
    for plan in ("R,1,8", "R,1,12", "R,1,16", "R,2,4", "R,2,6", "R,2,8"):
        monkeypatch.setenv("BITNET_B200_SMALLM_PLAN", plan)
        l0 = b200.launch_count()
        try:
            y, acc = ctx.linear_i8(w, dev(q), dev(s), want_acc=True)
        except b200.B200Error as ex:     # a forced plan that does not fit in shared memory is refused, not rerouted
            assert "not realisable" in str(ex)
            continue
        ran += 1
        assert b200.last_linear_kernel() == "gemm_small_resident"
        assert b200.launch_count() - l0 == 1, plan
        assert np.array_equal(acc.cpu().numpy(), racc), plan
        assert np.array_equal(y.cpu().numpy(), ry), plan
    assert ran >= 2


def test_linear_qk256_gemm_small_extremes_and_maps(ctx, b200, oracle):
    # all-(+/-127) activations against all-(+2) / all-(-2) codes: the largest |acc| the layout allows
    rows, cols, M = 96, 14336, 64
    for lut_name, byte in (("MAP_A", 0xFF), ("MAP_A", 0x00), ("MAP_B", 0xAA), ("MAP_C", 0x55)):
        qs = np.full(rows * cols // 4, byte, np.uint8)
        q = np.full((M, cols), 127, np.int8)
        q[1::2] = -127
        s = np.linspace(0.5, 2.0, M).astype(np.float32)
        lut = getattr(oracle, lut_name)
        w = ctx.weight(qs, b200.LAYOUT_QK256, rows, cols, lut=lut)
        y, acc = ctx.linear_i8(w, dev(q), dev(s), want_acc=True)
        racc, ry = oracle.linear_qk256_i8(qs, rows, cols, q, s, lut=lut)
        assert np.array_equal(acc.cpu().numpy(), racc)
        assert np.array_equal(y.cpu().numpy(), ry)


def test_linear_qk256_gemm_cold_weights_repeated(ctx, b200):
    # regression: with HBM-cold weights the decoder warps are the critical path; a packed-code stage
    # released before its shared loads had landed was overwritten by the next TMA box (wrong rows,
    # intermittently). Fresh weights + activations every repetition, BN = 256, several rounds per CTA.
    rows, cols, M = 13824, 2560, 1024
    rng = np.random.default_rng(7)
    lut = torch.tensor([-2, -1, 1, 2], dtype=torch.float32, device="cuda")
    for rep in range(6):
        qs = rng.integers(0, 256, rows * cols // 4, dtype=np.uint8)
        codes = torch.from_numpy(qs).cuda().view(rows, cols // 4)
        wd = torch.stack([lut[((codes >> (2 * j)) & 3).long()] for j in range(4)], dim=2).reshape(rows, cols)
        q = torch.from_numpy(rng.integers(-127, 128, (M, cols)).astype(np.int8)).cuda()
        ref = (q.float() @ wd.t()).round().to(torch.int32)   # exact in f32: |acc| < 2^24
        w = ctx.weight(qs, b200.LAYOUT_QK256, rows, cols)
        _, acc = ctx.linear_i8(w, q, torch.ones(M, device="cuda"), want_acc=True)
        assert torch.equal(acc, ref), "rep %d" % rep


# ---- configs[2] / configs[3] sizes: M = 4096 and M = 16384 on the four 2B-4T projections ------------------
# (the persistent GEMM wraps its rings the most here and split-K is off). The int32 reference is a device-side
# dense product in f32 over K halves (every partial |sum| < 2^24, so f32 is exact), added as int64 -> int32.
def _torch_acc_qk256(qs, rows, cols, q_dev, chunk=2048):
    import torch
    lut = torch.tensor([-2, -1, 1, 2], dtype=torch.float32, device="cuda")
    codes = torch.from_numpy(qs).cuda().view(rows, cols // 4)
    wd = torch.stack([lut[((codes >> (2 * j)) & 3).long()] for j in range(4)], dim=2).reshape(rows, cols)
    acc = torch.zeros(q_dev.shape[0], rows, dtype=torch.int64, device="cuda")
    for k0 in range(0, cols, chunk):   # |partial| <= 2048 * 127 * 2 < 2^24: exact in f32
        acc += (q_dev[:, k0:k0 + chunk].float() @ wd[:, k0:k0 + chunk].t()).round().to(torch.int64)
    return acc.to(torch.int32)


@pytest.mark.parametrize("rows,cols", [(3840, 2560), (2560, 2560), (13824, 2560), (2560, 6912)])
@pytest.mark.parametrize("M", [4096, 16384])
def test_linear_qk256_gemm_prefill_sizes(ctx, b200, rows, cols, M):
    rng = np.random.default_rng(rows + cols + M)
    qs = rng.integers(0, 256, rows * cols // 4, dtype=np.uint8)
    q = torch.from_numpy(rng.integers(-127, 128, (M, cols)).astype(np.int8)).cuda()
    s = torch.from_numpy(rng.uniform(0.5, 2.0, M).astype(np.float32)).cuda()
    w = ctx.weight(qs, b200.LAYOUT_QK256, rows, cols)
    y, acc = ctx.linear_i8(w, q, s, want_acc=True)
    ref = _torch_acc_qk256(qs, rows, cols, q)
    assert torch.equal(acc, ref)
    assert torch.equal(y, ref.float() * s[:, None])      # one exact int->f32 and one multiply
    del y, acc, ref
    y2 = ctx.linear_i8(w, q, s)                            # the no-accumulator path
    assert torch.equal(y2, _torch_acc_qk256(qs, rows, cols, q).float() * s[:, None])


@pytest.mark.parametrize("rows,cols,M", [(256, 512, 9), (13824, 2560, 64), (13824, 2560, 1000), (1024, 2560, 4096), (640, 1024, 300)])
def test_linear_qk256_gemm_swiglu_epilogue(ctx, b200, oracle, rows, cols, M):
    # SwiGLU in the GEMM epilogue == the plain GEMM followed by the stand-alone glue kernel, bit for bit
    qs, q, s = _qk256_case(oracle, rows, cols, M, rows + M + 5)
    w = ctx.weight(qs, b200.LAYOUT_QK256, rows, cols)
    y = ctx.linear_i8(w, dev(q), dev(s))
    h_ref, _ = ctx.row_absmax(y, swiglu=True)
    h = ctx.linear_i8_swiglu(w, dev(q), dev(s))
    assert torch.equal(h, h_ref)
    g, u = y[:, 0::2], y[:, 1::2]
    assert torch.allclose(h, torch.nn.functional.silu(g) * u, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("lut_name", ["MAP_B", "MAP_C"])
def test_linear_qk256_gemm_other_code_maps(ctx, b200, oracle, lut_name):
    rows, cols, M = 192, 1024, 70
    qs, q, s = _qk256_case(oracle, rows, cols, M, 9)
    lut = getattr(oracle, lut_name)
    w = ctx.weight(qs, b200.LAYOUT_QK256, rows, cols, lut=lut)
    _, acc = ctx.linear_i8(w, dev(q), dev(s), want_acc=True)
    racc, _ = oracle.linear_qk256_i8(qs, rows, cols, q, s, lut)
    assert np.array_equal(acc.cpu().numpy(), racc)


def test_linear_qk256_gemm_extreme_activations(ctx, b200, oracle):
    # every activation +-127 and every code +-2 over K = 14336: |acc| = 3 641 344, the F5 bound, in int32 TMEM
    rows, cols, M = 128, 14336, 64
    qs = np.where(np.arange(rows * cols // 4) % 2 == 0, 0xFF, 0x00).astype(np.uint8)
    q = np.where(np.arange(M * cols).reshape(M, cols) % 3 == 0, -127, 127).astype(np.int8)
    s = np.ones(M, np.float32)
    w = ctx.weight(qs, b200.LAYOUT_QK256, rows, cols)
    _, acc = ctx.linear_i8(w, dev(q), dev(s), want_acc=True)
    racc, _ = oracle.linear_qk256_i8(qs, rows, cols, q, s)
    assert np.array_equal(acc.cpu().numpy(), racc)


@pytest.mark.parametrize("lut_name", ["MAP_B", "MAP_C"])
def test_linear_qk256_other_code_maps(ctx, b200, oracle, lut_name):
    rows, cols, M = 64, 1024, 3
    qs, q, s = _qk256_case(oracle, rows, cols, M, 5)
    lut = getattr(oracle, lut_name)
    w = ctx.weight(qs, b200.LAYOUT_QK256, rows, cols, lut=lut)
    _, acc = ctx.linear_i8(w, dev(q), dev(s), want_acc=True)
    racc, _ = oracle.linear_qk256_i8(qs, rows, cols, q, s, lut)
    assert np.array_equal(acc.cpu().numpy(), racc)


def test_linear_qk256_extreme_activations(ctx, b200, oracle):
    # every activation +-127 and every code +-2: |acc| = 127*2*K, the F5 bound
    rows, cols = 16, 14336
    stride = cols // 4
    qs = np.zeros((rows, stride), np.uint8)
    qs[0::2] = 0xFF
    qs[1::2] = 0x00
    q = np.full((2, cols), 127, np.int8)
    q[1] = -127
    s = np.array([1.0, 0.5], np.float32)
    w = ctx.weight(qs, b200.LAYOUT_QK256, rows, cols)
    _, acc = ctx.linear_i8(w, dev(q), dev(s), want_acc=True)
    racc, _ = oracle.linear_qk256_i8(qs.ravel(), rows, cols, q, s)
    assert np.array_equal(acc.cpu().numpy(), racc)
    assert abs(int(racc[0, 0])) == 127 * 2 * cols


def test_linear_qk256_zero_row_and_padding(ctx, b200, oracle):
    rows, cols = 12, 512
    qs, q, s = _qk256_case(oracle, rows, cols, 3, 9)
    q[1] = 0
    s[1] = 0.0                                            # all-zero token: scale 0 -> outputs 0
    padded = np.concatenate([qs, np.zeros(96, np.uint8)])  # +-128 B alignment padding tolerated
    w = ctx.weight(padded, b200.LAYOUT_QK256, rows, cols)
    y, acc = ctx.linear_i8(w, dev(q), dev(s), want_acc=True)
    racc, ry = oracle.linear_qk256_i8(qs, rows, cols, q, s)
    assert np.array_equal(acc.cpu().numpy(), racc) and np.array_equal(y.cpu().numpy(), ry)
    assert not y[1].any().item()
    with pytest.raises(b200.B200Error, match="data size mismatch"):
        ctx.weight(np.zeros(rows * 128 + 129, np.uint8), b200.LAYOUT_QK256, rows, cols)


def test_linear_f32_device(ctx, b200, oracle):
    rows, cols, M = 640, 2560, 3
    qs = oracle.fill_bytes(rows * cols // 4, 3)
    x = oracle.fill_normal(M * cols, 4).reshape(M, cols)
    w = ctx.weight(qs, b200.LAYOUT_QK256, rows, cols)
    y = ctx.linear_f32(w, dev(x)).cpu().numpy()
    for i in range(M):
        ref = oracle.gemv_qk256(qs, x[i], rows, cols, cols // 4, "scalar")
        d = np.abs(y[i] - ref)
        assert np.all((d <= 1e-4 * np.sqrt(cols / 256)) | (d <= 1e-4 * np.abs(ref)))


# =========================================================================================
# the ternary linear, BitNet32-F16 -- per-block int32 bit-exact, scaled output within 1e-3
# =========================================================================================
def _rel(y, ref):
    return np.max(np.abs(y - ref)) / (np.max(np.abs(ref)) + 1e-30)


@pytest.mark.parametrize("rows,cols", [(4, 32), (9, 64), (33, 320), (17, 4096)] + SHAPES_2B)
@pytest.mark.parametrize("M", [1, 3, 16])
def test_linear_bitnet32(ctx, b200, oracle, rows, cols, M):
    if rows * cols > 4_000_000 and M == 3:
        pytest.skip("covered by M=1 and M=16")
    data = oracle.make_bitnet32(rows, cols, 0xB2000000 + rows)
    x = oracle.fill_normal(M * cols, 43 + rows).reshape(M, cols)
    q, s = oracle.quantize_rows_i8(x)
    w = ctx.weight(data, b200.LAYOUT_BITNET32_F16, rows, cols)
    assert w.nbytes == rows * cols // 32 * 10
    y, acc = ctx.linear_i8(w, dev(q), dev(s), want_acc=True)
    racc, ry = oracle.linear_bitnet32_i8(data, rows, cols, q, s, want_blocks=True)
    assert np.array_equal(acc.cpu().numpy(), racc)                 # per-32-block int32, bit exact
    y = y.cpu().numpy()
    assert _rel(y, ry) < 1e-5                                      # same math, different fp32 sum order
    assert _rel(y, oracle.linear_bitnet32_dequant(data, rows, cols, q, s)) < 1e-3   # the north_star bar
    # the no-accumulator path: same GEMV at M <= 8, the tcgen05 kind::f16 GEMM at M >= 9 (fp32 sum order differs)
    y2 = ctx.linear_i8(w, dev(q), dev(s)).cpu().numpy()
    assert _rel(y2, y) < (1e-6 if M <= 8 else 2e-5)


# ---- M >= 9: the tcgen05 kind::f16 GEMM on s * LUT[code] (gemm_f16_tcgen05.cu) -------------------
@pytest.mark.parametrize("rows,cols,M", [
    (128, 64, 16), (128, 256, 9), (100, 1024, 33), (640, 2560, 64), (2560, 6912, 64), (3840, 2560, 300),
    (6912, 2560, 1000), (2560, 2560, 2048), (4096, 512, 64), (130, 320, 129),
])
def test_linear_bitnet32_gemm_tcgen05(ctx, b200, oracle, rows, cols, M):
    data = oracle.make_bitnet32(rows, cols, 0xB2000040 + rows + M)
    x = oracle.fill_normal(M * cols, 47 + rows).reshape(M, cols)
    q, s = oracle.quantize_rows_i8(x)
    w = ctx.weight(data, b200.LAYOUT_BITNET32_F16, rows, cols)
    y = ctx.linear_i8(w, dev(q), dev(s)).cpu().numpy()
    ry = oracle.linear_bitnet32_i8(data, rows, cols, q, s)         # block-ordered int32 x scale
    assert _rel(y, ry) < 2e-5                                      # exact products, fp32 sum order differs
    assert _rel(y, oracle.linear_bitnet32_dequant(data, rows, cols, q, s)) < 1e-3   # the north_star bar
    assert np.array_equal(ctx.linear_i8(w, dev(q), dev(s)).cpu().numpy(), y) or rows * M < 148 * 128 * 16


@pytest.mark.parametrize("rows,cols", [(3840, 2560), (2560, 2560), (13824, 2560), (2560, 6912)])
@pytest.mark.parametrize("M", [4096, 16384])
def test_linear_bitnet32_gemm_prefill_sizes(ctx, b200, oracle, rows, cols, M):
    # configs[2] / configs[3] sizes on the BitNet32-F16 layout; reference = dequantise-then-f64 product on the device
    data = oracle.make_bitnet32(rows, cols, 0xB2000090 + rows + M)
    rng = np.random.default_rng(rows + cols + M + 1)
    q = torch.from_numpy(rng.integers(-127, 128, (M, cols)).astype(np.int8)).cuda()
    s = torch.from_numpy(rng.uniform(0.5, 2.0, M).astype(np.float32)).cuda()
    w = ctx.weight(data, b200.LAYOUT_BITNET32_F16, rows, cols)
    y = ctx.linear_i8(w, q, s)
    blk = torch.from_numpy(data).cuda().view(rows, cols // 32, 10)
    codes = blk[:, :, :8]
    lut = torch.tensor([-2, -1, 1, 2], dtype=torch.float64, device="cuda")
    wq = torch.stack([lut[((codes >> (2 * j)) & 3).long()] for j in range(4)], dim=3).reshape(rows, cols // 32, 32)
    sc = blk[:, :, 8:].contiguous().view(torch.float16).reshape(rows, cols // 32).double().abs().clamp(1e-3, 1e3)  # SCALE_ABS_CLAMP (quant/i2s.rs:93)
    wd = (wq * sc[:, :, None]).reshape(rows, cols)
    for m0 in range(0, M, 2048):
        ref = (q[m0:m0 + 2048].double() @ wd.t()) * s[m0:m0 + 2048, None].double()
        err = (y[m0:m0 + 2048].double() - ref).norm() / ref.norm()
        assert float(err) < 2e-5, "rows %d..%d rel %.3g" % (m0, m0 + 2048, float(err))


def test_linear_bitnet32_gemm_other_code_map(ctx, b200, oracle):
    rows, cols, M = 192, 1024, 70
    data = oracle.make_bitnet32(rows, cols, 77)
    x = oracle.fill_normal(M * cols, 3).reshape(M, cols)
    q, s = oracle.quantize_rows_i8(x)
    w = ctx.weight(data, b200.LAYOUT_BITNET32_F16, rows, cols, lut=oracle.MAP_C)
    y = ctx.linear_i8(w, dev(q), dev(s)).cpu().numpy()
    ry = oracle.linear_bitnet32_i8(data, rows, cols, q, s, oracle.MAP_C)
    assert _rel(y, ry) < 2e-5


@pytest.mark.parametrize("policy", ["SCALE_ABS_CLAMP", "SCALE_RAW"])
def test_linear_bitnet32_scale_edge_cases(ctx, b200, oracle, policy):
    # quant/i2s.rs:1007-1057: zero / negative / inf / tiny / max f16 scales
    rows, cols, M = 8, 64, 2
    data = oracle.make_bitnet32(rows, cols, 11).reshape(-1, 10)
    specials = [0x0000, 0xB800, 0x7C00, 0x0001, 0x7BFF, 0x8000, 0x3C00, 0x1400]   # 0,-0.5,inf,denorm,65504,-0,1,~1e-3
    for i, bits in enumerate(specials):
        data[i, 8] = bits & 0xFF
        data[i, 9] = bits >> 8
    if policy == "SCALE_RAW":
        data[2, 8:10] = [0x00, 0x3C]        # raw policy: keep inf out of the sum
    x = oracle.fill_normal(M * cols, 5).reshape(M, cols)
    q, s = oracle.quantize_rows_i8(x)
    pol = getattr(b200, policy)
    w = ctx.weight(data.ravel(), b200.LAYOUT_BITNET32_F16, rows, cols, scale_policy=pol)
    y = ctx.linear_i8(w, dev(q), dev(s)).cpu().numpy()
    ry = oracle.linear_bitnet32_i8(data.ravel(), rows, cols, q, s, scale_policy=pol)
    assert np.all(np.isfinite(y))
    assert _rel(y, ry) < 1e-5


@pytest.mark.parametrize("policy", ["SCALE_ABS_CLAMP", "SCALE_RAW"])
def test_linear_bitnet32_same_result_across_the_gemv_gemm_boundary(ctx, b200, oracle, policy):
    # The tcgen05 kind::f16 GEMM (M >= 9) folds the block scale into f16 weights. Clamped scales (1e-3 is not an f16)
    # and raw scales whose |s| * 2 overflows f16 would make M = 9 differ from M = 8: such weights are flagged at
    # upload and stay on the f32-scale GEMV, so a token's output does not depend on the batch it travels in.
    rows, cols = 64, 256
    data = oracle.make_bitnet32(rows, cols, 21).reshape(-1, 10)
    specials = [0x0001, 0x1418, 0x7BFF, 0xFBFF, 0x6400] if policy == "SCALE_ABS_CLAMP" else [0x7BFF, 0xFBFF, 0x7800]
    for i, bits in enumerate(specials):
        data[3 * i, 8] = bits & 0xFF
        data[3 * i, 9] = bits >> 8
    x = oracle.fill_normal(9 * cols, 6).reshape(9, cols)
    q, s = oracle.quantize_rows_i8(x)
    pol = getattr(b200, policy)
    w = ctx.weight(data.ravel(), b200.LAYOUT_BITNET32_F16, rows, cols, scale_policy=pol)
    y9 = ctx.linear_i8(w, dev(q), dev(s)).cpu().numpy()
    y8 = ctx.linear_i8(w, dev(q[:8]), dev(s[:8])).cpu().numpy()
    assert np.array_equal(y9[:8], y8)
    ry = oracle.linear_bitnet32_i8(data.ravel(), rows, cols, q, s, scale_policy=pol)
    assert np.all(np.isfinite(y9)) and _rel(y9, ry) < 1e-5


def test_linear_bitnet32_rejects_bad_sizes(ctx, b200):
    with pytest.raises(b200.B200Error, match="size mismatch"):
        ctx.weight(np.zeros(79, np.uint8), b200.LAYOUT_BITNET32_F16, 4, 64)
    with pytest.raises(b200.B200Error, match="cols % 32"):
        ctx.weight(np.zeros(80, np.uint8), b200.LAYOUT_BITNET32_F16, 4, 40)


# =========================================================================================
# launch_i2s_matmul fill (map C, f32 activations, per-block f32 scales)
# =========================================================================================
def test_i2s_matmul_f32_reference_vectors(b200, oracle):
    def run(act, packed, scales, m, n, k, bs):
        out = np.full(m * n, np.nan, np.float32)
        b200.i2s_matmul_f32(np.asarray(act, np.float32), np.asarray(packed, np.uint8),
                            np.asarray(scales, np.float32), out, m, n, k, bs)
        return out.tolist()
    p = oracle.pack_i2s([1, 1, 1, 1])
    assert run([1, 2, 3, 4], [p, p], [1, 1], 1, 2, 4, 4) == [10.0, 10.0]            # snapshot [10,10]
    assert run([1, 2, 3, 4], [oracle.pack_i2s([1, -1, 0, 1])], [2.0], 1, 1, 4, 4) == [6.0]   # snapshot [6]
    assert run(np.ones(64), np.full(16, 0b01010101), [1.0, 3.0], 1, 1, 64, 32) == [128.0]   # :589-613
    m, n, k = 5, 37, 512
    packed = oracle.fill_bytes(n * k // 4, 21)
    act = oracle.fill_normal(m * k, 22)
    scales = oracle.fill_uniform(n * (k // 32), 23, 0.5, 1.5)
    out = np.zeros(m * n, np.float32)
    b200.i2s_matmul_f32(act, packed, scales, out, m, n, k, 32)
    ref = oracle.i2s_matmul(act, packed, scales, m, n, k, 32, "f32")
    assert _rel(out.reshape(m, n), ref) < 1e-5
    with pytest.raises(b200.KernelError, match="scales too small"):
        b200.i2s_matmul_f32(np.ones(64, np.float32), np.zeros(16, np.uint8), np.ones(1, np.float32),
                            np.zeros(1, np.float32), 1, 1, 64, 32)


# =========================================================================================
# threading: every export is re-entrant
# =========================================================================================
def test_concurrent_host_threads(b200, oracle):
    import threading
    rows, cols = 64, 512
    qs = oracle.fill_bytes(rows * cols // 4, 31)
    errs = []

    def work(i):
        try:
            k = b200.B200Kernel()
            x = (oracle.fill_normal(cols, 100 + i) * 20).round().astype(np.float32)
            y = np.zeros(rows, np.float32)
            for _ in range(3):
                b200.gemv_qk256(qs, x, y, rows, cols, cols // 4)
                assert np.array_equal(y, oracle.gemv_qk256(qs, x, rows, cols, cols // 4, "scalar"))
            k.close()
        except Exception as e:  # noqa: BLE001
            errs.append(e)

    ts = [threading.Thread(target=work, args=(i,)) for i in range(4)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errs, errs


# =========================================================================================
# the decode fast path: [RMSNorm ->] act-quant -> linear -> epilogue in one launch
# (gemv_stream.cu). It must equal the composition of the pinned pieces bit for bit.
# =========================================================================================
def _silu_mul(g, u):
    g = g.astype(np.float32)
    return (g / (np.float32(1.0) + np.exp(-g, dtype=np.float32))) * u.astype(np.float32)


@pytest.mark.parametrize("rows,cols", [(640, 2560), (3840, 2560), (2560, 6912), (512, 256), (100, 1024), (64, 4096)])
@pytest.mark.parametrize("M", [1, 2, 3, 4, 5, 8])
@pytest.mark.parametrize("norm", [True, False])
def test_fused_linear_qk256(ctx, b200, oracle, rows, cols, M, norm):
    qs = oracle.fill_bytes(rows * cols // 4, 0xB2000000 + rows + M)
    x = oracle.fill_normal(M * cols, 77 + cols + M, 2.0).reshape(M, cols)
    g = oracle.fill_uniform(cols, 78, 0.5, 1.5)
    w = ctx.weight(qs, b200.LAYOUT_QK256, rows, cols)
    y, q, s = ctx.fused_linear(w, dev(x), dev(g) if norm else None, 1e-5, want_q=True)
    # the prologue is the two-step API's prologue, bit for bit
    if norm:
        q2, s2 = ctx.rmsnorm_actquant(dev(x), dev(g), 1e-5)
    else:
        q2, s2 = ctx.actquant_i8(dev(x))
        rq, rs = oracle.quantize_rows_i8(x)
        assert np.array_equal(q.cpu().numpy(), rq) and np.array_equal(s.cpu().numpy(), rs)
    assert torch.equal(q, q2) and torch.equal(s, s2)
    racc, ry = oracle.linear_qk256_i8(qs, rows, cols, q.cpu().numpy(), s.cpu().numpy())
    assert np.array_equal(y.cpu().numpy(), ry)
    # residual epilogue (in place) and SwiGLU epilogue
    res = oracle.fill_normal(M * rows, 5).reshape(M, rows)
    yr = dev(res)
    ctx.fused_linear(w, dev(x), dev(g) if norm else None, 1e-5, epilogue=b200.EPI_RESIDUAL, y=yr, residual=yr)
    assert np.array_equal(yr.cpu().numpy(), res + ry)
    h = ctx.fused_linear(w, dev(x), dev(g) if norm else None, 1e-5, epilogue=b200.EPI_SWIGLU).cpu().numpy()
    ref_h = _silu_mul(ry[:, 0::2], ry[:, 1::2])
    assert h.shape == (M, rows // 2)
    assert np.max(np.abs(h - ref_h)) <= 2e-6 * np.max(np.abs(ref_h))      # expf: device vs libm, <= 2 ulp


@pytest.mark.parametrize("rows,cols", [(640, 2560), (2560, 6912), (64, 256)])
@pytest.mark.parametrize("M", [1, 3])
def test_fused_linear_bitnet32(ctx, b200, oracle, rows, cols, M):
    data = oracle.make_bitnet32(rows, cols, 0xB2000000 + rows)
    x = oracle.fill_normal(M * cols, 99 + rows, 2.0).reshape(M, cols)
    g = oracle.fill_uniform(cols, 78, 0.5, 1.5)
    w = ctx.weight(data, b200.LAYOUT_BITNET32_F16, rows, cols)
    y, q, s = ctx.fused_linear(w, dev(x), dev(g), 1e-5, want_q=True)
    q2, s2 = ctx.rmsnorm_actquant(dev(x), dev(g), 1e-5)
    assert torch.equal(q, q2) and torch.equal(s, s2)
    ry = oracle.linear_bitnet32_i8(data, rows, cols, q.cpu().numpy(), s.cpu().numpy())
    assert _rel(y.cpu().numpy(), ry) < 1e-5
    assert _rel(y.cpu().numpy(), oracle.linear_bitnet32_dequant(data, rows, cols, q.cpu().numpy(), s.cpu().numpy())) < 1e-3


def test_fused_linear_edge_rows_and_errors(ctx, b200, oracle):
    rows, cols = 64, 512
    qs = oracle.fill_bytes(rows * cols // 4, 3)
    w = ctx.weight(qs, b200.LAYOUT_QK256, rows, cols)
    x = np.zeros((3, cols), np.float32)          # all-zero token -> scale 0 -> outputs 0
    x[1, 7] = -3.0                                # single spike
    x[2] = (np.arange(cols) % 255 - 127) * 0.5    # .5 rounding boundaries
    y, q, s = ctx.fused_linear(w, dev(x), want_q=True)
    rq, rs = oracle.quantize_rows_i8(x)
    assert np.array_equal(q.cpu().numpy(), rq) and np.array_equal(s.cpu().numpy(), rs)
    _, ry = oracle.linear_qk256_i8(qs, rows, cols, rq, rs)
    assert np.array_equal(y.cpu().numpy(), ry) and not y[0].any().item()
    with pytest.raises(b200.B200Error, match="1..8"):
        ctx.fused_linear(w, torch.zeros(9, cols, device="cuda"))
    w2 = ctx.weight(oracle.fill_bytes(4 * 128, 1), b200.LAYOUT_QK256, 4, 300)
    with pytest.raises(b200.B200Error, match="not covered"):
        ctx.fused_linear(w2, torch.zeros(1, 300, device="cuda"))


# =========================================================================================
# a12: the legacy QK256 entry points (crates/bitnet-quantization/src/qk256_dispatch.rs:20-99)
# =========================================================================================
@pytest.mark.parametrize("rows,cols", [(1, 256), (7, 512), (256, 256), (33, 2560), (300, 1024)])
def test_legacy_qk256_gemv_scalar_matches_reference(b200, oracle, rows, cols):
    packed = oracle.fill_bytes(rows * cols // 4, 0xB2000D00 + rows)
    scales = oracle.fill_uniform(rows * (cols // 256), 17 + rows, 0.25, 2.0)
    act = oracle.fill_normal(cols, 23 + cols)
    out = np.full(rows, np.nan, np.float32)
    b200.qk256_gemv_scalar(out, rows, cols, packed, scales, act)
    ref = oracle.qk256_gemv_scalar_legacy(rows, cols, packed, scales, act)     # map D, per-block f32 scales
    assert np.allclose(out, ref, rtol=1e-5, atol=1e-5 * np.abs(ref).max())      # f32 sum order inside a block differs


def test_legacy_qk256_gemv_forwards_to_gemv_qk256(b200, oracle):
    # qk256_dispatch.rs:20-43: the legacy signature validates `scales` and does not consume it (map A, no scale)
    rows, cols = 64, 768
    packed = oracle.fill_bytes(rows * cols // 4, 0xB2000D77)
    act = np.round(oracle.fill_normal(cols, 5) * 8).astype(np.float32)          # integer-valued: exact in f32
    out = np.zeros(rows, np.float32)
    b200.qk256_gemv(out, rows, cols, packed, np.full(rows * cols // 256, 123.0, np.float32), act)
    assert np.array_equal(out, oracle.gemv_qk256(packed, act, rows, cols, cols // 4, "scalar"))
    # the smoke test of the reference (qk256_dispatch.rs:106-124): all codes 1 -> -1, activations 0.5
    out = np.zeros(256, np.float32)
    b200.qk256_gemv(out, 256, 256, np.full(256 * 64, 0x55, np.uint8), np.ones(256, np.float32), np.full(256, 0.5, np.float32))
    assert np.all(out == -128.0)


@pytest.mark.parametrize("fn", ["qk256_gemv", "qk256_gemv_scalar"])
def test_legacy_qk256_contract_messages(b200, fn):
    f = getattr(b200, fn)
    ok = dict(rows=2, cols=256, packed=np.zeros(128, np.uint8), scales=np.ones(2, np.float32), act=np.zeros(256, np.float32))
    with pytest.raises(AssertionError, match="Output length mismatch"):
        f(np.zeros(3, np.float32), ok["rows"], ok["cols"], ok["packed"], ok["scales"], ok["act"])
    with pytest.raises(AssertionError, match="Activation length mismatch"):
        f(np.zeros(2, np.float32), 2, 256, ok["packed"], ok["scales"], np.zeros(255, np.float32))
    with pytest.raises(AssertionError, match="Cols must be multiple of QK256"):
        f(np.zeros(2, np.float32), 2, 128, np.zeros(64, np.uint8), ok["scales"], np.zeros(128, np.float32))
    with pytest.raises(AssertionError, match="Packed weight size mismatch"):
        f(np.zeros(2, np.float32), 2, 256, np.zeros(127, np.uint8), ok["scales"], ok["act"])
    with pytest.raises(AssertionError, match="Scales length mismatch"):
        f(np.zeros(2, np.float32), 2, 256, ok["packed"], np.ones(3, np.float32), ok["act"])


# =========================================================================================
# a16: QuantizedLinear, the in-tree consumer of the provider seam (quantized_linear.rs:704-802)
# =========================================================================================
@pytest.mark.parametrize("scale_kind", ["per_feature", "per_block"])
def test_quantized_linear_consumer_path(b200, kernel, oracle, scale_kind):
    batch, fin, fout, bs = 5, 96, 40, 32
    codes = (oracle.fill_bytes(fin * fout, 0xB2000A16) & 3).astype(np.int8) - 2        # i8 in [-2, 1], [in, out] row-major
    packed = oracle.pack_2bit_values(codes)
    n_scales = fout if scale_kind == "per_feature" else fin * fout // bs
    scales = oracle.fill_uniform(n_scales, 31, 0.5, 1.5)
    x = oracle.fill_normal(batch * fin, 9, 1.2).reshape(batch, fin)
    ql = b200.QuantizedLinear(packed, scales, fin, fout, block_size=bs, provider=kernel)
    assert np.array_equal(b200.unpack_2bit_values(packed, fin * fout), oracle.unpack_2bit_values(packed, fin * fout))
    assert np.array_equal(ql.prepare_quantized_weights_i2s(), (codes + 2).astype(np.uint8))
    y = ql.forward(x)
    # the same pipeline stated with the oracle's kernel: quantize_input_i2s -> FallbackKernel::matmul_i2s -> scales
    q = np.clip(x, -2.0, 1.0)
    q = (np.sign(q) * np.floor(np.abs(q) + 0.5)).astype(np.int8)
    ref = oracle.matmul_i2s(q.reshape(-1), (codes + 2).astype(np.uint8), batch, fout, fin, "fallback").reshape(batch, fout)
    if scale_kind == "per_feature":
        idx = np.arange(fout)
    else:
        idx = np.minimum(np.arange(fout) * fin // bs, n_scales - 1)                    # quantized_linear.rs:791-796
    assert np.array_equal(y, ref * scales[idx])
