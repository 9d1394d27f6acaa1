"""bitnet_b200_linear_i8_static_scratch: the linear for callers that must not allocate device memory. Collected last
(a failure here must not hide the other tests). The new entry point is called through the
C ABI and compared with bitnet_b200_linear_i8 -- bit for bit, the bar of the other QK256 linear tests -- on two calls in a
row (the static scratch keeps the previous call's contents), and checked to allocate nothing."""
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


def _case(rows, cols, M, seed):
    rng = np.random.default_rng(seed)
    qs = rng.integers(0, 256, rows * cols // 4, dtype=np.uint8)
    q = torch.from_numpy(rng.integers(-127, 128, (M, cols)).astype(np.int8)).cuda()
    s = torch.from_numpy(rng.uniform(0.5, 2.0, M).astype(np.float32)).cuda()
    return qs, q, s


# split-K shapes (few output tiles, long K: the scratch is in use) and single-pass ones; ragged rows and tokens
@pytest.mark.parametrize("rows,cols,M", [(2560, 6912, 64), (4096, 4096, 33), (13824, 2560, 64), (200, 14336, 9),
                                         (2560, 2560, 320), (1000, 1024, 130)])
def test_static_scratch_linear_matches_linear_i8_twice_in_a_row(ctx, b200, monkeypatch, rows, cols, M):
    monkeypatch.setenv("BITNET_B200_SMALLM_MAX", "0")   # the reference call takes the same tcgen05 GEMM, not gemm_small.cu
    qs, q, s = _case(rows, cols, M, rows + cols + M)
    w = ctx.weight(qs, b200.LAYOUT_QK256, rows, cols)
    y_ref = ctx.linear_i8(w, q, s)                       # existing entry point (cudaMalloc'ed split-K scratch)
    assert b200.last_linear_kernel() == "gemm_tcgen05"
    y = torch.full((M, rows), float("nan"), device="cuda")
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    ctx.linear_i8_static_scratch(w, q, s, y)
    ctx.synchronize()
    # nothing was allocated behind the caller's back (a cudaMalloc of the scratch would be 4 MiB)
    assert free_before - torch.cuda.mem_get_info()[0] < (1 << 20)
    assert torch.equal(y, y_ref)
    # second call straight after, other activations: the scratch holds what the first call left
    q2 = (-q).contiguous()
    s2 = (s * 0.5).contiguous()
    y2_ref = ctx.linear_i8(w, q2, s2)
    y2 = torch.full((M, rows), float("nan"), device="cuda")
    ctx.linear_i8_static_scratch(w, q2, s2, y2)
    ctx.linear_i8_static_scratch(w, q, s, y)             # and a third, back to back on the stream
    ctx.synchronize()
    assert torch.equal(y2, y2_ref)
    assert torch.equal(y, y_ref)


def test_static_scratch_linear_rejects_what_it_cannot_hold(ctx, b200):
    rows, cols = 13824, 2560
    qs, q, s = _case(rows, cols, 512, 3)
    w = ctx.weight(qs, b200.LAYOUT_QK256, rows, cols)
    with pytest.raises(b200.B200Error, match="too large for the 4 MiB static scratch"):
        ctx.linear_i8_static_scratch(w, q, s, torch.empty(512, rows, device="cuda"))
    with pytest.raises(b200.B200Error, match="M >= 9"):     # decode-regime M: not the tcgen05 GEMM's
        ctx.linear_i8_static_scratch(w, q[:4].contiguous(), s[:4].contiguous(), torch.empty(4, rows, device="cuda"))
    # the limit itself is usable: 13824 x 64 needs 3.4 MiB
    y = torch.empty(64, rows, device="cuda")
    ctx.linear_i8_static_scratch(w, q[:64].contiguous(), s[:64].contiguous(), y)
    ctx.synchronize()
    assert torch.isfinite(y).all()
