"""tests/golden/linear_golden.npz (written by tests/golden/make_golden.py): frozen inputs and outputs of the hot path
for both I2_S layouts. The CPU test holds the oracle to the file, the GPU test holds the CUDA path (through the C ABI) to
the same file, without the oracle in the loop: the two cannot drift together.

Bars as everywhere: quantiser and QK256 bit-exact; BitNet32-F16 within 1e-5 relative of the oracle's block-order sum
(fp32 summation order) and 1e-3 of dequantise-then-f32 (the north_star tolerance)."""
import os
import sys

import numpy as np
import pytest

torch = pytest.importorskip("torch")

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
sys.path.insert(0, GOLDEN)
from make_golden import CASES, compute  # noqa: E402

PATH = os.path.join(GOLDEN, "linear_golden.npz")


@pytest.fixture(scope="module")
def golden():
    z = np.load(PATH)
    return {name: {k.split(".", 1)[1]: z[k] for k in z.files if k.startswith(name + ".")} for name, *_ in CASES}


def _rel(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def test_fixture_holds_every_case(golden):
    for name, layout, rows, cols, M in CASES:
        g = golden[name]
        assert g["x"].shape == (M, cols) and g["q"].shape == (M, cols) and g["s"].shape == (M,) and g["y"].shape == (M, rows)
        assert g["w"].size == (rows * (-(-cols // 256) * 64) if layout == "qk256" else rows * cols // 32 * 10)
        assert np.isfinite(g["y"]).all()


@pytest.mark.parametrize("seed,case", list(enumerate(CASES)), ids=[c[0] for c in CASES])
def test_oracle_reproduces_the_fixture(oracle, golden, seed, case):
    g = golden[case[0]]
    now = compute(oracle, *case, seed)
    assert sorted(now) == sorted(g)
    for k in g:
        assert np.array_equal(now[k], g[k]), k


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_cuda_path_reproduces_the_fixture(b200, golden, case, monkeypatch):
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    name, layout, rows, cols, M = case
    g = golden[name]
    ctx = b200.Context(0)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    w = ctx.weight(g["w"], b200.LAYOUT_QK256 if layout == "qk256" else b200.LAYOUT_BITNET32_F16, rows, cols)
    q, s = ctx.actquant_i8(dev(g["x"]))
    assert np.array_equal(q.cpu().numpy(), g["q"]) and np.array_equal(s.cpu().numpy(), g["s"])
    ys = {}
    ys["linear_i8"] = ctx.linear_i8(w, dev(g["q"]), dev(g["s"])).cpu().numpy()
    if layout == "qk256":   # gemv_qk256 on f32 activations (no quantiser): fp32 sum order differs, the reference's 1e-4 bar
        d = np.abs(ctx.linear_f32(w, dev(g["x"])).cpu().numpy() - g["y_f32"])
        assert np.all((d <= 1e-4 * np.sqrt(-(-cols // 256))) | (d <= 1e-4 * np.abs(g["y_f32"])))
    if M <= 8 and cols % 256 == 0:
        ys["fused_linear"] = ctx.fused_linear(w, dev(g["x"])).cpu().numpy()         # the one-launch decode form
    monkeypatch.setenv("BITNET_B200_SMALLM_MAX", "0")       # the GEMV / tcgen05 GEMMs alone, without gemm_small.cu
    ys["linear_i8, no gemm_small"] = ctx.linear_i8(w, dev(g["q"]), dev(g["s"])).cpu().numpy()
    for how, y in ys.items():
        if layout == "qk256":
            assert np.array_equal(y, g["y"]), how
        else:
            assert _rel(y, g["y"]) < 2e-5, how
            assert _rel(y, g["y_dequant"]) < 1e-3, how
    if layout == "qk256":
        _, acc = ctx.linear_i8(w, dev(g["q"]), dev(g["s"]), want_acc=True)
        assert np.array_equal(acc.cpu().numpy(), g["acc"])
    ctx.close()
