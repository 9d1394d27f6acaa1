"""Tensor-parallel path on ONE GPU: the ranks' shards are stepped in lockstep on the same device and
the two exchanges (all-reduce MAX of the per-token absmax, all-reduce SUM of int32 partials) are
applied in place -- the kernels, shard layouts and epilogues are the ones the NCCL run uses.
Bar: bit-exact against tp = 1 (int32 partials add up to the unsharded accumulator)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(b200):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return b200.Context(0)


@pytest.mark.parametrize("tp_n", [2, 4, 8])
@pytest.mark.parametrize("rows,cols,M", [(4096, 4096, 64), (4096, 14336, 64), (2560, 2048, 3), (512, 1024, 16)])
def test_row_parallel_linear_bit_exact(ctx, b200, oracle, rows, cols, M, tp_n):
    from bitnet_rs_b200 import tp
    if (cols // tp_n) % 256:
        pytest.skip("K/tp is not a multiple of the QK256 block")
    qs = oracle.fill_bytes(rows * cols // 4, 0xB2000020 + rows)
    x = oracle.fill_normal(M * cols, 5 + M, 1.5).reshape(M, cols)
    q, s = oracle.quantize_rows_i8(x)
    racc, ry = oracle.linear_qk256_i8(qs, rows, cols, q, s)          # the unsharded reference
    xd = torch.from_numpy(x).cuda()
    kl = cols // tp_n
    ams, xs, ws = [], [], []
    for r in range(tp_n):
        xs.append(xd[:, r * kl:(r + 1) * kl].contiguous())
        ams.append(ctx.row_absmax(xs[r]))
        ws.append(ctx.weight(tp.shard_cols(qs, b200.LAYOUT_QK256, rows, cols, r * kl, (r + 1) * kl),
                             b200.LAYOUT_QK256, rows, kl))
    am = torch.stack(ams).amax(dim=0)                                 # all-reduce MAX
    accs, scales = [], []
    for r in range(tp_n):
        qr, sr = ctx.actquant_i8_absmax(xs[r], am)
        assert np.array_equal(qr.cpu().numpy(), q[:, r * kl:(r + 1) * kl])   # same int8 as the unsharded quantiser
        _, acc = ctx.linear_i8(ws[r], qr, sr, want_acc=True)
        accs.append(acc)
        scales.append(sr)
    total = torch.stack(accs).sum(dim=0, dtype=torch.int32)           # all-reduce SUM (int32)
    assert np.array_equal(total.cpu().numpy(), racc)
    assert np.array_equal(scales[0].cpu().numpy(), s)
    res = torch.from_numpy(oracle.fill_normal(M * rows, 9, 1.0).reshape(M, rows)).cuda()
    want = res + torch.from_numpy(ry).cuda()
    ctx.residual_acc(total, scales[0], res)
    assert torch.equal(res, want)


def test_swiglu_absmax_matches_fused_quantiser(ctx, b200, oracle):
    M, F = 7, 1792
    gu = torch.from_numpy(oracle.fill_normal(M * 2 * F, 11, 2.0).reshape(M, 2 * F)).cuda()
    h, am = ctx.row_absmax(gu, swiglu=True)
    q, s = ctx.actquant_i8_absmax(h, am)
    q2, s2 = ctx.actquant_i8(h)                                       # the row's own absmax: same thing at tp = 1
    assert torch.equal(q, q2) and torch.equal(s, s2)
    g, u = gu[:, 0::2], gu[:, 1::2]
    assert torch.allclose(h, torch.nn.functional.silu(g) * u, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("tp_n", [2, 4])
def test_decoder_tp_equals_tp1(ctx, b200, tp_n):
    from bitnet_rs_b200 import tp
    shape = b200.ModelShape(hidden=1024, ffn=2048, n_layers=2, n_heads=8, n_kv_heads=4, head_dim=128,
                            vocab=4096, max_ctx=16, batch=16)
    ref = b200.TPDecoder(ctx, shape, 0, 1)
    ranks = [b200.TPDecoder(ctx, shape, r, tp_n) for r in range(tp_n)]
    tok0 = (torch.arange(shape.batch, device="cuda") * 7919) % shape.vocab
    for d in [ref] + ranks:
        d.tokens.copy_(tok0)
    for step in range(4):
        tp.run_emulated([ref.step_gen()])
        tp.run_emulated([d.step_gen() for d in ranks])
        for d in ranks:
            assert torch.equal(d.last_hidden, ref.last_hidden), "hidden state differs at step %d" % step
            assert torch.equal(d.tokens, ref.tokens)
    assert int(ref.pos) == 4
