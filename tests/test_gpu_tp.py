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


# ---- fused exchanges: the consumer kernels pull their peers' buffers themselves (csrc/tp_xchg.cuh) ---------------
@pytest.mark.parametrize("tp_n", [1, 2, 4])
def test_decoder_fused_exchanges_equal_collective_path(ctx, b200, tp_n):
    # same shards, same kernels around the exchanges; the all-reduces are replaced by kernels that read the peers'
    # partial accumulators / absmax directly (here: N emulated ranks on one GPU, signal and work launches split).
    # Bar: hidden state and tokens bit-identical to the tp = 1 decoder that uses no exchange at all.
    from bitnet_rs_b200 import tp
    shape = b200.ModelShape(hidden=1024, ffn=2048, n_layers=2, n_heads=8, n_kv_heads=4, head_dim=128,
                            vocab=4096, max_ctx=16, batch=16)
    ref = b200.TPDecoder(ctx, shape, 0, 1)
    fabs = b200.TpFabric.emulated(torch, torch.device("cuda", 0), tp_n, shape.batch, shape.hidden)
    ranks = [b200.TPDecoder(ctx, shape, r, tp_n, fabric=fabs[r], emulated=True) for r in range(tp_n)]
    tok0 = (torch.arange(shape.batch, device="cuda") * 7919) % shape.vocab
    for d in [ref] + ranks:
        d.tokens.copy_(tok0)
    for step in range(5):
        tp.run_emulated([ref.step_gen()])
        tp.run_emulated([d.step_gen() for d in ranks])
        for d in ranks:
            assert torch.equal(d.last_hidden, ref.last_hidden), "hidden state differs at step %d" % step
            assert torch.equal(d.tokens, ref.tokens)
    assert int(ranks[0].fabric.counters[0]) == 5 * shape.n_layers       # one epoch per exchange and slot


def test_fused_exchange_kernels_match_the_three_step_form(ctx, b200, oracle):
    # tp_xchg_sum_rmsnorm_actquant == all-reduce(SUM) ; residual_acc ; rmsnorm_actquant, bit for bit, incl. `normed`
    M, H, n = 7, 4096, 4
    fabs = b200.TpFabric.emulated(torch, torch.device("cuda", 0), n, M, H)
    g = torch.Generator(device="cuda").manual_seed(3)
    parts = [torch.randint(-200000, 200000, (M, H), dtype=torch.int32, device="cuda", generator=g) for _ in range(n)]
    for r in range(n):
        fabs[r].acc_down.copy_(parts[r])
    sc = torch.rand(M, device="cuda", generator=g) * 1e-3
    gamma = torch.rand(H, device="cuda", generator=g) + 0.5
    x0 = torch.randn(M, H, device="cuda", generator=g)
    total = torch.stack(parts).sum(dim=0, dtype=torch.int32)
    xr = x0.clone()
    ctx.residual_acc(total, sc, xr)
    qr, sr, nr = ctx.rmsnorm_actquant(xr, gamma, 1e-5, want_normed=True)
    for r in range(n):
        ctx.tp_xchg_sum_rmsnorm_actquant(fabs[r].desc(fabs[r].SLOT_DOWN_SUM, 1), sc, x0, gamma, 1e-5)   # signal only
    for r in range(n):
        x = x0.clone()
        q, s, nm = ctx.tp_xchg_sum_rmsnorm_actquant(fabs[r].desc(fabs[r].SLOT_DOWN_SUM, 2), sc, x, gamma, 1e-5, want_normed=True)
        assert torch.equal(x, xr) and torch.equal(q, qr) and torch.equal(s, sr) and torch.equal(nm, nr)
    # absmax exchange
    K = 1536
    xs = [torch.randn(M, K, device="cuda", generator=g) * (r + 1) for r in range(n)]
    for r in range(n):
        ctx.row_absmax(xs[r], absmax=fabs[r].amax_o)
    am = torch.stack([f.amax_o for f in fabs]).amax(dim=0)
    for r in range(n):
        ctx.tp_xchg_max_actquant(fabs[r].desc(fabs[r].SLOT_O_MAX, 1), xs[r])
    for r in range(n):
        q, s = ctx.tp_xchg_max_actquant(fabs[r].desc(fabs[r].SLOT_O_MAX, 2), xs[r])
        q2, s2 = ctx.actquant_i8_absmax(xs[r], am)
        assert torch.equal(q, q2) and torch.equal(s, s2)


def test_tp1_decoder_matches_the_oracle_model(ctx, b200, oracle):
    # the decoder-level glue of the TP path (tp_rope_kv / tp_attn_decode kernels, vocabulary-sharded LM head with
    # lowest-index ties) against the CPU oracle model on a small GQA shape, batch > 1: same greedy tokens
    from bitnet_rs_b200 import tp
    shape = b200.ModelShape(hidden=512, ffn=1024, n_layers=2, n_heads=8, n_kv_heads=2, head_dim=64,
                            vocab=2048, max_ctx=16, batch=3)
    dec = b200.TPDecoder(ctx, shape, 0, 1)
    om = oracle.Model(shape, n_seqs=shape.batch)
    for l in range(shape.n_layers):
        om.set_layer(l, tp.fast_layer(shape, l))
    emb, fn = b200.synth_embedding(shape)
    om.set_embedding(emb, fn)
    om.set_rope(*b200.rope_tables(shape))
    toks = [5, 77, 1234]
    dec.tokens.copy_(torch.tensor(toks, device="cuda"))
    for step in range(6):
        tp.run_emulated([dec.step_gen()])
        toks = [om.step(t, seq=b) for b, t in enumerate(toks)]
        assert dec.tokens.cpu().tolist() == toks, "step %d" % step
    om.close()


def test_8b_geometry_two_layer_decoder_matches_the_oracle_model(ctx, b200, oracle):
    # configs[4] geometry (hidden 4096, FFN 14336, GQA 32/8, head 128), 2 layers, small vocabulary: the tp = 2 fused
    # path against the CPU oracle model
    from bitnet_rs_b200 import tp
    shape = b200.ModelShape(hidden=4096, ffn=14336, n_layers=2, n_heads=32, n_kv_heads=8, head_dim=128,
                            vocab=1024, max_ctx=8, batch=2)
    fabs = b200.TpFabric.emulated(torch, torch.device("cuda", 0), 2, shape.batch, shape.hidden)
    ranks = [b200.TPDecoder(ctx, shape, r, 2, fabric=fabs[r], emulated=True) for r in range(2)]
    om = oracle.Model(shape, n_seqs=shape.batch)
    for l in range(shape.n_layers):
        om.set_layer(l, tp.fast_layer(shape, l))
    emb, fn = b200.synth_embedding(shape)
    om.set_embedding(emb, fn)
    om.set_rope(*b200.rope_tables(shape))
    toks = [11, 900]
    for d in ranks:
        d.tokens.copy_(torch.tensor(toks, device="cuda"))
    for step in range(4):
        tp.run_emulated([d.step_gen() for d in ranks])
        toks = [om.step(t, seq=b) for b, t in enumerate(toks)]
        for d in ranks:
            assert d.tokens.cpu().tolist() == toks, "step %d" % step
    om.close()


# the fused RoPE + KV append + GQA attention kernel against the two-kernel form it replaces and a torch fp32 reference:
# caches and the rotated q bit-equal (same arithmetic), attention within fp32 summation-order noise. Head counts, head
# sizes and context bounds cover 1 / 2 / 4 position slices per head, the first position, and a long context.
@pytest.mark.parametrize("NH,NKV,D,max_ctx,pos", [(8, 2, 64, 16, 0), (8, 2, 64, 16, 9), (32, 8, 128, 128, 57), (20, 5, 128, 600, 333),
                                                   (4, 4, 128, 2048, 1500), (16, 2, 256, 300, 77), (4, 1, 128, 5000, 4100)])
def test_fused_rope_attention_matches_the_two_kernel_form(ctx, b200, monkeypatch, NH, NKV, D, max_ctx, pos):
    B = 3
    g = torch.Generator(device="cuda").manual_seed(NH * 1000 + pos)
    qkv0 = torch.randn(B, (NH + 2 * NKV) * D, device="cuda", generator=g)
    kc0 = torch.randn(B, NKV, max_ctx, D, device="cuda", generator=g)
    vc0 = torch.randn(B, NKV, max_ctx, D, device="cuda", generator=g)
    ang = torch.rand(max_ctx, D // 2, device="cuda", generator=g) * 6.28
    cos, sin = ang.cos().contiguous(), ang.sin().contiguous()
    posd = torch.tensor(pos, dtype=torch.int32, device="cuda")
    outs = []
    for split in (True, False):
        if split:
            monkeypatch.setenv("BITNET_B200_TP_ATTN_SPLIT", "1")
        else:
            monkeypatch.delenv("BITNET_B200_TP_ATTN_SPLIT")
        qkv, kc, vc = qkv0.clone(), kc0.clone(), vc0.clone()
        l0 = b200.launch_count()
        att = ctx.rope_attn_decode(qkv, cos, sin, kc, vc, posd, NH, NKV, D, max_ctx)
        torch.cuda.synchronize()
        outs.append((qkv, kc, vc, att, b200.launch_count() - l0))
    (q_a, k_a, v_a, att_a, n_a), (q_b, k_b, v_b, att_b, n_b) = outs
    assert (n_a, n_b) == (2, 1)
    assert torch.equal(k_a, k_b) and torch.equal(v_a, v_b) and torch.equal(q_a, q_b)
    # torch fp32 reference on the rotated q and the appended caches
    R = NH // NKV
    q = q_b[:, :NH * D].view(B, NKV, R, D)
    K, V = k_b[:, :, :pos + 1], v_b[:, :, :pos + 1]
    sc = torch.einsum("bgrd,bgpd->bgrp", q.double(), K.double()) / (D ** 0.5)
    ref = torch.einsum("bgrp,bgpd->bgrd", sc.softmax(-1), V.double()).reshape(B, NH * D).float()
    scale = ref.abs().max().item()
    assert (att_b - ref).abs().max().item() < 2e-5 * max(scale, 1.0)
    assert (att_a - ref).abs().max().item() < 2e-5 * max(scale, 1.0)


# several position slices share one q row and slice 0 writes the rotated row back in place: a slice that read the row
# after that write would rotate it twice. Many sequences, many repeats: the attention must come out the same every time.
@pytest.mark.parametrize("NH,NKV,D,max_ctx,pos", [(4, 4, 128, 2048, 1500), (8, 2, 128, 1024, 700), (2, 1, 64, 6000, 37)])
def test_fused_rope_attention_slices_read_q_before_it_is_rotated_in_place(ctx, b200, NH, NKV, D, max_ctx, pos):
    B = 48
    g = torch.Generator(device="cuda").manual_seed(NH + pos)
    qkv0 = torch.randn(B, (NH + 2 * NKV) * D, device="cuda", generator=g)
    kc = torch.randn(B, NKV, max_ctx, D, device="cuda", generator=g)
    vc = torch.randn(B, NKV, max_ctx, D, device="cuda", generator=g)
    ang = torch.rand(max_ctx, D // 2, device="cuda", generator=g) * 6.28
    cos, sin = ang.cos().contiguous(), ang.sin().contiguous()
    posd = torch.tensor(pos, dtype=torch.int32, device="cuda")
    first = None
    for rep in range(25):
        qkv = qkv0.clone()
        att = ctx.rope_attn_decode(qkv, cos, sin, kc, vc, posd, NH, NKV, D, max_ctx)   # the cache row at pos is rewritten alike
        torch.cuda.synchronize()
        if first is None:
            first = att.clone()
            R = NH // NKV
            q = qkv[:, :NH * D].view(B, NKV, R, D)
            sc = torch.einsum("bgrd,bgpd->bgrp", q.double(), kc[:, :, :pos + 1].double()) / (D ** 0.5)
            ref = torch.einsum("bgrp,bgpd->bgrd", sc.softmax(-1), vc[:, :, :pos + 1].double()).reshape(B, NH * D).float()
            assert (att - ref).abs().max().item() < 2e-5 * max(ref.abs().max().item(), 1.0)
        else:
            assert torch.equal(att, first), "repeat %d differs" % rep
