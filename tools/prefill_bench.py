"""BASELINE.json configs[3]: 2B-4T-shape prefill of a 2048-token synthetic prompt, batch 8 (M = 16384 token
rows), the tcgen05 kind::i8 GEMM path. Times the ternary-linear pipeline of every layer -- RMSNorm+act-quant,
qkv GEMM, act-quant, o GEMM, RMSNorm+act-quant, gate/up GEMM, SwiGLU+act-quant, down GEMM -- with CUDA events.
Prefill attention is outside the ternary path and is not run: the number is linears-only.

  python tools/prefill_bench.py [--tokens 2048] [--batch 8] [--layers 30] [--reps 3]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bitnet_rs_b200 as b2  # noqa: E402
from bitnet_rs_b200 import tp  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tokens", type=int, default=2048)
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--layers", type=int, default=30)
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()
    M = args.tokens * args.batch
    s = b2.ModelShape.bitnet_2b4t()
    ctx = b2.Context(0)
    H, F, QD = s.hidden, s.ffn, s.n_heads * s.head_dim
    layers = []
    for l in range(args.layers):
        w = tp.shard_layer(s, tp.fast_layer(s, l), 0, 1)          # tp = 1: the fused qkv / interleaved gate-up tensors
        ent = {k: ctx.weight(w[k][0], s.layout, w[k][1], w[k][2]) for k in ("wqkv", "wo", "wgu", "wdown")}
        ent["attn_norm"] = torch.from_numpy(w["attn_norm"]).cuda()
        ent["ffn_norm"] = torch.from_numpy(w["ffn_norm"]).cuda()
        layers.append(ent)
    x = torch.randn(M, H, device="cuda")
    q8 = torch.empty(M, F, dtype=torch.int8, device="cuda")
    qH, qQ, qF = q8[:, :H].contiguous(), torch.empty(M, QD, dtype=torch.int8, device="cuda"), q8
    sc = torch.empty(M, device="cuda")
    qkv = torch.empty(M, QD + 2 * s.n_kv_heads * s.head_dim, device="cuda")
    y = torch.empty(M, H, device="cuda")
    gu = torch.empty(M, 2 * F, device="cuda")

    def one_pass():
        for L in layers:
            ctx.rmsnorm_actquant(x, L["attn_norm"], s.rms_eps, q=qH, scale=sc)
            ctx.linear_i8(L["wqkv"], qH, sc, y=qkv)
            att = qkv[:, :QD].contiguous()                        # stand-in for the attention output (same shape)
            ctx.actquant_i8(att, q=qQ, scale=sc)
            ctx.linear_i8(L["wo"], qQ, sc, y=y)
            ctx.rmsnorm_actquant(y, L["ffn_norm"], s.rms_eps, q=qH, scale=sc)
            ctx.linear_i8(L["wgu"], qH, sc, y=gu)
            h, am = ctx.row_absmax(gu, swiglu=True)
            ctx.actquant_i8_absmax(h, am, q=qF, scale=sc)
            ctx.linear_i8(L["wdown"], qF, sc, y=y)

    one_pass()
    torch.cuda.synchronize()
    # GEMMs alone
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    best = 1e30
    for _ in range(args.reps):
        ev[0].record()
        one_pass()
        ev[1].record()
        torch.cuda.synchronize()
        best = min(best, ev[0].elapsed_time(ev[1]))
    gemm_ms = 1e30
    for _ in range(args.reps):
        ev[2].record()
        for L in layers:
            ctx.linear_i8(L["wqkv"], qH, sc, y=qkv)
            ctx.linear_i8(L["wo"], qQ, sc, y=y)
            ctx.linear_i8(L["wgu"], qH, sc, y=gu)
            ctx.linear_i8(L["wdown"], qF, sc, y=y)
        ev[3].record()
        torch.cuda.synchronize()
        gemm_ms = min(gemm_ms, ev[2].elapsed_time(ev[3]))
    per_layer = sum(r * c for r, c in s.proj_shapes().values())
    ops = 2.0 * M * per_layer * args.layers
    print(json.dumps({
        "workload": "2B-4T-shape prefill, %d tokens x batch %d (M = %d), %d layers, QK256, linears + act-quant only"
                    % (args.tokens, args.batch, M, args.layers),
        "ms_linears_and_quant": best, "ms_gemm_only": gemm_ms, "int8_TOP": ops / 1e12,
        "TOPS_gemm_only": ops / gemm_ms / 1e9, "TOPS_with_quant": ops / best / 1e9,
        "frac_of_4500_TOPS_nominal": ops / gemm_ms / 1e9 / 4500.0,
        "prefill_tok_per_s_linears_only": M / (best * 1e-3)}))


if __name__ == "__main__":
    main()
