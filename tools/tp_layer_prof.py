#!/usr/bin/env python
"""Where one layer of the 8B-class batch-64 tensor-parallel decode step goes (one GPU, tp = 1 shapes or the per-rank shapes
of --tp N): every op of TPDecoder.layers_fused_gen is replayed ALONE from a CUDA graph over all layers (its own weights
per layer, so the weight stream stays cold as in the step) and timed with CUDA events; the sum is compared with the graph of
the whole layer loop. Not a profiler run: these are in-graph, warm-instruction-cache times including the launch gap.

    python tools/tp_layer_prof.py [--tp 1] [--layers 32] [--batch 64]"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tp", type=int, default=1, help="time the per-rank shapes of this tensor-parallel degree (rank 0)")
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--ctx-len", type=int, default=48, help="tokens already in the KV cache")
    args = ap.parse_args()
    import torch
    import bitnet_rs_b200 as b2

    torch.cuda.set_device(0)
    shape = b2.ModelShape.llama_8b_class(batch=args.batch, max_ctx=args.ctx_len + 64)
    ctx = b2.Context(0)
    fabs = b2.TpFabric.emulated(torch, torch.device("cuda", 0), 1, shape.batch, shape.hidden)
    fab = fabs[0]
    dec = b2.TPDecoder(ctx, shape, 0, args.tp, n_layers=args.layers, fabric=fab)
    dec.pos.fill_(args.ctx_len)
    s, B = shape, shape.batch
    x = torch.randn(B, s.hidden, device="cuda")
    L0 = dec.layers[0]

    # one set of intermediate tensors, produced by running the ops once in order
    q8, s8 = ctx.rmsnorm_actquant(x, L0["attn_norm"], s.rms_eps)
    qkv = ctx.linear_i8(L0["wqkv"], q8, s8)
    att = ctx.rope_attn_decode(qkv, dec.cos, dec.sin, dec.kc[0], dec.vc[0], dec.pos, dec.nh, dec.nkv, s.head_dim, s.max_ctx)
    ctx.row_absmax(att, absmax=fab.amax_o)
    q8o, so = ctx.tp_xchg_max_actquant(fab.desc(fab.SLOT_O_MAX, 3), att)
    ctx.linear_i8(L0["wo"], q8o, so, acc_out=fab.acc_o)
    q8f, s8f = ctx.tp_xchg_sum_rmsnorm_actquant(fab.desc(fab.SLOT_O_SUM, 3), so, x, L0["ffn_norm"], s.rms_eps)
    gu = ctx.linear_i8(L0["wgu"], q8f, s8f)
    h, _ = ctx.row_absmax(gu, swiglu=True, absmax=fab.amax_down)
    q8d, sd = ctx.tp_xchg_max_actquant(fab.desc(fab.SLOT_DOWN_MAX, 3), h)
    ctx.linear_i8(L0["wdown"], q8d, sd, acc_out=fab.acc_down)
    torch.cuda.synchronize()

    ops = {
        "linear qkv": lambda l, L: ctx.linear_i8(L["wqkv"], q8, s8, y=qkv),
        "rope + attention": lambda l, L: ctx.rope_attn_decode(qkv, dec.cos, dec.sin, dec.kc[l], dec.vc[l], dec.pos, dec.nh,
                                                              dec.nkv, s.head_dim, s.max_ctx),
        "row_absmax (att)": lambda l, L: ctx.row_absmax(att, absmax=fab.amax_o),
        "xchg max + quant (o in)": lambda l, L: ctx.tp_xchg_max_actquant(fab.desc(fab.SLOT_O_MAX, 3), att),
        "linear o (acc)": lambda l, L: ctx.linear_i8(L["wo"], q8o, so, acc_out=fab.acc_o),
        "xchg sum + norm + quant (ffn in)": lambda l, L: ctx.tp_xchg_sum_rmsnorm_actquant(
            fab.desc(fab.SLOT_O_SUM, 3), so, x, L["ffn_norm"], s.rms_eps),
        "linear gate+up": lambda l, L: ctx.linear_i8(L["wgu"], q8f, s8f, y=gu),
        "swiglu + row_absmax": lambda l, L: ctx.row_absmax(gu, swiglu=True, absmax=fab.amax_down),
        "xchg max + quant (down in)": lambda l, L: ctx.tp_xchg_max_actquant(fab.desc(fab.SLOT_DOWN_MAX, 3), h),
        "linear down (acc)": lambda l, L: ctx.linear_i8(L["wdown"], q8d, sd, acc_out=fab.acc_down),
        "xchg sum + norm + quant (next in)": lambda l, L: ctx.tp_xchg_sum_rmsnorm_actquant(
            fab.desc(fab.SLOT_DOWN_SUM, 3), sd, x, L["attn_norm"], s.rms_eps),
    }

    def timed_graph(body, reps=5):
        body()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            ctx.use_torch_stream()
            body()
        ctx.use_torch_stream()
        g.replay()
        torch.cuda.synchronize()
        best = 1e30
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            g.replay()
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) * 1e3)
        return best

    nl = len(dec.layers)
    rows, total = [], 0.0
    for name, fn in ops.items():
        def body(fn=fn):
            for l, L in enumerate(dec.layers):
                fn(l, L)
        us = timed_graph(body) / nl
        total += us
        rows.append({"op": name, "us_per_layer": round(us, 2), "kernel": b2.last_linear_kernel() if name.startswith("linear") else None})
        print(json.dumps(rows[-1]), flush=True)

    def whole():
        for l, L in enumerate(dec.layers):
            for fn in ops.values():
                fn(l, L)
    us_whole = timed_graph(whole) / nl
    print(json.dumps({"tp": args.tp, "batch": B, "layers": nl, "sum_of_ops_us_per_layer": round(total, 2),
                      "whole_layer_us": round(us_whole, 2)}))


if __name__ == "__main__":
    main()
