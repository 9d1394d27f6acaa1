"""GEMV / act-quant micro-benchmark on one B200 (CUDA events, weights rotated through
copies larger than L2). Writes JSON lines to stdout. Not the judged bench (see bench.py)."""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bitnet_rs_b200 as b2  # noqa: E402

SHAPES = {"q/o": (2560, 2560), "kv": (640, 2560), "qkv": (3840, 2560), "gate/up": (6912, 2560),
          "gateup": (13824, 2560), "down": (2560, 6912), "8b_qkv": (6144, 4096), "8b_down": (4096, 14336)}


def rand_weight(ctx, layout, rows, cols, seed):
    rng = np.random.default_rng(seed)
    if layout == b2.LAYOUT_QK256:
        data = rng.integers(0, 256, rows * cols // 4, dtype=np.uint8)
    else:
        nb = rows * cols // 32
        codes = rng.integers(0, 256, (nb, 8), dtype=np.uint8)
        sc = (rng.uniform(0.5, 1.5, nb) / np.sqrt(cols)).astype(np.float16).view(np.uint8).reshape(nb, 2)
        data = np.concatenate([codes, sc], 1).ravel()
    return ctx.weight(data, layout, rows, cols)


def graph_time(ctx, fn, iters, reps=5):
    """Capture `iters` back-to-back launches in one CUDA graph; time replays with events."""
    for i in range(3):
        fn(i)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        ctx.use_torch_stream()
        for i in range(iters):
            fn(i)
    ctx.use_torch_stream()
    g.replay()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e3 / iters)
    return best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ms", default="1,2,4,8,16")
    ap.add_argument("--shapes", default="qkv,gateup,down,q/o,kv")
    ap.add_argument("--layouts", default="qk256,bitnet32")
    ap.add_argument("--iters", type=int, default=200)
    ap.add_argument("--l2mb", type=int, default=300)
    ap.add_argument("--no-actquant", action="store_true")
    args = ap.parse_args()
    ctx = b2.Context(0)
    peak = 6548.5
    try:
        peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    for lname in args.layouts.split(","):
        layout = b2.LAYOUT_QK256 if lname == "qk256" else b2.LAYOUT_BITNET32_F16
        for sname in args.shapes.split(","):
            rows, cols = SHAPES[sname]
            wbytes = rows * cols // 4 if layout == b2.LAYOUT_QK256 else rows * cols // 32 * 10
            ncopies = max(2, (args.l2mb * 1024 * 1024) // wbytes + 1)
            ws = [rand_weight(ctx, layout, rows, cols, i) for i in range(min(ncopies, 64))]
            for M in [int(m) for m in args.ms.split(",")]:
                q = torch.randint(-127, 128, (M, cols), dtype=torch.int8, device="cuda")
                s = torch.rand(M, device="cuda") + 0.5
                y = torch.empty(M, rows, device="cuda")
                us = graph_time(ctx, lambda i: ctx.linear_i8(ws[i % len(ws)], q, s, y=y), args.iters)
                alg = wbytes + M * cols + 4 * M + 4 * M * rows
                print(json.dumps({"kernel": "linear_i8", "layout": lname, "shape": sname, "N": rows, "K": cols,
                                  "M": M, "us": round(us, 3), "alg_bytes": alg,
                                  "TOPS": round(2.0 * M * rows * cols / us / 1e6, 1), "GBs": round(alg / us / 1e3, 1), "frac_of_measured_peak": round(alg / us / 1e3 / peak, 4),
                                  "weight_copies": len(ws)}), flush=True)
            del ws
            torch.cuda.empty_cache()
    if args.no_actquant:
        return
    # act-quant
    for K in (2560, 6912):
        for M in (1, 16, 4096):
            x = torch.randn(M, K, device="cuda")
            g = torch.rand(K, device="cuda") + 0.5
            for fused in (False, True):
                qb = torch.empty(M, K, dtype=torch.int8, device="cuda")
                sb = torch.empty(M, device="cuda")
                fn = (lambda i: ctx.rmsnorm_actquant(x, g, 1e-5, q=qb, scale=sb)) if fused else \
                    (lambda i: ctx.actquant_i8(x, q=qb, scale=sb))
                us = graph_time(ctx, fn, 50)
                alg = M * K * 5 + 4 * M
                print(json.dumps({"kernel": "rmsnorm_actquant" if fused else "actquant_i8", "M": M, "K": K,
                                  "us": round(us, 3), "GBs": round(alg / us / 1e3, 1)}), flush=True)


if __name__ == "__main__":
    main()
