#!/usr/bin/env python
"""Time bitnet_b200_linear_i8 in the batched-decode regime (M = 8..128) on the 2B-4T and 8B-class
projection shapes (whole and tp = 2 shards). One JSON line per (shape, M); weights rotate through
more than the L2 so that every call streams its weights from HBM.

    python tools/smallm_bench.py [--layout qk256|bitnet32] [--out file.jsonl] [--ms 8,16,32,64,128]

BITNET_B200_NO_SMALLM=1 disables the mma.sync small-M kernel (A/B against the tcgen05 GEMM)."""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

SHAPES = {
    "2b.qkv": (3840, 2560), "2b.o": (2560, 2560), "2b.gateup": (13824, 2560), "2b.down": (2560, 6912),
    "8b.qkv": (6144, 4096), "8b.o": (4096, 4096), "8b.gateup": (28672, 4096), "8b.down": (4096, 14336),
    "8b/2.qkv": (3072, 4096), "8b/2.o": (4096, 2048), "8b/2.gateup": (14336, 4096), "8b/2.down": (4096, 7168),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layout", default="qk256")
    ap.add_argument("--ms", default="8,16,32,64,128")
    ap.add_argument("--shapes", default="")
    ap.add_argument("--out", default="")
    ap.add_argument("--acc", action="store_true", help="also request the int32 accumulators (TP row-parallel form)")
    args = ap.parse_args()
    import torch
    import bitnet_rs_b200 as b2
    torch.cuda.set_device(0)
    ctx = b2.Context(0)
    layout = b2.LAYOUT_QK256 if args.layout == "qk256" else b2.LAYOUT_BITNET32_F16
    from bench import _sweep_weight_bytes
    ms = [int(x) for x in args.ms.split(",")]
    names = [s for s in args.shapes.split(",") if s] or list(SHAPES)
    out = open(args.out, "w") if args.out else None

    def graph_time(fn, iters, reps=5):
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
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            g.replay()
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) * 1e3 / iters)
        return best

    for name in names:
        rows, cols = SHAPES[name]
        wbytes = rows * cols // 4 if args.layout == "qk256" else rows * cols // 32 * 10
        ncopies = min(48, max(2, (300 << 20) // wbytes + 1))
        ws = [ctx.weight(_sweep_weight_bytes(np.random.default_rng(i), args.layout, rows, cols), layout, rows, cols)
              for i in range(ncopies)]
        for M in ms:
            q = torch.randint(-127, 128, (M, cols), dtype=torch.int8, device="cuda")
            s = torch.rand(M, device="cuda") + 0.5
            y = torch.empty(M, rows, device="cuda")
            acc = torch.empty(M, rows, dtype=torch.int32, device="cuda") if args.acc else None
            l0 = b2.launch_count()
            us = graph_time(lambda i: ctx.linear_i8(ws[i % len(ws)], q, s, y=y, acc_out=acc), 100)
            row = {"shape": name, "N": rows, "K": cols, "M": M, "layout": args.layout, "us": round(us, 2),
                   "GBs_weights": round(wbytes / us / 1e3, 1), "TOPS": round(2.0 * M * rows * cols / us / 1e6, 1),
                   "kernel": b2.last_linear_kernel(), "acc_out": bool(args.acc)}
            line = json.dumps(row)
            print(line, flush=True)
            if out:
                out.write(line + "\n")
        del ws
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
