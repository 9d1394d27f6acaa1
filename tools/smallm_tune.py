#!/usr/bin/env python
"""Exhaustive timing of the gemm_small.cu plan space (row warps, k warps, words per lane, cluster K split) per shape and M,
next to what the built-in heuristic picks and to the tcgen05 GEMM. Used to fit the cost model in gemm_small.cu.

    python tools/smallm_tune.py [--shapes 8b.qkv,8b.o] [--ms 16,32,64] [--out file.jsonl]"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from smallm_bench import SHAPES  # noqa: E402


def derive(N, K, M, RW, KW, U, ks):
    """mirror of small_derive (gemm_small.cu): None when the plan is not realisable"""
    if RW * KW > 16:
        return None
    nt = (M + 7) // 8
    NT = 1 if nt <= 1 else 2 if nt <= 2 else 3 if nt <= 3 else 4 if nt <= 4 else 6 if nt <= 6 else 8
    CK = 128 * U * KW
    if CK % 256:
        return None
    nch = -(-K // CK)
    if nch % ks:
        return None
    npr = nch // ks
    stage = CK * NT * 8 + CK // 256 * 32 * RW * 64
    jobs = RW * 2 * NT
    ring_off = (128 + (jobs * 512 if ks > 1 else 0) + 1023) & ~1023
    part = jobs * KW * 512
    budget = 220 * 1024
    if ring_off + part > budget:
        return None
    ns = min((budget - ring_off) // stage, 8, npr)
    if ns < 1 or (ns < 2 and npr > 1):
        return None
    return {"NS": ns, "npr": npr, "CK": CK, "ctas": -(-N // (32 * RW)) * ks}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", default="")
    ap.add_argument("--ms", default="16,32,64")
    ap.add_argument("--out", default="")
    ap.add_argument("--iters", type=int, default=40)
    args = ap.parse_args()
    import torch
    import bitnet_rs_b200 as b2
    from bench import _sweep_weight_bytes
    torch.cuda.set_device(0)
    ctx = b2.Context(0)
    out = open(args.out, "w") if args.out else None
    names = [s for s in args.shapes.split(",") if s] or list(SHAPES)

    def graph_time(fn, iters, reps=3):
        for i in range(2):
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
        wbytes = rows * cols // 4
        ncopies = min(40, max(2, (300 << 20) // wbytes + 1))
        ws = [ctx.weight(_sweep_weight_bytes(np.random.default_rng(i), "qk256", rows, cols), b2.LAYOUT_QK256, rows, cols)
              for i in range(ncopies)]
        for M in [int(x) for x in args.ms.split(",")]:
            q = torch.randint(-127, 128, (M, cols), dtype=torch.int8, device="cuda")
            s = torch.rand(M, device="cuda") + 0.5
            y = torch.empty(M, rows, device="cuda")
            fn = lambda i: ctx.linear_i8(ws[i % len(ws)], q, s, y=y)  # noqa: E731
            os.environ.pop("BITNET_B200_SMALLM_PLAN", None)
            os.environ["BITNET_B200_SMALLM_MAX"] = "0"
            t_tc = graph_time(fn, args.iters)
            os.environ.pop("BITNET_B200_SMALLM_MAX")
            t_h = graph_time(fn, args.iters)
            res = []
            for RW in (1, 2):
                for KW in range(2, 16 // RW + 1):
                    for U in (1, 2, 4):
                        for ks in range(1, 9):
                            d = derive(rows, cols, M, RW, KW, U, ks)
                            if d is None:
                                continue
                            if d["ctas"] < 100 or d["ctas"] > 148 * 8:
                                continue
                            os.environ["BITNET_B200_SMALLM_PLAN"] = "%d,%d,%d,%d" % (RW, KW, U, ks)
                            res.append((graph_time(fn, args.iters, reps=2), (RW, KW, U, ks), d))
            os.environ.pop("BITNET_B200_SMALLM_PLAN", None)
            res.sort(key=lambda r: r[0])
            row = {"shape": name, "N": rows, "K": cols, "M": M, "tcgen05_us": round(t_tc, 2), "heuristic_us": round(t_h, 2),
                   "best": [{"us": round(t, 2), "plan": pl, **d} for t, pl, d in res[:6]],
                   "worst_us": round(res[-1][0], 2) if res else None, "n_plans": len(res),
                   "all": [[round(t, 2), list(pl), d["NS"], d["npr"], d["ctas"]] for t, pl, d in res]}
            line = json.dumps(row)
            print(line, flush=True)
            if out:
                out.write(line + "\n")
                out.flush()
        del ws
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
