"""Fixed-cost probe of the tcgen05 GEMM: tiny-K problems, graph-replayed back to back."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bitnet_rs_b200 as b2  # noqa: E402

ctx = b2.Context(0)
rng = np.random.default_rng(0)
for (N, K, M) in [(4096, 256, 16), (4096, 256, 64), (4096, 512, 64), (4096, 2048, 64), (16384, 256, 64), (18944, 512, 64),
                  (4096, 4096, 64), (4096, 4096, 256)]:
    w = ctx.weight(rng.integers(0, 256, N * K // 4, dtype=np.uint8), b2.LAYOUT_QK256, N, K)
    q = torch.randint(-127, 128, (M, K), dtype=torch.int8, device="cuda")
    s = torch.ones(M, device="cuda")
    y = torch.empty(M, N, device="cuda")
    for _ in range(3):
        ctx.linear_i8(w, q, s, y=y)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        ctx.use_torch_stream()
        for _ in range(50):
            ctx.linear_i8(w, q, s, y=y)
    ctx.use_torch_stream()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    print("N=%d K=%d M=%d: %.2f us per launch (hot L2)" % (N, K, M, e0.elapsed_time(e1) * 1e3 / 50), flush=True)
