"""Run a few linear_i8 launches of one shape (for ncu): python tools/gemm_prof.py N K M [reps]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bitnet_rs_b200 as b2  # noqa: E402

N, K, M = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 4
ctx = b2.Context(0)
rng = np.random.default_rng(0)
w = ctx.weight(rng.integers(0, 256, N * K // 4, dtype=np.uint8), b2.LAYOUT_QK256, N, K)
q = torch.randint(-127, 128, (M, K), dtype=torch.int8, device="cuda")
s = torch.rand(M, device="cuda") + 0.5
y = torch.empty(M, N, device="cuda")
for _ in range(reps):
    ctx.linear_i8(w, q, s, y=y)
torch.cuda.synchronize()
print("ok")
