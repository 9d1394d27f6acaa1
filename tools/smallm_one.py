#!/usr/bin/env python
"""Run linear_i8 a few times on one shape (for ncu captures of gemm_small.cu): python tools/smallm_one.py SHAPE M [reps]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from smallm_bench import SHAPES  # noqa: E402
import torch  # noqa: E402
import bitnet_rs_b200 as b2  # noqa: E402
from bench import _sweep_weight_bytes  # noqa: E402

name, M = sys.argv[1], int(sys.argv[2])
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 4
rows, cols = SHAPES[name]
ctx = b2.Context(0)
ws = [ctx.weight(_sweep_weight_bytes(np.random.default_rng(i), "qk256", rows, cols), b2.LAYOUT_QK256, rows, cols)
      for i in range(reps)]
q = torch.randint(-127, 128, (M, cols), dtype=torch.int8, device="cuda")
s = torch.rand(M, device="cuda") + 0.5
y = torch.empty(M, rows, device="cuda")
for w in ws:
    ctx.linear_i8(w, q, s, y=y)
torch.cuda.synchronize()
print("ok", name, M, float(y.abs().sum()))
