"""Repeat one linear_i8 GEMM shape and report where results differ from a torch int reference."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bitnet_rs_b200 as b2  # noqa: E402

N, K, M = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 10
ctx = b2.Context(0)
rng = np.random.default_rng(0)
qs = rng.integers(0, 256, N * K // 4, dtype=np.uint8)
w = ctx.weight(qs, b2.LAYOUT_QK256, N, K)
# reference: decode codes (LSB-first 2-bit, map A) and matmul in float64 chunks on the GPU
codes = torch.from_numpy(qs).cuda().view(N, K // 4)
lut = torch.tensor([-2, -1, 1, 2], dtype=torch.float32, device="cuda")
wd = torch.stack([lut[((codes >> (2 * j)) & 3).long()] for j in range(4)], dim=2).reshape(N, K)
q = torch.randint(-127, 128, (M, K), dtype=torch.int8, device="cuda")
s = torch.ones(M, device="cuda")
ref = (q.float() @ wd.t()).round().to(torch.int32)  # exact: |acc| < 2^24
bad = 0
fresh = os.environ.get("FRESH") == "1"
for r in range(reps):
    if fresh:  # mimic the parity tests: new weight + activations right before the launch, a split-K launch in between
        w2 = ctx.weight(rng.integers(0, 256, 4096 * 512 // 4, dtype=np.uint8), b2.LAYOUT_QK256, 4096, 512)
        q2 = torch.randint(-127, 128, (64, 512), dtype=torch.int8, device="cuda")
        ctx.linear_i8(w2, q2, torch.ones(64, device="cuda"))
        w = ctx.weight(qs, b2.LAYOUT_QK256, N, K)
        q = torch.from_numpy(rng.integers(-127, 128, (M, K)).astype(np.int8)).cuda()
        ref = (q.float() @ wd.t()).round().to(torch.int32)
    y, acc = ctx.linear_i8(w, q, s, want_acc=True)
    torch.cuda.synchronize()
    d = (acc != ref)
    nb = int(d.sum())
    if nb:
        bad += 1
        idx = d.nonzero()
        ms, ns = idx[:, 0], idx[:, 1]
        print("rep", r, "mismatches", nb, "m range", int(ms.min()), int(ms.max()), "n range", int(ns.min()), int(ns.max()),
              "distinct m-tiles(256)", sorted(set((ms // 256).tolist()))[:8], "distinct n-tiles(128)", sorted(set((ns // 128).tolist()))[:8])
        dd = (acc - ref)[d]
        print("   diff sample", dd[:8].tolist(), "ref sample", ref[d][:4].tolist())
print("bad reps", bad, "of", reps)
