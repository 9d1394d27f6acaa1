#!/usr/bin/env python
"""Randomised soak of gemm_small.cu: random (rows, cols, M), the cost model's plan and random forced plans, int32
accumulators against a torch reference (exact: |partial| < 2^24 per 2048-k slice). python tools/smallm_stress.py [cases]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch  # noqa: E402
import bitnet_rs_b200 as b2  # noqa: E402
from smallm_tune import derive  # noqa: E402

cases = int(sys.argv[1]) if len(sys.argv) > 1 else 150
rng = np.random.default_rng(int(os.environ.get("SEED", "1")))
ctx = b2.Context(0)
lut = torch.tensor([-2, -1, 1, 2], dtype=torch.float32, device="cuda")
os.environ["BITNET_B200_SMALLM_AUTOTUNE"] = "0"
bad = 0
for c in range(cases):
    rows = int(rng.integers(1, 6000))
    cols = 256 * int(rng.integers(1, 40))
    M = int(rng.integers(1, 65))
    qs = rng.integers(0, 256, rows * cols // 4, dtype=np.uint8)
    codes = torch.from_numpy(qs).cuda().view(rows, cols // 4)
    wd = torch.stack([lut[((codes >> (2 * j)) & 3).long()] for j in range(4)], dim=2).reshape(rows, cols)
    q = torch.from_numpy(rng.integers(-127, 128, (M, cols)).astype(np.int8)).cuda()
    ref = torch.zeros(M, rows, dtype=torch.int64, device="cuda")
    for k0 in range(0, cols, 2048):
        ref += (q[:, k0:k0 + 2048].float() @ wd[:, k0:k0 + 2048].t()).round().to(torch.int64)
    ref = ref.to(torch.int32)
    w = ctx.weight(qs, b2.LAYOUT_QK256, rows, cols)
    s = torch.ones(M, device="cuda")
    plans = [None]
    for _ in range(3):
        pl = (int(rng.integers(1, 3)), int(rng.integers(2, 17)), int(rng.choice([1, 2, 4])), int(rng.integers(1, 9)))
        if derive(rows, cols, M, *pl) is not None:
            plans.append("%d,%d,%d,%d" % pl)
    if rows > 148 * 32:
        plans.append("R,%d,%d" % (int(rng.integers(1, 3)), int(rng.choice([4, 6, 8]))))
    for pl in plans:
        if pl is None:
            os.environ.pop("BITNET_B200_SMALLM_PLAN", None)
        else:
            os.environ["BITNET_B200_SMALLM_PLAN"] = pl
        try:
            _, acc = ctx.linear_i8(w, q, s, want_acc=True)
        except b2.B200Error as ex:
            if "not realisable" in str(ex):
                continue
            raise
        if not torch.equal(acc, ref):
            bad += 1
            print("MISMATCH", rows, cols, M, pl, b2.last_linear_kernel(), int((acc != ref).sum()))
    del w
print("cases", cases, "mismatches", bad)
sys.exit(1 if bad else 0)
