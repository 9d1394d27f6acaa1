"""Measured int8 tensor peak of this GPU: a cuBLASLt int8 GEMM (torch._int_mm, 8192^3) timed with CUDA
events -- the library denominator SURVEY 8(d) asks for beside the 4 500 TOPS nominal figure."""
import json

import torch

n = 8192
a = torch.randint(-127, 128, (n, n), dtype=torch.int8, device="cuda")
b = torch.randint(-127, 128, (n, n), dtype=torch.int8, device="cuda").t().contiguous().t()  # column-major B
for _ in range(5):
    torch._int_mm(a, b)
torch.cuda.synchronize()
best = 1e30
for _ in range(10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    torch._int_mm(a, b)
    e1.record()
    torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(200):
    torch._int_mm(a, b)
e1.record()
torch.cuda.synchronize()
sus = e0.elapsed_time(e1) / 200
ops = 2.0 * n ** 3
print(json.dumps({"what": "cuBLASLt int8 GEMM 8192^3 (torch._int_mm)", "burst_TOPS": ops / best / 1e9,
                  "sustained_TOPS": ops / sus / 1e9, "nominal_TOPS": 4500.0}))
