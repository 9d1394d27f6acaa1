"""Diagnostics: where the host-buffer decode call (bench e2e) spends its time per step."""
import os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bitnet_rs_b200 as b2

shape = b2.ModelShape.bitnet_2b4t(max_ctx=512)
ctx = b2.Context(0)
m = b2.Model(ctx, shape)
m.load_synthetic()
tok = np.array([1], np.int32)
for _ in range(16):
    tok = m.decode(tok)
N = 64
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(N):
    tok = m.decode(tok)
t1 = time.perf_counter()
print("decode() host-buffer call: %.1f us/step" % ((t1 - t0) / N * 1e6))
t0 = time.perf_counter()
for _ in range(N):
    m.decode_device(1)
    torch.cuda.synchronize()
t1 = time.perf_counter()
print("decode_device(1) + sync:   %.1f us/step" % ((t1 - t0) / N * 1e6))
t0 = time.perf_counter()
m.decode_device(N)
torch.cuda.synchronize()
t1 = time.perf_counter()
print("decode_device(%d) queued:  %.1f us/step" % (N, (t1 - t0) / N * 1e6))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ctx.use_torch_stream()
ts = []
for _ in range(8):
    e0.record()
    m.decode_device(1)
    e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1) * 1e3)
print("single launch, events around it: %s us" % ["%.0f" % t for t in ts])
m.close()
