"""Stress: the persistent decode kernel must be deterministic -- the same greedy chain on every repetition
(a lost flag traps; a torn or stale pair would change the chain). usage: mega_stress.py [steps] [reps] [max_ctx]"""
import os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bitnet_rs_b200 as b2

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 1100
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
max_ctx = int(sys.argv[3]) if len(sys.argv) > 3 else steps + 16
shape = b2.ModelShape.bitnet_2b4t(max_ctx=max_ctx)
ctx = b2.Context(0)
m = b2.Model(ctx, shape)
m.load_synthetic()
ref = None
bad = 0
for r in range(reps):
    m.reset()
    m.set_tokens(np.array([1 + 0 * r], np.int32))
    t0 = time.perf_counter()
    chain = []
    done = 0
    while done < steps:
        n = min(97, steps - done)       # odd chunking: host syncs fall at different positions
        m.decode_device(n)
        torch.cuda.synchronize()
        chain.append(int(m.get_tokens()[0]))
        done += n
    dt = time.perf_counter() - t0
    if ref is None:
        ref = chain
    same = chain == ref
    bad += 0 if same else 1
    first = next((i for i, (a, b) in enumerate(zip(chain, ref)) if a != b), None)
    print("rep %d: %d steps in %.2f s, chain %s%s" % (r, steps, dt, "identical" if same else "DIFFERS",
                                                     "" if same else " from checkpoint %d" % first), flush=True)
m.close()
print("stress", "FAILED" if bad else "ok")
sys.exit(1 if bad else 0)
