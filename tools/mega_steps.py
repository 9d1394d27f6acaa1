"""A few decode steps of the 2B-4T-shape model (for ncu captures of the persistent kernel)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bitnet_rs_b200 as b2
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 8
shape = b2.ModelShape.bitnet_2b4t(max_ctx=128)
ctx = b2.Context(0)
m = b2.Model(ctx, shape)
m.load_synthetic()
tok = np.array([1], np.int32)
for i in range(steps):
    tok = m.decode(tok)
print("tokens ok", int(tok[0]), "launches/step", m.launches_per_step)
if len(sys.argv) > 2:
    print("linears ms", m.time_linears(3))
m.close()
