"""Diagnostics: per-phase time breakdown of the persistent decode kernel (CTA 0 stamps)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["BITNET_B200_MEGA_PROF"] = out = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/megaprof.txt"
import bitnet_rs_b200 as b2
shape = b2.ModelShape.bitnet_2b4t(max_ctx=128)
ctx = b2.Context(0)
m = b2.Model(ctx, shape)
m.load_synthetic()
tok = np.array([1], np.int32)
for i in range(40):
    tok = m.decode(tok)
m.close()
t = np.array([float(l) for l in open(out) if l[0] not in '#@'])
extra = [l.split()[1:] for l in open(out) if l.startswith('#')]
L = shape.n_layers
per = 18
names = ["qkv.pro", "qkv.tiles", "qkv.epi", "bar>attn", "attn", "bar>o", "o.pro", "o.tiles", "o.epi", "bar>gu", "gu.pro",
         "gu.tiles", "gu.epi", "bar>down", "down.pro", "down.tiles", "down.epi", "bar>next"]
d = np.diff(t)
lay = d[: L * per].reshape(L, per)  # interval k = stamp k -> k+1
print("stamps", len(t), "total us", t[-1] / 1e3)
for k, nme in enumerate(names):
    print("%-10s mean %.2f us  (layers 5..25: %.2f)" % (nme, lay[:, k].mean() / 1e3, lay[5:25, k].mean() / 1e3))
print("layer total mean us", lay.sum(1).mean() / 1e3)
print("tail (logits etc) us", d[L * per:] / 1e3)

ex = {int(a): (int(b), int(c)) for a, b, c in extra}
for w, base in (("warp0", 0), ("warp5", 4096)):
    for k, nme in ((2, "qkv"), (8, "o"), (12, "gu"), (16, "down")):
        v = [ex[base + l * per + k] for l in range(5, 25) if base + l * per + k in ex]
        if v:
            print(w, nme, "wait cycles %.0f of total %.0f" % (np.mean([a for a, b in v]), np.mean([b for a, b in v])))

print("gu phase layer 10 warp 5: per-tile clock stamps (loop top, after wait, after steps, after arrive)")
for l in open(out):
    if l.startswith('@'):
        print(l.strip())
