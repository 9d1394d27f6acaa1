"""Diagnostics: per-phase time breakdown of the persistent decode kernel.

Every CTA stamps the phase boundaries (globaltimer); the warps of one CTA (BITNET_B200_MEGA_PROF_CTA) add
finer stamps. The dump holds the LAST step run (the stamps of each launch overwrite the previous ones).
usage: python tools/megaprof.py [out.bin] [steps]
"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["BITNET_B200_MEGA_PROF"] = out = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/megaprof.bin"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 40
import bitnet_rs_b200 as b2
shape = b2.ModelShape.bitnet_2b4t(max_ctx=128)
ctx = b2.Context(0)
m = b2.Model(ctx, shape)
m.load_synthetic()
tok = np.array([1], np.int32)
lin = len(sys.argv) > 3 and sys.argv[3] == "lin"
for i in range(steps):
    tok = m.decode(tok)
if lin:
    print("linears only: ms per pass", m.time_linears(3))
m.close()
raw = np.fromfile(out, dtype=np.uint64)
G = (len(raw) - 16 * 4096) // 1024
st = raw[: G * 1024].reshape(G, 1024).astype(np.int64)
wt = raw[G * 1024:].reshape(16, 4096)
L = shape.n_layers
if lin:
    per = 16
    names = ["qkv.pro", "qkv.tiles", "qkv.epi", ">o", "o.pro", "o.tiles", "o.epi", ">gu", "gu.pro", "gu.tiles", "gu.epi", ">down",
             "down.pro", "down.tiles", "down.epi", ">next"]
    st = st[:, : L * per + 1]
    t0 = st[:, 0].min()
    st = st - t0
    d = np.diff(st, axis=1).reshape(G, L, per)
    print("LINEARS ONLY: CTAs", G, "pass us %.1f" % (st[:, -1].max() / 1e3))
    print("interval        CTA0   median   max-CTA   (us, mean over layers 5..25)")
    for k, nme in enumerate(names):
        x = d[:, 5:25, k].mean(1)
        print("%-12s %7.2f  %7.2f  %7.2f" % (nme, x[0] / 1e3, np.median(x) / 1e3, x.max() / 1e3))
    print("layer total us (CTA0) %.2f" % (d[0].sum(1).mean() / 1e3))
if not lin:
    per = 18
    names = ["qkv.pro", "qkv.tiles", "qkv.epi", ">attn", "attn", ">o", "o.pro", "o.tiles", "o.epi", ">gu", "gu.pro",
             "gu.tiles", "gu.epi", ">down", "down.pro", "down.tiles", "down.epi", ">next"]
    nst = L * per + 3
    t0 = st[:, 0].min()
    st = st[:, :nst] - t0
    print("CTAs", G, "stamps/CTA", nst, "step us %.1f" % (st[:, -1].max() / 1e3))
    d = np.diff(st, axis=1)[:, : L * per].reshape(G, L, per)
    print("interval        CTA0   median   max-CTA   (us, mean over layers 5..25)")
    for k, nme in enumerate(names):
        x = d[:, 5:25, k].mean(1)
        print("%-12s %7.2f  %7.2f  %7.2f" % (nme, x[0] / 1e3, np.median(x) / 1e3, x.max() / 1e3))
    print("layer total us (CTA0) %.2f" % (d[0].sum(1).mean() / 1e3))
    # end-of-phase skew: time of 'epilogue done' per CTA relative to the earliest CTA, per phase kind
    for nme, k in (("qkv", 3), ("attn", 5), ("o", 9), ("gu", 13), ("down", 17)):
        e = st[:, [l * per + k for l in range(5, 25)]]
        sk = e - e.min(0)
        print("end-of-%-5s skew over CTAs: median %.2f us  p90 %.2f  max %.2f" % (nme, np.median(sk) / 1e3, np.percentile(sk, 90) / 1e3, sk.max(0).mean() / 1e3))
    # critical path per phase: latest 'epilogue done' of phase n -> latest 'epilogue done' of phase n+1
    ends = [3, 5, 9, 13, 17]
    lat = np.array([[st[:, l * per + k].max() for k in ends] for l in range(L)]).reshape(-1)
    dl = np.diff(lat).reshape(-1)
    dl = np.concatenate([[0], dl]).reshape(L, 5)
    print("critical path (last CTA done -> last CTA done), us: qkv %.2f attn %.2f o %.2f gu %.2f down %.2f" % tuple(dl[5:25].mean(0) / 1e3))

# warp trace of the selected CTA: all consumer warps, layer 10
ids = {1: "start", 2: "valid", 3: "bfrag", 4: "pair", 5: "tiles", 6: "epi", 7: "bar1", 8: "bar2", 9: "bar3", 10: "WAITCYC"}
tr = []
for w in range(15):
    v = wt[w]
    v = v[v != 0]
    tid = (v >> np.uint64(48)).astype(np.int64)
    tt = (v & np.uint64(0xffffffffffff)).astype(np.int64)
    tr.append((tid, tt))
nph = 4
for ph, nme in zip(range(40, 44), ("qkv", "o", "gu", "down")):
    print("phase %s (layer 10), per warp: us since the CTA's earliest phase start" % nme)
    rows = []
    for w in range(15):
        tid, tt = tr[w]
        starts = np.where(tid == 1)[0]
        if len(starts) <= ph + 1:
            continue
        a, b = starts[ph], starts[ph + 1]
        rows.append((w, tid[a:b], tt[a:b]))
    if not rows:
        continue
    base = min(r[2][0] for r in rows)
    for w, i, x in rows:
        print("  w%-2d " % w + " ".join(("%s %d" % (ids[int(k)], y)) if k == 10 else ("%s %.2f" % (ids[int(k)], (y - base) / 1e3)) for k, y in zip(i, x)))

v = wt[15]
v = v[v != 0]
print("producer warp: slots / free-at-first-look / cycles slept:", [(int(x >> np.uint64(48)), int(x & np.uint64(0xffffffffffff))) for x in v[:3]])
