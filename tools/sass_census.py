"""Per-kernel census of the SASS mnemonics that prove the Blackwell-native paths (B200_PROFILING.md):
UTC*MMA (tcgen05.mma), LDTM/STTM (tcgen05.ld/st), UTMALDG / UBLKCP (TMA), IMMA / HMMA (legacy mma.sync), IDP (dp4a).
usage: python tools/sass_census.py [lib.so] > profiles/r02_sass_opcodes.txt"""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "bitnet-rs_b200/libbitnet_b200.so"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
want = ["UTCIMMA", "UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTCBAR", "IMMA", "HMMA", "IDP", "SYNCS", "LDG", "STG",
        "PRMT"]
cur, counts, total = None, collections.OrderedDict(), collections.Counter()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        counts[cur] = collections.Counter()
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and cur:
        op = m.group(1)
        total[cur] += 1
        for w in want:
            if op.startswith(w):
                counts[cur][w] += 1
print("# SASS opcode census of %s (cuobjdump -sass), per kernel: total instructions, then the mnemonics of interest" % lib)
for k, c in counts.items():
    print("%-60s %6d  %s" % (k[-60:], total[k], "  ".join("%s x%d" % (w, c[w]) for w in want if c[w])))
