#!/bin/bash
# time the prefill GEMM shapes for several ring configurations: tools/gemm_rings.sh "5,2" "4,4" ...
for cfg in "$@"; do
  echo "rings $cfg"
  BITNET_B200_GEMM_RINGS=$cfg python tools/smallm_bench.py --ms 4096,16384 --shapes 2b.qkv,2b.o,2b.gateup,2b.down 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('  ',d['shape'],d['M'],d['us'],d['TOPS'])"
done
