#!/bin/bash
# A/B/C... of several builds of libbitnet_b200.so on ONE box: bench.py per library (BITNET_B200_LIB), print tok/s, e2e, roofline us.
# usage: tools/ab_many.sh lib1.so lib2.so ... ; REPS=n for repetitions; LAYOUT=qk256|bitnet32
REPS=${REPS:-2}
for i in $(seq $REPS); do
  for L in "$@"; do
    BITNET_B200_LIB=$PWD/$L timeout 200 python bench.py --no-cpu-baseline --no-gemm-record --no-bitnet32-record --layout ${LAYOUT:-qk256} 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$L', 'tok/s %.1f' % d['value'], 'e2e %.1f' % d['e2e']['value'], 'lin us %.1f' % d['roofline']['us_per_launch'], 'frac %.3f' % d['roofline']['frac'])"
  done
done
