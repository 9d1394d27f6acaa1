#!/bin/bash
# bench.py tok/s for several builds of the library on ONE box, interleaved. usage: tools/ab_many.sh reps A.so B.so ...
REPS=$1; shift; L=bitnet-rs_b200/libbitnet_b200.so; cp $L /tmp/_keep.so
for i in $(seq $REPS); do for f in "$@"; do cp "$f" $L; echo "run $i $f $(timeout 200 python bench.py --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('%.1f tok/s  e2e %.1f' % (d['value'], d['e2e']['value']))")"; done; done
cp /tmp/_keep.so $L
