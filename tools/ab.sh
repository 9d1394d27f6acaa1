#!/bin/bash
# A/B of an environment knob on ONE box: alternate runs of bench.py, print tok/s per run.
# usage: tools/ab.sh KNOB [reps]
KNOB=$1; REPS=${2:-3}
for i in $(seq $REPS); do
  a=$(timeout 200 python bench.py --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('%.1f' % d['value'])")
  b=$(env $KNOB=1 timeout 200 python bench.py --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('%.1f' % d['value'])")
  echo "run $i: default $a tok/s   $KNOB=1 $b tok/s"
done
