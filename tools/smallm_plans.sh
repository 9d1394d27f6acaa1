#!/bin/bash
# time forced gemm_small plans on one shape: tools/smallm_plans.sh SHAPE M plan1 plan2 ...
shape=$1; m=$2; shift 2
for pl in "$@"; do
  echo -n "$shape M=$m plan=$pl: "
  BITNET_B200_SMALLM_PLAN=$pl python tools/smallm_bench.py --shapes $shape --ms $m 2>&1 | python -c "import sys,json; [print(json.loads(l)['us']) for l in sys.stdin if l.startswith('{')]"
done
