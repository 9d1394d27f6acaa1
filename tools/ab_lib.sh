#!/bin/bash
# A/B of two builds of libbitnet_b200.so on ONE box: alternate bench.py runs, print tok/s.
# usage: tools/ab_lib.sh A.so B.so [reps]   (the in-tree library is restored to A at the end)
A=$1; B=$2; REPS=${3:-2}; L=bitnet-rs_b200/libbitnet_b200.so
one() { cp "$1" $L; timeout 200 python bench.py --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('%.1f' % d['value'])"; }
for i in $(seq $REPS); do echo "run $i: $A $(one $A) tok/s   $B $(one $B) tok/s"; done
cp "$A" $L
