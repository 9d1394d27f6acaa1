# usage: bash tools/diag1.sh name[:ENV=VAL,...] ...   -- short bench runs under env knobs, one summary line each
mkdir -p gpurun_out
for spec in "$@"; do
  name=${spec%%:*}; envs=""
  if [ "$spec" != "$name" ]; then envs=$(echo "${spec#*:}" | tr ',' ' '); fi
  env $envs timeout 300 python bench.py --steps ${STEPS:-16} --warmup 3 --prompt ${PROMPT:-4} --no-cpu-baseline > gpurun_out/d_$name.json 2> gpurun_out/d_$name.err
  python - "$name" <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.load(open("gpurun_out/d_%s.json"%n))
    print(n, "tok/s=%.0f e2e=%.0f ms=%.3f launches=%d lin_us=%.2f GBs=%.0f"%(d["value"],d["e2e"]["value"],d["ms_per_step"],d["launches_per_step"],d["roofline"]["us_per_launch"],d["roofline"]["achieved"]))
except Exception as e: print(n,"FAILED",e)
PY
done
