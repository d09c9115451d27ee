#!/bin/bash
# usage: ab_env.sh VAR "v1 v2 ...": bench with VAR=v for each value, twice
for rep in 1 2; do for v in $2; do
  env $1=$v timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --gym-steps 0 --no-secondary > gpurun_out/ab_tmp.log 2>&1
  python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/ab_tmp.log").read().strip().splitlines()[-1])
    print("$1=$v", "hops/s %.4e"%j["value"], "ms/step %.3f"%j["ms_per_step"])
except Exception as ex:
    print("$1=$v", "FAILED", ex, open("gpurun_out/ab_tmp.log").read()[-400:])
PY
done; done
