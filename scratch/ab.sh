#!/bin/bash
# usage: ab.sh libA.so libB.so ...: bench each library 2x (alternating), print hops/s
for rep in 1 2; do for lib in "$@"; do
  PRB_LIB=$PWD/$lib timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --gym-steps 0 --no-secondary > gpurun_out/ab_tmp.log 2>&1
  python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/ab_tmp.log").read().strip().splitlines()[-1])
    print("$lib", "hops/s %.4e"%j["value"], "ms/step %.3f"%j["ms_per_step"], j.get("engine",{}).get("events_per_round"))
except Exception as ex:
    print("$lib", "FAILED", ex, open("gpurun_out/ab_tmp.log").read()[-500:])
PY
done; done
