#!/bin/bash
# usage: sweep_libs.sh "<lib names under scratch/libs, or 'default'>" "<team list>" "<env list>"
for lib in $1; do for t in $2; do for e in $3; do
  if [ "$lib" = default ]; then unset PRB_LIB; else export PRB_LIB=$PWD/scratch/libs/$lib; fi
  timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --gym-steps 0 --team $t --envs $e > gpurun_out/swl.log 2>&1
  python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/swl.log").read().strip().splitlines()[-1])
    print("$lib team",$t,"envs",$e,"hops/s %.3e"%j["value"],"ms/step %.2f"%j["ms_per_step"],j.get("engine",{}).get("events_per_round"))
except Exception as ex:
    print("$lib team",$t,"envs",$e,"FAILED",ex)
PY
done; done; done
