#!/bin/bash
# usage: sweep.sh "<team list>" "<env list>"
for t in $1; do for e in $2; do
  timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --team $t --envs $e > gpurun_out/sw_t${t}_e${e}.log 2>&1
  python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/sw_t${t}_e${e}.log").read().strip().splitlines()[-1])
    print("team",$t,"envs",$e,"hops/s %.3e"%j["value"],"ms/step %.2f"%j["ms_per_step"],j.get("engine",{}).get("events_per_round"))
except Exception as ex:
    print("team",$t,"envs",$e,"FAILED",ex)
PY
done; done
