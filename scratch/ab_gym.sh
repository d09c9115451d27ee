#!/bin/bash
# usage: ab_gym.sh libA.so libB.so ...: the Gym-mode leg (prb_step_qnet, device-side dqn_buffer agents) of bench.py per library
for rep in 1 2; do for lib in "$@"; do
  PRB_LIB=$PWD/$lib timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ab_tmp.log 2>&1
  python - <<PY
import json
try:
    j=json.loads(open("gpurun_out/ab_tmp.log").read().strip().splitlines()[-1])
    print("$lib", "agent-steps/s %.4e"%j["gym_dqn_buffer"]["value"], "hops/s %.4e"%j["value"])
except Exception as ex:
    print("$lib", "FAILED", ex, open("gpurun_out/ab_tmp.log").read()[-500:])
PY
done; done
