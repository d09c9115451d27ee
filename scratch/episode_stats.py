"""One full episode on the GPU in one-second steps with a read-out per step: wall time, hand-overs to the serial engine
and the device time of their replays.  usage: episode_stats.py <topology> <load> <envs> [seconds] [engine]"""
import json, sys, time
sys.path.insert(0, "/root/repo")
import numpy as np
import torch
from tests import common
from prisma_b200.env import BatchedRoutingEnv, sp_table

topo, lf, envs = sys.argv[1], float(sys.argv[2]), int(sys.argv[3])
T = int(sys.argv[4]) if len(sys.argv) > 4 else 60
engine = sys.argv[5] if len(sys.argv) > 5 else "auto"
cfg = common.env_config(topo, num_envs=envs, simTime=float(T), activateUnderlayTraffic=1, load_factor=lf, pingAsObs=0)
arrays = cfg.load()
env = BatchedRoutingEnv(cfg, arrays=arrays)
env.set_engine(engine)
env.set_policy("table", sp_table(arrays))
env.run(until_s=1.0)
env.reset()
torch.cuda.synchronize()
t0 = time.perf_counter()
per = []
for t in range(1, T + 1):
    a = time.perf_counter()
    env.run(until_s=float(t))
    n = env.handovers() if env.engine == "parallel" else 0
    env.sync()
    per.append((time.perf_counter() - a, n))
wall = time.perf_counter() - t0
c = env.counters()
env.check_errors(c)
n, ms = env.handover_stats() if env.engine == "parallel" else (0, 0.0)
rounds, executed = env.round_stats() if env.engine == "parallel" else (1, 1)
hops = int(c["packet_hops"].sum())
print(json.dumps({"topology": topo, "load_factor": lf, "envs": envs, "sim_s": T, "engine": env.engine, "wall_s": wall,
                  "packet_hops": hops, "packet_hops_per_s": hops / wall, "handovers": n, "serial_replay_ms": ms,
                  "replay_share_of_wall": ms / 1000.0 / wall, "events_per_round": executed / max(rounds, 1),
                  "loss_overlay": float(c["overlay_lost"].sum()) / max(float(c["overlay_injected"].sum()), 1.0),
                  "slowest_steps_ms": sorted([round(1000 * x, 1) for x, _ in per])[-3:]}))
