"""Throughput of prb_run on the other fixture topologies (not bench lines: DESIGN.md table)."""
import sys, time
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from tests import common
from prisma_b200.env import BatchedRoutingEnv, sp_table
cases = [("abilene", dict(load_factor=0.9, pingAsObs=0), 4096),
         ("geant", dict(load_factor=0.9, pingAsObs=0), 4096),
         ("geant", dict(load_factor=1.5, pingAsObs=0), 4096),
         ("overlay_full_mesh_3n_abilene", dict(load_factor=0.9, pingAsObs=1), 16384),
         ("abilene", dict(load_factor=0.9, pingAsObs=0, train=1, signaling_type="target", agent_type="dqn_buffer"), 4096),
         # BASELINE.json configs[4]: 100-node random underlay, 4-packet buffers (drop-heavy), control packets
         ("synthetic100", dict(load_factor=1.3, pingAsObs=0, max_out_buffer_size=2168, train=1, signaling_type="NN",
                               agent_type="dqn_buffer"), 4096)]
only = sys.argv[1:]
for topo, over, E in cases:
    if only and topo not in only:
        continue
    for eng in ("parallel", "serial"):
        cfg = common.env_config(topo, num_envs=E, simTime=8.0, activateUnderlayTraffic=1, **over)
        arrays = cfg.load()
        env = BatchedRoutingEnv(cfg, arrays=arrays)
        env.set_engine(eng)
        env.set_policy("table", sp_table(arrays))
        env.run(until_s=2.0)
        torch.cuda.synchronize()
        c0 = env.counters()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); env.run(until_s=4.0); b.record(); torch.cuda.synchronize()
        c1 = env.counters(); env.check_errors(c1)
        dt = a.elapsed_time(b) / 1000
        rs = env.round_stats() if eng == "parallel" else (1, 0)
        print("%-30s %s E=%d %-8s hops/s %.3e steps/s %.3e ev/round %.2f" % (topo, over, E, eng,
              (c1["packet_hops"].sum() - c0["packet_hops"].sum()) / dt, (c1["env_steps"].sum() - c0["env_steps"].sum()) / dt, rs[1] / max(rs[0], 1)), flush=True)
        env.close()
