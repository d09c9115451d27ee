"""BASELINE.json configs[2..4] on one GPU (not bench lines: the DESIGN.md table).  configs[2]: overlay topology, dqn
agents on the device (Gym loop, pingAsObs=1), 16k envs; configs[3]: GEANT load sweep, 8192 envs (= 64k / 8 GPUs);
configs[4]: synthetic100, drop-heavy + control packets."""
import sys, time
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from tests import common
from prisma_b200 import qnet
from prisma_b200.env import BatchedRoutingEnv, sp_table

def timed(fn):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); fn(); b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / 1000

# configs[2]
cfg = common.env_config("overlay_full_mesh_3n_abilene", num_envs=16384, simTime=8.0, load_factor=0.9, pingAsObs=1,
                        activateUnderlayTraffic=1, agent_type="dqn_buffer")
arrays = cfg.load()
env = BatchedRoutingEnv(cfg, arrays=arrays)
deg = [len(n) for n in env.neighbors]
env.qnet_load(qnet.pack_weights(qnet.init_weights(len(deg), deg, max(deg), seed=1)))
env.replay_enable(rows_per_node=1 << 18)
env.step_qnet(256, eps=0.05, seed=0)
c0 = env.counters()
dt = timed(lambda: env.step_qnet(512, eps=0.05, seed=1))
c1 = env.counters(); env.check_errors(c1)
print("configs[2] overlay 16384 envs dqn_buffer on device, pingAsObs=1: agent-steps/s %.3e  packet-hops/s %.3e" % (
    (c1["env_steps"].sum() - c0["env_steps"].sum()) / dt, (c1["packet_hops"].sum() - c0["packet_hops"].sum()) / dt), flush=True)
env.close()
# configs[3]
for lf in (0.5, 0.7, 0.9, 1.1, 1.3, 1.5):
    cfg = common.env_config("geant", num_envs=8192, simTime=6.0, load_factor=lf, pingAsObs=0, activateUnderlayTraffic=1)
    arrays = cfg.load()
    env = BatchedRoutingEnv(cfg, arrays=arrays)
    env.set_policy("table", sp_table(arrays))
    env.run(until_s=2.0)
    c0 = env.counters()
    dt = timed(lambda: env.run(until_s=3.0))
    c1 = env.counters(); env.check_errors(c1)
    lost = c1["overlay_lost"].sum() / max(c1["overlay_injected"].sum(), 1)
    print("configs[3] geant 8192 envs lf %.1f engine %s: packet-hops/s %.3e env-steps/s %.3e loss %.3f handovers %d" % (
        lf, env.engine, (c1["packet_hops"].sum() - c0["packet_hops"].sum()) / dt,
        (c1["env_steps"].sum() - c0["env_steps"].sum()) / dt, lost, env.handovers()), flush=True)
    env.close()
