"""Where a batched Gym step goes: prb_step alone, prb_qnet_act alone (warp / mma), per batch size."""
import sys, json
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from tests import common
from prisma_b200 import qnet
from prisma_b200.env import BatchedRoutingEnv
def timed(fn, n):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) * 1000 / n
for envs in [int(x) for x in sys.argv[1:]] or [4096, 65536]:
    cfg = common.env_config("abilene", num_envs=envs, simTime=60.0, load_factor=0.9, pingAsObs=0, activateUnderlayTraffic=1, agent_type="dqn_buffer")
    arrays = cfg.load()
    env = BatchedRoutingEnv(cfg, arrays=arrays)
    deg = [len(n) for n in env.neighbors]
    env.qnet_load(qnet.pack_weights(qnet.init_weights(len(deg), deg, max(deg), seed=1)))
    env.step_qnet(128, eps=0.05, seed=1)
    acts = torch.zeros(envs, dtype=torch.int32, device="cuda")
    out = {"envs": envs}
    out["step_us"] = timed(lambda: env.step(acts), 100)
    for mode in ("warp", "mma"):
        env.qnet_set_mode(mode)
        env.qnet_act()
        out["act_%s_us" % mode] = timed(lambda: env.qnet_act(), 100)
    print(json.dumps(out), flush=True)
    env.close()
