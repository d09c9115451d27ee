"""agent-steps/s of prb_step_qnet by Q-network mode and batch size.  usage: qnet_modes.py [envs ...]"""
import sys, json
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from tests import common
from prisma_b200 import qnet
from prisma_b200.env import BatchedRoutingEnv
for envs in [int(x) for x in sys.argv[1:]] or [4096, 16384, 65536]:
    for mode in ("fused", "mma"):
        cfg = common.env_config("abilene", num_envs=envs, simTime=60.0, load_factor=0.9, pingAsObs=0, activateUnderlayTraffic=1,
                                agent_type="dqn_buffer")
        arrays = cfg.load()
        env = BatchedRoutingEnv(cfg, arrays=arrays)
        deg = [len(n) for n in env.neighbors]
        env.qnet_load(qnet.pack_weights(qnet.init_weights(len(deg), deg, max(deg), seed=1)))
        env.qnet_set_mode(mode)
        env.replay_enable(rows_per_node=1 << 16)
        env.step_qnet(64, eps=0.05, seed=1)
        torch.cuda.synchronize()
        s0 = int(env.counters()["env_steps"].sum())
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); env.step_qnet(256, eps=0.05, seed=2); b.record(); torch.cuda.synchronize()
        c = env.counters(); env.check_errors(c)
        dt = a.elapsed_time(b) / 1000
        print(json.dumps({"envs": envs, "mode": mode, "agent_steps_per_s": (int(c["env_steps"].sum()) - s0) / dt, "ms_per_call": 1000 * dt / 256}), flush=True)
        env.close()
