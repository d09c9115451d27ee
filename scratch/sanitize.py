"""Small workloads of both engines and the Gym / replay / Q-network paths for compute-sanitizer runs."""
import sys
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from tests import common
from prisma_b200 import qnet
from prisma_b200.env import BatchedRoutingEnv, sp_table
for topo, over in (("abilene", dict(load_factor=1.6, max_out_buffer_size=2168, pingAsObs=1, train=1, signaling_type="NN",
                                    agent_type="dqn_buffer")),
                   ("overlay_full_mesh_3n_abilene", dict(load_factor=2.0, pingAsObs=1)),
                   ("geant", dict(load_factor=1.2, pingAsObs=0))):
    cfg = common.env_config(topo, num_envs=24, simTime=0.6, activateUnderlayTraffic=1, **over)
    arrays = cfg.load()
    for eng in ("serial", "parallel"):
        env = BatchedRoutingEnv(cfg, arrays=arrays)
        env.set_engine(eng)
        env.set_policy("table", sp_table(arrays))
        env.run(until_s=0.3)
        env.run()
        c = env.counters(); env.check_errors(c)
        print(topo, eng, int(c["packet_hops"].sum()), flush=True)
        env.close()
    env = BatchedRoutingEnv(cfg, arrays=arrays)
    deg = [len(n) for n in env.neighbors]
    env.qnet_load(qnet.pack_weights(qnet.init_weights(len(deg), deg, max(deg), seed=1)))
    env.replay_enable(rows_per_node=512, pending_slots=1024)
    env.replay_set_target_mode(True, 0.9)
    env.replay_enable_prioritized(1.0)
    b = env.step(None)
    for _ in range(20):
        a = env.qnet_act(eps=0.1, seed=1); env.replay_record(a); b = env.step(a)
    env.step_qnet(150, eps=0.1, seed=2)
    env.replay_sample(0, 64, seed=1, weighted=True)
    torch.cuda.synchronize()
    env.check_errors()
    print(topo, "gym", env.replay_counts()[0].sum(), flush=True)
    env.close()
print("done")
