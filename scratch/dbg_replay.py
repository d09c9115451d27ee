import numpy as np, torch, sys
sys.path.insert(0, "/root/repo")
from tests import common
from tests.test_gpu_replay import expected_rows, canon
from prisma_b200 import qnet
from prisma_b200.env import BatchedRoutingEnv
cfg = common.env_config("abilene", num_envs=48, simTime=1.5, load_factor=2.0, pingAsObs=0, activateUnderlayTraffic=1, max_out_buffer_size=2168)
arrays = cfg.load()
env = BatchedRoutingEnv(cfg, arrays=arrays)
degrees = [len(n) for n in env.neighbors]
k_max, m = max(degrees), len(degrees)
env.qnet_load(qnet.pack_weights(qnet.init_weights(m, degrees, k_max, seed=3)))
env.replay_enable(rows_per_node=1 << 16, pending_slots=4096)
history = []
batch = env.step(None)
for it in range(1500):
    actions = env.qnet_act(eps=0.3, seed=5)
    env.replay_record(actions)
    torch.cuda.synchronize()
    history.append((batch.numpy().copy(), actions.cpu().numpy().copy()))
    batch = env.step(actions)
exp = expected_rows(history, k_max, degrees, cfg.loss_penalty, cfg.num_envs, m)
w = 2 * (1 + k_max) + 3
for node in range(2):
    a, b = canon(exp[node], w), canon(env.replay_rows(node), w)
    sa = {r.tobytes() for r in a}; sb = {r.tobytes() for r in b}
    oa = [np.frombuffer(x, dtype=np.float32) for x in sa - sb][:5]
    ob = [np.frombuffer(x, dtype=np.float32) for x in sb - sa][:5]
    print(node, len(sa - sb), len(sb - sa))
    for x in oa: print(" exp", x.tolist())
    for x in ob: print(" got", x.tolist())
