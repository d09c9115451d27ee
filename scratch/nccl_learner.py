"""Data-parallel DQN training over NCCL: every rank owns a shard of the environments on its own GPU (no collective in
the environment step), the learners exchange ONE gradient bucket per training step (torch.distributed all_reduce over
NCCL / NVLink) and must hold identical weights afterwards.  Launch:
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P scratch/nccl_learner.py"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
from tests import common
from prisma_b200 import learner
from prisma_b200.env import BatchedRoutingEnv

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
cfg = common.env_config("abilene", num_envs=512, simTime=3.0, load_factor=1.1, pingAsObs=0, activateUnderlayTraffic=1, device=local)
first, n = cfg.shard(rank, world)
cfg.num_envs = n
arrays = cfg.load()
env = BatchedRoutingEnv(cfg, arrays=arrays, env_id_base=first, device=local)
u2o = {int(u): i for i, u in enumerate(env.overlay_to_underlay)}
neigh = [[u2o[int(u)] for u in nb] for nb in env.neighbors]
L = learner.DQNLearner(neigh, lr=1e-3, seed=2, device="cuda:%d" % local)       # same seed: same initial weights
env.qnet_load(L.weights())
env.replay_enable(rows_per_node=1 << 15)
DEVICE = os.environ.get("PRB_LEARNER", "device") != "torch"
if DEVICE:                              # the gradient step as kernels of the library (prb_learner_*); PRB_LEARNER=torch: autograd
    L = learner.DeviceLearner(env, lr=1e-3)
t_ar = 0.0
for it in range(10):
    env.step_qnet(150, eps=0.2, seed=100 * rank + it)            # different traffic (env ids) and exploration per rank
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    def ar(t):
        a.record(); dist.all_reduce(t); b.record()
    mse = L.train_on_env(env, batch_size=256, seed=1000 * rank + it, world_size=world, all_reduce=ar)
    assert any(x is not None for x in mse)
    torch.cuda.synchronize()
    t_ar += a.elapsed_time(b)
    if it % 4 == 3:
        L.sync_targets()
w = torch.from_numpy(L.weights()).cuda() if DEVICE else L.W.detach()
ref = w.clone()
dist.broadcast(ref, 0)
same = torch.equal(w, ref)
flags = torch.tensor([1.0 if same else 0.0, float(env.counters()["packet_hops"].sum())], device="cuda")
dist.all_reduce(flags)
env.check_errors()
if rank == 0:
    print("learner %s ranks %d identical_weights %s total_packet_hops %d bucket_floats %d mean_all_reduce_ms %.3f" % (
        "device" if DEVICE else "torch", world, flags[0].item() == world, int(flags[1].item()), w.numel() + L.m, t_ar / 10), flush=True)
assert same
env.close()
dist.destroy_process_group()
