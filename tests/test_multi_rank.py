"""Multi-GPU host logic on CPU: world_size-2 gloo processes shard the environments the way bench.py and
`EnvConfig.shard` do (contiguous env-id ranges, no collective in the step), each rank advances its own
shard, and the report is reduced with max (time) / sum (counts).  The step itself is the host build of
the kernel body (tests/host_harness) because this container has no GPU; what is under test is the
sharding: env ids are disjoint and complete, and a sharded job gives the same per-env results as one
rank running every env."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, total_envs, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch
    import torch.distributed as dist
    from tests import common
    from tests.host_harness import host_sim
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cfg = common.env_config("abilene", num_envs=total_envs, simTime=0.5, load_factor=0.9, pingAsObs=0,
                            activateUnderlayTraffic=1)
    arrays = cfg.load()
    table = common.sp_table(arrays)
    first, n = cfg.shard(rank, world)
    hops = np.zeros(total_envs, dtype=np.int64)
    for i in range(n):
        hs = host_sim.HostSim(cfg.abi_params(arrays, env_id_base=first), env_index=i)
        hs.set_policy("table", table)
        hs.run()
        c = hs.counters()
        assert c["errors"] == 0
        hops[first + i] = c["packet_hops"]
    ids = torch.zeros(total_envs, dtype=torch.int64)
    ids[first:first + n] = 1
    dist.all_reduce(ids, op=dist.ReduceOp.SUM)
    assert bool((ids == 1).all()), "every env id must be owned by exactly one rank"
    t = torch.from_numpy(hops.copy())
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    el = torch.tensor([0.1 * (rank + 1)], dtype=torch.float64)
    dist.all_reduce(el, op=dist.ReduceOp.MAX)
    assert abs(float(el) - 0.1 * world) < 1e-12
    if rank == 0:
        np.save(os.path.join(out_dir, "hops.npy"), t.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_env_sharding_world2_gloo(tmp_path):
    import torch.multiprocessing as mp
    from tests import common
    from tests.host_harness import host_sim
    total = 5          # odd on purpose: ranks own 3 and 2 envs
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, total, str(tmp_path)), nprocs=2, join=True)
    sharded = np.load(os.path.join(str(tmp_path), "hops.npy"))
    cfg = common.env_config("abilene", num_envs=total, simTime=0.5, load_factor=0.9, pingAsObs=0, activateUnderlayTraffic=1)
    arrays = cfg.load()
    table = common.sp_table(arrays)
    single = []
    for e in range(total):
        hs = host_sim.HostSim(cfg.abi_params(arrays, env_id_base=0), env_index=e)
        hs.set_policy("table", table)
        hs.run()
        single.append(int(hs.counters()["packet_hops"]))
    assert sharded.tolist() == single
    assert len(set(single)) > 1, "envs draw different traffic"


def test_shard_ranges():
    from prisma_b200.config import EnvConfig
    for total in (1, 7, 4096, 65536):
        for world in (1, 2, 3, 8):
            cfg = EnvConfig(num_envs=total)
            seen = []
            for r in range(world):
                first, n = cfg.shard(r, world)
                seen.extend(range(first, first + n))
            assert seen == list(range(total))
