import sys
sys.path.insert(0, "/root/repo")
import numpy as np
from tests import common
from prisma_b200.env import BatchedRoutingEnv, sp_table
cfg = common.env_config("geant", num_envs=4096, simTime=8.0, activateUnderlayTraffic=1, load_factor=1.5, pingAsObs=0)
arrays = cfg.load()
env = BatchedRoutingEnv(cfg, arrays=arrays)
env.set_engine("parallel"); env.set_policy("table", sp_table(arrays))
for t in (0.5, 1.0, 2.0, 4.0):
    env.run(until_s=t)
    c = env.counters()
    bad = np.nonzero(c["errors"])[0]
    print(t, len(bad), bad[:10].tolist(), c["errors"][bad[:10]].tolist(), flush=True)
