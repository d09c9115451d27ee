import sys
sys.path.insert(0, "/root/repo")
import numpy as np
from tests import common
from prisma_b200.env import BatchedRoutingEnv, sp_table
cfg = common.env_config("abilene", num_envs=4096, simTime=8.0, activateUnderlayTraffic=1, load_factor=0.9, pingAsObs=0,
                        train=1, signaling_type="target", agent_type="dqn_buffer")
arrays = cfg.load()
for eng in ("parallel", "serial"):
    env = BatchedRoutingEnv(cfg, arrays=arrays)
    env.set_engine(eng); env.set_policy("table", sp_table(arrays))
    for t in (0.25, 0.5, 1.0, 2.0, 4.0):
        env.run(until_s=t)
        c = env.counters()
        bad = np.nonzero(c["errors"])[0]
        print(eng, t, len(bad), bad[:8].tolist(), c["errors"][bad[:8]].tolist(), flush=True)
        if len(bad): break
    env.close()
