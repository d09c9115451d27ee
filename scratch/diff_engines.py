import sys; sys.path.insert(0,'.')
import numpy as np
from tests import common
from prisma_b200.env import BatchedRoutingEnv
cfg = common.env_config("abilene", num_envs=1024, simTime=2.0, load_factor=0.9, pingAsObs=0, activateUnderlayTraffic=1)
arrays = cfg.load(); table = common.sp_table(arrays)
res = {}
for mode in ("serial", "parallel"):
    env = BatchedRoutingEnv(cfg, arrays=arrays); env.set_engine(mode); env.set_policy("table", table)
    env.run(until_s=0.75); env.run(); res[mode] = env.counters(); env.close()
a, b = res["serial"], res["parallel"]
for k in common.COUNTER_FLOATS:
    d = np.nonzero(a[k].view(np.uint32) != b[k].view(np.uint32))[0]
    print(k, "differs in envs", d[:10].tolist(), [(float(a[k][i]), float(b[k][i])) for i in d[:3]])
