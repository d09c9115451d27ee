import sys, time
sys.path.insert(0, "/root/repo")
from tests import common
from tests.host_harness import host_sim
import importlib.util
spec = importlib.util.spec_from_file_location("bench", "/root/repo/bench.py"); bench = importlib.util.module_from_spec(spec); spec.loader.exec_module(bench)
cfg = bench.workload_cfg(4096, 14.0)
arrays = cfg.load()
from prisma_b200.env import sp_table
table = sp_table(arrays)
base = int(sys.argv[1]); lo = int(sys.argv[2]); hi = int(sys.argv[3]); T = float(sys.argv[4])
t0 = time.time()
for e in range(lo, hi):
    hp = host_sim.HostParSim(cfg.abi_params(arrays, env_id_base=base), env_index=e, team=16)
    hp.set_policy("table", table)
    hp.run(until_ns=int(T * 1e9), log=False)
    c = hp.counters()
    if c["errors"]:
        print("env", base + e, "errors", c["errors"], flush=True)
print("done", time.time() - t0)
