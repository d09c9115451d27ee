import sys
sys.path.insert(0, "/root/repo")
from tests import common
from tests.host_harness import host_sim
topo, lf, e, T, team = sys.argv[1], float(sys.argv[2]), int(sys.argv[3]), float(sys.argv[4]), int(sys.argv[5])
cfg = common.env_config(topo, simTime=8.0, activateUnderlayTraffic=1, load_factor=lf, pingAsObs=0)
arrays = cfg.load()
hp = host_sim.HostParSim(cfg.abi_params(arrays, env_id_base=0), env_index=e, team=team)
hp.set_policy("table", common.sp_table(arrays))
hp.run(until_ns=int(T * 1e9), log=False)
print("errors", hp.counters()["errors"])
