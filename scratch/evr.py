import sys, time
sys.path.insert(0, "/root/repo")
from tests import common
from tests.host_harness import host_sim
cfg = common.env_config("abilene", simTime=2.0, load_factor=0.9, pingAsObs=0, activateUnderlayTraffic=1)
arrays = cfg.load()
table = common.sp_table(arrays)
tot_r = tot_e = 0
for e in range(4):
    hp = host_sim.HostParSim(cfg.abi_params(arrays, env_id_base=0), env_index=e, team=16)
    hp.set_policy("table", table)
    hp.run(log=False)
    assert hp.counters()["errors"] == 0
    tot_r += hp.rounds; tot_e += hp.executed
print("events/round %.3f  rounds %d" % (tot_e / tot_r, tot_r))
