"""Host-side census of undecided ties: runs the node-parallel engine's host build on a range of environments of a
bench-like workload (one process per core) and prints the environments that end with an error flag.
usage: tie_census.py <topology> <load> <first env> <count> <seconds> [team]"""
import sys, time
sys.path.insert(0, "/root/repo")
from multiprocessing import Pool
from tests import common
from tests.host_harness import host_sim
from prisma_b200.env import sp_table

topo, lf, lo, cnt, T = sys.argv[1], float(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), float(sys.argv[5])
team = int(sys.argv[6]) if len(sys.argv) > 6 else 16


def one(e):
    cfg = common.env_config(topo, simTime=T, activateUnderlayTraffic=1, load_factor=lf, pingAsObs=0)
    arrays = cfg.load()
    hp = host_sim.HostParSim(cfg.abi_params(arrays, env_id_base=0), env_index=e, team=team)
    hp.set_policy("table", sp_table(arrays))
    hp.run(until_ns=int(T * 1e9), log=False)
    c = hp.counters()
    return e, c["errors"], c["packet_hops"]


if __name__ == "__main__":
    host_sim.lib()
    t0 = time.time()
    bad = []
    with Pool(8) as p:
        for e, err, hops in p.imap_unordered(one, range(lo, lo + cnt)):
            if err:
                bad.append((e, err))
                print("env", e, "errors", err, flush=True)
    print("done", cnt, "envs,", len(bad), "flagged,", round(time.time() - t0, 1), "s")
