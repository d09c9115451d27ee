"""The node-parallel engine of prb_run (k_run_par, prisma_b200/csrc/par_core.cuh) on the GPU, through the C
ABI: against the committed outputs of the reference's ns-3 program, against the CPU restatement, and against
the event-serial engine at batch sizes.  The engine executes independent events of an environment
concurrently, so event logs are compared as multisets of complete records; counters, float32 statistics and
queue occupancies bit for bit."""
import struct

import numpy as np
import pytest

from oracle import oracle
from tests import common, golden_util
from tests.test_par_host import assert_same_multiset, oracle_uid_map

pytestmark = pytest.mark.gpu


def make_env(cfg, **kw):
    from prisma_b200.env import BatchedRoutingEnv
    return BatchedRoutingEnv(cfg, **kw)


def node_trace(g):
    from tests.host_harness.host_sim import node_trace as nt
    return nt(g.dec["node"], g.dec["action"], g.arrays.overlay_to_underlay())


@pytest.mark.parametrize("team", [0, 32])
@pytest.mark.parametrize("name", golden_util.names(parallel=True))
def test_parallel_engine_replays_reference_golden(name, team):
    g = golden_util.Golden(name)
    g.cfg.num_envs = 5
    env = make_env(g.cfg, arrays=g.arrays)
    inj, policy, table, seed = common.replay_inputs(g)
    env.set_inject_trace(*inj)
    env.set_engine("parallel", team)
    env.enable_log(2 * len(g.events), 0)
    env.reset()
    env.set_policy("node_trace", node_trace(g))
    env.run()
    assert env.engine == "parallel"
    cs = env.counters()
    assert (cs["errors"] == 0).all()
    for e in (0, 4):
        assert_same_multiset(g.events, env.events(e))
        c = cs[e]
        for key, v in g.counters.items():
            assert int(c[key]) == v, key
        for key, v in g.float_bits.items():
            assert struct.unpack("I", struct.pack("f", float(c[key])))[0] == v, key
    env.close()


CASES = [
    ("abilene", dict(simTime=3, load_factor=0.9, pingAsObs=0, activateUnderlayTraffic=1)),
    ("abilene", dict(simTime=3, load_factor=1.6, max_out_buffer_size=2168, pingAsObs=1, activateUnderlayTraffic=1)),
    ("overlay_full_mesh_3n_abilene", dict(simTime=8, load_factor=2.5, pingAsObs=1, activateUnderlayTraffic=1, train=1,
                                          signaling_type="target", agent_type="dqn_buffer", max_out_buffer_size=3000)),
    ("geant", dict(simTime=1.5, load_factor=1.3, pingAsObs=0, activateUnderlayTraffic=1)),
    # BASELINE.json configs[4]: 100-node random underlay, 4-packet buffers (drop-heavy), control packets
    ("synthetic100", dict(simTime=1.5, load_factor=1.4, max_out_buffer_size=2168, pingAsObs=0, train=1,
                          signaling_type="NN", agent_type="dqn_buffer", activateUnderlayTraffic=1)),
]


@pytest.mark.parametrize("topo,over", CASES)
def test_parallel_engine_matches_oracle_synthetic_traffic(topo, over):
    cfg = common.env_config(topo, num_envs=130, **over)
    arrays = cfg.load()
    table = common.sp_table(arrays)
    env = make_env(cfg, arrays=arrays, env_id_base=1000)
    env.set_engine("parallel")
    env.enable_log(400000, 0)
    env.reset()
    env.set_policy("table", table)
    env.run()
    cs = env.counters()
    env.check_errors(cs)
    for e in (0, 37, 129):
        osim = common.oracle_run(cfg, arrays, oracle, policy="table", table=table, env_id=1000 + e)
        common.assert_counters_equal(osim.counters(), cs[e])
        assert np.array_equal(osim.queue_bytes(), env.queue_bytes(e))
        umap, _ = oracle_uid_map(osim.events(), arrays)
        assert_same_multiset(osim.events(), env.events(e), umap)
    rounds, executed = env.round_stats()
    assert executed > rounds
    env.close()


def test_parallel_and_serial_engines_agree_on_a_batch():
    """1024 Abilene environments, 2 simulated seconds in two calls: every counter of every environment."""
    cfg = common.env_config("abilene", num_envs=1024, simTime=2.0, load_factor=0.9, pingAsObs=0, activateUnderlayTraffic=1)
    arrays = cfg.load()
    table = common.sp_table(arrays)
    res = {}
    for mode in ("serial", "parallel"):
        env = make_env(cfg, arrays=arrays)
        env.set_engine(mode)
        env.set_policy("table", table)
        env.run(until_s=0.75)
        env.run()
        assert env.engine == mode
        cs = env.counters()
        env.check_errors(cs)
        res[mode] = (cs, np.stack([env.queue_bytes(e) for e in (0, 511, 1023)]))
        env.close()
    a, b = res["serial"][0], res["parallel"][0]
    for k in common.COUNTER_INTS:
        assert np.array_equal(a[k], b[k]), k
    # float32 statistics: the sums are taken in the reference's order; deliveries at DIFFERENT nodes in the same
    # nanosecond that were also scheduled in the same nanosecond (lock-step transmissions after a ping round)
    # are folded in lane order, which can move a sum by one ulp.  Specification: 1e-6 relative.
    for k in common.COUNTER_FLOATS:
        x, y = a[k].astype(np.float64), b[k].astype(np.float64)
        assert np.all(np.abs(x - y) <= 1e-6 * np.maximum(np.abs(x), 1e-30)), k
        assert np.mean(a[k].view(np.uint32) != b[k].view(np.uint32)) < 0.02, k
    assert np.array_equal(res["serial"][1], res["parallel"][1])
    assert int(a["packet_hops"].sum()) > 3_000_000


def test_engine_rules():
    from prisma_b200 import _abi
    cfg = common.env_config("abilene", num_envs=8, simTime=0.5, load_factor=0.9, pingAsObs=0, activateUnderlayTraffic=1)
    arrays = cfg.load()
    env = make_env(cfg, arrays=arrays)
    env.set_policy("table", common.sp_table(arrays))
    env.run(until_s=0.25)
    assert env.engine == "parallel"                      # auto: table agent, no ordered log
    with pytest.raises(_abi.PrbError):
        env.step_host(None)                              # an episode stays on its engine
    env.reset()
    assert env.engine == "none"
    recs = env.step_host(None)
    assert recs["valid"].all() and env.engine == "serial"
    env.run()                                            # continues on the serial engine
    assert env.engine == "serial"
    env.check_errors()
    env.reset()
    env.set_policy("random", seed=3)
    env.run()
    assert env.engine == "serial"                        # agents with global state run event by event
    env.close()


def test_drop_heavy_100_nodes_four_nodes_per_lane():
    """A 100-node random underlay with 4-packet buffers (BASELINE.json configs[4] shape): the parallel engine
    (4 nodes per lane) and the serial engine give identical counters."""
    rng = np.random.default_rng(100)
    n = 100
    adj = np.zeros((n, n), dtype=np.uint8)
    for i in range(1, n):
        j = int(rng.integers(0, i))
        adj[i, j] = adj[j, i] = 1
    for _ in range(70):
        i, j = (int(x) for x in rng.integers(0, n, 2))
        if i != j:
            adj[i, j] = adj[j, i] = 1
    tm = rng.integers(1000, 12000, size=(n, n)).astype(np.uint64)
    np.fill_diagonal(tm, 0)
    from prisma_b200.config import EnvConfig
    cfg = EnvConfig(physical_adjacency_matrix=adj, overlay_adjacency_matrix=adj, map_overlay=np.arange(n),
                    traffic_matrix_bps=tm, num_envs=16, simTime=1.0, load_factor=1.0, max_out_buffer_size=2168,
                    pingAsObs=0, activateUnderlayTraffic=1, train=1, signaling_type="target", agent_type="dqn_buffer")
    arrays = cfg.load()
    from prisma_b200.env import sp_table
    table = sp_table(arrays)
    res = {}
    for mode in ("serial", "parallel"):
        env = make_env(cfg, arrays=arrays)
        env.set_engine(mode)
        env.set_policy("table", table)
        env.run()
        cs = env.counters()
        env.check_errors(cs)
        res[mode] = cs
        env.close()
    for k in common.COUNTER_INTS:
        assert np.array_equal(res["serial"][k], res["parallel"][k]), k
    for k in common.COUNTER_FLOATS:
        x, y = res["serial"][k].astype(np.float64), res["parallel"][k].astype(np.float64)
        assert np.all(np.abs(x - y) <= 1e-6 * np.maximum(np.abs(x), 1e-30)), k
    assert int(res["parallel"]["overlay_lost"].sum()) > 0


HANDOVER = [
    # GEANT at load 1.5 (full queues, long lock-step chains after the ping rounds) with the lineage walk cut short
    # (PRB_LINEAGE_WALK_MAX=4): about one environment in twelve meets a tie it cannot decide within four generations
    ("geant", dict(load_factor=1.5), 1600, (0.5, 1.0, 1.5), 1558),
    # small signalling packets: in environment 2853 a queue holds more records than the link ring has slots
    ("abilene", dict(load_factor=0.9, train=1, signaling_type="target", agent_type="dqn_buffer"), 2900, (1.0, 2.0), 2853),
]


@pytest.mark.parametrize("topo,over,num_envs,times,probe", HANDOVER)
def test_handover_to_the_serial_engine(topo, over, num_envs, times, probe, monkeypatch):
    """What the parallel engine cannot finish exactly (an undecided order of simultaneous events, a full link
    ring) is replayed by the serial engine: every counter of every environment still equals the serial run, at
    every read-out."""
    monkeypatch.setenv("PRB_LINEAGE_WALK_MAX", "4")
    cfg = common.env_config(topo, num_envs=num_envs, simTime=times[-1], pingAsObs=0, activateUnderlayTraffic=1, **over)
    arrays = cfg.load()
    table = common.sp_table(arrays)
    res = {}
    for mode in ("serial", "parallel"):
        env = make_env(cfg, arrays=arrays)
        env.set_engine(mode)
        env.set_policy("table", table)
        out = []
        for t in times:
            env.run(until_s=t)
            cs = env.counters()
            env.check_errors(cs)
            out.append((cs, env.queue_bytes(probe), env.queue_bytes(7)))
        if mode == "parallel":
            assert env.engine == "parallel"
            assert env.handovers() >= 1
        res[mode] = out
        env.close()
    for (a, qa, ra), (b, qb, rb) in zip(res["serial"], res["parallel"]):
        for k in common.COUNTER_INTS:
            assert np.array_equal(a[k], b[k]), k
        for k in common.COUNTER_FLOATS:
            x, y = a[k].astype(np.float64), b[k].astype(np.float64)
            assert np.all(np.abs(x - y) <= 1e-6 * np.maximum(np.abs(x), 1e-30)), k
        assert np.array_equal(qa, qb) and np.array_equal(ra, rb)


def test_bench_size_batch_engines_agree_and_conserve_packets():
    """BASELINE.json configs[1] at its full size (4096 Abilene environments, load 0.9, next-hop table): the two engines
    agree on every integer counter of every environment after one simulated second, and every environment obeys
    the conservation law of compute-stats-v2.cc (injected = arrived + lost + still in the network)."""
    cfg = common.env_config("abilene", num_envs=4096, simTime=1.0, load_factor=0.9, pingAsObs=0, activateUnderlayTraffic=1,
                            agent_type="dqn_routing")
    arrays = cfg.load()
    table = common.sp_table(arrays)
    res = {}
    for mode in ("serial", "parallel"):
        env = make_env(cfg, arrays=arrays)
        env.set_engine(mode)
        env.set_policy("table", table)
        env.run()
        cs = env.counters()
        env.check_errors(cs)
        res[mode] = cs
        env.close()
    a, b = res["serial"], res["parallel"]
    for k in common.COUNTER_INTS:
        assert np.array_equal(a[k], b[k]), k
    for c in (a, b):
        for kind in ("overlay", "underlay"):
            left = c[kind + "_injected"] - c[kind + "_arrived"] - c[kind + "_lost"]
            assert (left >= 0).all() and (left < 400).all(), kind
    assert int(b["packet_hops"].sum()) > 4_000_000 and len(np.unique(b["packet_hops"])) > 300
    # env 4095 against the restatement
    osim = common.oracle_run(cfg, arrays, oracle, policy="table", table=table, env_id=4095)
    common.assert_counters_equal(osim.counters(), b[4095])


def test_run_to_an_earlier_time_is_a_noop_on_both_engines():
    """prb_run(until) with until at or before the time already reached does nothing (times inside a parallel launch are
    unsigned offsets from the frontier, which never moves backwards): run(1.0); run(0.5); run(2.0) equals the
    restatement at 2.0 on both engines."""
    cfg = common.env_config("abilene", num_envs=8, simTime=3.0, load_factor=1.0, pingAsObs=0, activateUnderlayTraffic=1)
    arrays = cfg.load()
    table = common.sp_table(arrays)
    osim = common.oracle_run(cfg, arrays, oracle, policy="table", table=table, env_id=5, until_ns=2_000_000_000)
    for mode in ("serial", "parallel"):
        env = make_env(cfg, arrays=arrays)
        env.set_engine(mode)
        env.set_policy("table", table)
        env.run(until_s=1.0)
        mid = env.counters()
        env.run(until_s=0.5)
        again = env.counters()
        for k in common.COUNTER_INTS:
            assert np.array_equal(mid[k], again[k]), (mode, k)
        env.run(until_s=2.0)
        cs = env.counters()
        env.check_errors(cs)
        assert env.engine == mode
        common.assert_counters_equal(osim.counters(), cs[5])
        assert np.array_equal(osim.queue_bytes(), env.queue_bytes(5))
        env.close()


def test_handover_replays_every_stretch_with_the_table_it_ran_with(monkeypatch):
    """prb_set_policy between prb_run calls of one episode (a periodically refreshed dq_routing table): an environment
    handed over AFTER the change is replayed by the serial engine with the old table for the old stretch and the new one
    from then on (GEANT load 1.5 with the lineage walk cut to four generations, so that environments do hand over; the table changes
    at 0.4 s and at 0.8 s)."""
    monkeypatch.setenv("PRB_LINEAGE_WALK_MAX", "4")
    cfg = common.env_config("geant", num_envs=1600, simTime=1.2, pingAsObs=0, activateUnderlayTraffic=1, load_factor=1.5)
    arrays = cfg.load()
    t0 = np.asarray(common.sp_table(arrays), dtype=np.int32)
    deg = np.array([len(x) for x in arrays.overlay_neighbors()], dtype=np.int32)
    t1 = (t0 + 1) % deg[:, None]          # every node deflects to its next neighbour: legal, and very different
    t2 = t0.copy()
    t2[::2] = t1[::2]
    res = {}
    for mode in ("serial", "parallel"):
        env = make_env(cfg, arrays=arrays)
        env.set_engine(mode)
        out = []
        for until, tab in ((0.4, t0), (0.8, t1), (1.2, t2)):
            env.set_policy("table", tab)
            env.run(until_s=until)
            cs = env.counters()
            env.check_errors(cs)
            out.append((cs, env.queue_bytes(1558)))
        if mode == "parallel":
            n, ms = env.handover_stats()
            assert n >= 1 and n == env.handovers() and ms > 0.0
        res[mode] = out
        env.close()
    for (a, qa), (b, qb) in zip(res["serial"], res["parallel"]):
        for k in common.COUNTER_INTS:
            assert np.array_equal(a[k], b[k]), k
        assert np.array_equal(qa, qb)


@pytest.mark.parametrize("topo,lf,envs", [("abilene", 0.9, 4096), ("geant", 1.5, 2048)])
def test_full_length_episodes_need_no_handover(topo, lf, envs):
    """A whole 60-second episode (the reference's simTime, argument_parser.py:35-37) of the headline configuration
    (BASELINE.json configs[1]: 4096 Abilene environments at load 0.9) and of GEANT at load 1.5 on the node-parallel engine:
    no environment is handed over to the serial engine (round 1: 2.4 % of Abilene's environments by t = 60 s), packets are
    conserved, and the saturated second half of the episode really drops packets."""
    cfg = common.env_config(topo, num_envs=envs, simTime=60.0, load_factor=lf, pingAsObs=0, activateUnderlayTraffic=1,
                            agent_type="dqn_routing", seed=100)
    arrays = cfg.load()
    env = make_env(cfg, arrays=arrays)
    env.set_policy("table", common.sp_table(arrays))
    for t in range(10, 61, 10):
        env.run(until_s=float(t))
        assert env.handovers() == 0, t
    assert env.engine == "parallel"
    n, ms = env.handover_stats()
    assert n == 0 and ms == 0.0
    c = env.counters()
    env.check_errors(c)
    inflight = c["overlay_injected"] - c["overlay_arrived"] - c["overlay_lost"]
    assert (inflight >= 0).all() and (inflight < 2000).all()
    assert int(c["overlay_lost"].sum()) > 0 and (c["end_ns"] == 60_000_000_000).all()
    env.close()
