"""Node-parallel engine (prisma_b200/csrc/par_core.cuh) on the CPU: the lanes of a team are emulated one
after the other, in a shuffled order inside each round (tests/host_harness).  The engine executes
independent events of one environment out of global order, so its event log is compared with the
reference / oracle log as a MULTISET of complete records (every enqueue with its queue bytes, every drop,
every tx start and every delivery with its nanosecond timestamp), and counters, float32 statistics and
final queue occupancies bit for bit."""
import struct

import numpy as np
import pytest

from oracle import oracle
from tests import common, golden_util
from tests.host_harness import host_sim

FIELDS = ["t_ns", "kind", "node", "link", "uid", "size", "type", "src", "dst", "last_hop", "next_hop", "valable", "qbytes"]


def canon(ev, uid_map=None):
    """Sorted (n, 13) int64 matrix of complete event records."""
    a = np.stack([ev[f].astype(np.int64) for f in FIELDS], axis=1)
    if uid_map is not None:
        u = a[:, 4]
        a[:, 4] = np.where(u >= 0, np.array([uid_map.get(int(x), -2) for x in u], dtype=np.int64), u)
    order = np.lexsort(a.T[::-1])
    return a[order]


def oracle_uid_map(ev, arrays, underlay_traffic=True):
    """Synthetic traffic: the oracle numbers packets in global injection order, the parallel engine as
    ordinal * F + flow.  Map the former onto the latter through the injection records.  Without underlay
    traffic only the pairs of overlay nodes have a source (sim.cc:505-507)."""
    flows = {}
    n = arrays.n_nodes
    f = 0
    m = arrays.map_overlay
    act = arrays.traffic_bps > 0
    if not underlay_traffic:
        act = act & (m[:, None] >= 0) & (m[None, :] >= 0)
    for i in range(n):
        for j in range(n):
            if i != j and act[i, j]:
                flows[(i, j)] = f
                f += 1
    inj = ev[ev["kind"] == 0]
    count = {}
    out = {}
    for s, d, u in zip(inj["src"], inj["dst"], inj["uid"]):
        k = flows[(int(s), int(d))]
        count[k] = count.get(k, 0) + 1
        out[int(u)] = count[k] * f + k
    return out, f


def assert_same_multiset(ref_ev, par_ev, uid_map=None):
    a, b = canon(ref_ev, uid_map), canon(par_ev)
    assert a.shape == b.shape, (a.shape, b.shape)
    bad = np.nonzero((a != b).any(axis=1))[0]
    assert len(bad) == 0, (int(bad[0]), a[bad[0]].tolist(), b[bad[0]].tolist())


@pytest.mark.parametrize("team", [16, 32])
@pytest.mark.parametrize("name", golden_util.names(parallel=True))
def test_parallel_engine_replays_reference_golden(name, team):
    """Real ns-3 outputs: same injections, the agents' actions replayed per node."""
    g = golden_util.Golden(name)
    hp = host_sim.HostParSim(g.cfg.abi_params(g.arrays), team=team)
    inj, policy, table, seed = common.replay_inputs(g)
    hp.set_inject_trace(*inj)
    o2u = g.arrays.overlay_to_underlay()
    hp.set_policy("node_trace", host_sim.node_trace(g.dec["node"], g.dec["action"], o2u))
    hp.run()
    c = hp.counters()
    assert c["errors"] == 0
    assert_same_multiset(g.events, hp.events())
    for key, v in g.counters.items():
        assert c[key] == v, key
    for key, v in g.float_bits.items():
        assert struct.unpack("I", struct.pack("f", c[key]))[0] == v, key


CASES = [
    ("abilene", dict(simTime=3, load_factor=0.9, pingAsObs=0, activateUnderlayTraffic=1), 0, 16),
    ("abilene", dict(simTime=3, load_factor=1.6, max_out_buffer_size=2168, pingAsObs=1, activateUnderlayTraffic=1), 5, 16),
    ("overlay_full_mesh_3n_abilene", dict(simTime=8, load_factor=2.5, pingAsObs=1, activateUnderlayTraffic=1,
                                          train=1, signaling_type="target", agent_type="dqn_buffer",
                                          max_out_buffer_size=3000), 123456789012, 8),
    ("geant", dict(simTime=1.5, load_factor=1.3, pingAsObs=0, activateUnderlayTraffic=1), 77, 32),
    ("geant", dict(simTime=1.0, load_factor=1.0, pingAsObs=0, activateUnderlayTraffic=1), 3, 8),   # 3 nodes per lane
    # BASELINE.json configs[4]: 100-node random underlay (4 nodes per lane), 4-packet buffers, control packets;
    # and the same construction at 56 nodes, the largest the reference itself can address (pinned in test_oracle_vs_ref)
    ("synthetic100", dict(simTime=1.5, load_factor=1.4, max_out_buffer_size=2168, pingAsObs=0, train=1,
                          signaling_type="NN", agent_type="dqn_buffer", activateUnderlayTraffic=1), 11, 32),
    ("synthetic56", dict(simTime=1.5, load_factor=1.4, max_out_buffer_size=2168, pingAsObs=0, train=1,
                          signaling_type="NN", agent_type="dqn_buffer", activateUnderlayTraffic=1), 4, 32),
    ("synthetic100", dict(simTime=1.2, load_factor=1.0, pingAsObs=1, activateUnderlayTraffic=1), 5, 32),
    # a source whose next packet follows 4 ns after the previous one (the access link's serialisation time): the
    # arrival at the node and the next injection fall into the same nanosecond (found by the 8-GPU bench)
    ("abilene", dict(simTime=1.3, load_factor=0.9, pingAsObs=0, activateUnderlayTraffic=1, agent_type="dqn_routing",
                     seed=100), 32284, 16),
]


@pytest.mark.parametrize("topo,over,env_id,team", CASES)
def test_parallel_engine_matches_oracle_synthetic_traffic(topo, over, env_id, team):
    cfg = common.env_config(topo, **over)
    arrays = cfg.load()
    table = common.sp_table(arrays)
    osim = common.oracle_run(cfg, arrays, oracle, policy="table", table=table, env_id=env_id)
    hp = host_sim.HostParSim(cfg.abi_params(arrays, env_id_base=env_id), team=team)
    hp.set_policy("table", table)
    hp.run()
    pc = hp.counters()
    common.assert_counters_equal(osim.counters(), pc)
    assert np.array_equal(osim.queue_bytes(), hp.queue_bytes())
    umap, _ = oracle_uid_map(osim.events(), arrays)
    assert_same_multiset(osim.events(), hp.events(), umap)
    assert hp.executed / hp.rounds > 1.0       # the nodes of one environment really advance concurrently


def test_parallel_engine_chunked_run_equals_single_run():
    cfg = common.env_config("abilene", simTime=2, load_factor=1.0, pingAsObs=0, activateUnderlayTraffic=1)
    arrays = cfg.load()
    table = common.sp_table(arrays)
    a = host_sim.HostParSim(cfg.abi_params(arrays), team=16)
    a.set_policy("table", table)
    a.run(log=False)
    b = host_sim.HostParSim(cfg.abi_params(arrays), team=16)
    b.set_policy("table", table)
    for k in range(1, 9):
        b.run(until_ns=k * 250_000_000, log=False)
        osim = common.oracle_run(cfg, arrays, oracle, policy="table", table=table, until_ns=k * 250_000_000)
        common.assert_counters_equal(osim.counters(), b.counters())
    ca, cb = a.counters(), b.counters()
    assert ca == cb
    assert np.array_equal(a.queue_bytes(), b.queue_bytes())


def test_parallel_engine_many_envs_against_oracle():
    """Different environments draw different traffic: 24 of them, counters and queues bit for bit."""
    cfg = common.env_config("abilene", simTime=1.0, load_factor=1.2, pingAsObs=0, activateUnderlayTraffic=1)
    arrays = cfg.load()
    table = common.sp_table(arrays)
    for e in range(24):
        osim = common.oracle_run(cfg, arrays, oracle, policy="table", table=table, env_id=1000 + e)
        hp = host_sim.HostParSim(cfg.abi_params(arrays, env_id_base=1000), env_index=e, team=16)
        hp.set_policy("table", table)
        hp.run(log=False)
        common.assert_counters_equal(osim.counters(), hp.counters())
        assert np.array_equal(osim.queue_bytes(), hp.queue_bytes())


def test_parallel_engine_rebases_time_offsets():
    """Times inside a launch are 32-bit offsets from its start: 37 ms launches (an odd size, so that packets
    are in flight and queued across every boundary) against one launch, on the real ns-3 trace."""
    g = golden_util.Golden("abilene_random_drops_3s")
    o2u = g.arrays.overlay_to_underlay()
    inj, policy, table, seed = common.replay_inputs(g)
    hp = host_sim.HostParSim(g.cfg.abi_params(g.arrays), team=16)
    hp.set_chunk(37_000_001)
    hp.set_inject_trace(*inj)
    hp.set_policy("node_trace", host_sim.node_trace(g.dec["node"], g.dec["action"], o2u))
    hp.run()
    c = hp.counters()
    assert c["errors"] == 0
    assert_same_multiset(g.events, hp.events())
    for key, v in g.counters.items():
        assert c[key] == v, key
    for key, v in g.float_bits.items():
        assert struct.unpack("I", struct.pack("f", c[key]))[0] == v, key
