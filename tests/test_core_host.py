"""Kernel body (prisma_b200/csrc/sim_core.cuh) compiled for the host and single-stepped against the
committed outputs of the reference and against the CPU restatement.  The host build is a debugging aid
of the test-suite for the GPU-less container (tests/host_harness); the GPU parity tests proper are in
test_gpu_parity.py and call the CUDA library through the C ABI."""
import struct

import numpy as np
import pytest

from oracle import oracle
from tests import common, golden_util
from tests.host_harness import host_sim


@pytest.mark.parametrize("name", golden_util.names())
def test_core_replays_reference_golden(name):
    g = golden_util.Golden(name)
    hs = host_sim.HostSim(g.cfg.abi_params(g.arrays))
    inj, policy, table, seed = common.replay_inputs(g)
    hs.set_inject_trace(*inj)
    hs.set_policy(policy, table, seed=seed)
    hs.run()
    assert common.first_event_mismatch(g.events, hs.events()) is None
    st = hs.steps()
    assert np.array_equal(st["t_ns"], g.dec["t"]) and np.array_equal(st["node"], g.dec["node"])
    assert np.array_equal(st["ptype"], g.dec["ptype"]) and np.array_equal(st["done"], g.dec["done"])
    assert np.array_equal(st["action"], g.dec["action"])
    data = g.dec["ptype"] == 0
    assert np.array_equal(st["uid"][data], g.dec["uid"][data])
    k = g.dec["obs"].shape[1]
    mask = np.arange(k)[None, :] < g.dec["nobs"][:, None]
    assert np.array_equal(np.where(mask, st["obs"][:, :k], -1), g.dec["obs"])
    if "big" in g.dec:
        common.assert_big_signaling_fields(g.dec["big"], st["uid"], st["prev_node"], g.arrays.map_overlay)
    c = hs.counters()
    assert c["errors"] == 0
    for key, v in g.counters.items():
        assert c[key] == v, key
    for key, v in g.float_bits.items():
        assert struct.unpack("I", struct.pack("f", c[key]))[0] == v, key


PHILOX_CASES = [
    ("abilene", dict(simTime=3, load_factor=0.9, pingAsObs=0, activateUnderlayTraffic=1), "table", 0),
    ("abilene", dict(simTime=3, load_factor=1.6, max_out_buffer_size=2168, pingAsObs=1, activateUnderlayTraffic=1), "random", 5),
    ("overlay_full_mesh_3n_abilene", dict(simTime=8, load_factor=2.5, pingAsObs=1, activateUnderlayTraffic=1,
                                          train=1, signaling_type="target", agent_type="dqn_buffer",
                                          max_out_buffer_size=3000), "random", 123456789012),
    ("geant", dict(simTime=1.5, load_factor=1.3, pingAsObs=0, activateUnderlayTraffic=1), "table", 77),
    # BASELINE.json configs[4]: 100-node random underlay, 4-packet buffers (drop-heavy), control packets
    ("synthetic100", dict(simTime=1.5, load_factor=1.4, max_out_buffer_size=2168, pingAsObs=0, train=1,
                          signaling_type="NN", agent_type="dqn_buffer", activateUnderlayTraffic=1), "random", 9),
    # NN-weight segment sources (big signalling) next to Philox traffic
    ("abilene", dict(simTime=2.5, load_factor=0.8, pingAsObs=0, activateUnderlayTraffic=1, train=1, signaling_type="NN",
                     signalingSim=1, sync_step=0.25, bigSignalingSize=4096, agent_type="dqn_buffer"), "random", 2),
]


@pytest.mark.parametrize("topo,over,policy,env_id", PHILOX_CASES)
def test_core_matches_oracle_synthetic_traffic(topo, over, policy, env_id):
    cfg = common.env_config(topo, **over)
    arrays = cfg.load()
    table = common.sp_table(arrays)
    osim = common.oracle_run(cfg, arrays, oracle, policy=policy, table=table, seed=11 + env_id, env_id=env_id)
    hs = host_sim.HostSim(cfg.abi_params(arrays, env_id_base=env_id - 3 if env_id >= 3 else 0),
                          env_index=3 if env_id >= 3 else env_id)
    # the random agent of env e is seeded with seed + e
    hs.set_policy(policy, table, seed=11 + env_id - (3 if env_id >= 3 else env_id))
    hs.run()
    oe, he = osim.events(), hs.events()
    assert len(oe) > 1000
    assert common.first_event_mismatch(oe, he) is None
    common.assert_steps_equal(osim.decisions(), hs.steps())
    common.assert_step_sources(oe, hs.steps())
    common.assert_counters_equal(osim.counters(), hs.counters())
    assert np.array_equal(osim.queue_bytes(), hs.queue_bytes())


def test_core_gym_mode_equals_run_mode():
    """Stepping Notify() by Notify() with the table's actions supplied from outside == fused run."""
    cfg = common.env_config("abilene", simTime=2, load_factor=1.3, pingAsObs=0, activateUnderlayTraffic=1,
                            max_out_buffer_size=4000, train=1, signaling_type="NN", agent_type="dqn_buffer")
    arrays = cfg.load()
    table = common.sp_table(arrays)
    a = host_sim.HostSim(cfg.abi_params(arrays))
    a.set_policy("table", table)
    a.run()
    b = host_sim.HostSim(cfg.abi_params(arrays))
    recs = []
    action = 0
    while True:
        r = b.step(action)
        if not r.valid:
            break
        action = int(table[r.overlay, r.obs[0]]) if r.need_action else 0
        recs.append((r.t_ns, r.overlay, r.uid, r.ptype, r.done, action if r.need_action else -1))
    sa = a.steps()
    assert len(recs) == len(sa) > 1000
    got = np.array(recs, dtype=np.int64)
    assert np.array_equal(got[:, 0], sa["t_ns"]) and np.array_equal(got[:, 1], sa["overlay"])
    assert np.array_equal(got[:, 2], sa["uid"]) and np.array_equal(got[:, 5], sa["action"])
    assert common.first_event_mismatch(a.events(), b.events()) is None
    ca, cb = a.counters(), b.counters()
    assert ca == cb


def test_core_chunked_run_equals_single_run():
    cfg = common.env_config("abilene", simTime=2, load_factor=1.0, pingAsObs=1, activateUnderlayTraffic=1)
    arrays = cfg.load()
    table = common.sp_table(arrays)
    a = host_sim.HostSim(cfg.abi_params(arrays))
    a.set_policy("table", table)
    a.run()
    b = host_sim.HostSim(cfg.abi_params(arrays))
    b.set_policy("table", table)
    for t in (1, 250_000_000, 250_000_001, 900_000_000, 1_999_999_999, 2_000_000_000, 5_000_000_000):
        b.run(t)
    assert common.first_event_mismatch(a.events(), b.events()) is None
    assert a.counters() == b.counters()


def test_time_arithmetic_matches_oracle():
    """ns-3 Time/DataRate arithmetic of the product (host side and the integer emulation the kernels use)
    against the restatement that is pinned to the reference binary (test_oracle_vs_ref.py)."""
    L = host_sim.lib()
    rng = np.random.default_rng(3)
    hi, lo = (host_sim.C.c_uint64(), host_sim.C.c_uint64())
    L.hs_invert_1e9(host_sim.C.byref(hi), host_sim.C.byref(lo))
    ohi, olo = (host_sim.C.c_uint64(), host_sim.C.c_uint64())
    oracle.lib().po_invert_1e9(host_sim.C.byref(ohi), host_sim.C.byref(olo))
    assert (hi.value, lo.value) == (ohi.value, olo.value)
    for d in list(rng.random(500)) + list(rng.random(100) * 1e-3) + list(rng.random(100) * 70) + \
            [0.001, 0.2, float(np.float32(0.2)), 60.0, 0.0001, 1.0001, 0.0, 1e-9, 2.60]:
        assert L.hs_seconds_to_ns(d) == oracle.seconds_to_ns(d), d
    for size in (38, 46, 54, 62, 110, 200, 230, 542, 1000, 1500):
        for bps in (500000, 1000000, 155520000, 5 * 10 ** 11, 15 * 10 ** 11, 45 * 10 ** 11, 123457):
            assert L.hs_tx_time_ns(size, bps) == oracle.tx_time_ns(size, bps)
    cfg = common.env_config("abilene")
    hs = host_sim.HostSim(cfg.abi_params())
    ns = [int(x) for x in rng.integers(0, 70 * 10 ** 9, 20000)] + [0, 1, 999999999, 10 ** 9, 10 ** 9 - 1, 60 * 10 ** 9] + \
         [k * 10 ** 9 + d for k in range(0, 70, 7) for d in (-1, 0, 1) if k * 10 ** 9 + d >= 0] + \
         [int(x) for x in rng.integers(0, 2 ** 62, 2000)]
    for v in ns:
        a = L.hs_ns_to_seconds(hs.h, v)
        b = oracle.ns_to_seconds(v)
        assert struct.pack("d", a) == struct.pack("d", b), v


@pytest.mark.parametrize("topo", common.TOPOLOGIES)
def test_global_routes_match_oracle(topo):
    arrays = common.env_config(topo).load()
    n = arrays.n_nodes
    out = np.zeros((n, n), dtype=np.int32)
    host_sim.lib().hs_global_routes(n, arrays.phys_adj.ctypes.data_as(host_sim.C.POINTER(host_sim.C.c_uint8)),
                                    out.ctypes.data_as(host_sim.C.POINTER(host_sim.C.c_int32)))
    assert np.array_equal(out, oracle.global_routes(arrays.phys_adj))


def test_core_record_free_fast_path_equals_logged_run():
    """Fused runs that nobody logs skip the Gym step record (sim_core.cuh Sim::REC == false): counters,
    queues and clock must not change."""
    for policy in ("table", "random"):
        cfg = common.env_config("abilene", simTime=2, load_factor=1.4, max_out_buffer_size=3000, pingAsObs=1,
                                activateUnderlayTraffic=1, train=1, signaling_type="NN", agent_type="dqn_buffer")
        arrays = cfg.load()
        table = common.sp_table(arrays)
        a = host_sim.HostSim(cfg.abi_params(arrays))
        a.set_policy(policy, table, seed=5)
        a.run()
        b = host_sim.HostSim(cfg.abi_params(arrays))
        b.set_policy(policy, table, seed=5)
        b.run_nolog()
        assert a.counters() == b.counters()
        assert np.array_equal(a.queue_bytes(), b.queue_bytes())


def test_core_sparse_timers_beyond_key_range():
    """Scan keys are 32-bit offsets from an epoch (4.29 s range).  With a 7 s ping interval and a trickle of
    traffic every armed timer is regularly farther away than that: the exact 64-bit fallback must agree
    with the restatement event by event."""
    cfg = common.env_config("overlay_full_mesh_3n_abilene", simTime=60, load_factor=0.004, pingAsObs=1,
                            activateUnderlayTraffic=1, pingPacketIntervalTime=7.0)
    arrays = cfg.load()
    table = common.sp_table(arrays)
    osim = common.oracle_run(cfg, arrays, oracle, policy="table", table=table, env_id=2)
    hs = host_sim.HostSim(cfg.abi_params(arrays), env_index=2)
    hs.set_policy("table", table)
    hs.run()
    oe = osim.events()
    assert 50 < len(oe) < 3000
    assert common.first_event_mismatch(oe, hs.events()) is None
    common.assert_steps_equal(osim.decisions(), hs.steps())
    common.assert_counters_equal(osim.counters(), hs.counters())


def test_whole_seconds_of_a_timestamp_is_the_integer_quotient():
    """`uint64_t (Simulator::Now ().GetSeconds ())` (poisson-application.cc:310) goes through the Q64.64 product
    with Invert (1e9) and a double; the parallel engine uses ns / 10^9.  Equal at every second boundary +-3 ns
    and on random values (2e8 random values were checked once offline)."""
    L = host_sim.lib()
    rng = np.random.default_rng(5)
    vals = [k * 1_000_000_000 + d for k in range(0, 4000, 7) for d in (-3, -2, -1, 0, 1, 2, 3) if k * 1_000_000_000 + d >= 0]
    vals += [int(v) for v in rng.integers(0, 1 << 44, 20000)]
    for ns in vals:
        assert int(L.hs_ns_to_seconds(None, ns)) == ns // 1_000_000_000, ns
