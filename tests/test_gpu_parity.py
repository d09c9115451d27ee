"""Parity tests proper: the CUDA path (libprisma_b200.so through the C ABI) against the committed outputs
of the reference's ns-3 program (tests/golden), against the CPU restatement on seeded synthetic traffic,
and size-independent properties at the full batch sizes of BASELINE.json.  Integer / index / timestamp
results are compared bit for bit; float32 statistics bit for bit as well (the kernel executes one
environment's events in the reference's order, so even the float sums agree), which is tighter than the
1e-6 relative the specification asks for."""
import struct

import numpy as np
import pytest

from oracle import oracle
from tests import common, golden_util

pytestmark = pytest.mark.gpu


def make_env(cfg, **kw):
    from prisma_b200.env import BatchedRoutingEnv
    return BatchedRoutingEnv(cfg, **kw)


@pytest.mark.parametrize("lanes", [32, 8])
@pytest.mark.parametrize("name", golden_util.names())
def test_cuda_replays_reference_golden(name, lanes):
    g = golden_util.Golden(name)
    g.cfg.num_envs = 3
    env = make_env(g.cfg, arrays=g.arrays, lanes_per_env=lanes)
    inj, policy, table, seed = common.replay_inputs(g)
    env.set_inject_trace(*inj)
    env.enable_log(2 * len(g.events), 2 * len(g.dec["t"]))   # envs 1, 2 of a random agent draw other actions
    env.reset()
    env.set_policy(policy, table, seed=seed)
    env.run()
    cs = env.counters()
    assert (cs["errors"] == 0).all()
    ev = env.events(0)
    assert common.first_event_mismatch(g.events, ev) is None
    st = env.steps(0)
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
    c = cs[0]
    for key, v in g.counters.items():
        assert int(c[key]) == v, key
    for key, v in g.float_bits.items():
        assert struct.unpack("I", struct.pack("f", float(c[key])))[0] == v, key
    if policy == "table":
        # every env replays the same trace with the same table: identical logs
        assert np.array_equal(env.events(2), ev)
    env.close()


PHILOX_CASES = [
    ("abilene", dict(simTime=3, load_factor=0.9, pingAsObs=0, activateUnderlayTraffic=1), "table"),
    ("abilene", dict(simTime=3, load_factor=1.6, max_out_buffer_size=2168, pingAsObs=1, activateUnderlayTraffic=1), "random"),
    ("overlay_full_mesh_3n_abilene", dict(simTime=8, load_factor=2.5, pingAsObs=1, activateUnderlayTraffic=1, train=1,
                                          signaling_type="target", agent_type="dqn_buffer", max_out_buffer_size=3000), "random"),
    ("geant", dict(simTime=1.5, load_factor=1.3, pingAsObs=0, activateUnderlayTraffic=1), "table"),
    # BASELINE.json configs[4]: 100-node random underlay, 4-packet buffers (drop-heavy), control packets
    ("synthetic100", dict(simTime=1.5, load_factor=1.4, max_out_buffer_size=2168, pingAsObs=0, train=1,
                          signaling_type="NN", agent_type="dqn_buffer", activateUnderlayTraffic=1), "random"),
    # NN-weight segment sources (big signalling) next to Philox traffic
    ("abilene", dict(simTime=2.5, load_factor=0.8, pingAsObs=0, activateUnderlayTraffic=1, train=1, signaling_type="NN",
                     signalingSim=1, sync_step=0.25, bigSignalingSize=4096, agent_type="dqn_buffer"), "random"),
]


@pytest.mark.parametrize("topo,over,policy", PHILOX_CASES)
def test_cuda_matches_oracle_synthetic_traffic(topo, over, policy):
    """Seeded Philox traffic: envs 0, 37 and 129 of a 130-env batch against the restatement run with the
    same (seed, env id): event logs, Gym steps (obs, reward, lost uids), counters, final queues."""
    cfg = common.env_config(topo, num_envs=130, **over)
    arrays = cfg.load()
    table = common.sp_table(arrays)
    env = make_env(cfg, arrays=arrays, env_id_base=1000)
    env.enable_log(400000, 120000)
    env.reset()
    env.set_policy(policy, table, seed=11)
    env.run()
    cs = env.counters()
    assert (cs["errors"] == 0).all()
    for e in (0, 37, 129):
        osim = common.oracle_run(cfg, arrays, oracle, policy=policy, table=table, seed=11 + e, env_id=1000 + e)
        oe = osim.events()
        assert len(oe) > 1000
        assert common.first_event_mismatch(oe, env.events(e)) is None
        common.assert_steps_equal(osim.decisions(), env.steps(e))
        common.assert_step_sources(oe, env.steps(e))
        common.assert_counters_equal(osim.counters(), cs[e])
        assert np.array_equal(osim.queue_bytes(), env.queue_bytes(e))
    # different env ids draw different traffic
    assert len({int(x) for x in cs["packet_hops"]}) > 10
    env.close()


def test_gym_mode_matches_oracle_and_run_mode():
    """prb_step: one Notify() per call for every env; actions from outside (sp table looked up on the
    host) must reproduce the fused run and the restatement's step sequence."""
    cfg = common.env_config("abilene", num_envs=5, simTime=1.0, load_factor=1.2, pingAsObs=0, activateUnderlayTraffic=1,
                            max_out_buffer_size=4000, train=1, signaling_type="NN", agent_type="dqn_buffer")
    arrays = cfg.load()
    table = common.sp_table(arrays)
    env = make_env(cfg, arrays=arrays)
    env.enable_log(100000, 0)
    env.reset()
    per_env = [[] for _ in range(5)]
    actions = np.zeros(5, dtype=np.int32)
    for _ in range(20000):
        recs = env.step_host(actions)
        if not recs["valid"].any():
            break
        for e in range(5):
            r = recs[e]
            if r["valid"]:
                actions[e] = table[r["overlay"], r["obs"][0]] if r["need_action"] else 0
                per_env[e].append((int(r["t_ns"]), int(r["overlay"]), int(r["uid"]), int(r["ptype"]), int(r["done"]),
                                   int(actions[e]) if r["need_action"] else -1, float(r["hop_time"]), int(r["n_lost"])))
    else:
        raise AssertionError("simulation did not end")
    cs = env.counters()
    assert (cs["errors"] == 0).all()
    for e in (0, 4):
        osim = common.oracle_run(cfg, arrays, oracle, policy="table", table=table, env_id=e)
        od = osim.decisions()
        assert len(od) == len(per_env[e]) > 500
        got = per_env[e]
        assert [x[0] for x in got] == list(od["t_ns"]) and [x[1] for x in got] == list(od["overlay"])
        assert [x[2] for x in got] == list(od["uid"]) and [x[5] for x in got] == list(od["action"])
        assert [x[6] for x in got] == list(od["hop_time"]) and [x[7] for x in got] == list(od["n_lost"])
        assert common.first_event_mismatch(osim.events(), env.events(e)) is None
        common.assert_counters_equal(osim.counters(), cs[e])
    env.close()


def test_gym_mode_device_tensors():
    """prb_step on device tensors: the agent (a table lookup written in torch) never leaves the GPU."""
    import torch
    cfg = common.env_config("abilene", num_envs=512, simTime=0.5, load_factor=1.0, pingAsObs=0, activateUnderlayTraffic=1)
    arrays = cfg.load()
    table = common.sp_table(arrays)
    env = make_env(cfg, arrays=arrays)
    tab = torch.tensor(table, dtype=torch.int32, device="cuda")
    acts = torch.zeros(512, dtype=torch.int32, device="cuda")
    n = 0
    while True:
        b = env.step(acts)
        if not bool(b.valid.any()):
            break
        acts = torch.where(b.need_action == 1, tab[b.overlay.long().clamp(min=0), b.obs[:, 0].long().clamp(0, 10)], 0).to(torch.int32).contiguous()
        n += 1
        assert n < 10000
    cs = env.counters()
    env2 = make_env(cfg, arrays=arrays)
    env2.set_policy("table", table)
    env2.run()
    cs2 = env2.counters()
    for k in common.COUNTER_INTS:
        assert np.array_equal(cs[k], cs2[k]), k
    assert np.array_equal(cs["sum_e2e"].view(np.uint32), cs2["sum_e2e"].view(np.uint32))
    assert int(cs["env_steps"].sum()) > 100000
    env.close()
    env2.close()


def test_chunked_run_and_reset_are_deterministic():
    cfg = common.env_config("abilene", num_envs=64, simTime=2, load_factor=0.9, pingAsObs=1, activateUnderlayTraffic=1)
    arrays = cfg.load()
    table = common.sp_table(arrays)
    env = make_env(cfg, arrays=arrays)
    env.set_policy("table", table)
    env.run()
    a = env.counters()
    env.reset()
    for t in (0.3, 0.30000001, 1.1, 1.999999999, 2.0, 3.0):
        env.run(until_s=t)
    b = env.counters()
    assert a.tobytes() == b.tobytes()
    env.close()


def test_full_batch_properties_config2():
    """BASELINE.json configs[1] at full size (Abilene, 4096 envs, load 0.9): conservation laws that hold
    for every env whatever the traffic, and agreement of a sample of envs with the restatement."""
    cfg = common.env_config("abilene", num_envs=4096, simTime=2, load_factor=0.9, pingAsObs=0, activateUnderlayTraffic=1)
    arrays = cfg.load()
    table = common.sp_table(arrays)
    env = make_env(cfg, arrays=arrays)
    env.set_policy("table", table)
    env.run()
    c = env.counters()
    assert (c["errors"] == 0).all()
    assert (c["end_ns"] == 2_000_000_000).all()
    buffered = c["overlay_injected"] - c["overlay_arrived"] - c["overlay_lost"]
    assert (buffered >= 0).all() and (buffered < 400).all()
    assert (c["n_e2e"] == c["overlay_arrived"]).all() and (c["n_cost"] == c["overlay_arrived"] + c["overlay_lost"]).all()
    assert (c["data_hops"] <= c["packet_hops"]).all() and (c["decisions"] <= c["env_steps"]).all()
    assert (c["env_steps"] == c["decisions"] + c["overlay_arrived"]).all()
    assert (c["vanished"] == 0).all()
    assert 1100 < c["overlay_injected"].mean() < 1230      # 860 pkt/s * 0.9 * (2 s - mean start offset 0.5 s)
    for e in (0, 1777, 4095):
        osim = common.oracle_run(cfg, arrays, oracle, policy="table", table=table, env_id=e)
        common.assert_counters_equal(osim.counters(), c[e])
        assert np.array_equal(osim.queue_bytes(), env.queue_bytes(e))
    env.close()


def test_drop_heavy_small_buffers():
    """Small max_out_buffer_size: drops are the common path; lost uids reach the right agents."""
    cfg = common.env_config("geant", num_envs=40, simTime=1.0, load_factor=1.5, max_out_buffer_size=1100, pingAsObs=1,
                            activateUnderlayTraffic=1, train=1, signaling_type="NN", agent_type="dqn_buffer")
    arrays = cfg.load()
    table = common.sp_table(arrays)
    env = make_env(cfg, arrays=arrays)
    env.enable_log(200000, 60000)
    env.reset()
    env.set_policy("random", seed=3)
    env.run()
    cs = env.counters()
    assert (cs["errors"] == 0).all() and (cs["overlay_lost"] > 50).all()
    for e in (0, 39):
        osim = common.oracle_run(cfg, arrays, oracle, policy="random", seed=3 + e, env_id=e)
        assert common.first_event_mismatch(osim.events(), env.events(e)) is None
        common.assert_steps_equal(osim.decisions(), env.steps(e))
        common.assert_counters_equal(osim.counters(), cs[e])
    env.close()


def test_sparse_timers_beyond_key_range():
    """32-bit scan keys saturate when every timer is > 4.29 s away: exact fallback on the device."""
    cfg = common.env_config("overlay_full_mesh_3n_abilene", num_envs=33, simTime=60, load_factor=0.004, pingAsObs=1,
                            activateUnderlayTraffic=1, pingPacketIntervalTime=7.0)
    arrays = cfg.load()
    table = common.sp_table(arrays)
    env = make_env(cfg, arrays=arrays)
    env.enable_log(5000, 2000)
    env.reset()
    env.set_policy("table", table)
    env.run()
    cs = env.counters()
    for e in (0, 32):
        osim = common.oracle_run(cfg, arrays, oracle, policy="table", table=table, env_id=e)
        assert common.first_event_mismatch(osim.events(), env.events(e)) is None
        common.assert_steps_equal(osim.decisions(), env.steps(e))
        common.assert_counters_equal(osim.counters(), cs[e])
    env.close()


def test_gym_mode_surfaces_big_signaling_steps():
    """prb_step with NN-weight segment sources: every segment that reaches its neighbour is a Gym step with obs
    [1000] whose record carries NN Index / segment Index / the signalling node, and the reference's info string
    holds them in the positions Forwarder.treat_info reads (forwarder.py:251-254)."""
    from prisma_b200.env import Ns3EnvGroup
    cfg = common.env_config("abilene", num_envs=4, simTime=0.6, load_factor=0.8, pingAsObs=0, activateUnderlayTraffic=1,
                            train=1, signaling_type="NN", signalingSim=1, sync_step=0.1, bigSignalingSize=2048,
                            agent_type="dqn_buffer")
    arrays = cfg.load()
    env = make_env(cfg, arrays=arrays)
    osim = common.oracle_run(cfg, arrays, oracle, policy="zero", env_id=2)
    od = osim.decisions()
    grp = Ns3EnvGroup(env, 2)
    acts = np.zeros(cfg.num_envs, dtype=np.int32)
    n = n_big = 0
    while True:
        rec = env.step_host(acts)[2]
        if not rec["valid"]:
            break
        o = od[n]
        assert rec["t_ns"] == o["t_ns"] and rec["ptype"] == o["ptype"] and rec["node"] == o["node"], (n, rec, o)
        if rec["ptype"] == 1:
            assert rec["uid"] == o["uid"] and rec["prev_node"] == o["prev_node"] and rec["obs"][0] == 1000
            tok = grp.info_string(rec, env.counters()[2]).split(",") if n_big < 5 else None
            if tok:
                assert int(tok[4].split("=")[-1]) == 1
                assert int(tok[18].split("=")[-1]) == rec["uid"] >> 32
                assert int(tok[19].split("=")[-1]) == rec["uid"] & 0xFFFFFFFF
                assert int(tok[20].split("=")[-1]) == arrays.map_overlay[rec["prev_node"]]
            n_big += 1
        n += 1
    assert n == len(od) and n_big == int((od["ptype"] == 1).sum()) and n_big > 100
    # segments per NN = 2048 / 512: the segment index wraps at 4 and the NN index advances
    big = od[od["ptype"] == 1]
    assert (big["uid"] & 0xFFFFFFFF).max() == 3 and (big["uid"] >> 32).max() >= 1
    common.assert_counters_equal(osim.counters(), env.counters()[2])
    env.close()


def test_compat_per_node_view():
    """Ns3EnvCompat: the reference's per-node reset/step surface over env 0, driven the way
    prisma/main.py:117-127 does — one thread per overlay node running a Forwarder-style loop."""
    import threading
    from prisma_b200.env import Ns3EnvGroup
    cfg = common.env_config("overlay_full_mesh_3n_abilene", num_envs=2, simTime=1.0, load_factor=2.0, pingAsObs=1,
                            activateUnderlayTraffic=1)
    arrays = cfg.load()
    table = common.sp_table(arrays)
    env = make_env(cfg, arrays=arrays)
    grp = Ns3EnvGroup(env, 0)
    assert [n.action_space.n for n in grp.nodes] == [len(x) for x in arrays.overlay_neighbors()]
    assert grp.nodes[0].observation_space.shape == (1 + len(arrays.overlay_neighbors()[0]),)
    counts = [0] * arrays.n_overlay
    errors = []

    def forwarder(m):
        try:
            node = grp.nodes[m]
            obs = node.reset(timeout=120)
            while obs[0] != -1:
                a = int(table[m, obs[0]]) if obs[0] not in (m, 1000) else 0     # forwarder.py:149-152
                obs, reward, done, info = node.step(a, timeout=120)
                counts[m] += 1
                if obs[0] != -1:
                    assert reward == 1.0 and info.startswith("End to End Delay=") and len(info.split(",")) == 22
        except Exception as ex:   # pragma: no cover
            errors.append(ex)
            with grp._cv:
                grp._ended = True
                grp._cv.notify_all()

    threads = [threading.Thread(target=forwarder, args=(m,)) for m in range(arrays.n_overlay)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(300)
    assert not errors, errors
    osim = common.oracle_run(cfg, arrays, oracle, policy="table", table=table, env_id=0)
    od = osim.decisions()
    assert sum(counts) == len(od)
    assert counts == [int((od["overlay"] == m).sum()) for m in range(arrays.n_overlay)]
    common.assert_counters_equal(osim.counters(), env.counters()[0])
    env.close()


def test_c1_full_length_episode_against_the_restatement():
    """BASELINE.json configs[0] at its full length: Abilene, sp agent, one 60-second episode.  The event-serial engine's
    complete ordered log (every injection, enqueue, drop, transmission and reception with nanosecond timestamps and
    queue bytes), every Gym step and all counters against the restatement; the node-parallel engine's counters and
    final queues on the same episode; and the conservation law of compute-stats-v2.cc (injected = arrived + lost +
    still buffered) on both."""
    cfg = common.env_config("abilene", num_envs=2, simTime=60, load_factor=1.0, pingAsObs=0, activateUnderlayTraffic=1)
    arrays = cfg.load()
    table = common.sp_table(arrays)
    osim = common.oracle_run(cfg, arrays, oracle, policy="table", table=table, env_id=1)
    oe, oc = osim.events(), osim.counters()
    assert len(oe) > 400_000 and oc["overlay_injected"] > 40_000
    env = make_env(cfg, arrays=arrays)
    env.enable_log(len(oe) * 5 // 4, len(osim.decisions()) * 5 // 4)      # env 0 draws other traffic
    env.reset()
    env.set_policy("table", table)
    env.run()
    cs = env.counters()
    env.check_errors(cs)
    assert env.engine == "serial"
    assert common.first_event_mismatch(oe, env.events(1)) is None
    common.assert_steps_equal(osim.decisions(), env.steps(1))
    common.assert_counters_equal(oc, cs[1])
    env.close()
    par = make_env(cfg, arrays=arrays)
    par.set_engine("parallel")
    par.set_policy("table", table)
    par.run()
    cp = par.counters()
    par.check_errors(cp)
    assert par.engine == "parallel"
    common.assert_counters_equal(oc, cp[1])
    assert np.array_equal(osim.queue_bytes(), par.queue_bytes(1))
    for c in (cs[1], cp[1]):
        buffered = int(c["overlay_injected"] - c["overlay_arrived"] - c["overlay_lost"])
        assert 0 <= buffered < 400
    par.close()
