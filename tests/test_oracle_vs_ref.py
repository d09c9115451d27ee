"""Pins the CPU restatement (oracle/prisma_oracle.c) to the reference's own ns-3 program
(oracle/_ref/prisma_ref): same seeded injections, same policy -> identical event logs, decisions,
counters and float32 statistics.  Needs /root/reference-built oracle/_ref, i.e. runs in the build
container; the same comparison against committed outputs lives in test_golden.py."""
import os
import struct

import numpy as np
import pytest

from oracle import oracle, run_ref
from tests import common

pytestmark = pytest.mark.skipif(not run_ref.available(), reason="oracle/_ref/prisma_ref not built")

CASES = {
    # name: (topology, cfg overrides, policy, ref kwargs)
    "c1_abilene_sp_60s": ("abilene", dict(simTime=60, load_factor=1.0, pingAsObs=0, activateUnderlayTraffic=1), "table"),
    "abilene_random_drop_heavy": ("abilene", dict(simTime=10, load_factor=1.4, max_out_buffer_size=2168, pingAsObs=0,
                                                  activateUnderlayTraffic=1), "random"),
    "overlay_ping_obs_underlay": ("overlay_full_mesh_3n_abilene", dict(simTime=20, load_factor=1.5, pingAsObs=1,
                                                                      activateUnderlayTraffic=1), "random"),
    "overlay_no_underlay_sp": ("overlay_full_mesh_3n_abilene", dict(simTime=10, load_factor=2.0, pingAsObs=1,
                                                                    activateUnderlayTraffic=0), "table"),
    "geant_sp": ("geant", dict(simTime=5, load_factor=1.2, pingAsObs=0, activateUnderlayTraffic=1), "table"),
    "abilene_train_small_signaling": ("abilene", dict(simTime=5, load_factor=1.2, pingAsObs=0, train=1,
                                                      signaling_type="NN", agent_type="dqn_buffer",
                                                      activateUnderlayTraffic=1), "random"),
    "abilene_train_nn_big_signaling": ("abilene", dict(simTime=4, load_factor=1.1, pingAsObs=0, train=1,
                                                       signaling_type="NN", signalingSim=1, sync_step=1.0,
                                                       bigSignalingSize=35328, agent_type="dqn_buffer",
                                                       activateUnderlayTraffic=1), "random"),
    # BASELINE.json configs[4] (random underlay, 4-packet buffers = drop-heavy, control packets) at the largest size the
    # reference can run: it numbers all switch devices out of one /24 (sim.cc:446-447), so 100 nodes abort in ns-3
    "c5_synthetic56_drop_heavy_signaling": ("synthetic56", dict(simTime=1.5, load_factor=1.3, max_out_buffer_size=2168,
                                                                  pingAsObs=0, train=1, signaling_type="NN",
                                                                  agent_type="dqn_buffer", activateUnderlayTraffic=1),
                                             "random"),
    "overlay_train_target": ("overlay_full_mesh_3n_abilene", dict(simTime=10, load_factor=3.0, pingAsObs=1, train=1,
                                                                  signaling_type="target", agent_type="dqn_buffer",
                                                                  activateUnderlayTraffic=1,
                                                                  max_out_buffer_size=4000), "random"),
}


def run_case(name, tmp):
    topo, over, policy = CASES[name]
    cfg = common.env_config(topo, **over)
    arrays = cfg.load()
    table = common.sp_table(arrays)
    files = common.ref_files(cfg, tmp)
    lp = cfg.derived_loss_penalty(arrays.n_overlay)
    out = os.path.join(tmp, "ref_out")
    run_ref.run(files, out, sim_time=cfg.simTime, load_factor=cfg.load_factor, link_cap=cfg.link_cap,
                link_delay_ms=cfg.link_delay, packet_size=cfg.packet_size,
                max_out_buffer_size=cfg.max_out_buffer_size, agent_type=cfg.agent_type,
                signaling_type=cfg.signaling_type, train=bool(cfg.train), loss_penalty=lp,
                moving_average=cfg.movingAverageObsSize, activate_underlay_traffic=bool(cfg.activateUnderlayTraffic),
                ping_as_obs=bool(cfg.pingAsObs), ping_interval=cfg.pingPacketIntervalTime,
                policy=policy, policy_data=table, policy_seed=7, seed=cfg.seed,
                signaling=cfg.effective_signaling(), sync_step=cfg.sync_step, big_signaling_size=cfg.bigSignalingSize,
                log_info=True)
    summ = run_ref.parse_summary(out)
    fr, to = oracle.build_links(arrays.phys_adj)
    _, route = run_ref.link_maps(summ, fr, to)
    ev = run_ref.parse_events(out, summ, fr, to)
    dec = [d for d in run_ref.parse_decisions(out, with_info=True) if d["ptype"] >= 0]
    sim = oracle.OracleSim(common.oracle_params(cfg, arrays, oracle))
    inj = ev[ev["kind"] == 0]
    sim.set_inject_trace(inj["t_ns"], inj["src"], inj["dst"], inj["uid"], inj["valable"], inj["rejected"], inj["type"])
    sim.set_policy(policy, table, seed=7)
    sim.run()
    return cfg, arrays, summ, route, ev, dec, sim


@pytest.mark.parametrize("name", list(CASES))
def test_restatement_matches_reference(name, tmp_path):
    cfg, arrays, summ, route, ev, dec, sim = run_case(name, str(tmp_path))
    # routing tables: po_global_routes vs Ipv4GlobalRouting as dumped by the reference
    mine = oracle.global_routes(arrays.phys_adj)
    assert np.array_equal(mine, route)
    # every INJ/ARR/ENQ/DRP/TXB/RCV record, in order
    oe = sim.events()
    bad = common.first_event_mismatch(ev, oe)
    assert bad is None, (bad, ev[max(0, bad - 2):bad + 2], oe[max(0, bad - 2):bad + 2])
    assert len(ev) == len(oe) and len(ev) > 1000
    # every Gym step
    od = sim.decisions()
    assert len(od) == len(dec)
    for i, (r, o) in enumerate(zip(dec, od)):
        assert r["t_ns"] == o["t_ns"] and r["overlay"] == o["overlay"] and r["node"] == o["node"], i
        assert r["ptype"] == o["ptype"] and r["done"] == o["done"] and r["action"] == o["action"], (i, r, o)
        assert r["obs"] == list(o["obs"][:o["nobs"]]), (i, r, o)
        if r["ptype"] == 0:
            assert r["uid"] == o["uid"]
        if r["ptype"] == 1:
            # info fields 18-20 of the reference against the record fields that carry them
            f = dict(t.strip().split("=") for t in r["info"].split(",") if "=" in t)
            assert int(f["NN Index"]) == o["uid"] >> 32 and int(f["segment Index"]) == o["uid"] & 0xFFFFFFFF, (i, r, o)
            assert int(f["NodeId Signaled"]) == arrays.map_overlay[o["prev_node"]], (i, r, o)
    # counters and the float32 statistics, bit for bit
    c = sim.counters()
    for k in ("overlay_injected", "overlay_arrived", "overlay_lost", "underlay_injected", "underlay_arrived",
              "underlay_lost", "env_steps", "decisions", "n_e2e", "n_cost"):
        assert summ[k] == c[k], k
    bits = lambda f: struct.unpack("I", struct.pack("f", f))[0]
    assert summ["avg_e2e_delay_bits"] == bits(c["avg_e2e"])
    assert summ["avg_cost_bits"] == bits(c["avg_cost"])
    assert summ["avg_underlay_e2e_delay_bits"] == bits(c["avg_underlay_e2e"])
    assert summ["avg_underlay_cost_bits"] == bits(c["avg_underlay_cost"])
    assert summ["signaling_overhead_bits"] == bits(c["signaling_overhead"])


def test_time_arithmetic_kats():
    """ns-3 Time / DataRate known answers, asked of the reference binary itself."""
    rng = np.random.default_rng(5)
    q = []
    sizes = [38, 46, 54, 62, 110, 200, 230, 542, 1000, 1500]
    rates = [500000, 1000000, 155520000, 5 * 10 ** 11, 15 * 10 ** 11, 123457]
    for r in rates:
        for s in sizes:
            q.append("txbps %d %d" % (r, s))
    ds = list(rng.random(200)) + list(rng.random(50) * 1e-3) + [0.001, 0.2, float(np.float32(0.2)), 60.0, 0.0001, 1.0001]
    for d in ds:
        q.append("s2ns %016x" % struct.unpack("Q", struct.pack("d", float(d)))[0])
    ns = [int(x) for x in rng.integers(0, 60 * 10 ** 9, 300)] + [0, 1, 999999999, 10 ** 9, 60 * 10 ** 9]
    for v in ns:
        q.append("ns2s %d" % v)
    ans = run_ref.kat(q)
    assert len(ans) == len(q)
    for a in ans:
        if a[0] == "txbps":
            assert oracle.tx_time_ns(int(a[2]), int(a[1])) == int(a[3]), a
        elif a[0] == "s2ns":
            d = struct.unpack("d", struct.pack("Q", int(a[1], 16)))[0]
            assert oracle.seconds_to_ns(d) == int(a[2]), a
        elif a[0] == "ns2s":
            got = struct.unpack("Q", struct.pack("d", oracle.ns_to_seconds(int(a[1]))))[0]
            assert got == int(a[2], 16), a


def _skeleton(info):
    """An info string with every value removed: the keys, the separators and the irregular blanks of the reference."""
    import re
    return re.sub(r"=[^,]*", "=", info)


def test_info_string_is_byte_compatible_with_the_reference(tmp_path):
    """format_info (prisma_b200/env.py) against the strings the reference's getInfo() produced in the same run
    (packet-manager.cc:119-176, data-packet-manager.cc:229-248, small-/big-signaling-packet-manager.cc): keys,
    separators and blanks byte for byte for every packet type; for data packets also every value that belongs to the
    step (fields 0-4 and the manager's own fields).  The running averages (fields 5-17) are compared by the counter
    tests; fields 0, 1 and 3 of signalling records hold an uninitialised member / ns-3's global Packet uid."""
    from prisma_b200.env import format_info
    cfg, arrays, summ, route, ev, dec, sim = run_case("abilene_train_nn_big_signaling", str(tmp_path))
    od = sim.decisions()
    zero = {k: 0 for k in ("overlay_injected", "overlay_arrived", "overlay_lost", "underlay_injected", "underlay_arrived",
                           "underlay_lost", "avg_e2e", "avg_cost", "avg_underlay_e2e", "avg_underlay_cost",
                           "signaling_overhead")}
    seen = set()
    n_data = 0
    inj = ev[(ev["kind"] == 0) & (ev["type"] == 0)]
    src_of = {int(u): int(s) for u, s in zip(inj["uid"], inj["src"])}      # the restatement's record has no source field
    for r, o in zip(dec, od):
        o = {k: o[k] for k in o.dtype.names}
        o["src"] = src_of.get(int(o["uid"]), 0)
        mine = format_info(o, zero, arrays.map_overlay, cfg.packet_size)
        assert _skeleton(mine) == _skeleton(r["info"]), (mine, r["info"])
        seen.add(int(o["ptype"]))
        if o["ptype"] == 0:
            a, b = mine.split(","), r["info"].split(",")
            assert a[:5] == b[:5] and a[18:] == b[18:], (mine, r["info"])
            n_data += 1
        elif o["ptype"] == 1:
            assert mine.split(",")[18:] == r["info"].split(",")[18:] and mine.split(",")[4] == r["info"].split(",")[4]
    assert seen == {0, 1, 2} and n_data > 500


def test_compat_view_has_the_bridge_forwarder_closes():
    """Forwarder.run ends with self.env.ns3ZmqBridge.send_close_command() (forwarder.py:331)."""
    from prisma_b200.env import _NoBridge
    assert _NoBridge().send_close_command() is True
