"""Shared helpers of the test-suite: fixture topologies, the sp table, oracle parameter dicts."""
import os

import numpy as np

from prisma_b200.config import EnvConfig

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FIXTURES = os.path.join(ROOT, "tests", "fixtures")   # copies of the reference's example inputs (data, P17)
REF_EXAMPLES = "/root/reference/prisma/examples"

TOPOLOGIES = ["abilene", "geant", "overlay_full_mesh_3n_abilene"]


def topo_dir(name):
    d = os.path.join(FIXTURES, name)
    return d


def env_config(name="abilene", **kw):
    d = topo_dir(name)
    return EnvConfig(physical_adjacency_matrix_path=os.path.join(d, "topology_files", "physical_adjacency_matrix.txt"),
                     overlay_adjacency_matrix_path=os.path.join(d, "topology_files", "overlay_adjacency_matrix.txt"),
                     map_overlay_path=os.path.join(d, "topology_files", "map_overlay.txt"),
                     traffic_matrix_root_path=os.path.join(d, "traffic_matrices"), **kw)


def ref_files(cfg, workdir):
    """Input file paths for oracle/run_ref.run."""
    n = cfg.load().n_nodes
    rej = os.path.join(workdir, "rejected.txt")
    os.makedirs(workdir, exist_ok=True)
    np.savetxt(rej, np.zeros((n, n)), fmt="%.6f")
    return {"adj": cfg.physical_adjacency_matrix_path, "overlay": cfg.overlay_adjacency_matrix_path,
            "map": cfg.map_overlay_path,
            "tm": os.path.join(cfg.traffic_matrix_root_path, f"node_intensity_normalized_{cfg.traffic_matrix_index}.txt"),
            "rej": rej}


def sp_table(arrays):
    """The sp agent as a table: neighbors.index(nx.shortest_path(G, me, dst)[1]) (forwarder.py:190-191),
    G = DiGraph of the OVERLAY adjacency (argument_parser.py:127-130)."""
    import networkx as nx
    g = nx.DiGraph(nx.from_numpy_array(arrays.overlay_adj.astype(int)))
    nbrs = arrays.overlay_neighbors()
    o2u = arrays.overlay_to_underlay()
    m = arrays.n_overlay
    t = np.zeros((m, m), dtype=np.int32)
    for a in range(m):
        for d in range(m):
            if a != d:
                t[a, d] = nbrs[a].index(int(o2u[nx.shortest_path(g, a, d)[1]]))
    return t


def oracle_params(cfg, arrays, oracle_mod, **over):
    m = arrays.n_overlay
    p = dict(n_nodes=arrays.n_nodes, n_overlay=m, phys_adj=arrays.phys_adj, overlay_adj=arrays.overlay_adj,
             map_overlay=arrays.map_overlay, traffic_bps=arrays.traffic_bps,
             link_rate_bps=cfg.link_cap, link_delay_ns=oracle_mod.seconds_to_ns(cfg.link_delay * 0.001),
             packet_size=cfg.packet_size, max_out_buffer_size=cfg.max_out_buffer_size,
             sim_time_ns=oracle_mod.seconds_to_ns(float(cfg.simTime)),
             ping_interval_ns=oracle_mod.seconds_to_ns(float(np.float32(cfg.pingPacketIntervalTime))),
             moving_average_size=cfg.movingAverageObsSize, ping_as_obs=cfg.pingAsObs,
             activate_underlay_traffic=cfg.activateUnderlayTraffic, train=cfg.train,
             signaling_type=cfg.signaling_type, opt_agent=1 if cfg.agent_type == "opt" else 0,
             loss_penalty=cfg.derived_loss_penalty(m), seed=cfg.seed,
             signaling=int(cfg.effective_signaling()), big_signaling_size=cfg.bigSignalingSize,
             sync_step=cfg.sync_step)
    p.update(over)
    return p


EVENT_FIELDS = ["t_ns", "kind", "node", "link", "uid", "size", "type", "src", "dst", "last_hop", "next_hop", "qbytes"]


def first_event_mismatch(a, b, fields=EVENT_FIELDS):
    n = min(len(a), len(b))
    for f in fields:
        d = np.nonzero(a[f][:n] != b[f][:n])[0]
        if len(d):
            n = min(n, int(d[0]) + 1)
            first = int(d[0])
        else:
            first = None
        if first is not None:
            best = first if "best" not in locals() else min(best, first)
    if "best" in locals():
        return best
    if len(a) != len(b):
        return min(len(a), len(b))
    return None


# ---- comparisons between the oracle (po_*) and the product / host-harness records (prb_*) ----

STEP_FIELDS = ["t_ns", "node", "overlay", "uid", "ptype", "done", "action", "nobs", "is_new", "prev_node", "n_lost"]


def assert_steps_equal(oracle_dec, steps, check_action=True):
    """po_decision[] (oracle) vs prb_step_rec[] (CUDA path / host harness): every Gym step, bit for bit."""
    assert len(oracle_dec) == len(steps), (len(oracle_dec), len(steps))
    for f in STEP_FIELDS:
        if f == "action" and not check_action:
            continue
        a, b = oracle_dec[f], steps[f]
        if f == "uid":
            keep = oracle_dec["ptype"] != 1          # big signalling carries no uid
            a, b = a[keep], b[keep]
        bad = np.nonzero(np.asarray(a) != np.asarray(b))[0]
        assert len(bad) == 0, (f, int(bad[0]), oracle_dec[bad[0]], steps[bad[0]])
    k = steps["obs"].shape[1]
    mask = np.arange(k)[None, :] < steps["nobs"][:, None]
    assert np.array_equal(np.where(mask, oracle_dec["obs"][:, :k], 0), np.where(mask, steps["obs"], 0))
    # reward (hop time): exact ns difference scaled once -> identical doubles; e2e delay: identical float32
    assert np.array_equal(oracle_dec["hop_time"].view(np.uint64), steps["hop_time"].view(np.uint64))
    assert np.array_equal(oracle_dec["e2e_delay"].view(np.uint32), steps["e2e_delay"].view(np.uint32))
    nl = np.minimum(steps["n_lost"], 8)
    lm = np.arange(8)[None, :] < nl[:, None]
    assert np.array_equal(np.where(lm, oracle_dec["lost_uid"], 0), np.where(lm, steps["lost_uid"], 0))


COUNTER_INTS = ["overlay_injected", "overlay_arrived", "overlay_lost", "underlay_injected", "underlay_arrived",
                "underlay_lost", "bytes_data", "bytes_signaling", "n_e2e", "n_cost", "n_underlay_e2e",
                "n_underlay_cost", "packet_hops", "data_hops", "env_steps", "decisions", "events", "vanished", "end_ns"]
COUNTER_FLOATS = ["sum_e2e", "sum_cost", "sum_underlay_e2e", "sum_underlay_cost", "avg_e2e", "avg_cost",
                  "avg_underlay_e2e", "avg_underlay_cost", "signaling_overhead"]


def assert_counters_equal(oc, pc):
    """oracle counters dict vs product counters (dict or numpy record): integers and float32 bits."""
    import struct
    for k in COUNTER_INTS:
        assert int(oc[k]) == int(pc[k]), (k, oc[k], pc[k])
    for k in COUNTER_FLOATS:
        a = struct.unpack("I", struct.pack("f", float(oc[k])))[0]
        b = struct.unpack("I", struct.pack("f", float(pc[k])))[0]
        assert a == b, (k, oc[k], pc[k])
    assert int(pc["errors"]) == 0


def assert_big_signaling_fields(big, uid, prev_node, map_overlay):
    """Golden (NN Index, segment Index, NodeId Signaled) of every big-signalling step against the record fields
    that carry them: uid = NNIndex << 32 | segIndex, prev_node = underlay id of the signalling node."""
    sel = big[:, 0] >= 0
    assert sel.any()
    assert np.array_equal(np.asarray(uid)[sel] >> 32, big[sel, 0])
    assert np.array_equal(np.asarray(uid)[sel] & 0xFFFFFFFF, big[sel, 1])
    assert np.array_equal(np.asarray(map_overlay)[np.asarray(prev_node)[sel]], big[sel, 2])


def assert_step_sources(events, steps):
    """prb_step_rec.src of every data step = the source its packet was injected with (INJ record of that uid)."""
    inj = events[(events["kind"] == 0) & (events["type"] == 0)]
    src_of = dict(zip(inj["uid"].tolist(), inj["src"].tolist()))
    data = steps[steps["ptype"] == 0]
    assert len(data) > 0
    assert all(src_of[int(u)] == int(s) for u, s in zip(data["uid"], data["src"]))


def replay_inputs(g):
    """(injection arrays, policy kind, policy data, seed) of a golden fixture."""
    inj = g.injections
    return ((inj["t_ns"], inj["src"], inj["dst"], inj["uid"], inj["valable"], inj["rejected"], inj["type"]),
            g.policy, g.table, g.policy_seed)


def oracle_run(cfg, arrays, oracle_mod, policy="table", table=None, seed=7, env_id=0, until_ns=None, inject=None, **over):
    """Run the CPU restatement on one environment (Philox traffic unless `inject` is given)."""
    p = oracle_params(cfg, arrays, oracle_mod, env_id=env_id, inject_mode=1, **over)
    sim = oracle_mod.OracleSim(p)
    if inject is not None:
        sim.set_inject_trace(*inject)
    sim.set_policy(policy, table, seed=seed)
    sim.run(until_ns)
    return sim
