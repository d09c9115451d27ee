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
             loss_penalty=cfg.derived_loss_penalty(m), seed=cfg.seed)
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
