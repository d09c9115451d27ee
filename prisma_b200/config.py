"""Host-side configuration of the batched routing environment.

Mirrors the parameter surface of the reference (names kept so that existing launch
scripts map one to one):

* flags and defaults: prisma/source/argument_parser.py:33-98
* derived values (loss_penalty, traffic_matrix_path): prisma/source/argument_parser.py:101-163
* how the flags reach ns-3: prisma/source/run_ns3.py:41-73
* file formats and topology build: prisma/ns3/sim.cc:274-514,726-1012

The arrays produced here are what the C ABI (`include/prisma_b200.h: prb_config`) takes.
"""
from __future__ import annotations

import dataclasses
import math
import os
from typing import Optional

import numpy as np

AGENT_TYPES = ["dqn_buffer", "dqn_routing", "dqn_buffer_fp", "dqn_buffer_lite", "dqn_buffer_lighter",
               "dqn_buffer_lighter_2", "dqn_buffer_lighter_3", "dqn_buffer_ff",
               "dqn_buffer_with_throughputs", "sp", "opt"]
SIGNALING_TYPES = {"ideal": 0, "NN": 1, "target": 2, "digital_twin": 3}

_RATE_UNITS = [  # ns3-gym/src/network/utils/data-rate.cc:33-200 (DataRate::DoParse)
    ("bps", 1.0), ("b/s", 1.0), ("Bps", 8.0), ("B/s", 8.0),
    ("kbps", 1e3), ("kb/s", 1e3), ("Kbps", 1e3), ("Kb/s", 1e3),
    ("kBps", 8e3), ("kB/s", 8e3), ("KBps", 8e3), ("KB/s", 8e3),
    ("Kib/s", 1024.0), ("KiB/s", 8192.0),
    ("Mbps", 1e6), ("Mb/s", 1e6), ("MBps", 8e6), ("MB/s", 8e6),
    ("Mib/s", 1048576.0), ("MiB/s", 8388608.0),
    ("Gbps", 1e9), ("Gb/s", 1e9), ("GBps", 8e9), ("GB/s", 8e9),
    ("Gib/s", 1073741824.0), ("GiB/s", 8589934592.0),
]


def parse_data_rate(s: str) -> int:
    """ns-3 DataRate string -> bit/s, truncating like `(uint64_t)(r * unit)` does."""
    s = s.strip()
    n = 0
    while n < len(s) and (s[n].isdigit() or s[n] == "."):
        n += 1
    if n == len(s):
        return int(float(s))
    r = float(s[:n])
    unit = s[n:]
    for u, mul in _RATE_UNITS:
        if unit == u:
            return int(r * mul)
    raise ValueError(f"cannot parse data rate {s!r}")


def read_matrix(path: str) -> np.ndarray:
    """Square 0/1 matrix, whitespace separated, blank line ends it (sim.cc:726-790)."""
    rows = []
    with open(path) as f:
        for line in f:
            if line.strip() == "":
                break
            rows.append([int(float(x)) for x in line.split()])
    a = np.array(rows, dtype=np.uint8)
    if a.ndim != 2 or a.shape[0] != a.shape[1]:
        raise ValueError(f"{path}: not a square matrix: {a.shape}")
    return a


def read_map_overlay(path: str) -> np.ndarray:
    """One integer per line: overlay index of underlay node i or -1 (sim.cc:281-282,792-830)."""
    vals = []
    with open(path) as f:
        for line in f:
            if line.strip() == "":
                break
            vals.append(int(line.split()[0]))
    return np.array(vals, dtype=np.int32)


def read_traffic_matrix(path: str) -> np.ndarray:
    """N x N rate strings such as `46.83Kbps` (sim.cc:952-1012) -> bit/s (uint64)."""
    rows = []
    with open(path) as f:
        for line in f:
            if line.strip() == "":
                break
            rows.append([parse_data_rate(x) for x in line.split()])
    return np.array(rows, dtype=np.uint64)


def scale_traffic(tm_bps: np.ndarray, load_factor: float) -> np.ndarray:
    """`DataRate(ceil(rate * load_factor))` per flow (sim.cc:623); 0 stays 0 (sim.cc:604)."""
    out = np.zeros_like(tm_bps, dtype=np.uint64)
    it = np.nditer(tm_bps, flags=["multi_index"])
    for v in it:
        v = int(v)
        if v > 0:
            out[it.multi_index] = int(math.ceil(float(v) * float(load_factor)))
    return out


@dataclasses.dataclass
class EnvConfig:
    """Parameters of the reference that reach the simulator, plus the batch size.

    Field names follow prisma/source/argument_parser.py; units follow the flags
    (link_delay in ms, link_cap in bit/s, sizes in bytes, simTime in s).
    """
    # topology / traffic files (argument_parser.py:52-58)
    physical_adjacency_matrix_path: Optional[str] = None
    overlay_adjacency_matrix_path: Optional[str] = None
    map_overlay_path: Optional[str] = None
    traffic_matrix_root_path: Optional[str] = None
    traffic_matrix_index: int = 0
    # or in-memory arrays (used by synthetic topologies)
    physical_adjacency_matrix: Optional[np.ndarray] = None
    overlay_adjacency_matrix: Optional[np.ndarray] = None
    map_overlay: Optional[np.ndarray] = None
    traffic_matrix_bps: Optional[np.ndarray] = None
    opt_rejected: Optional[np.ndarray] = None
    # network parameters
    load_factor: float = 1.0
    link_cap: int = 500000
    link_delay: int = 1
    packet_size: int = 512
    max_out_buffer_size: int = 16260
    # simulation
    simTime: float = 60
    seed: int = 100
    train: int = 0
    agent_type: str = "sp"
    signaling_type: str = "ideal"
    signalingSim: int = 0
    sync_step: float = 1.0
    bigSignalingSize: int = 35328
    movingAverageObsSize: int = 5
    activateUnderlayTraffic: int = 0
    pingAsObs: int = 1
    pingPacketIntervalTime: float = 0.2
    loss_penalty: Optional[float] = None
    # batch
    num_envs: int = 1
    device: int = 0

    @classmethod
    def from_params(cls, params: dict, **over) -> "EnvConfig":
        """Builds the configuration from the reference's `params` dict (the output of
        prisma/source/argument_parser.py:11-164): same keys, unknown keys ignored."""
        names = {f.name for f in dataclasses.fields(cls)}
        kw = {k: v for k, v in params.items() if k in names and v is not None}
        kw.update(over)
        return cls(**kw)

    def shard(self, rank: int, world: int) -> "tuple[int, int]":
        """(first env id, env count) of rank `rank` when `num_envs` environments are sharded over `world`
        ranks: contiguous ranges, remainder to the low ranks.  Environments are independent, so this is the
        whole multi-GPU story of the step loop (no collective)."""
        base, rem = divmod(int(self.num_envs), int(world))
        n = base + (1 if rank < rem else 0)
        first = rank * base + min(rank, rem)
        return first, n

    # ---- derived ----
    def load(self) -> "EnvArrays":
        adj = self.physical_adjacency_matrix
        if adj is None:
            adj = read_matrix(self.physical_adjacency_matrix_path)
        adj = np.ascontiguousarray(adj, dtype=np.uint8)
        n = adj.shape[0]
        ov = self.overlay_adjacency_matrix
        if ov is None:
            ov = read_matrix(self.overlay_adjacency_matrix_path) if self.overlay_adjacency_matrix_path else adj
        ov = np.ascontiguousarray(ov, dtype=np.uint8)
        mp = self.map_overlay
        if mp is None:
            mp = read_map_overlay(self.map_overlay_path) if self.map_overlay_path else np.arange(n)
        mp = np.ascontiguousarray(mp, dtype=np.int32)
        tm = self.traffic_matrix_bps
        if tm is None:
            # argument_parser.py:119: <root>/node_intensity_normalized_<index>.txt
            p = os.path.join(self.traffic_matrix_root_path, f"node_intensity_normalized_{self.traffic_matrix_index}.txt")
            tm = read_traffic_matrix(p)
        tm = np.ascontiguousarray(tm, dtype=np.uint64)
        if tm.shape != (n, n):
            raise ValueError("traffic matrix and adjacency matrix sizes differ (sim.cc:310-313)")
        if mp.shape[0] != n:
            raise ValueError("map_overlay must have one entry per underlay node")
        m = ov.shape[0]
        if sorted(int(x) for x in mp if x >= 0) != list(range(m)):
            raise ValueError("map_overlay must name every overlay index exactly once")
        rej = self.opt_rejected
        if rej is None:
            rej = np.zeros((n, n))   # argument_parser.py:155-156
        return EnvArrays(n_nodes=n, n_overlay=m, phys_adj=adj, overlay_adj=ov, map_overlay=mp,
                         traffic_bps=scale_traffic(tm, self.load_factor),
                         opt_rejected=np.ascontiguousarray(rej, dtype=np.float64))

    @property
    def num_overlay_nodes(self) -> int:
        return self.load().n_overlay

    def derived_loss_penalty(self, num_nodes: int) -> float:
        """argument_parser.py:163 (numNodes = number of OVERLAY nodes, :127-131)."""
        if self.loss_penalty is not None:
            return float(self.loss_penalty)
        return (((self.max_out_buffer_size + 512 + 30) * 8) / self.link_cap + 0.001) * num_nodes

    def abi_params(self, arrays: "Optional[EnvArrays]" = None, **over) -> dict:
        """The prb_config fields (include/prisma_b200.h) as a dict: integer-ns times through the library's
        ns-3 `Seconds()` arithmetic, the float32 `pingPacketIntervalTime` of sim.cc:201, the derived loss
        penalty of argument_parser.py:163."""
        from . import _abi
        L = _abi.lib()
        a = arrays if arrays is not None else self.load()
        p = dict(num_envs=self.num_envs, n_nodes=a.n_nodes, n_overlay=a.n_overlay, phys_adj=a.phys_adj,
                 overlay_adj=a.overlay_adj, map_overlay=a.map_overlay, traffic_bps=a.traffic_bps,
                 opt_rejected=a.opt_rejected if self.agent_type == "opt" else None,
                 link_cap_bps=int(self.link_cap), link_delay_ns=L.prb_seconds_to_ns(float(self.link_delay) * 0.001),
                 packet_size=int(self.packet_size), max_out_buffer_size=int(self.max_out_buffer_size),
                 sim_time_ns=L.prb_seconds_to_ns(float(self.simTime)),
                 ping_interval_ns=L.prb_seconds_to_ns(float(np.float32(self.pingPacketIntervalTime))),
                 moving_average_size=int(self.movingAverageObsSize), ping_as_obs=int(self.pingAsObs),
                 activate_underlay_traffic=int(self.activateUnderlayTraffic), train=int(self.train),
                 signaling_type=self.signaling_type, opt_agent=1 if self.agent_type == "opt" else 0,
                 loss_penalty=self.derived_loss_penalty(a.n_overlay), seed=int(self.seed), env_id_base=0,
                 signaling=int(self.effective_signaling()), big_signaling_size=int(self.bigSignalingSize),
                 sync_step=float(self.sync_step))
        p.update(over)
        return p

    def effective_signaling(self) -> bool:
        """sim.cc:373-375: sp / opt agents and ideal signalling switch the control plane off."""
        return bool(self.signalingSim) and self.agent_type not in ("sp", "opt") and self.signaling_type != "ideal"


@dataclasses.dataclass
class EnvArrays:
    n_nodes: int
    n_overlay: int
    phys_adj: np.ndarray
    overlay_adj: np.ndarray
    map_overlay: np.ndarray
    traffic_bps: np.ndarray
    opt_rejected: np.ndarray

    def overlay_to_underlay(self) -> np.ndarray:
        o2u = np.zeros(self.n_overlay, dtype=np.int32)
        for i, m in enumerate(self.map_overlay):
            if m >= 0:
                o2u[m] = i
        return o2u

    def overlay_neighbors(self):
        """Action order of every overlay node, as underlay ids (sim.cc:469-476)."""
        o2u = self.overlay_to_underlay()
        nbrs = [[] for _ in range(self.n_overlay)]
        for i in range(self.n_overlay):
            for j in range(i, self.n_overlay):
                if self.overlay_adj[i, j] == 1:
                    nbrs[i].append(int(o2u[j]))
                    nbrs[j].append(int(o2u[i]))
        return nbrs
