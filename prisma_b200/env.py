"""Batched, device-resident packet-routing environment — the Python face of libprisma_b200.so.

`BatchedRoutingEnv` holds E independent copies of the network that PRISMA simulates with one ns-3 process
(prisma/ns3/sim.cc) and steps them with the sm_100a kernels of csrc/.  Two ways to drive it:

* `run(until)` — fused-agent mode.  The agent hook (Forwarder.take_action, prisma/source/forwarder.py:171-195)
  is a table evaluated inside the kernel (`sp`, `opt`-like tables, frozen `dq_routing` policies whose
  input is the destination only), so no host round trip happens per packet hop.
* `step(actions)` — Gym mode, the reference's contract (Ns3Env.step, prisma/ns3_model/ns3env.py:417-420):
  every call advances EVERY environment to its next valid Notify() (one packet arrival at one overlay
  node) and returns the record of that step; the actions answer the records of the previous call.

`Ns3EnvCompat` is the per-node view with the reference's `reset/step/close/action_space/
observation_space/connected` surface for one environment of the batch.

All computation happens in CUDA kernels; PyTorch is used for device buffers and streams only.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np

from . import _abi
from .config import EnvArrays, EnvConfig


def sp_table(arrays: EnvArrays) -> np.ndarray:
    """The `sp` agent as a table [M, M]: row = stepping overlay node, column = destination overlay index,
    value = neighbors.index(nx.shortest_path(G, me, dst)[1]) with G the DiGraph of the OVERLAY adjacency
    (prisma/source/forwarder.py:190-191, prisma/source/argument_parser.py:127-130)."""
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


class StepBatch:
    """Device-resident records of one `step` call: torch views over E prb_step_rec structs."""

    def __init__(self, buf):
        import torch
        self.raw = buf                                  # uint8 [E, 208]
        i32 = buf.view(torch.int32)                     # [E, 52]
        i64 = buf.view(torch.int64)                     # [E, 26]
        self.t_ns = i64[:, 0]
        self.uid = i64[:, 1]
        self.hop_time = buf.view(torch.float64)[:, 2]
        self.node, self.overlay, self.ptype, self.done = i32[:, 6], i32[:, 7], i32[:, 8], i32[:, 9]
        self.action, self.need_action, self.nobs, self.is_new = i32[:, 10], i32[:, 11], i32[:, 12], i32[:, 13]
        self.prev_node, self.n_lost = i32[:, 14], i32[:, 15]
        self.e2e_delay = buf.view(torch.float32)[:, 16]
        self.valid = i32[:, 17]
        self.obs = i32[:, 18:18 + 1 + _abi.PRB_MAX_K]
        self.lost_uid = i64[:, 18:26]

    def numpy(self):
        return np.frombuffer(self.raw.cpu().numpy().tobytes(), dtype=_abi.STEP_DTYPE)


class BatchedRoutingEnv:
    def __init__(self, cfg: EnvConfig, device: Optional[int] = None, stream=None, lanes_per_env: int = 0,
                 pool_capacity: int = 0, env_id_base: int = 0, arrays: Optional[EnvArrays] = None):
        import torch
        self._torch = torch
        self.lib = _abi.lib()
        if not torch.cuda.is_available():
            raise RuntimeError("prisma_b200 needs a CUDA device: there is no CPU execution path")
        self.cfg = cfg
        self.arrays = arrays if arrays is not None else cfg.load()
        self.device = int(cfg.device if device is None else device)
        self.num_envs = int(cfg.num_envs)
        self._keep = []
        params = cfg.abi_params(self.arrays, env_id_base=env_id_base, pool_capacity=pool_capacity,
                                lanes_per_env=lanes_per_env)
        self._cfg_struct = _abi.fill_config(_abi.PrbConfig(), self._keep, params)
        self.sim_time_ns = int(self._cfg_struct.sim_time_ns)
        with torch.cuda.device(self.device):
            self.stream = stream if stream is not None else torch.cuda.current_stream()
        h = C.c_void_p()
        _abi.check(self.lib.prb_create(C.byref(self._cfg_struct), self.device, C.c_void_p(self.stream.cuda_stream),
                                       C.byref(h)))
        self.h = h
        self.num_links = self.lib.prb_num_links(self.h)
        dev = torch.device("cuda", self.device)
        self._out = torch.zeros((self.num_envs, 208), dtype=torch.uint8, device=dev)
        self._counters = torch.zeros((self.num_envs, 192), dtype=torch.uint8, device=dev)
        self._zero_actions = torch.zeros(self.num_envs, dtype=torch.int32, device=dev)
        self.last = StepBatch(self._out)
        self.neighbors = self.arrays.overlay_neighbors()
        self.overlay_to_underlay = self.arrays.overlay_to_underlay()

    # ---- lifecycle ----
    def close(self):
        if getattr(self, "h", None):
            self.lib.prb_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self):
        _abi.check(self.lib.prb_reset(self.h))
        self._out.zero_()

    def sync(self):
        _abi.check(self.lib.prb_sync(self.h))

    @property
    def launch_count(self) -> int:
        return int(self.lib.prb_launch_count(self.h))

    def state_bytes(self):
        a, b = C.c_int64(), C.c_int64()
        _abi.check(self.lib.prb_state_bytes(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    # ---- traffic / agents ----
    def set_inject_trace(self, t_ns, src, dst, uid, valable, rejected=None, ptype=None):
        """Recorded injections shared by every environment; ptype: 0 data (default), 1 big-signalling segment."""
        a = [np.ascontiguousarray(t_ns, dtype=np.int64), np.ascontiguousarray(src, dtype=np.int32),
             np.ascontiguousarray(dst, dtype=np.int32), np.ascontiguousarray(uid, dtype=np.int64),
             np.ascontiguousarray(valable, dtype=np.uint8),
             np.ascontiguousarray(rejected if rejected is not None else np.zeros(len(t_ns)), dtype=np.uint8)]
        a.append(np.ascontiguousarray(ptype if ptype is not None else np.zeros(len(a[0])), dtype=np.uint8))
        _abi.check(self.lib.prb_set_inject_trace_typed(self.h, len(a[0]), _abi.ptr(a[0], C.c_int64),
                                                       _abi.ptr(a[1], C.c_int32), _abi.ptr(a[2], C.c_int32),
                                                       _abi.ptr(a[3], C.c_int64), _abi.ptr(a[4], C.c_uint8),
                                                       _abi.ptr(a[5], C.c_uint8), _abi.ptr(a[6], C.c_uint8)))

    def set_policy(self, kind: str, data=None, seed: int = 1):
        d = np.ascontiguousarray(data if data is not None else [], dtype=np.int32).ravel()
        _abi.check(self.lib.prb_set_policy(self.h, _abi.POLICY[kind], _abi.ptr(d, C.c_int32) if d.size else None,
                                           d.size, int(seed)))

    def set_agent(self, agent_type: Optional[str] = None, table=None):
        """Select the device-side agent hook: `sp` builds the shortest-path table; any agent whose decision
        depends on the destination only (dq_routing, opt without rejection) can be passed as `table`."""
        agent_type = agent_type or self.cfg.agent_type
        if table is None:
            if agent_type != "sp":
                raise ValueError("only the sp agent has a built-in table; pass table= for the others")
            table = sp_table(self.arrays)
        self.set_policy("table", table)

    def set_engine(self, mode: str = "auto", team: int = 0):
        """Engine of `run`: "auto" (the node-parallel engine for table / per-node-trace agents without ordered
        logs, else the event-serial one), "serial" or "parallel"; `team` = lanes per environment (0: auto)."""
        _abi.check(self.lib.prb_set_engine(self.h, _abi.ENGINE[mode], int(team)))

    @property
    def engine(self) -> str:
        """Engine that has advanced the environments since the last reset."""
        return {0: "none", 1: "serial", 2: "parallel"}[self.lib.prb_engine(self.h)]

    def round_stats(self):
        """(rounds, events executed) of the node-parallel engine so far: their ratio is the number of events
        of one environment that ran concurrently."""
        a, b = C.c_int64(), C.c_int64()
        _abi.check(self.lib.prb_round_stats(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def handovers(self) -> int:
        """Environments the node-parallel engine handed over to the event-serial engine in this episode."""
        n = self.lib.prb_handovers(self.h)
        if n < 0:
            _abi.check(n)
        return int(n)

    def handover_stats(self):
        """(environments handed over, device milliseconds the serial engine spent replaying them) in this episode."""
        n, ms = C.c_int64(), C.c_double()
        _abi.check(self.lib.prb_handover_stats(self.h, C.byref(n), C.byref(ms)))
        return int(n.value), float(ms.value)

    # ---- stepping ----
    def run(self, until_s: Optional[float] = None, until_ns: Optional[int] = None):
        if until_ns is None:
            until_ns = self.sim_time_ns if until_s is None else int(self.lib.prb_seconds_to_ns(float(until_s)))
        _abi.check(self.lib.prb_run(self.h, int(until_ns)))

    def step(self, actions=None) -> StepBatch:
        """Gym mode on device tensors: `actions` int32 [E] (cuda) answers the previous records."""
        a = self._zero_actions if actions is None else actions
        if a.dtype != self._torch.int32 or not a.is_cuda or a.numel() != self.num_envs or not a.is_contiguous():
            raise ValueError("actions must be a contiguous cuda int32 tensor with one entry per environment")
        _abi.check(self.lib.prb_step(self.h, C.c_void_p(a.data_ptr()), C.c_void_p(self._out.data_ptr())))
        return self.last

    def step_host(self, actions: Optional[np.ndarray] = None) -> np.ndarray:
        """Gym mode through host buffers (actions up, records down inside the call)."""
        out = np.empty(self.num_envs, dtype=_abi.STEP_DTYPE)
        ap = None
        if actions is not None:
            actions = np.ascontiguousarray(actions, dtype=np.int32)
            ap = actions.ctypes.data_as(C.c_void_p)
        _abi.check(self.lib.prb_step_host(self.h, ap, out.ctypes.data_as(C.c_void_p)))
        return out

    # ---- dqn_buffer agents on the device ----
    def qnet_load(self, weights: np.ndarray):
        """Per-node Q-networks (prisma_b200.qnet.pack_weights layout), float32."""
        w = np.ascontiguousarray(weights, dtype=np.float32).ravel()
        _abi.check(self.lib.prb_qnet_load(self.h, w.ctypes.data_as(C.c_void_p), w.size))

    def qnet_set_arch(self, hidden_layers: int = 2):
        """Hidden layers after the concatenation: 2 (dqn_buffer, dqn_buffer_lite, dqn_routing) or 1 (dqn_buffer_lighter*)."""
        _abi.check(self.lib.prb_qnet_set_arch(self.h, int(hidden_layers)))

    def qnet_set_mode(self, mode: str = "fused"):
        """'fused': Q-networks inside the simulator kernel (one warp per environment, fp32); 'warp': the same as its own
        kernel; 'mma': decisions of all environments bucketed by node, tiles of 16 on the tensor cores."""
        _abi.check(self.lib.prb_qnet_set_mode(self.h, {"fused": 0, "warp": 1, "mma": 2}[mode]))

    def qnet_act(self, eps: float = 0.0, seed: int = 0, actions=None, q_out=None):
        """argmin-Q actions (eps-greedy) for the records of the last `step`; returns the int32 [E] cuda tensor."""
        torch = self._torch
        if actions is None:
            if not hasattr(self, "_q_actions"):
                self._q_actions = torch.zeros(self.num_envs, dtype=torch.int32, device=self._out.device)
            actions = self._q_actions
        _abi.check(self.lib.prb_qnet_act(self.h, C.c_void_p(self._out.data_ptr()), C.c_void_p(actions.data_ptr()),
                                         C.c_void_p(q_out.data_ptr()) if q_out is not None else None,
                                         C.c_float(eps), int(seed)))
        return actions

    def step_qnet(self, n_steps: int = 1, eps: float = 0.0, seed: int = 0) -> StepBatch:
        """n_steps x (Q-network forward, argmin, env step) without leaving the device."""
        _abi.check(self.lib.prb_step_qnet(self.h, int(n_steps), C.c_float(eps), int(seed), C.c_void_p(self._out.data_ptr())))
        return self.last

    # ---- replay buffers in HBM ----
    def replay_enable(self, rows_per_node: int, pending_slots: int = 4096):
        _abi.check(self.lib.prb_replay_enable(self.h, int(rows_per_node), int(pending_slots)))
        self._replay_k = max(len(n) for n in self.neighbors)

    def replay_set_target_mode(self, on: bool = True, gamma: float = 1.0):
        """`target` signalling: transit transitions store the TD target computed by the next node's network
        (hop time + gamma * min Q over the actions that do not lead back) instead of the hop time."""
        _abi.check(self.lib.prb_replay_set_target_mode(self.h, 1 if on else 0, C.c_float(gamma)))

    def replay_record(self, actions):
        """Feed the records of the last `step` and the actions that answer them (before the next `step`)."""
        _abi.check(self.lib.prb_replay_record(self.h, C.c_void_p(self._out.data_ptr()), C.c_void_p(actions.data_ptr())))

    def replay_counts(self):
        n = (C.c_int64 * len(self.neighbors))()
        missed = C.c_int64()
        _abi.check(self.lib.prb_replay_counts(self.h, n, C.byref(missed)))
        return np.array(list(n), dtype=np.int64), int(missed.value)

    def replay_rows(self, node: int) -> np.ndarray:
        """Stored transitions of one node: float32 [n, 2*(1+K)+3] = obs, action, reward, next_obs, done."""
        n = C.c_int64()
        _abi.check(self.lib.prb_replay_read(self.h, int(node), None, 0, C.byref(n)))
        out = np.empty((n.value, 2 * (1 + self._replay_k) + 3), dtype=np.float32)
        if n.value:
            _abi.check(self.lib.prb_replay_read(self.h, int(node), out.ctypes.data_as(C.c_void_p), n.value, C.byref(n)))
        return out

    def replay_enable_prioritized(self, alpha: float = 1.0):
        """PrioritizedReplayBuffer (replay_buffer.py:393-534): rows remember the gradient step of the network that
        produced their target; `replay_sample (..., weighted=True)` returns the importance weights."""
        _abi.check(self.lib.prb_replay_enable_prioritized(self.h, C.c_float(alpha)))
        self._replay_prioritized = True

    def replay_set_gradient_steps(self, steps):
        """Gradient steps taken so far by every node's learner (int [M])."""
        s = np.ascontiguousarray(steps, dtype=np.int32)
        assert s.size == len(self.neighbors)
        _abi.check(self.lib.prb_replay_set_gradient_steps(self.h, _abi.ptr(s, C.c_int32)))

    def replay_priorities(self, node: int):
        """(gradient step of every stored row of `node`, latest step seen per action)."""
        n = len(self.replay_rows(node))
        g = np.zeros(max(n, 1), dtype=np.float32)
        lat = np.zeros(self._replay_k, dtype=np.int32)
        _abi.check(self.lib.prb_replay_read_priorities(self.h, int(node), g.ctypes.data_as(C.c_void_p), n, _abi.ptr(lat, C.c_int32)))
        return g[:n], lat

    def replay_sample(self, node: int, batch: int, seed: int = 0, weighted: bool = False):
        """ReplayBuffer.sample on the device: (obs [B,1+K], action [B], reward [B], next_obs [B,1+K], done [B]);
        weighted: also the importance weights [B] of PrioritizedReplayBuffer.sample (ones without the prioritized variant)."""
        if weighted:
            return self._replay_sample_weighted(node, batch, seed)
        torch = self._torch
        dev, ow = self._out.device, 1 + self._replay_k
        obs = torch.zeros((batch, ow), dtype=torch.float32, device=dev)
        nxt = torch.zeros((batch, ow), dtype=torch.float32, device=dev)
        act = torch.zeros(batch, dtype=torch.int32, device=dev)
        rew = torch.zeros(batch, dtype=torch.float32, device=dev)
        done = torch.zeros(batch, dtype=torch.float32, device=dev)
        _abi.check(self.lib.prb_replay_sample(self.h, int(node), int(batch), int(seed), C.c_void_p(obs.data_ptr()),
                                              C.c_void_p(act.data_ptr()), C.c_void_p(rew.data_ptr()),
                                              C.c_void_p(nxt.data_ptr()), C.c_void_p(done.data_ptr())))
        return obs, act, rew, nxt, done

    def _replay_sample_weighted(self, node, batch, seed):
        torch = self._torch
        dev, ow = self._out.device, 1 + self._replay_k
        obs = torch.zeros((batch, ow), dtype=torch.float32, device=dev)
        nxt = torch.zeros((batch, ow), dtype=torch.float32, device=dev)
        act = torch.zeros(batch, dtype=torch.int32, device=dev)
        rew, done, wts = (torch.zeros(batch, dtype=torch.float32, device=dev) for _ in range(3))
        _abi.check(self.lib.prb_replay_sample_weighted(self.h, int(node), int(batch), int(seed), C.c_void_p(obs.data_ptr()),
                                                       C.c_void_p(act.data_ptr()), C.c_void_p(rew.data_ptr()),
                                                       C.c_void_p(nxt.data_ptr()), C.c_void_p(done.data_ptr()),
                                                       C.c_void_p(wts.data_ptr())))
        return obs, act, rew, nxt, done, wts

    # ---- read-out ----
    def counters(self) -> np.ndarray:
        out = np.empty(self.num_envs, dtype=_abi.COUNTERS_DTYPE)
        _abi.check(self.lib.prb_counters_host(self.h, out.ctypes.data_as(C.c_void_p)))
        return out

    def counters_device(self):
        _abi.check(self.lib.prb_counters_dev(self.h, C.c_void_p(self._counters.data_ptr())))
        return self._counters

    def check_errors(self, counters=None):
        c = self.counters() if counters is None else counters
        bits = int(np.bitwise_or.reduce(c["errors"]))
        if bits:
            raise _abi.PrbError("environment error flags: " + ", ".join(v for k, v in _abi.ERR_BITS.items() if bits & k))

    def queue_bytes(self, e: int = 0) -> np.ndarray:
        out = np.zeros(self.num_links, dtype=np.int32)
        _abi.check(self.lib.prb_queue_bytes(self.h, int(e), _abi.ptr(out, C.c_int32)))
        return out

    def queue_bytes_noalloc(self, e: int = 0) -> np.ndarray:
        """queue_bytes through the allocation-free entry point (static device scratch; include/prisma_b200.h)."""
        L = self.lib.prb_num_links(self.h)
        out = np.zeros(L, dtype=np.int32)
        _abi.check(self.lib.prb_queue_bytes_noalloc(self.h, int(e), _abi.ptr(out, C.c_int32), L))
        return out

    def enable_log(self, event_capacity: int, step_capacity: int):
        _abi.check(self.lib.prb_enable_log(self.h, int(event_capacity), int(step_capacity)))
        self._ev_cap, self._st_cap = int(event_capacity), int(step_capacity)

    def events(self, e: int = 0) -> np.ndarray:
        n = C.c_int64()
        _abi.check(self.lib.prb_read_events(self.h, int(e), None, 0, C.byref(n)))
        out = np.empty(min(n.value, self._ev_cap), dtype=_abi.EVENT_DTYPE)
        if len(out):
            _abi.check(self.lib.prb_read_events(self.h, int(e), out.ctypes.data_as(C.c_void_p), len(out), C.byref(n)))
        return out

    def steps(self, e: int = 0) -> np.ndarray:
        n = C.c_int64()
        _abi.check(self.lib.prb_read_steps(self.h, int(e), None, 0, C.byref(n)))
        out = np.empty(min(n.value, self._st_cap), dtype=_abi.STEP_DTYPE)
        if len(out):
            _abi.check(self.lib.prb_read_steps(self.h, int(e), out.ctypes.data_as(C.c_void_p), len(out), C.byref(n)))
        return out


class _Discrete:
    def __init__(self, n):
        self.n = int(n)

    def contains(self, a):
        return 0 <= int(a) < self.n

    def __repr__(self):
        return f"Discrete({self.n})"


class _Box:
    def __init__(self, low, high, shape, dtype):
        self.low, self.high, self.shape, self.dtype = low, high, tuple(shape), dtype

    def __repr__(self):
        return f"Box({self.low}, {self.high}, {self.shape}, {self.dtype})"


class Ns3EnvGroup:
    """One environment of the batch seen through the reference's per-node interface.

    In the reference every overlay node owns an `Ns3Env` whose `step` blocks until ns-3 delivers that
    node's next packet, one Python thread per node interleaving (prisma/main.py:117-127).  The same
    contract holds here: `nodes[m].step(action)` hands the action to the kernel, and blocks until the
    environment's next valid Notify() belongs to node m (other nodes' threads consume theirs meanwhile).
    Single-threaded drivers use `next()` to learn whose turn it is.
    """

    def __init__(self, benv: BatchedRoutingEnv, env_index: int = 0):
        import threading
        self.benv = benv
        self.e = int(env_index)
        self._rec = None
        self._ended = False
        self._cv = threading.Condition()
        self.nodes = [Ns3EnvCompat(self, m) for m in range(benv.arrays.n_overlay)]

    def _advance(self, action=0):
        acts = np.zeros(self.benv.num_envs, dtype=np.int32)
        acts[self.e] = int(action)
        self._rec = self.benv.step_host(acts)[self.e].copy()
        if not self._rec["valid"]:
            self._ended = True

    def next(self) -> int:
        """Overlay index of the node whose step is ready, or -1 when the simulation has ended."""
        with self._cv:
            if self._rec is None and not self._ended:
                self._advance()
                self._cv.notify_all()
            return -1 if self._ended else int(self._rec["overlay"])

    def info_string(self, rec, c) -> str:
        return format_info(rec, c, self.benv.arrays.map_overlay, self.benv.cfg.packet_size)


def format_info(rec, c, map_overlay, packet_size) -> str:
    """The reference's info string, key for key and separator for separator: PacketManager::getInfo
    (prisma/ns3/packet-manager.cc:119-176, fields 0-17: note the irregular blanks around '=' and after ','), followed by
    the fields of the packet's own manager — data: Packet Lost / Source / Destination / node
    (data-packet-manager.cc:229-248); small signalling: PacketIdSignaled / Arrived at final dest
    (small-signaling-packet-manager.cc:109-119); big signalling: NN Index / segment Index / NodeId Signaled
    (big-signaling-packet-manager.cc:110-122).  Numbers are std::to_string's: "%f" for doubles and floats.
    Forwarder.treat_info reads the fields by position (prisma/source/forwarder.py:197-289).
    rec: one prb_step_rec (numpy record or mapping), c: the environment's prb_counters."""
    ptype = int(rec["ptype"])
    # m_packetSize is the size on the wire: payload + 30 (data; a source pads payloads below 200 bytes,
    # poisson-application.cc:277-278), 512 + 30 for an NN segment.  Not reproduced for signalling records, and never
    # read there by Forwarder.treat_info: field 0 (an uninitialised member in the reference), field 1 of small
    # signalling (the sender's degree decides it) and field 3 (ns-3's global Packet uid).
    if ptype == 0:
        size = (max(packet_size, 200) if int(rec["is_new"]) and int(rec["prev_node"]) < 0 else packet_size) + 30
    else:
        size = 542 if ptype == 1 else 0
    mo = map_overlay
    uid = int(rec["uid"])
    buffered_o = int(c["overlay_injected"]) - int(c["overlay_arrived"]) - int(c["overlay_lost"])
    buffered_u = int(c["underlay_injected"]) - int(c["underlay_arrived"]) - int(c["underlay_lost"])
    # field 0 is a double in the reference: Now().GetSeconds() - m_sourceTimeStamp, the tag's start time being whole
    # seconds (poisson-application.cc:310); the record carries it as float32, from which the integer is recovered
    now_s = int(rec["t_ns"]) / 1e9
    start_s = round(now_s - float(rec["e2e_delay"]))
    s = "End to End Delay=%f" % (now_s - start_s if ptype == 0 else float(rec["e2e_delay"]))
    s += ", Packet Size=%d" % size
    s += ", Current sim time =%f" % now_s
    s += ", Pkt ID =%d" % (0 if ptype == 1 else uid)
    s += ", packetType =%d" % ptype
    s += ", Avg End to End Delay =%f" % float(c["avg_e2e"])
    s += ", Avg Cost =%f" % float(c["avg_cost"])
    s += ", Avg Underlay End to End Delay =%f" % float(c["avg_underlay_e2e"])
    s += ", Avg Underlay Cost =%f" % float(c["avg_underlay_cost"])
    s += ", Packets dropped =%d" % int(c["overlay_lost"])
    s += ", Packets delivered =%d" % int(c["overlay_arrived"])
    s += ", Packets injected =%d" % int(c["overlay_injected"])
    s += ",Packets Buffered =%d" % buffered_o
    s += ", Packets dropped Underlay =%d" % int(c["underlay_lost"])
    s += ", Packets delivered Underlay=%d" % int(c["underlay_arrived"])
    s += ", Packets injected Underlay=%d" % int(c["underlay_injected"])
    s += ",Packets Buffered Underlay=%d" % buffered_u
    s += ",Signaling overhead =%f" % float(c["signaling_overhead"])
    if ptype == 1:
        s += ", NN Index=%d, segment Index=%d, NodeId Signaled=%d" % (uid >> 32, uid & 0xFFFFFFFF, int(mo[int(rec["prev_node"])]))
    elif ptype == 2:
        s += ", PacketIdSignaled=%d, Arrived at final dest=%d" % (uid, 1)
    else:
        lost = "".join("%d;" % int(u) for u in rec["lost_uid"][:min(int(rec["n_lost"]), 8)])
        s += ", Packet Lost=%s, Source=%d, Destination=%d, node=%d" % (lost, int(mo[int(rec["src"])]), int(rec["obs"][0]),
                                                                      int(rec["overlay"]))
    return s


class _NoBridge:
    """Stands where Ns3ZmqBridge (prisma/ns3_model/ns3env.py:33-375) stood: there is no socket to close."""

    def send_close_command(self):
        return True

    def close(self):
        return None


class Ns3EnvCompat:
    """Per-node view with the surface of prisma/ns3_model/ns3env.py:378-455 (reset/step/close,
    action_space, observation_space, connected)."""

    def __init__(self, group: Ns3EnvGroup, overlay_index: int):
        self.group = group
        self.index = int(overlay_index)
        k = len(group.benv.neighbors[self.index])
        self.action_space = _Discrete(k)
        self.observation_space = _Box(0, 16260, (1 + k,), np.uint32)   # data-packet-manager.cc:139-141
        self.connected = True
        self.ns3ZmqBridge = _NoBridge()       # Forwarder.run ends with env.ns3ZmqBridge.send_close_command() (forwarder.py:331)

    def _wait_mine(self, timeout=None):
        """Called with the group's condition held: blocks until the ready step is this node's."""
        g = self.group
        while True:
            if g._rec is None and not g._ended:
                g._advance()
                g._cv.notify_all()
            if g._ended:
                return [-1], -1.0, True, ""      # simulation-end sentinel (packet-routing-gym.cc:158-160)
            if int(g._rec["overlay"]) == self.index:
                r = g._rec
                c = g.benv.counters()[g.e]
                obs = [int(x) for x in r["obs"][:int(r["nobs"])]]
                return obs, 1.0, bool(r["done"]), g.info_string(r, c)
            if not g._cv.wait(timeout):
                raise TimeoutError(f"node {self.index}: the ready step belongs to node {int(g._rec['overlay'])}")

    def reset(self, timeout=None):
        with self.group._cv:
            return self._wait_mine(timeout)[0]

    def step(self, action, timeout=None):
        g = self.group
        with g._cv:
            if g._rec is not None and not g._ended and int(g._rec["overlay"]) == self.index:
                g._advance(action)
                g._cv.notify_all()
            return self._wait_mine(timeout)

    def close(self):
        self.connected = False
