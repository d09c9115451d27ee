"""TEST INFRASTRUCTURE — ctypes front end of oracle/libprisma_oracle.so (the CPU restatement).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product (prisma_b200/) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libprisma_oracle.so")

PO_MAX_K = 16
EV_KINDS = ["INJ", "ARR", "ENQ", "DRP", "TXB", "RCV"]
POLICY = {"zero": 0, "table": 1, "trace": 2, "random": 3}


class PoConfig(C.Structure):
    _fields_ = [
        ("n_nodes", C.c_int32), ("n_overlay", C.c_int32),
        ("phys_adj", C.POINTER(C.c_uint8)), ("overlay_adj", C.POINTER(C.c_uint8)),
        ("map_overlay", C.POINTER(C.c_int32)), ("traffic_bps", C.POINTER(C.c_uint64)),
        ("route_out", C.POINTER(C.c_int32)), ("opt_rejected", C.POINTER(C.c_double)),
        ("link_rate_bps", C.c_uint64), ("link_delay_ns", C.c_int64),
        ("packet_size", C.c_uint32), ("max_out_buffer_size", C.c_uint32),
        ("sim_time_ns", C.c_int64), ("ping_interval_ns", C.c_int64),
        ("moving_average_size", C.c_uint32), ("ping_as_obs", C.c_int32),
        ("activate_underlay_traffic", C.c_int32), ("train", C.c_int32),
        ("signaling_type", C.c_int32), ("opt_agent", C.c_int32),
        ("loss_penalty", C.c_double), ("inject_mode", C.c_int32),
        ("seed", C.c_uint64), ("env_id", C.c_uint64),
        ("signaling", C.c_int32), ("big_signaling_size", C.c_uint32), ("sync_step", C.c_double),
    ]


class PoEvent(C.Structure):
    _fields_ = [("t_ns", C.c_int64), ("kind", C.c_int32), ("node", C.c_int32), ("link", C.c_int32),
                ("uid", C.c_int64), ("size", C.c_int32), ("type", C.c_int32), ("src", C.c_int32),
                ("dst", C.c_int32), ("last_hop", C.c_int32), ("next_hop", C.c_int32),
                ("valable", C.c_int32), ("qbytes", C.c_int32)]


class PoDecision(C.Structure):
    _fields_ = [("t_ns", C.c_int64), ("node", C.c_int32), ("overlay", C.c_int32), ("uid", C.c_int64),
                ("ptype", C.c_int32), ("done", C.c_int32), ("action", C.c_int32), ("nobs", C.c_int32),
                ("obs", C.c_int32 * (1 + PO_MAX_K)), ("is_new", C.c_int32), ("prev_node", C.c_int32),
                ("hop_time", C.c_double), ("e2e_delay", C.c_float), ("n_lost", C.c_int32),
                ("lost_uid", C.c_int64 * 8)]


class PoCounters(C.Structure):
    _fields_ = [(n, C.c_int64) for n in
                ("overlay_injected", "overlay_arrived", "overlay_lost", "underlay_injected",
                 "underlay_arrived", "underlay_lost", "bytes_data", "bytes_signaling",
                 "n_e2e", "n_cost", "n_underlay_e2e", "n_underlay_cost")] + \
               [(n, C.c_float) for n in
                ("sum_e2e", "sum_cost", "sum_underlay_e2e", "sum_underlay_cost", "avg_e2e", "avg_cost",
                 "avg_underlay_e2e", "avg_underlay_cost", "signaling_overhead")] + \
               [(n, C.c_int64) for n in
                ("packet_hops", "data_hops", "env_steps", "decisions", "events", "vanished", "end_ns")]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", HERE, "libprisma_oracle.so"])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(os.path.join(HERE, "prisma_oracle.c")):
            build()
        L = C.CDLL(LIB)
        L.po_seconds_to_ns.restype = C.c_int64
        L.po_seconds_to_ns.argtypes = [C.c_double]
        L.po_ns_to_seconds.restype = C.c_double
        L.po_ns_to_seconds.argtypes = [C.c_int64]
        L.po_tx_time_ns.restype = C.c_int64
        L.po_tx_time_ns.argtypes = [C.c_uint32, C.c_uint64]
        L.po_invert_1e9.argtypes = [C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.po_global_routes.restype = C.c_int32
        L.po_global_routes.argtypes = [C.c_int32, C.POINTER(C.c_uint8), C.POINTER(C.c_int32)]
        L.po_build_links.restype = C.c_int32
        L.po_build_links.argtypes = [C.c_int32, C.POINTER(C.c_uint8), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
        L.po_num_links.restype = C.c_int32
        L.po_num_links.argtypes = [C.c_int32, C.POINTER(C.c_uint8)]
        L.po_create.restype = C.c_void_p
        L.po_create.argtypes = [C.POINTER(PoConfig)]
        L.po_destroy.argtypes = [C.c_void_p]
        L.po_set_event_log.argtypes = [C.c_void_p, C.c_int32]
        L.po_set_inject_trace.restype = C.c_int32
        L.po_set_inject_trace.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_int32),
                                          C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.POINTER(C.c_uint8),
                                          C.POINTER(C.c_uint8)]
        L.po_set_inject_trace_typed.restype = C.c_int32
        L.po_set_inject_trace_typed.argtypes = L.po_set_inject_trace.argtypes + [C.POINTER(C.c_uint8)]
        L.po_set_policy.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_int32), C.c_int64, C.c_uint64]
        L.po_run.argtypes = [C.c_void_p, C.c_int64]
        L.po_get_counters.argtypes = [C.c_void_p, C.POINTER(PoCounters)]
        L.po_num_events.restype = C.c_int64
        L.po_num_events.argtypes = [C.c_void_p]
        L.po_events.restype = C.POINTER(PoEvent)
        L.po_events.argtypes = [C.c_void_p]
        L.po_num_decisions.restype = C.c_int64
        L.po_num_decisions.argtypes = [C.c_void_p]
        L.po_decisions.restype = C.POINTER(PoDecision)
        L.po_decisions.argtypes = [C.c_void_p]
        L.po_queue_bytes.restype = C.c_int32
        L.po_queue_bytes.argtypes = [C.c_void_p, C.c_int32]
        L.po_num_links_of.restype = C.c_int32
        L.po_num_links_of.argtypes = [C.c_void_p]
        L.po_philox4x32.argtypes = [C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
        _lib = L
    return _lib


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def seconds_to_ns(s):
    return int(lib().po_seconds_to_ns(float(s)))


def ns_to_seconds(ns):
    return float(lib().po_ns_to_seconds(int(ns)))


def tx_time_ns(nbytes, bps):
    return int(lib().po_tx_time_ns(int(nbytes), int(bps)))


def global_routes(adj):
    adj = np.ascontiguousarray(adj, dtype=np.uint8)
    n = adj.shape[0]
    out = np.zeros((n, n), dtype=np.int32)
    lib().po_global_routes(n, _p(adj, C.c_uint8), _p(out, C.c_int32))
    return out


def build_links(adj):
    adj = np.ascontiguousarray(adj, dtype=np.uint8)
    n = adj.shape[0]
    nl = lib().po_num_links(n, _p(adj, C.c_uint8))
    fr = np.zeros(nl + 1, dtype=np.int32)
    to = np.zeros(nl + 1, dtype=np.int32)
    lib().po_build_links(n, _p(adj, C.c_uint8), _p(fr, C.c_int32), _p(to, C.c_int32))
    return fr[:nl], to[:nl]


def philox(ctr, key):
    c = (C.c_uint32 * 4)(*ctr)
    k = (C.c_uint32 * 2)(*key)
    o = (C.c_uint32 * 4)()
    lib().po_philox4x32(c, k, o)
    return list(o)


SIGNALING = {"ideal": 0, "NN": 1, "target": 2, "digital_twin": 3}


class OracleSim:
    """One environment of the CPU restatement.  `p` is a dict with the arrays/scalars of po_config."""

    def __init__(self, p, log_events=True):
        self._keep = []
        cfg = PoConfig()
        n = int(p["n_nodes"])
        m = int(p["n_overlay"])

        def arr(key, dt, ct, shape=None, optional=False):
            v = p.get(key)
            if v is None:
                if optional:
                    return None
                raise KeyError(key)
            a = np.ascontiguousarray(np.asarray(v), dtype=dt)
            if shape is not None:
                assert a.size == int(np.prod(shape)), (key, a.shape, shape)
            self._keep.append(a)
            return _p(a, ct)

        cfg.n_nodes, cfg.n_overlay = n, m
        cfg.phys_adj = arr("phys_adj", np.uint8, C.c_uint8, (n, n))
        cfg.overlay_adj = arr("overlay_adj", np.uint8, C.c_uint8, (m, m))
        cfg.map_overlay = arr("map_overlay", np.int32, C.c_int32, (n,))
        cfg.traffic_bps = arr("traffic_bps", np.uint64, C.c_uint64, (n, n))
        r = arr("route_out", np.int32, C.c_int32, (n, n), optional=True)
        if r is not None:
            cfg.route_out = r
        r = arr("opt_rejected", np.float64, C.c_double, (n, n), optional=True)
        if r is not None:
            cfg.opt_rejected = r
        for k in ("link_rate_bps", "link_delay_ns", "packet_size", "max_out_buffer_size", "sim_time_ns",
                  "ping_interval_ns", "moving_average_size", "ping_as_obs", "activate_underlay_traffic",
                  "train", "opt_agent", "inject_mode", "seed", "env_id", "signaling", "big_signaling_size"):
            setattr(cfg, k, int(p.get(k, 0)))
        st = p.get("signaling_type", 0)
        cfg.signaling_type = SIGNALING[st] if isinstance(st, str) else int(st)
        cfg.loss_penalty = float(p.get("loss_penalty", 0.0))
        cfg.sync_step = float(p.get("sync_step", 0.0))
        self.cfg = cfg
        self.h = lib().po_create(C.byref(cfg))
        lib().po_set_event_log(self.h, 1 if log_events else 0)

    def close(self):
        if self.h:
            lib().po_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_inject_trace(self, t_ns, src, dst, uid, valable, rejected, ptype=None):
        """ptype: packet type of every injection (0 data, 1 big signalling); None = all data."""
        a = [np.ascontiguousarray(t_ns, dtype=np.int64), np.ascontiguousarray(src, dtype=np.int32),
             np.ascontiguousarray(dst, dtype=np.int32), np.ascontiguousarray(uid, dtype=np.int64),
             np.ascontiguousarray(valable, dtype=np.uint8), np.ascontiguousarray(rejected, dtype=np.uint8)]
        if ptype is None:
            ptype = np.zeros(len(a[0]), dtype=np.uint8)
        a.append(np.ascontiguousarray(ptype, dtype=np.uint8))
        rc = lib().po_set_inject_trace_typed(self.h, len(a[0]), _p(a[0], C.c_int64), _p(a[1], C.c_int32),
                                             _p(a[2], C.c_int32), _p(a[3], C.c_int64), _p(a[4], C.c_uint8),
                                             _p(a[5], C.c_uint8), _p(a[6], C.c_uint8))
        if rc != 0:
            raise ValueError("injection trace names a flow that the configuration does not have")

    def set_policy(self, kind, data=None, seed=1):
        d = np.ascontiguousarray(data if data is not None else [], dtype=np.int32).ravel()
        lib().po_set_policy(self.h, POLICY[kind], _p(d, C.c_int32), d.size, int(seed))

    def run(self, until_ns=None):
        lib().po_run(self.h, int(until_ns if until_ns is not None else self.cfg.sim_time_ns))

    def counters(self):
        c = PoCounters()
        lib().po_get_counters(self.h, C.byref(c))
        return c.as_dict()

    def events(self):
        n = lib().po_num_events(self.h)
        if n == 0:
            return np.zeros(0, dtype=np.dtype(PoEvent))
        buf = C.cast(lib().po_events(self.h), C.POINTER(PoEvent * n)).contents
        return np.frombuffer(buf, dtype=np.dtype(PoEvent)).copy()

    def decisions(self):
        n = lib().po_num_decisions(self.h)
        if n == 0:
            return np.zeros(0, dtype=np.dtype(PoDecision))
        buf = C.cast(lib().po_decisions(self.h), C.POINTER(PoDecision * n)).contents
        return np.frombuffer(buf, dtype=np.dtype(PoDecision)).copy()

    def queue_bytes(self):
        nl = lib().po_num_links_of(self.h)
        return np.array([lib().po_queue_bytes(self.h, l) for l in range(nl)], dtype=np.int32)
