"""TEST INFRASTRUCTURE — ctypes front end of tests/host_harness/libhost_sim.so: the kernel body of
prisma_b200 compiled with g++ for one environment (see host_sim.cpp).  Only tests import this."""
import ctypes as C
import os
import subprocess

import numpy as np

from prisma_b200 import _abi

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libhost_sim.so")
SRC = [os.path.join(HERE, "host_sim.cpp"),
       os.path.join(HERE, "..", "..", "prisma_b200", "csrc", "sim_core.cuh"),
       os.path.join(HERE, "..", "..", "prisma_b200", "csrc", "par_core.cuh"),
       os.path.join(HERE, "..", "..", "prisma_b200", "csrc", "model.h"),
       os.path.join(HERE, "..", "..", "include", "prisma_b200.h")]
_lib = None


def build():
    subprocess.check_call(["g++", "-O2", "-g", "-std=c++17", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-shared",
                           "-Wall", "-Wno-unused-function", "-Wno-unknown-pragmas"]
                          + ["-D" + x for x in os.environ.get("PRB_HOST_DEFINES", "").split()] + ["-o", LIB, SRC[0]])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB) or any(os.path.getmtime(LIB) < os.path.getmtime(s) for s in SRC):
            build()
        L = C.CDLL(LIB)
        vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
        L.hs_create.restype = vp
        L.hs_create.argtypes = [C.POINTER(_abi.PrbConfig), C.c_int, i64, i64]
        L.hs_destroy.argtypes = [vp]
        L.hs_set_inject_trace.argtypes = [vp, i64, C.POINTER(i64), C.POINTER(i32), C.POINTER(i32), C.POINTER(i64),
                                          C.POINTER(C.c_uint8), C.POINTER(C.c_uint8), C.POINTER(C.c_uint8)]
        L.hs_set_policy.argtypes = [vp, C.c_int, C.POINTER(i32), i64, C.c_uint64]
        L.hs_run.argtypes = [vp, i64]
        L.hs_run_nolog.argtypes = [vp, i64]
        L.hs_step.argtypes = [vp, i32, C.POINTER(_abi.PrbStepRec)]
        L.hs_counters.argtypes = [vp, C.POINTER(_abi.PrbCounters)]
        L.hs_num_events.restype = i64
        L.hs_num_events.argtypes = [vp]
        L.hs_events.restype = vp
        L.hs_events.argtypes = [vp]
        L.hs_num_steps.restype = i64
        L.hs_num_steps.argtypes = [vp]
        L.hs_steps.restype = vp
        L.hs_steps.argtypes = [vp]
        L.hs_num_links.argtypes = [vp]
        L.hs_queue_bytes.argtypes = [vp, i32]
        L.hs_seconds_to_ns.restype = i64
        L.hs_seconds_to_ns.argtypes = [C.c_double]
        L.hs_ns_to_seconds.restype = C.c_double
        L.hs_ns_to_seconds.argtypes = [vp, i64]
        L.hs_tx_time_ns.restype = i64
        L.hs_tx_time_ns.argtypes = [C.c_uint32, C.c_uint64]
        L.hs_invert_1e9.argtypes = [C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.hs_global_routes.argtypes = [i32, C.POINTER(C.c_uint8), C.POINTER(i32)]
        L.hp_create.restype = vp
        L.hp_create.argtypes = [C.POINTER(_abi.PrbConfig), C.c_int, i64, C.c_int]
        L.hp_destroy.argtypes = [vp]
        L.hp_set_inject_trace.argtypes = L.hs_set_inject_trace.argtypes[:-1]
        L.hp_set_policy.argtypes = [vp, C.c_int, C.POINTER(i32), i64]
        L.hp_run.argtypes = [vp, i64, C.c_int]
        L.hp_counters.argtypes = [vp, C.POINTER(_abi.PrbCounters)]
        for f in ("hp_num_events", "hp_rounds", "hp_executed", "hp_lookahead_ns"):
            getattr(L, f).restype = i64
            getattr(L, f).argtypes = [vp]
        L.hp_events.restype = vp
        L.hp_events.argtypes = [vp]
        L.hp_num_links.argtypes = [vp]
        L.hp_queue_bytes.argtypes = [vp, i32]
        L.hp_wire_slots.argtypes = [vp]
        L.hp_state_bytes.argtypes = [vp]
        L.hp_set_chunk.argtypes = [vp, i64]
        _lib = L
    return _lib


def node_trace(decisions_node, decisions_action, o2u):
    """Per-node action lists for PRB_POLICY_NODE_TRACE from a global decision log: [M+1 offsets][actions].
    Only the steps where the agent was consulted (action >= 0) consume an action."""
    m = len(o2u)
    u2o = {int(u): i for i, u in enumerate(o2u)}
    per = [[] for _ in range(m)]
    for n, a in zip(decisions_node, decisions_action):
        if a >= 0:
            per[u2o[int(n)]].append(int(a))
    off = np.zeros(m + 1, dtype=np.int32)
    off[1:] = np.cumsum([len(x) for x in per])
    flat = np.array([a for x in per for a in x], dtype=np.int32)
    return np.concatenate([off, flat]).astype(np.int32)


class HostSim:
    def __init__(self, params, env_index=0, ev_cap=1 << 21, st_cap=1 << 20):
        self._keep = []
        self.cfg = _abi.fill_config(_abi.PrbConfig(), self._keep, params)
        self.h = lib().hs_create(C.byref(self.cfg), env_index, ev_cap, st_cap)
        if not self.h:
            raise ValueError("hs_create failed")

    def close(self):
        if self.h:
            lib().hs_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_inject_trace(self, t_ns, src, dst, uid, valable, rejected, ptype=None):
        a = [np.ascontiguousarray(t_ns, dtype=np.int64), np.ascontiguousarray(src, dtype=np.int32),
             np.ascontiguousarray(dst, dtype=np.int32), np.ascontiguousarray(uid, dtype=np.int64),
             np.ascontiguousarray(valable, dtype=np.uint8), np.ascontiguousarray(rejected, dtype=np.uint8),
             np.ascontiguousarray(ptype if ptype is not None else np.zeros(len(t_ns)), dtype=np.uint8)]
        rc = lib().hs_set_inject_trace(self.h, len(a[0]), _abi.ptr(a[0], C.c_int64), _abi.ptr(a[1], C.c_int32),
                                       _abi.ptr(a[2], C.c_int32), _abi.ptr(a[3], C.c_int64),
                                       _abi.ptr(a[4], C.c_uint8), _abi.ptr(a[5], C.c_uint8), _abi.ptr(a[6], C.c_uint8))
        if rc != 0:
            raise ValueError("bad injection trace")

    def set_policy(self, kind, data=None, seed=1):
        d = np.ascontiguousarray(data if data is not None else [], dtype=np.int32).ravel()
        lib().hs_set_policy(self.h, _abi.POLICY[kind], _abi.ptr(d, C.c_int32), d.size, int(seed))

    def run(self, until_ns=None):
        lib().hs_run(self.h, int(until_ns if until_ns is not None else self.cfg.sim_time_ns))

    def run_nolog(self, until_ns=None):
        lib().hs_run_nolog(self.h, int(until_ns if until_ns is not None else self.cfg.sim_time_ns))

    def step(self, action=0):
        r = _abi.PrbStepRec()
        lib().hs_step(self.h, int(action), C.byref(r))
        return r

    def counters(self):
        c = _abi.PrbCounters()
        lib().hs_counters(self.h, C.byref(c))
        return {n: getattr(c, n) for n, _ in c._fields_}

    def events(self):
        n = lib().hs_num_events(self.h)
        if n == 0:
            return np.zeros(0, dtype=_abi.EVENT_DTYPE)
        buf = (C.c_char * (n * 64)).from_address(lib().hs_events(self.h))
        return np.frombuffer(buf, dtype=_abi.EVENT_DTYPE).copy()

    def steps(self):
        n = lib().hs_num_steps(self.h)
        if n == 0:
            return np.zeros(0, dtype=_abi.STEP_DTYPE)
        buf = (C.c_char * (n * 208)).from_address(lib().hs_steps(self.h))
        return np.frombuffer(buf, dtype=_abi.STEP_DTYPE).copy()

    def queue_bytes(self):
        nl = lib().hs_num_links(self.h)
        return np.array([lib().hs_queue_bytes(self.h, l) for l in range(nl)], dtype=np.int32)


class HostParSim:
    """The node-parallel engine (prisma_b200/csrc/par_core.cuh) with the lanes of a team emulated in turn."""

    def __init__(self, params, env_index=0, ev_cap=1 << 21, team=32):
        self._keep = []
        self.cfg = _abi.fill_config(_abi.PrbConfig(), self._keep, params)
        self.h = lib().hp_create(C.byref(self.cfg), env_index, ev_cap, team)
        if not self.h:
            raise ValueError("hp_create failed (configuration not supported by the parallel engine?)")

    def close(self):
        if self.h:
            lib().hp_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_inject_trace(self, t_ns, src, dst, uid, valable, rejected, ptype=None):
        assert ptype is None or not np.any(ptype), "data injections only"
        a = [np.ascontiguousarray(t_ns, dtype=np.int64), np.ascontiguousarray(src, dtype=np.int32),
             np.ascontiguousarray(dst, dtype=np.int32), np.ascontiguousarray(uid, dtype=np.int64),
             np.ascontiguousarray(valable, dtype=np.uint8), np.ascontiguousarray(rejected, dtype=np.uint8)]
        rc = lib().hp_set_inject_trace(self.h, len(a[0]), _abi.ptr(a[0], C.c_int64), _abi.ptr(a[1], C.c_int32),
                                       _abi.ptr(a[2], C.c_int32), _abi.ptr(a[3], C.c_int64),
                                       _abi.ptr(a[4], C.c_uint8), _abi.ptr(a[5], C.c_uint8))
        if rc != 0:
            raise ValueError("bad injection trace")

    def set_policy(self, kind, data=None):
        d = np.ascontiguousarray(data if data is not None else [], dtype=np.int32).ravel()
        lib().hp_set_policy(self.h, _abi.POLICY[kind], _abi.ptr(d, C.c_int32), d.size)

    def run(self, until_ns=None, log=True):
        lib().hp_run(self.h, int(until_ns if until_ns is not None else self.cfg.sim_time_ns), 1 if log else 0)

    def counters(self):
        c = _abi.PrbCounters()
        lib().hp_counters(self.h, C.byref(c))
        return {n: getattr(c, n) for n, _ in c._fields_}

    def events(self):
        n = lib().hp_num_events(self.h)
        if n == 0:
            return np.zeros(0, dtype=_abi.EVENT_DTYPE)
        buf = (C.c_char * (n * 64)).from_address(lib().hp_events(self.h))
        return np.frombuffer(buf, dtype=_abi.EVENT_DTYPE).copy()

    def queue_bytes(self):
        nl = lib().hp_num_links(self.h)
        return np.array([lib().hp_queue_bytes(self.h, l) for l in range(nl)], dtype=np.int32)

    def set_chunk(self, ns):
        """Longest stretch of simulated time per emulated launch (each one re-bases the 32-bit time offsets)."""
        lib().hp_set_chunk(self.h, int(ns))

    @property
    def state_bytes(self):
        return lib().hp_state_bytes(self.h)

    @property
    def rounds(self):
        return lib().hp_rounds(self.h)

    @property
    def executed(self):
        return lib().hp_executed(self.h)
