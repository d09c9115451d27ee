"""ctypes mirror of include/prisma_b200.h (structs, enums) and the loader of libprisma_b200.so.

There is no Python or CPU implementation behind this module: if the CUDA library has not been built
(`python -c "import __graft_entry__ as g; g.build()"` or `python -m prisma_b200.build`) loading fails
with an error instead of falling back.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PRB_LIB") or os.path.join(HERE, "libprisma_b200.so")   # PRB_LIB: kernel experiments

PRB_MAX_K = 16
PRB_MAX_LOST = 8
PRB_LASTHOP_NONE = 1000

PRB_DATA, PRB_BIG_SIGN, PRB_SMALL_SIGN, PRB_PING_FWD, PRB_PING_BACK = range(5)
EV_KINDS = ["INJ", "ARR", "ENQ", "DRP", "TXB", "RCV"]
POLICY = {"zero": 0, "table": 1, "trace": 2, "random": 3, "node_trace": 4}
ENGINE = {"auto": 0, "serial": 1, "parallel": 2}
SIGNALING = {"ideal": 0, "NN": 1, "target": 2, "digital_twin": 3}
ERR_BITS = {1: "packet pool exhausted", 2: "log capacity exceeded", 4: "ping window too short",
            8: "action trace exhausted", 16: "scheduler uid overflow",
            32: "order of simultaneous events not determined (parallel engine)"}


class PrbConfig(C.Structure):
    _fields_ = [
        ("num_envs", C.c_int32), ("n_nodes", C.c_int32), ("n_overlay", C.c_int32),
        ("phys_adj", C.POINTER(C.c_uint8)), ("overlay_adj", C.POINTER(C.c_uint8)),
        ("map_overlay", C.POINTER(C.c_int32)), ("traffic_bps", C.POINTER(C.c_uint64)),
        ("route_out", C.POINTER(C.c_int32)), ("opt_rejected", C.POINTER(C.c_double)),
        ("link_cap_bps", C.c_uint64), ("link_delay_ns", C.c_int64),
        ("packet_size", C.c_uint32), ("max_out_buffer_size", C.c_uint32),
        ("sim_time_ns", C.c_int64), ("ping_interval_ns", C.c_int64),
        ("moving_average_size", C.c_uint32), ("ping_as_obs", C.c_int32),
        ("activate_underlay_traffic", C.c_int32), ("train", C.c_int32),
        ("signaling_type", C.c_int32), ("opt_agent", C.c_int32),
        ("loss_penalty", C.c_double), ("seed", C.c_uint64), ("env_id_base", C.c_uint64),
        ("pool_capacity", C.c_int32), ("lanes_per_env", C.c_int32),
        ("signaling", C.c_int32), ("big_signaling_size", C.c_uint32), ("sync_step", C.c_double),
    ]


class PrbEvent(C.Structure):
    _fields_ = [("t_ns", C.c_int64), ("kind", C.c_int32), ("node", C.c_int32), ("link", C.c_int32),
                ("pad_", C.c_int32), ("uid", C.c_int64), ("size", C.c_int32), ("type", C.c_int32),
                ("src", C.c_int32), ("dst", C.c_int32), ("last_hop", C.c_int32), ("next_hop", C.c_int32),
                ("valable", C.c_int32), ("qbytes", C.c_int32)]


class PrbStepRec(C.Structure):
    _fields_ = [("t_ns", C.c_int64), ("uid", C.c_int64), ("hop_time", C.c_double),
                ("node", C.c_int32), ("overlay", C.c_int32), ("ptype", C.c_int32), ("done", C.c_int32),
                ("action", C.c_int32), ("need_action", C.c_int32), ("nobs", C.c_int32), ("is_new", C.c_int32),
                ("prev_node", C.c_int32), ("n_lost", C.c_int32), ("e2e_delay", C.c_float), ("valid", C.c_int32),
                ("obs", C.c_int32 * (1 + PRB_MAX_K)), ("src", C.c_int32), ("lost_uid", C.c_int64 * PRB_MAX_LOST)]


class PrbCounters(C.Structure):
    _fields_ = [(n, C.c_int64) for n in
                ("overlay_injected", "overlay_arrived", "overlay_lost", "underlay_injected",
                 "underlay_arrived", "underlay_lost", "bytes_data", "bytes_signaling",
                 "n_e2e", "n_cost", "n_underlay_e2e", "n_underlay_cost")] + \
               [(n, C.c_float) for n in
                ("sum_e2e", "sum_cost", "sum_underlay_e2e", "sum_underlay_cost", "avg_e2e", "avg_cost",
                 "avg_underlay_e2e", "avg_underlay_cost", "signaling_overhead")] + \
               [("errors", C.c_int32)] + \
               [(n, C.c_int64) for n in
                ("packet_hops", "data_hops", "env_steps", "decisions", "events", "vanished", "end_ns")]


EVENT_DTYPE = np.dtype(PrbEvent)
STEP_DTYPE = np.dtype(PrbStepRec)
COUNTERS_DTYPE = np.dtype(PrbCounters)
assert C.sizeof(PrbEvent) == 64 and C.sizeof(PrbStepRec) == 208 and C.sizeof(PrbCounters) == 192


def ptr(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def fill_config(cfg_struct, keep, p):
    """Fill a PrbConfig from a dict of arrays/scalars named like the struct fields; `keep` collects the
    numpy arrays that must outlive the struct."""
    n, m = int(p["n_nodes"]), int(p["n_overlay"])

    def arr(key, dt, ct, size, optional=False):
        v = p.get(key)
        if v is None:
            if optional:
                return None
            raise KeyError(key)
        a = np.ascontiguousarray(np.asarray(v), dtype=dt)
        if a.size != size:
            raise ValueError(f"{key}: expected {size} elements, got {a.shape}")
        keep.append(a)
        return ptr(a, ct)

    cfg_struct.num_envs = int(p.get("num_envs", 1))
    cfg_struct.n_nodes, cfg_struct.n_overlay = n, m
    cfg_struct.phys_adj = arr("phys_adj", np.uint8, C.c_uint8, n * n)
    cfg_struct.overlay_adj = arr("overlay_adj", np.uint8, C.c_uint8, m * m)
    cfg_struct.map_overlay = arr("map_overlay", np.int32, C.c_int32, n)
    cfg_struct.traffic_bps = arr("traffic_bps", np.uint64, C.c_uint64, n * n)
    r = arr("route_out", np.int32, C.c_int32, n * n, optional=True)
    if r is not None:
        cfg_struct.route_out = r
    r = arr("opt_rejected", np.float64, C.c_double, n * n, optional=True)
    if r is not None:
        cfg_struct.opt_rejected = r
    for k in ("link_cap_bps", "link_delay_ns", "packet_size", "max_out_buffer_size", "sim_time_ns",
              "ping_interval_ns", "moving_average_size", "ping_as_obs", "activate_underlay_traffic", "train",
              "opt_agent", "seed", "env_id_base", "pool_capacity", "lanes_per_env"):
        setattr(cfg_struct, k, int(p.get(k, 0)))
    stype = p.get("signaling_type", 0)
    cfg_struct.signaling_type = SIGNALING[stype] if isinstance(stype, str) else int(stype)
    cfg_struct.loss_penalty = float(p.get("loss_penalty", 0.0))
    cfg_struct.signaling = int(p.get("signaling", 0))
    cfg_struct.big_signaling_size = int(p.get("big_signaling_size", 0))
    cfg_struct.sync_step = float(p.get("sync_step", 0.0))
    return cfg_struct


_lib = None


def lib():
    """libprisma_b200.so, loaded once.  Raises if it has not been built — there is no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: build the CUDA library first (python -m prisma_b200.build); "
                           "prisma_b200 has no CPU or pure-Python execution path")
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, u64 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64
    sig = {
        "prb_last_error": (C.c_char_p, []),
        "prb_version": (C.c_char_p, []),
        "prb_seconds_to_ns": (i64, [C.c_double]),
        "prb_ns_to_seconds": (C.c_double, [i64]),
        "prb_tx_time_ns": (i64, [C.c_uint32, u64]),
        "prb_build_links": (i32, [i32, C.POINTER(C.c_uint8), C.POINTER(i32), C.POINTER(i32)]),
        "prb_global_routes": (i32, [i32, C.POINTER(C.c_uint8), C.POINTER(i32)]),
        "prb_create": (C.c_int, [C.POINTER(PrbConfig), C.c_int, vp, C.POINTER(vp)]),
        "prb_destroy": (C.c_int, [vp]),
        "prb_reset": (C.c_int, [vp]),
        "prb_num_links": (C.c_int, [vp]),
        "prb_state_bytes": (C.c_int, [vp, C.POINTER(i64), C.POINTER(i64)]),
        "prb_set_inject_trace": (C.c_int, [vp, i64, C.POINTER(i64), C.POINTER(i32), C.POINTER(i32), C.POINTER(i64),
                                           C.POINTER(C.c_uint8), C.POINTER(C.c_uint8)]),
        "prb_set_inject_trace_typed": (C.c_int, [vp, i64, C.POINTER(i64), C.POINTER(i32), C.POINTER(i32), C.POINTER(i64),
                                                 C.POINTER(C.c_uint8), C.POINTER(C.c_uint8), C.POINTER(C.c_uint8)]),
        "prb_set_policy": (C.c_int, [vp, C.c_int, C.POINTER(i32), i64, u64]),
        "prb_enable_log": (C.c_int, [vp, i64, i64]),
        "prb_read_events": (C.c_int, [vp, i32, vp, i64, C.POINTER(i64)]),
        "prb_read_steps": (C.c_int, [vp, i32, vp, i64, C.POINTER(i64)]),
        "prb_run": (C.c_int, [vp, i64]),
        "prb_step": (C.c_int, [vp, vp, vp]),
        "prb_step_host": (C.c_int, [vp, vp, vp]),
        "prb_counters_host": (C.c_int, [vp, vp]),
        "prb_counters_dev": (C.c_int, [vp, vp]),
        "prb_queue_bytes": (C.c_int, [vp, i32, C.POINTER(i32)]),
        "prb_sync": (C.c_int, [vp]),
        "prb_launch_count": (i64, [vp]),
        "prb_set_engine": (C.c_int, [vp, C.c_int, C.c_int]),
        "prb_qnet_num_floats": (i64, [vp]),
        "prb_qnet_load": (C.c_int, [vp, vp, i64]),
        "prb_qnet_act": (C.c_int, [vp, vp, vp, vp, C.c_float, u64]),
        "prb_step_qnet": (C.c_int, [vp, i32, C.c_float, u64, vp]),
        "prb_replay_enable": (C.c_int, [vp, i64, i32]),
        "prb_replay_record": (C.c_int, [vp, vp, vp]),
        "prb_replay_set_target_mode": (C.c_int, [vp, C.c_int, C.c_float]),
        "prb_replay_counts": (C.c_int, [vp, C.POINTER(i64), C.POINTER(i64)]),
        "prb_replay_read": (C.c_int, [vp, i32, vp, i64, C.POINTER(i64)]),
        "prb_replay_sample": (C.c_int, [vp, i32, i32, u64, vp, vp, vp, vp, vp]),
        "prb_replay_sample_weighted": (C.c_int, [vp, i32, i32, u64, vp, vp, vp, vp, vp, vp]),
        "prb_replay_enable_prioritized": (C.c_int, [vp, C.c_float]),
        "prb_replay_set_gradient_steps": (C.c_int, [vp, C.POINTER(i32)]),
        "prb_replay_read_priorities": (C.c_int, [vp, i32, vp, i64, C.POINTER(i32)]),
        "prb_engine": (C.c_int, [vp]),
        "prb_handovers": (C.c_int, [vp]),
        "prb_handover_stats": (C.c_int, [vp, C.POINTER(i64), C.POINTER(C.c_double)]),
        "prb_queue_bytes_noalloc": (C.c_int, [vp, i32, C.POINTER(i32), i32]),
        "prb_qnet_set_mode": (C.c_int, [vp, C.c_int]),
        "prb_qnet_set_arch": (C.c_int, [vp, C.c_int]),
        "prb_replay_sample_all": (C.c_int, [vp, i32, u64, i64, vp, vp, vp, vp, vp, vp, vp]),
        "prb_learner_init": (C.c_int, [vp, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int]),
        "prb_learner_sync_targets": (C.c_int, [vp]),
        "prb_learner_grad": (C.c_int, [vp, i32, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
        "prb_learner_apply": (C.c_int, [vp, vp, i32]),
        "prb_learner_read": (C.c_int, [vp, i32, vp, i64]),
        "prb_round_stats": (C.c_int, [vp, C.POINTER(i64), C.POINTER(i64)]),
    }
    for name, (res, args) in sig.items():
        f = getattr(L, name)   # raises AttributeError if the library does not export a declared symbol
        f.restype = res
        f.argtypes = args
    _lib = L
    return L


EXPORTS = ["prb_last_error", "prb_version", "prb_seconds_to_ns", "prb_ns_to_seconds", "prb_tx_time_ns",
           "prb_build_links", "prb_global_routes", "prb_create", "prb_destroy", "prb_reset", "prb_num_links",
           "prb_state_bytes", "prb_set_inject_trace", "prb_set_inject_trace_typed", "prb_set_policy", "prb_enable_log", "prb_read_events",
           "prb_read_steps", "prb_run", "prb_step", "prb_step_host", "prb_counters_host", "prb_counters_dev",
           "prb_queue_bytes", "prb_sync", "prb_launch_count", "prb_set_engine", "prb_engine", "prb_round_stats", "prb_handovers", "prb_handover_stats", "prb_queue_bytes_noalloc", "prb_qnet_set_mode", "prb_qnet_set_arch", "prb_replay_sample_all", "prb_learner_init", "prb_learner_sync_targets",
           "prb_learner_grad", "prb_learner_apply", "prb_learner_read",
           "prb_qnet_num_floats", "prb_qnet_load", "prb_qnet_act", "prb_step_qnet",
           "prb_replay_enable", "prb_replay_set_target_mode", "prb_replay_record", "prb_replay_counts", "prb_replay_read", "prb_replay_sample",
           "prb_replay_sample_weighted", "prb_replay_enable_prioritized", "prb_replay_set_gradient_steps",
           "prb_replay_read_priorities"]


class PrbError(RuntimeError):
    pass


def check(rc):
    if rc != 0:
        raise PrbError("prisma_b200: %s (status %d)" % (lib().prb_last_error().decode(), rc))
