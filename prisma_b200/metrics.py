"""Session statistics, TensorBoard scalars and model snapshots with the reference's names and folder layout
(SURVEY.md §8 f4).

* `SessionStats.from_counters`: what `Forwarder.treat_info` (prisma/source/forwarder.py:270-289) keeps in the
  `Agent.sim_*` class variables, computed from the per-environment `prb_counters` of a batch.  A batch holds E
  independent simulations of the same network, so packet counts are sums over the environments and delays /
  costs are delivery-weighted means; with E = 1 the values are the reference's.
* `StatsWriter`: the scalars of `stats_writer_train` / `stats_writer_test`
  (prisma/source/tb_logger.py:70-164), same tags, same sub-folders (`session`, `nb_arrived_pkts`, `nb_lost_pkts`,
  `nb_new_pkts`, `test_results/<model_version>`), same step conventions (`*_over_iterations`: training
  iteration; `*_over_time`: simulated microseconds; test scalars: `int (load_factor * 100)`).
* `save_all_models` / `load_model`: the folder layout of prisma/source/utils.py:14-99
  (`<root>/<path>/final/node<i>` or `episode_<n>_step_<t>/node<i>`); a node's network is one `.npz` with the
  tensors of prisma_b200/qnet.py (Keras SavedModel needs TensorFlow, which this package does not use).
"""
import dataclasses
import os
import time
from typing import Dict, List, Optional

import numpy as np

from . import qnet


@dataclasses.dataclass
class SessionStats:
    sim_avg_e2e_delay: float = -1.0
    sim_cost: float = -1.0
    sim_global_avg_e2e_delay: float = -1.0
    sim_global_cost: float = -1.0
    sim_dropped_packets: int = 0
    sim_delivered_packets: int = 0
    sim_injected_packets: int = 0
    sim_buffered_packets: int = 0
    sim_global_dropped_packets: int = 0
    sim_global_delivered_packets: int = 0
    sim_global_injected_packets: int = 0
    sim_global_buffered_packets: int = 0
    sim_signaling_overhead: float = 0.0
    total_hops: int = 0
    env_steps: int = 0
    curr_time: float = 0.0

    @classmethod
    def from_counters(cls, c: np.ndarray) -> "SessionStats":
        """c: prb_counters records of the batch (BatchedRoutingEnv.counters ())."""
        i = lambda k: int(np.asarray(c[k], dtype=np.int64).sum())
        f = lambda k: float(np.asarray(c[k], dtype=np.float64).sum())
        s = cls()
        s.sim_dropped_packets, s.sim_delivered_packets, s.sim_injected_packets = (
            i("overlay_lost"), i("overlay_arrived"), i("overlay_injected"))
        s.sim_buffered_packets = s.sim_injected_packets - s.sim_delivered_packets - s.sim_dropped_packets
        ul, ua, ui = i("underlay_lost"), i("underlay_arrived"), i("underlay_injected")
        # forwarder.py:280-283: the "global" counters are underlay + overlay
        s.sim_global_dropped_packets = ul + s.sim_dropped_packets
        s.sim_global_delivered_packets = ua + s.sim_delivered_packets
        s.sim_global_injected_packets = ui + s.sim_injected_packets
        s.sim_global_buffered_packets = (ui - ua - ul) + s.sim_buffered_packets
        ne, nc = i("n_e2e"), i("n_cost")
        s.sim_avg_e2e_delay = f("sum_e2e") / ne if ne else -1.0
        s.sim_cost = f("sum_cost") / nc if nc else -1.0
        nue, nuc = i("n_underlay_e2e"), i("n_underlay_cost")
        # forwarder.py:285-288: delivery-weighted mean of the underlay and overlay averages
        s.sim_global_avg_e2e_delay = (f("sum_underlay_e2e") + f("sum_e2e")) / (nue + ne) if nue + ne else -1.0
        s.sim_global_cost = (f("sum_underlay_cost") + f("sum_cost")) / (nuc + nc) if nuc + nc else -1.0
        bd, bs = i("bytes_data"), i("bytes_signaling")
        s.sim_signaling_overhead = bs / bd if bd else 0.0     # compute-stats-v2.cc:getSignalingOverhead
        s.total_hops = i("data_hops")
        s.env_steps = i("env_steps")
        s.curr_time = float(np.asarray(c["end_ns"]).max()) * 1e-9
        return s

    def results(self) -> Dict[str, float]:
        """The dictionary prisma/main.py prints at the end of a session."""
        inj, ginj = self.sim_injected_packets, self.sim_global_injected_packets
        return {"total_data_pkts": inj, "total_lost_pkts": self.sim_dropped_packets,
                "total_arrived_pkts": self.sim_delivered_packets, "loss_ratio": self.sim_dropped_packets / inj if inj else -1,
                "avg_e2e_delay": self.sim_avg_e2e_delay, "avg_cost": self.sim_cost,
                "avg_hops": self.total_hops / self.sim_delivered_packets if self.sim_delivered_packets else -1,
                "global_loss_ratio": self.sim_global_dropped_packets / ginj if ginj else -1,
                "global_avg_e2e_delay": self.sim_global_avg_e2e_delay, "global_avg_cost": self.sim_global_cost,
                "signaling_overhead": self.sim_signaling_overhead}


class StatsWriter:
    """TensorBoard event files under `<logs_parent_folder>/<session_name>/` with the reference's sub-folders."""

    def __init__(self, logs_parent_folder: str, session_name: str):
        from torch.utils.tensorboard import SummaryWriter
        self._SW = SummaryWriter
        self.root = os.path.join(logs_parent_folder, session_name)
        self._w = {}
        self.start_time = time.time()

    def _writer(self, sub: str):
        if sub not in self._w:
            self._w[sub] = self._SW(log_dir=os.path.join(self.root, sub))
        return self._w[sub]

    def write_train(self, s: SessionStats, iteration: int, base_time: float = 0.0, total_e2e_delay: float = 0.0,
                    total_rewards_with_loss: float = 0.0, ma_hops: Optional[float] = None,
                    ma_delay: Optional[float] = None, td_errors: Optional[List[Optional[float]]] = None,
                    epsilon: Optional[float] = None, replay_lengths: Optional[List[int]] = None):
        """stats_writer_train (tb_logger.py:70-143) + the per-node training scalars of trainer.py:90-99."""
        t_us = int((base_time + s.curr_time) * 1e6)
        loss_ratio = s.sim_dropped_packets / s.sim_injected_packets if s.sim_injected_packets > 0 else -1
        delivered = s.sim_delivered_packets > 0
        avg_delay = s.sim_avg_e2e_delay if delivered else -1
        avg_cost = s.sim_cost if delivered else -1
        avg_hops = s.total_hops / s.sim_delivered_packets if delivered else -1
        w = self._writer("session")
        for tag, v in (("total_e2e_delay", total_e2e_delay), ("total_rewards_with_loss", total_rewards_with_loss),
                       ("loss_ratio", loss_ratio), ("total_hops", s.total_hops), ("avg_hops", avg_hops),
                       ("ma_avg_hops", avg_hops if ma_hops is None else ma_hops),
                       ("nb_buffered_pkts", s.sim_buffered_packets), ("avg_cost", avg_cost), ("avg_delay", avg_delay),
                       ("ma_delays", avg_delay if ma_delay is None else ma_delay)):
            w.add_scalar(tag + "_over_iterations", v, iteration)
            w.add_scalar(tag + "_over_time", v, t_us)
        w.add_scalar("signalling ratio", s.sim_signaling_overhead, t_us)
        sim_t = base_time + s.curr_time
        if sim_t > 0:
            w.add_scalar("sim_second_per_real_seconds", (time.time() - self.start_time) / sim_t, t_us)
        for sub, v in (("nb_arrived_pkts", s.sim_delivered_packets), ("nb_lost_pkts", s.sim_dropped_packets),
                       ("nb_new_pkts", s.sim_injected_packets)):
            ww = self._writer(sub)
            ww.add_scalar("pkts_over_iterations", v, iteration)
            ww.add_scalar("pkts_over_time", v, t_us)
        if td_errors is not None:
            for i, v in enumerate(td_errors):
                if v is not None:
                    self._writer("td_error/node_%d" % i).add_scalar("MSE_loss_over_time", v, t_us)
        if epsilon is not None:
            w.add_scalar("exploaration_value_over_time", epsilon, t_us)
        if replay_lengths is not None:
            for i, v in enumerate(replay_lengths):
                self._writer("replay_buffer_length/node_%d" % i).add_scalar("replay_buffer_length_over_time", v, t_us)

    def write_test(self, s: SessionStats, load_factor: float, model_version: str = "final"):
        """stats_writer_test (tb_logger.py:145-164): one point per load factor."""
        w = self._writer(os.path.join("test_results", model_version))
        step = int(load_factor * 100)
        ginj, inj = s.sim_global_injected_packets, s.sim_injected_packets
        for tag, v in (("test_global_injected_pkts", ginj), ("test_overlay_injected_pkts", inj),
                       ("test_global_lost_pkts", s.sim_global_dropped_packets),
                       ("test_overlay_lost_pkts", s.sim_dropped_packets),
                       ("test_global_arrived_pkts", s.sim_global_delivered_packets),
                       ("test_overlay_arrived_pkts", s.sim_delivered_packets),
                       # the reference swaps the two delays (tb_logger.py:158-159): kept, the tags are the interface
                       ("test_global_e2e_delay", s.sim_avg_e2e_delay), ("test_overlay_e2e_delay", s.sim_global_avg_e2e_delay),
                       ("test_global_loss_rate", s.sim_global_dropped_packets / ginj if ginj else -1),
                       ("test_overlay_loss_rate", s.sim_dropped_packets / inj if inj else -1),
                       ("test_global_cost", s.sim_global_cost), ("test_overlay_cost", s.sim_cost)):
            w.add_scalar(tag, v, step)

    def close(self):
        for w in self._w.values():
            w.flush()
            w.close()
        self._w = {}


class LinearSchedule:
    """prisma/source/utils.py:101-124."""

    def __init__(self, schedule_timesteps: int, final_p: float, initial_p: float = 1.0):
        self.schedule_timesteps, self.final_p, self.initial_p = schedule_timesteps, final_p, initial_p

    def value(self, t):
        fraction = min(float(t) / self.schedule_timesteps, 1.0)
        return self.initial_p + fraction * (self.final_p - self.initial_p)


# ---- model snapshots (utils.py:14-99) ----

def _folder(root: str, path: str, t: int, num_episodes: int, snapshot: bool) -> str:
    return os.path.join(root, path.rstrip("/"), "episode_%d_step_%d" % (num_episodes, t) if snapshot else "final")


def save_all_models(weights: np.ndarray, m: int, k_max: int, path: str, t: int = 0, num_episodes: int = 0,
                    root: str = "saved_models/", snapshot: bool = False) -> str:
    """weights: the flat [M * P] array of DQNLearner.weights (); one file per overlay node."""
    folder = _folder(root, path, t, num_episodes, snapshot)
    os.makedirs(folder, exist_ok=True)
    p = qnet.floats_per_node(m, k_max)
    w = np.asarray(weights, dtype=np.float32).reshape(m, p)
    for i in range(m):
        out, o = {}, 0
        for name, shape in qnet.shapes(m, k_max).items():
            n = int(np.prod(shape))
            out[name] = w[i, o:o + n].reshape(shape)
            o += n
        np.savez(os.path.join(folder, "node%d" % i), **out)
    return folder


def load_model(path: str, node_index: int = -1) -> List[Optional[Dict[str, np.ndarray]]]:
    """The networks stored in `path` (a `final` or `episode_*` folder), indexed by node; with node_index >= 0 only
    that entry is filled (utils.py:69-99)."""
    items = [f for f in os.listdir(path) if f.startswith("node")]
    nets: List[Optional[Dict[str, np.ndarray]]] = [None] * len(items)
    for item in items:
        index = int(os.path.splitext(item)[0][4:])
        if node_index >= 0 and node_index != index:
            continue
        with np.load(os.path.join(path, item)) as z:
            nets[index] = {k: z[k] for k in qnet.LAYERS}
    return nets
