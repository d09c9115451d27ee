"""Loads tests/golden/*.npz (outputs of the reference's ns-3 program, see tests/golden/make_golden.py)."""
import ast
import glob
import os

import numpy as np

from tests import common

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
COUNTERS = ("overlay_injected", "overlay_arrived", "overlay_lost", "underlay_injected", "underlay_arrived",
            "underlay_lost", "env_steps", "decisions", "n_e2e", "n_cost")
FLOATS = ("avg_e2e", "avg_cost", "avg_underlay_e2e", "avg_underlay_cost", "signaling_overhead")


def names(parallel=False):
    """parallel: only the fixtures the node-parallel engine takes (NN-weight segment sources run on the
    event-serial engine)."""
    n = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(HERE, "*.npz")))
    return [x for x in n if not (parallel and "bigsig" in x)]


class Golden:
    def __init__(self, name):
        z = np.load(os.path.join(HERE, name + ".npz"), allow_pickle=False)
        self.name = name
        self.topology = str(z["topology"])
        self.overrides = ast.literal_eval(str(z["overrides"]))
        self.policy = str(z["policy"])
        self.policy_seed = int(z["policy_seed"])
        self.table = z["table"]
        self.route = z["route"]
        self.events = z["events"]
        self.dec = {k[4:]: z[k] for k in z.files if k.startswith("dec_")}
        self.counters = dict(zip(COUNTERS, (int(x) for x in z["counters"])))
        self.float_bits = dict(zip(FLOATS, (int(x) for x in z["float_bits"])))
        self.cfg = common.env_config(self.topology, **self.overrides)
        self.arrays = self.cfg.load()

    @property
    def injections(self):
        return self.events[self.events["kind"] == 0]
