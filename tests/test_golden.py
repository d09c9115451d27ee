"""CPU restatement vs the committed outputs of the reference (tests/golden/*.npz): bit-exact event
logs, Gym steps, counters and float32 statistics.  Runs without /root/reference."""
import struct

import numpy as np
import pytest

from oracle import oracle
from tests import common, golden_util


@pytest.mark.parametrize("name", golden_util.names())
def test_oracle_reproduces_golden(name):
    g = golden_util.Golden(name)
    assert np.array_equal(oracle.global_routes(g.arrays.phys_adj), g.route)
    sim = oracle.OracleSim(common.oracle_params(g.cfg, g.arrays, oracle))
    inj = g.injections
    sim.set_inject_trace(inj["t_ns"], inj["src"], inj["dst"], inj["uid"], inj["valable"], inj["rejected"], inj["type"])
    sim.set_policy(g.policy, g.table, seed=g.policy_seed)
    sim.run()
    oe = sim.events()
    assert common.first_event_mismatch(g.events, oe) is None
    od = sim.decisions()
    assert len(od) == len(g.dec["t"])
    assert np.array_equal(od["t_ns"], g.dec["t"])
    assert np.array_equal(od["node"], g.dec["node"])
    assert np.array_equal(od["ptype"], g.dec["ptype"])
    assert np.array_equal(od["done"], g.dec["done"])
    assert np.array_equal(od["action"], g.dec["action"])
    data = g.dec["ptype"] == 0
    assert np.array_equal(od["uid"][data], g.dec["uid"][data])
    k = g.dec["obs"].shape[1]
    mask = np.arange(k)[None, :] < g.dec["nobs"][:, None]
    assert np.array_equal(np.where(mask, od["obs"][:, :k], -1), g.dec["obs"])
    if "big" in g.dec:
        common.assert_big_signaling_fields(g.dec["big"], od["uid"], od["prev_node"], g.arrays.map_overlay)
    c = sim.counters()
    for key, v in g.counters.items():
        assert c[key] == v, key
    for key, v in g.float_bits.items():
        assert struct.unpack("I", struct.pack("f", c[key]))[0] == v, key
