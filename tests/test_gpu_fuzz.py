"""The random networks of tests/test_fuzz_engines.py through the C ABI on the GPU: the event-serial engine's ordered
log, Gym steps, counters and queues against the restatement; the node-parallel engine's counters and queues — hand-overs
to the serial engine included: an environment whose tie could not be decided is replayed and must come out exact."""
import numpy as np
import pytest

from oracle import oracle
from tests import common
from tests.test_fuzz_engines import random_case

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("seed", [0, 3, 7, 11, 38, 161, 64, 99])
def test_random_network_on_the_gpu(seed):
    from prisma_b200.env import BatchedRoutingEnv
    cfg, env_id = random_case(seed)
    cfg.num_envs = 6
    arrays = cfg.load()
    table = common.sp_table(arrays)
    e = 4
    # event-serial engine, random agent (env e draws seed + e)
    osim = common.oracle_run(cfg, arrays, oracle, policy="random", table=table, seed=5 + e, env_id=env_id + e)
    env = BatchedRoutingEnv(cfg, arrays=arrays, env_id_base=env_id)
    env.enable_log(2 * len(osim.events()) + 1000, 2 * len(osim.decisions()) + 1000)
    env.reset()
    env.set_policy("random", table, seed=5)
    env.run()
    cs = env.counters()
    env.check_errors(cs)
    assert common.first_event_mismatch(osim.events(), env.events(e)) is None
    common.assert_steps_equal(osim.decisions(), env.steps(e))
    common.assert_counters_equal(osim.counters(), cs[e])
    assert np.array_equal(osim.queue_bytes(), env.queue_bytes(e))
    env.close()
    # node-parallel engine, next-hop table; hand-overs are resolved before the read-out
    env = BatchedRoutingEnv(cfg, arrays=arrays, env_id_base=env_id)
    env.set_policy("table", table)
    try:
        env.set_engine("parallel")
        env.run()
    except Exception as ex:
        env.close()
        pytest.skip("parallel engine refuses: %s" % ex)
    cs = env.counters()
    env.check_errors(cs)
    for k in range(cfg.num_envs):
        osim = common.oracle_run(cfg, arrays, oracle, policy="table", table=table, env_id=env_id + k)
        common.assert_counters_equal(osim.counters(), cs[k])
        assert np.array_equal(osim.queue_bytes(), env.queue_bytes(k))
    env.close()
