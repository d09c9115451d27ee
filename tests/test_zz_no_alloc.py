"""prb_queue_bytes_noalloc (the allocation-free variant of prb_queue_bytes: its device-side staging buffer is a
file-scope static __device__ array) through the C ABI, against prb_queue_bytes, bit for bit (integers), on two calls in
a row and on both engines.  No counterpart in the reference; collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from tests import common

pytestmark = pytest.mark.gpu


def both(env, e):
    from prisma_b200 import _abi
    L = env.lib.prb_num_links(env.h)
    a = np.full(L, -1, dtype=np.int32)
    b = np.full(L + 3, -7, dtype=np.int32)
    _abi.check(env.lib.prb_queue_bytes(env.h, e, _abi.ptr(a, C.c_int32)))
    _abi.check(env.lib.prb_queue_bytes_noalloc(env.h, e, _abi.ptr(b, C.c_int32), b.size))
    assert (b[L:] == -7).all()            # nothing is written past L
    return a, b[:L]


@pytest.mark.parametrize("engine", ["parallel", "serial"])
def test_noalloc_queue_bytes_equals_queue_bytes_twice_in_a_row(engine):
    from prisma_b200 import _abi
    from prisma_b200.env import BatchedRoutingEnv
    cfg = common.env_config("abilene", num_envs=32, simTime=3.0, load_factor=1.4, pingAsObs=0, activateUnderlayTraffic=1)
    arrays = cfg.load()
    env = BatchedRoutingEnv(cfg, arrays=arrays)
    env.set_engine(engine)
    env.set_policy("table", common.sp_table(arrays))
    env.run(until_s=1.5)
    assert env.handovers() == 0
    a0, b0 = both(env, 3)
    a1, b1 = both(env, 17)                # second call: the shared array still holds env 3's values until it is cleared
    assert np.array_equal(a0, b0) and np.array_equal(a1, b1)
    assert a0.sum() > 0 and not np.array_equal(a0, a1)
    env.run()
    assert env.handovers() == 0
    a2, b2 = both(env, 3)
    assert np.array_equal(a2, b2)
    # bad arguments are refused with a clear error and nothing is written
    small = np.zeros(2, dtype=np.int32)
    rc = env.lib.prb_queue_bytes_noalloc(env.h, 3, _abi.ptr(small, C.c_int32), small.size)
    assert rc != 0 and b"fewer than L" in env.lib.prb_last_error()
    rc = env.lib.prb_queue_bytes_noalloc(env.h, 99, _abi.ptr(b2.copy(), C.c_int32), 64)
    assert rc != 0
    env.close()
