"""The C-ABI library: loads on a CPU-only box, exports every symbol include/prisma_b200.h declares, struct
layouts agree with the ctypes mirror, and nothing runs without a CUDA device (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from prisma_b200 import _abi, build
from tests import common

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    build.build()
    return _abi.lib()


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "prisma_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(prb_[a-z0-9_]+)\s*\(", txt)))


def test_exports_every_declared_symbol(lib):
    names = declared_symbols()
    assert len(names) >= 20
    raw = C.CDLL(_abi.LIB_PATH)
    for n in names:
        assert hasattr(raw, n), n
    assert sorted(_abi.EXPORTS) == names


def test_struct_layouts(lib):
    hdr = open(os.path.join(ROOT, "include", "prisma_b200.h")).read()
    assert "PRB_MAX_K 16" in hdr and "PRB_MAX_LOST 8" in hdr
    assert C.sizeof(_abi.PrbEvent) == 64
    assert C.sizeof(_abi.PrbStepRec) == 208
    assert C.sizeof(_abi.PrbCounters) == 192
    assert _abi.PrbStepRec.obs.offset == 72 and _abi.PrbStepRec.lost_uid.offset == 144
    assert _abi.PrbStepRec.e2e_delay.offset == 64 and _abi.PrbStepRec.hop_time.offset == 16


def test_host_arithmetic_entry_points(lib):
    assert lib.prb_seconds_to_ns(0.001) == 1000000
    assert lib.prb_tx_time_ns(542, 500000) == 8671999       # not 8672000: Q64.64 rounding of ns-3
    assert lib.prb_seconds_to_ns(float(np.float32(0.2))) == 200000002
    arrays = common.env_config("abilene").load()
    n = arrays.n_nodes
    fr = np.zeros(64, dtype=np.int32)
    to = np.zeros(64, dtype=np.int32)
    nl = lib.prb_build_links(n, _abi.ptr(arrays.phys_adj, C.c_uint8), _abi.ptr(fr, C.c_int32), _abi.ptr(to, C.c_int32))
    assert nl == 28
    route = np.zeros((n, n), dtype=np.int32)
    assert lib.prb_global_routes(n, _abi.ptr(arrays.phys_adj, C.c_uint8), _abi.ptr(route, C.c_int32)) == 28
    assert (np.diag(route) == -1).all() and (route[~np.eye(n, dtype=bool)] >= 0).all()


def test_no_cpu_fallback(lib):
    """Without a CUDA device the product refuses to run instead of computing on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    keep = []
    cfg = _abi.fill_config(_abi.PrbConfig(), keep, common.env_config("abilene").abi_params())
    h = C.c_void_p()
    rc = lib.prb_create(C.byref(cfg), 0, None, C.byref(h))
    assert rc == -2 and not h.value
    assert b"no CUDA device" in lib.prb_last_error()
    from prisma_b200.env import BatchedRoutingEnv
    with pytest.raises(RuntimeError):
        BatchedRoutingEnv(common.env_config("abilene"))


def test_product_does_not_import_oracle():
    for root, _, files in os.walk(os.path.join(ROOT, "prisma_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(root, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt and "prisma_oracle" not in txt, f
                assert "host_harness" not in txt or f in ("sim_core.cuh", "par_core.cuh"), f   # mentioned in a comment only
