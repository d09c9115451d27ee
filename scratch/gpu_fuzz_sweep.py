"""Offline sweep of tests/test_gpu_fuzz.py over many seeds (prints failures)."""
import sys
sys.path.insert(0, "/root/repo")
import pytest
from tests import test_gpu_fuzz as G
lo, hi = int(sys.argv[1]), int(sys.argv[2])
ok = skip = fail = 0
for seed in range(lo, hi):
    try:
        G.test_random_network_on_the_gpu(seed)
        ok += 1
    except pytest.skip.Exception as ex:
        skip += 1
    except Exception as ex:
        fail += 1
        print("FAIL", seed, type(ex).__name__, str(ex)[:400], flush=True)
print("ok", ok, "skip", skip, "fail", fail)
