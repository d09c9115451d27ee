"""Builds prisma_b200/libprisma_b200.so (the CUDA kernels + C ABI) in-tree for sm_100a.

    python -m prisma_b200.build [--force] [--verbose]

nvcc cross-compiles without a GPU.  `-fmad=false`: the simulator's floating-point statistics and the
synthetic-traffic generator are specified operation by operation, so the compiler must not
contract a*b+c; the explicit fma of the generator uses __fma_rn.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libprisma_b200.so")
SOURCES = [os.path.join(CSRC, "prisma_b200.cu")]
DEPS = SOURCES + [os.path.join(CSRC, "sim_core.cuh"), os.path.join(CSRC, "par_core.cuh"), os.path.join(CSRC, "learner_core.cuh"), os.path.join(CSRC, "model.h"),
                  os.path.join(HERE, "..", "include", "prisma_b200.h")]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-fmad=false",
         "-Xcompiler", "-fPIC,-O2,-ffp-contract=off", "-shared", "-cudart", "static", "--expt-relaxed-constexpr"]


def up_to_date():
    return os.path.exists(OUT) and all(os.path.getmtime(OUT) >= os.path.getmtime(d) for d in DEPS)


def build(force=False, verbose=False, defines=(), out=None):
    """defines/out: build an experimental variant next to the product library (kernel A/B tests)."""
    if out is None and not force and up_to_date():
        return OUT
    out = out or OUT
    cmd = [NVCC] + FLAGS + ["-D" + d for d in defines] + (["-Xptxas", "-v"] if verbose else []) + ["-o", out] + SOURCES
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout)
    if verbose:
        print(r.stdout)
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
