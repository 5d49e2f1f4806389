"""Builds libpillowfight_b200.so in-tree with nvcc for sm_100a.

    python libpillowfight_b200/build.py [--force] [--verbose]

(run it as a script, or load it by path as __graft_entry__.load_build_module
does: importing the package binds the .so and refuses a missing or stale one)

One shared library, CUDA runtime linked statically, no Python or torch
dependency: the C-ABI in include/pillowfight_b200.h is all it exports.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SO = os.path.join(HERE, "libpillowfight_b200.so")
SOURCES = ["pf_capi.cu", "pf_gaussian.cu", "pf_edges.cu", "pf_ace.cu", "pf_matrix.cu"]
HEADERS = [os.path.join(CSRC, "pf_common.cuh"),
           os.path.join(HERE, "..", "include", "pillowfight_b200.h")]


def nvcc_command(verbose=False):
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc, "-O3", "-std=c++17",
           "-gencode", "arch=compute_100a,code=sm_100a",
           "-lineinfo",
           "-Xcompiler", "-fPIC,-ffp-contract=off",
           "-cudart", "static", "-shared", "-o", SO]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    cmd += os.environ.get("PF_B200_NVCC_EXTRA", "").split()     # experiments, e.g. -DPF_ACE_INPLACE=1
    cmd += [os.path.join(CSRC, s) for s in SOURCES]
    return cmd


def is_stale():
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    deps = [os.path.join(CSRC, s) for s in SOURCES] + HEADERS
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if force or is_stale():
        cmd = nvcc_command(verbose)
        subprocess.check_call(cmd)
    return SO


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(SO)
