"""Builds libasvrl.so for sm_100a with nvcc, in-tree (so the .so travels to the GPU box).

    python -m distributional_rl_decision_and_control_b200.csrc.build [--force] [--verbose]
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(HERE)
SO = os.path.join(PKG, "libasvrl.so")
SOURCES = ["api.cu", "env_step.cu", "env_reset.cu", "replay.cu", "iqn.cu", "trunk.cu", "tower.cu", "optim.cu", "linear_umma.cu", "policy_umma.cu", "critic_umma.cu", "wgrad_umma.cu", "critic_bwd_umma.cu", "c51.cu", "small_ops.cu", "c51_head_umma.cu"]
HEADERS = ["common.cuh", "umma.cuh", "env_consts.cuh", "env_reset.cuh", os.path.join("..", "..", "include", "asvrl.h")]
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "--shared", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
         "--expt-relaxed-constexpr", "-cudart", "static"]


def nvcc():
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.sep not in c or os.path.exists(c)):
            return c
    raise RuntimeError("nvcc not found")


def stale():
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    deps = [os.path.join(HERE, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not (force or stale()):
        return SO
    srcs = [os.path.join(HERE, s) for s in SOURCES if os.path.exists(os.path.join(HERE, s))]
    extra = os.environ.get("ASVRL_NVCC_FLAGS", "").split()  # e.g. -DCF_TRACE for profiles/probes/k10_trace.py
    cmd = [nvcc()] + FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-o", SO] + srcs
    r = subprocess.run(cmd, cwd=HERE, capture_output=True, text=True)
    if verbose or r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed building libasvrl.so")
    return SO


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
