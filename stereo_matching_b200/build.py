"""Builds libsgm_b200.so (the C-ABI engine + sm_100a kernels) in-tree with nvcc.

    python -m stereo_matching_b200.build        # or: from stereo_matching_b200.build import build; build()

nvcc cross-compiles for sm_100a without a GPU.  -fmad=false keeps every float op separately rounded
(the reference's float recurrences must be reproduced bit for bit); -lineinfo lets ncu map SASS to source.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libsgm_b200.so")
SOURCES = ["engine.cu", "cost_kernels.cu", "aggregate_kernels.cu", "sweep_kernels.cu", "post_kernels.cu", "view_kernels.cu"]
HEADERS = ["common.cuh", "kernels.h", "sgm_device.cuh", os.path.join("..", "..", "include", "sgm_b200.h")]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-fmad=false", "--cudart", "static",
    "-Xcompiler", "-fPIC,-O2,-fvisibility=hidden", "-shared",
]


def _stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not _stale():
        return LIB
    nvcc = os.environ.get("NVCC", "nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB] + [os.path.join(CSRC, s) for s in SOURCES]
    env = dict(os.environ)
    env.pop("CXX", None)  # the image exports a g++ wrapper that nvcc must not pick up
    env.pop("CC", None)
    subprocess.check_call(cmd, env=env)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
