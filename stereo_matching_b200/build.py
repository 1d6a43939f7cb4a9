"""Builds libsgm_b200.so (the C-ABI engine + sm_100a kernels) in-tree with nvcc.

    python -m stereo_matching_b200.build [--force] [-v] [--tuning]

nvcc cross-compiles for sm_100a without a GPU.  -fmad=false keeps every float op separately rounded
(the reference's float recurrences must be reproduced bit for bit); -lineinfo lets ncu map SASS to source.
The translation units are compiled in parallel (one object each) and linked into the shared library.

--tuning builds libsgm_b200_tuning.so with -DSGM_B200_TUNING: the same code plus the environment knobs used for
performance experiments and test hooks (SGM_B200_SWEEP_DEBUG, *_CFG, *_LEGACY, SGM_B200_HALO_SEQ_START, ...).  The
product library reads no environment variable.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libsgm_b200.so")
LIB_TUNING = os.path.join(HERE, "libsgm_b200_tuning.so")
OBJ_DIR = os.path.join(HERE, "build")
SOURCES = sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))
HEADERS = sorted(f for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))) + [os.path.join("..", "..", "include", "sgm_b200.h")]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-fmad=false",
    "-Xcompiler", "-fPIC,-O2,-fvisibility=hidden,-pthread",
]
LINK_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "--cudart", "static", "-shared", "-Xcompiler", "-pthread"]


def _stale(lib: str) -> bool:
    if not os.path.exists(lib):
        return True
    t = os.path.getmtime(lib)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False, tuning: bool = False) -> str:
    lib = LIB_TUNING if tuning else LIB
    if not force and not _stale(lib):
        return lib
    nvcc = os.environ.get("NVCC", "nvcc")
    env = dict(os.environ)
    env.pop("CXX", None)  # the image exports a g++ wrapper that nvcc must not pick up
    env.pop("CC", None)
    obj_dir = os.path.join(OBJ_DIR, "tuning" if tuning else "product")
    os.makedirs(obj_dir, exist_ok=True)
    extra = (["-DSGM_B200_TUNING"] if tuning else []) + (["-Xptxas", "-v"] if verbose else [])
    newest_common = max([os.path.getmtime(os.path.join(CSRC, h)) for h in HEADERS] + [os.path.getmtime(os.path.abspath(__file__))])

    def compile_one(src: str) -> str:
        obj = os.path.join(obj_dir, src.replace(".cu", ".o"))
        src_path = os.path.join(CSRC, src)
        if force or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(src_path), newest_common):
            subprocess.check_call([nvcc] + NVCC_FLAGS + extra + ["-c", "-o", obj, src_path], env=env)
        return obj

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as pool:
        objs = list(pool.map(compile_one, SOURCES))
    subprocess.check_call([nvcc] + LINK_FLAGS + ["-o", lib] + objs, env=env)
    return lib


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, tuning="--tuning" in sys.argv))
