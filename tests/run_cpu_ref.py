"""Times the compiled reference (oracle/_ref/libsgm_ref.so = unmodified /root/reference sources) or, if it did not travel,
the restatement, in a process of its own so that bench.py can bound it with a timeout (the reference's speckle filter races
under OpenMP, src/Solver.cpp:525-553).  Helper of bench.py, not a pytest:
    python tests/run_cpu_ref.py H W D SCALE iters threads   -> one JSON line {"kind", "cores", "ms": [...]}"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
h, w, d, scale, iters, threads = (int(x) for x in sys.argv[1:7])
os.environ["OMP_NUM_THREADS"] = str(threads)

from oracle import refbind  # noqa: E402
from stereo_matching_b200 import synth  # noqa: E402

L, R, _ = synth.random_dot_pair(h, w, seed=1234, bg=12, fg=40)
if refbind.available():
    refbind.set_threads(threads)
    ms = [float(x) for x in refbind.time_process("sgm", L, R, scale, d, iters)]
    kind, cores = "reference", threads
else:
    from oracle import restate

    ms = []
    for _ in range(iters):
        t0 = time.perf_counter()
        restate.process(L, R, scale, d, "sgm", taps=("f_final",))
        ms.append((time.perf_counter() - t0) * 1e3)
    kind, cores = "port", 1
print(json.dumps({"kind": kind, "cores": cores, "ms": ms}), flush=True)
