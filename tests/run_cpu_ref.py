"""Times the compiled reference (oracle/_ref/libsgm_ref.so = unmodified /root/reference sources) or, if it did not travel,
the restatement, in a process of its own, frame by frame, so that bench.py can watch every frame with a deadline (the
reference's speckle filter races under OpenMP, src/Solver.cpp:525-553, and a frame can hang).  Helper of bench.py, not a pytest:

    python tests/run_cpu_ref.py H W D SCALE first_frame n_frames threads

prints {"kind", "cores"} first, then one line {"frame": k, "ms": t} per frame as soon as it is done (flushed).  Frame k is pair
k mod 8 of the generator bench.py uses (tests/golden/make_golden_bench.py: bench_pair), through ONE solver kept across frames
like node.cpp:49,93 keeps it."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
h, w, d, scale, first, count, threads = (int(x) for x in sys.argv[1:8])
os.environ["OMP_NUM_THREADS"] = str(threads)

from oracle import refbind  # noqa: E402
from stereo_matching_b200 import synth  # noqa: E402

N_PAIRS = 8


def pair(k):
    k %= N_PAIRS
    return synth.random_dot_pair(h, w, seed=1234 + k, bg=12, fg=40, noise=4 if k else 0)[:2]


pairs = {}
if refbind.available():
    refbind.set_threads(threads)
    session = refbind.Session("sgm", h, w, scale, d)
    print(json.dumps({"kind": "reference", "cores": threads}), flush=True)
    for f in range(first, first + count):
        if f % N_PAIRS not in pairs:
            pairs[f % N_PAIRS] = pair(f)
        L, R = pairs[f % N_PAIRS]
        print(json.dumps({"frame": f, "ms": float(session.run(L, R))}), flush=True)
    session.close()
else:
    from oracle import restate

    print(json.dumps({"kind": "port", "cores": 1}), flush=True)
    for f in range(first, first + count):
        L, R = pair(f)
        t0 = time.perf_counter()
        restate.process(L, R, scale, d, "sgm", taps=("f_final",))
        print(json.dumps({"frame": f, "ms": (time.perf_counter() - t0) * 1e3}), flush=True)
