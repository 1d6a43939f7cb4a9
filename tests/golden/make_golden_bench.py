"""Golden hashes of the 8 stereo pairs bench.py cycles (BASELINE.json configs[2] / configs[4]: 1242x375, D=128), produced by
the REFERENCE ITSELF (oracle/_ref/libsgm_ref.so = the unmodified /root/reference CPU sources, one thread because its speckle
filter races under OpenMP).  bench.py hashes the first 8 results of its frame stream against these outside the timed region
("results of the first 8 frames parity-checked").  Run in the build container:

    python tests/golden/make_golden_bench.py      # -> tests/golden/bench_pairs.json
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import refbind, restate  # noqa: E402
from stereo_matching_b200 import synth  # noqa: E402

H, W, D, N_PAIRS = 375, 1242, 128, 8


def bench_pair(k):
    """The generator bench.py uses for pair k (kept in one place: bench.py imports it from here)."""
    return synth.random_dot_pair(H, W, seed=1234 + k, bg=12, fg=40, noise=4 if k else 0)[:2]


def main():
    refbind.set_threads(1)
    recs = []
    for k in range(N_PAIRS):
        L, R = bench_pair(k)
        out = refbind.process("sgm", L, R, 1, D)
        recs.append(dict(pair=k, seed=1234 + k, noise=4 if k else 0, left=restate.fnv1a64(L), right=restate.fnv1a64(R),
                         f_final=restate.fnv1a64(out["fdisp"]), disp=restate.fnv1a64(out["disp"]),
                         valid=int((out["fdisp"] <= D - 1).sum())))
        print("golden (bench pair):", recs[-1], flush=True)
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "bench_pairs.json"), "w") as f:
        json.dump(dict(generator="tests/golden/make_golden_bench.py", source="oracle/_ref/libsgm_ref.so (reference, 1 thread)",
                       hash="FNV-1a-64 over raw little-endian float32 / uint8 buffers", h=H, w=W, d=D, pairs=recs), f, indent=1)


if __name__ == "__main__":
    main()
