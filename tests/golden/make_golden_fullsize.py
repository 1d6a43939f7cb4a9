"""Golden hashes at FULL size for the two BASELINE.json configurations the reference itself cannot run:

  * configs[3]: 1920x1080, 8-path SGM, D=256 (the reference asserts d in {32, 64, 128}, src/Solver.cpp:10) — float mode
    (recursive cost filter) and the integer / DPX mode (cost filter off), clean and noisy pair;
  * configs[1]: SAD 9x9 block matching with WTA + LR-check at 1242x375, D=64 (the reference's BM has neither SAD nor an
    LR check; the volume is the reference's own SAD(), src/cost.cpp:4-30, in a build_dsi-shaped loop).

Source: the widened restatement oracle/sgm_oracle.c, which tests/test_oracle_cpu.py pins bit for bit against the compiled
reference at the sizes / modes the reference can run (and its SAD volume against the reference's SAD()).  The vectors are
therefore "widened restatement", not "reference" (SURVEY 8(c)).  Run anywhere (needs ~25 GB of RAM, a few minutes):

    python tests/golden/make_golden_fullsize.py            # -> tests/golden/kat_hashes_fullsize.json

Hash: FNV-1a-64 over raw little-endian buffers, like kat_hashes.json.
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import restate  # noqa: E402
from stereo_matching_b200 import synth  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
H = restate.fnv1a64

SGM_TAPS = ("census", "cost_hv", "S", "disp", "fsub", "f_lr", "f_median", "f_final")
BM_TAPS = ("cost_raw", "disp", "fsub", "f_lr", "f_median", "f_final")

CASES = [
    # configs[3]: background d=24, foreground d=200
    dict(name="hd_sgm_d256_clean", algo="sgm", h=1080, w=1920, d=256, scale=1, seed=31, bg=24, fg=200, noise=0, integer=False),
    dict(name="hd_sgm_d256_noisy", algo="sgm", h=1080, w=1920, d=256, scale=1, seed=32, bg=24, fg=200, noise=8, integer=False),
    dict(name="hd_int_d256_noisy", algo="sgm", h=1080, w=1920, d=256, scale=1, seed=32, bg=24, fg=200, noise=8, integer=True),
    # configs[1]
    dict(name="kitti_bmsad_d64_clean", algo="bm_sad", h=375, w=1242, d=64, scale=1, seed=1234, bg=12, fg=40, noise=0, integer=False),
    dict(name="kitti_bmsad_d64_noisy", algo="bm_sad", h=375, w=1242, d=64, scale=1, seed=99, bg=7, fg=33, noise=8, integer=False),
]


def run_case(c):
    L, R, _ = synth.random_dot_pair(c["h"], c["w"], seed=c["seed"], bg=c["bg"], fg=c["fg"], noise=c["noise"])
    over = dict(box_w=c["scale"], box_h=c["scale"]) if c["integer"] else {}
    taps = SGM_TAPS if c["algo"] == "sgm" else BM_TAPS
    out = restate.process(L, R, c["scale"], c["d"], c["algo"], taps=taps, **over)
    hs = dict(left=H(L), right=H(R))
    for k in taps:
        if isinstance(out[k], list):
            hs[k + "_l"], hs[k + "_r"] = H(out[k][0]), H(out[k][1])
        else:
            hs[k] = H(out[k])
    f = out["f_final"]
    rec = dict(c)
    rec["hashes"] = hs
    rec["stats"] = dict(valid=int((f <= c["d"] - 1).sum()), holes_filled=int((out["f_median"] != out["f_lr"]).sum()))
    return rec


def main():
    only = set(sys.argv[1:])
    path = os.path.join(HERE, "kat_hashes_fullsize.json")
    recs = {}
    if only and os.path.exists(path):  # regenerate a subset, keep the rest
        with open(path) as f:
            recs = {c["name"]: c for c in json.load(f)["cases"]}
    for c in CASES:
        if only and c["name"] not in only:
            continue
        t0 = time.time()
        recs[c["name"]] = run_case(c)
        print(f"golden (full size): {c['name']} {time.time() - t0:.0f} s  valid={recs[c['name']]['stats']['valid']}", flush=True)
    with open(path, "w") as f:
        json.dump(dict(generator="tests/golden/make_golden_fullsize.py",
                       source="oracle/sgm_oracle.c (widened restatement; pinned against the compiled reference in tests/test_oracle_cpu.py)",
                       hash="FNV-1a-64 over raw little-endian buffers", cases=[recs[c["name"]] for c in CASES if c["name"] in recs]), f, indent=1)


if __name__ == "__main__":
    main()
