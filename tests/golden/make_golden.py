"""Generates tests/golden/*.json|*.npz from the REFERENCE ITSELF (oracle/_ref/libsgm_ref.so = the unmodified
/root/reference CPU sources compiled by oracle/Makefile).  Run in the build container, where /root/reference
exists:  python tests/golden/make_golden.py

Outputs
  kat_hashes.json : FNV-1a-64 hashes of every reference tap for the BASELINE.json configurations the
                    reference can run (KITTI-size 1242x375, D=64/128 SGM, D=64 BM), clean / noisy / sky-masked.
  small_sgm.npz   : full input + output arrays of a tiny SGM case (for diagnosing a mismatch).
The reference runs single-threaded here (its speckle filter races under OpenMP, src/Solver.cpp:525-553).
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import refbind, restate  # noqa: E402
from stereo_matching_b200 import synth  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
H = restate.fnv1a64


def case_inputs(c):
    L, R, _ = synth.random_dot_pair(c["h"], c["w"], seed=c["seed"], bg=c["bg"], fg=c["fg"], noise=c["noise"])
    sky = synth.sky_masks(c["h"] // c["scale"], c["w"] // c["scale"]) if c["sky"] else (None, None)
    return L, R, sky


CASES = [
    dict(name="kitti_sgm_d128_clean", algo="sgm", h=375, w=1242, d=128, scale=1, seed=1234, bg=12, fg=40, noise=0, sky=False),
    dict(name="kitti_sgm_d64_clean", algo="sgm", h=375, w=1242, d=64, scale=1, seed=1234, bg=12, fg=40, noise=0, sky=False),
    dict(name="kitti_sgm_d128_noisy_sky", algo="sgm", h=375, w=1242, d=128, scale=1, seed=4321, bg=12, fg=40, noise=8, sky=True),
    dict(name="kitti_sgm_d64_noisy", algo="sgm", h=375, w=1242, d=64, scale=1, seed=99, bg=7, fg=33, noise=8, sky=False),
    dict(name="kitti_sgm_d64_scale2", algo="sgm", h=375, w=1242, d=64, scale=2, seed=5, bg=12, fg=40, noise=4, sky=False),
    dict(name="kitti_bm_d64_clean", algo="bm", h=375, w=1242, d=64, scale=1, seed=1234, bg=12, fg=40, noise=0, sky=False),
    dict(name="kitti_bm_d64_noisy", algo="bm", h=375, w=1242, d=64, scale=1, seed=99, bg=7, fg=33, noise=8, sky=False),
    dict(name="odd_sgm_d32", algo="sgm", h=374, w=1241, d=32, scale=1, seed=3, bg=4, fg=20, noise=6, sky=False),
    dict(name="tall_sgm_d32", algo="sgm", h=300, w=200, d=32, scale=1, seed=8, bg=4, fg=20, noise=6, sky=True),
]


def run_case(c):
    L, R, (sl, sr) = case_inputs(c)
    out = refbind.process(c["algo"], L, R, c["scale"], c["d"], sl, sr, want_cost=True)
    rec = dict(c)
    hashes = dict(left=H(L), right=H(R), disp=H(out["disp"]))
    st = refbind.stages(c["algo"], L, R, c["scale"], c["d"], beta=False, sky=sl)
    hashes.update(census_l=H(st["ct_l"]), census_r=H(st["ct_r"]), cost_raw_l=H(st["cost_raw"]), cost_h_l=H(st["cost_h"]),
                  cost_hv_l=H(st["cost_hv"]))
    if c["algo"] == "sgm":
        stb = refbind.stages(c["algo"], L, R, c["scale"], c["d"], beta=True, sky=sr)
        hashes.update(cost_raw_r=H(stb["cost_raw"]), cost_hv_r=H(stb["cost_hv"]), S_r=H(out["cost_last"]),
                      disp_beta=H(out["disp_beta"]), fsub_r=H(out["fdisp_beta"]), f_final=H(out["fdisp"]))
        f = out["fdisp"]
        rec["stats"] = dict(valid=int((f <= c["d"] - 1).sum()), invalid_disp=int((out["disp"] > c["d"] - 1).sum()))
    rec["hashes"] = hashes
    return rec


def main():
    refbind.set_threads(1)
    recs = []
    for c in CASES:
        print("golden:", c["name"], flush=True)
        recs.append(run_case(c))
    with open(os.path.join(HERE, "kat_hashes.json"), "w") as f:
        json.dump(dict(generator="tests/golden/make_golden.py", source="oracle/_ref/libsgm_ref.so (reference, 1 thread)",
                       hash="FNV-1a-64 over raw little-endian buffers", cases=recs), f, indent=1)

    # tiny full-array fixture
    h, w, d = 40, 64, 32
    L, R, _ = synth.random_dot_pair(h, w, seed=11, bg=3, fg=9, noise=6)
    out = refbind.process("sgm", L, R, 1, d, want_cost=True)
    st = refbind.stages("sgm", L, R, 1, d, beta=False)
    stb = refbind.stages("sgm", L, R, 1, d, beta=True)
    np.savez_compressed(os.path.join(HERE, "small_sgm.npz"), left=L, right=R, d=d, census_l=st["ct_l"], census_r=st["ct_r"],
                        cost_hv_l=st["cost_hv"], cost_hv_r=stb["cost_hv"], S_r=out["cost_last"], disp=out["disp"],
                        disp_beta=out["disp_beta"], fsub_r=out["fdisp_beta"], f_final=out["fdisp"])
    bm = refbind.process("bm", L, R, 1, d)
    np.savez_compressed(os.path.join(HERE, "small_bm.npz"), left=L, right=R, d=d, disp=bm["disp"])


if __name__ == "__main__":
    main()
