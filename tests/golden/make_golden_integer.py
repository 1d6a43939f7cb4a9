"""Golden hashes for the engine's integer / DPX mode (SGM_FLAG_INTEGER) at BASELINE.json's KITTI size.

The reference cannot run this configuration itself (its cost-filter windows are compile-time constants,
inc/Solver.h:12-13), so the vectors come from the restatement (oracle/sgm_oracle.c, pinned bit for bit against the compiled
reference for the default windows) with the cost filter window at 1x1, after checking that its 16-bit integer form
(orc_aggregate_u16) gives the same path sums.  Run anywhere:  python tests/golden/make_golden_integer.py
Output: tests/golden/kat_hashes_integer.json (FNV-1a-64 over raw little-endian buffers, like kat_hashes.json).
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

CASES = [
    dict(name="kitti_int_d128_noisy", h=375, w=1242, d=128, scale=1, seed=1234, bg=12, fg=40, noise=8),
    dict(name="kitti_int_d64_clean", h=375, w=1242, d=64, scale=1, seed=1234, bg=12, fg=40, noise=0),
    dict(name="kitti_int_d64_scale2", h=375, w=1242, d=64, scale=2, seed=5, bg=12, fg=40, noise=4),
]


def run_case(c):
    L, R, _ = synth.random_dot_pair(c["h"], c["w"], seed=c["seed"], bg=c["bg"], fg=c["fg"], noise=c["noise"])
    s = c["scale"]
    out = restate.process(L, R, s, c["d"], "sgm", taps=restate.ALL_TAPS, box_w=s, box_h=s)
    for v in (0, 1):  # the claim the mode rests on: the float path computes integers, the 16-bit form agrees
        Si = restate.aggregate_u16(out["cost_raw"][v].astype(np.uint16), 10, 100, 8)
        assert np.array_equal(Si.astype(np.float32), out["S"][v]), "16-bit restatement differs from the float restatement"
    # direct pin where the reference can run it (scale 1): its own sources compiled with COST_WIN_H = COST_WIN_W = 1
    # (make -C oracle ref_nofilter) must give the same path sums, disparities and final image
    if s == 1 and refbind.available("nofilter"):
        with refbind.use("nofilter"):
            refbind.set_threads(1)
            ref = refbind.process("sgm", L, R, 1, c["d"], want_cost=True)
        assert np.array_equal(ref["cost_last"], out["S"][1]) and np.array_equal(ref["disp"], out["disp"][0])
        assert np.array_equal(ref["disp_beta"], out["disp"][1]) and np.array_equal(ref["fdisp"].view(np.uint32), out["f_final"].view(np.uint32))
        c = dict(c, pinned_by="oracle/_ref/libsgm_ref_nofilter.so (reference sources, COST_WIN_H = COST_WIN_W = 1)")
    f = out["f_final"]
    rec = dict(c)
    rec["hashes"] = dict(left=H(L), right=H(R), cost_l=H(out["cost_hv"][0]), cost_r=H(out["cost_hv"][1]), S_l=H(out["S"][0]),
                         S_r=H(out["S"][1]), disp=H(out["disp"][0]), disp_beta=H(out["disp"][1]), fsub_l=H(out["fsub"][0]),
                         fsub_r=H(out["fsub"][1]), f_lr=H(out["f_lr"]), f_median=H(out["f_median"]), f_final=H(f))
    rec["stats"] = dict(valid=int((f <= c["d"] - 1).sum()), S_max=int(out["S"][0].max()))
    return rec


def main():
    recs = []
    for c in CASES:
        print("golden (integer mode):", c["name"], flush=True)
        recs.append(run_case(c))
    with open(os.path.join(HERE, "kat_hashes_integer.json"), "w") as f:
        json.dump(dict(generator="tests/golden/make_golden_integer.py",
                       source="oracle/sgm_oracle.c with box_w = box_h = scale (window 1x1), cross-checked with orc_aggregate_u16",
                       hash="FNV-1a-64 over raw little-endian buffers", cases=recs), f, indent=1)


if __name__ == "__main__":
    main()
