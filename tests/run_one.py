"""Dev helper: run a few frames of one configuration (for ncu captures).  usage: python tests/run_one.py sgm|bm|bmsad H W D [frames] [integer]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereo_matching_b200 import capi, synth  # noqa: E402

algo, h, w, d = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
frames = int(sys.argv[5]) if len(sys.argv) > 5 else 4
cls = {"sgm": capi.SGM, "bm": capi.BM, "bmsad": capi.BMSad}[algo]
L, R, _ = synth.random_dot_pair(h, w, seed=7, bg=12, fg=min(40, d // 2), noise=4)
integer = len(sys.argv) > 6 and sys.argv[6] == "integer"
e = cls(h, w, 1, d, flags=capi.FLAG_STAGE_TIMING | (capi.FLAG_INTEGER if integer else 0), **(dict(box_w=1, box_h=1) if integer else {}))
for _ in range(frames):
    e.process(L, R)
print(e.stage_ms())
e.close()
