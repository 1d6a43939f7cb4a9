"""Small frames through every kernel, written for `compute-sanitizer --tool memcheck|racecheck python tests/sanitize_small.py`
(dev helper, not a pytest; one tool per GPU call, see B200_PROFILING.md).  compute-sanitizer was closed on the GPU pool during
round 2 (profiles/r02_summary.md), so the script only serves as a quick run of every kernel on small and on wide frames."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereo_matching_b200 import capi, synth  # noqa: E402

for cls, h, w, s, d, sky in ((capi.SGM, 48, 100, 1, 32, True), (capi.SGM, 40, 90, 1, 128, False), (capi.BM, 48, 100, 2, 32, False),
                             (capi.BMSad, 40, 64, 1, 32, False)):
    L, R, _ = synth.random_dot_pair(h, w, seed=h + d, bg=3, fg=9, noise=20)
    masks = synth.sky_masks(h // s, w // s) if sky else (None, None)
    sv = cls(h, w, s, d)
    f = sv.process(L, R, *masks).get_disp()
    view = sv.debug_view()
    xyz, inten = sv.point_cloud(300.0, 300.0, w / 2, h / 2)
    print(cls.__name__, h, w, d, "valid", float((f <= d - 1).mean()), "points", len(xyz), "launches", sv.launches)
    sv.close()
# wide rows: one block per SM with 13 / 17 / 19 columns and the hardware warp -> column permutation (sweep_kernels.cu: wide_shape)
for h, w, d, flags, over in ((16, 760, 64, 0, {}), (16, 1250, 32, 0, {}), (16, 1290, 32, 0, {}), (16, 1500, 256, 0, {}),
                             (16, 1250, 64, capi.FLAG_INTEGER, dict(box_w=1, box_h=1))):
    L, R, _ = synth.random_dot_pair(h, w, seed=h + d, bg=3, fg=9, noise=20)
    sv = capi.SGM(h, w, 1, d, flags=flags, **over)
    f = sv.process(L, R).get_disp()
    print("SGM wide", h, w, d, flags, "valid", float((f <= d - 1).mean()), "launches", sv.launches)
    sv.close()
for h, w, s, d in ((48, 100, 1, 64), (40, 90, 1, 128), (48, 100, 2, 64)):  # integer / DPX mode
    L, R, _ = synth.random_dot_pair(h, w, seed=h + d, bg=3, fg=9, noise=20)
    sv = capi.SGM(h, w, s, d, flags=capi.FLAG_INTEGER, box_w=s, box_h=s)
    f = sv.process(L, R).get_disp()
    print("SGM integer", h, w, d, "valid", float((f <= d - 1).mean()), "launches", sv.launches)
    sv.close()
print("sanitize_small done")
