"""Times the reference's own CUDA class GPU_SGM (oracle/_ref/libgpu_sgm_ref.so, built by `make -C oracle gpu_sgm`) on
one B200: a reported baseline ("the repo's gpu_sgm path"), not an oracle - it is an approximate algorithm (no pre-blur,
no right-view pass / LR check, truncated diagonal paths in `short`).  Dev helper, not a pytest:
    timeout 120 python tests/run_gpu_sgm_ref.py [H W D iters]   -> one JSON line"""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from stereo_matching_b200 import synth  # noqa: E402

h, w, d, iters = (int(x) for x in (sys.argv[1:5] + ["375", "1242", "128", "5"][len(sys.argv) - 1:]))
lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libgpu_sgm_ref.so"))
lib.gpu_sgm_time.argtypes = [C.c_int] * 4 + [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
L, R, gt = synth.random_dot_pair(h, w, seed=1234, bg=12, fg=40)
ms = np.zeros(iters, np.float64)
out = np.zeros((h, w), np.float32)
sys.stdout.flush()
rc = lib.gpu_sgm_time(h, w, 1, d, L.ctypes.data, R.ctypes.data, iters, ms.ctypes.data, out.ctypes.data)
valid = out <= d - 1
region = np.broadcast_to(np.arange(w)[None, :] >= d, (h, w))
good = valid & region & (np.abs(out - gt) <= 1)
print(json.dumps({"impl": "reference gpu_sgm (GPU_SGM::process, approximate algorithm, left view only)", "rc": rc, "h": h, "w": w, "d": d,
                  "ms_per_frame": [round(float(x), 3) for x in ms],
                  "frames_per_s_steady": float(1e3 / np.median(ms[1:])) if iters > 1 else None,
                  "valid_fraction_cols_ge_d": float((valid & region).sum() / region.sum()),
                  "within_1px_of_gt_cols_ge_d": float(good.sum() / region.sum())}))
