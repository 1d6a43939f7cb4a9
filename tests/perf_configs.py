"""Times every BASELINE.json config on one GPU (dev helper, not a pytest; bench.py stays the contract for configs[2]).

    python tests/perf_configs.py [--frames N] > gpurun_out/perf_configs.jsonl

One JSON line per config: frames/s with 3 frames in flight through the host C ABI (sgm_enqueue / sgm_wait, H2D and D2H
inside the timed region, wall clock around a synchronised region), Gdisp.px/s, per-stage device times of a frame run
alone, device memory per handle.  Config 4 (1080p, D=256) cannot run through the reference at all (src/Solver.cpp:10).
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereo_matching_b200 import capi, stream, synth  # noqa: E402

def _variant(**kw):
    """SGM with extra sgm_params / flags (e.g. the integer mode), as a class-like factory."""
    def make(h, w, s, d, flags=0):
        over = dict(kw)
        return capi.SGM(h, w, s, d, flags=flags | over.pop("flags", 0), **over)
    return make


CONFIGS = [
    ("configs[0] KITTI 1242x375 census + 8-path SGM D=64", capi.SGM, 375, 1242, 64, dict(bg=12, fg=40)),
    ("configs[1] KITTI 1242x375 SAD 9x9 block matching D=64 + LR-check", capi.BMSad, 375, 1242, 64, dict(bg=12, fg=40)),
    ("configs[1b] KITTI 1242x375 reference BM (census) D=64", capi.BM, 375, 1242, 64, dict(bg=12, fg=40)),
    ("configs[2] KITTI 1242x375 census + 8-path SGM D=128", capi.SGM, 375, 1242, 128, dict(bg=12, fg=40)),
    ("configs[3] 1920x1080 8-path SGM D=256", capi.SGM, 1080, 1920, 256, dict(bg=24, fg=200)),
    # the same SGM with the cost filter off (window 1x1): float arithmetic vs the integer / DPX mode (identical results)
    ("configs[2] no cost filter, float32 volumes D=128", _variant(box_w=1, box_h=1), 375, 1242, 128, dict(bg=12, fg=40)),
    ("configs[2] no cost filter, INTEGER mode (s16x2, DPX) D=128", _variant(box_w=1, box_h=1, flags=capi.FLAG_INTEGER), 375, 1242, 128,
     dict(bg=12, fg=40)),
    ("configs[0] no cost filter, INTEGER mode D=64", _variant(box_w=1, box_h=1, flags=capi.FLAG_INTEGER), 375, 1242, 64, dict(bg=12, fg=40)),
    ("configs[3] no cost filter, INTEGER mode 1920x1080 D=256", _variant(box_w=1, box_h=1, flags=capi.FLAG_INTEGER), 1080, 1920, 256,
     dict(bg=24, fg=200)),
]


def run(name, cls, h, w, d, gen, frames):
    pairs = [synth.random_dot_pair(h, w, seed=100 + k, noise=4 if k else 0, **gen)[:2] for k in range(4)]
    inflight = 5 if h * w * d < 3e8 else 3
    engines = [cls(h, w, 1, d) for _ in range(inflight)]
    probe = cls(h, w, 1, d, flags=capi.FLAG_STAGE_TIMING)
    fs = stream.FrameStream(engines)
    list(fs.run((k, pairs[k % 4]) for k in range(inflight * 2)))  # warm-up
    t0 = time.perf_counter()
    n = sum(1 for _ in fs.run((k, pairs[k % 4]) for k in range(frames)))
    dt = time.perf_counter() - t0
    stage = {}
    for k in range(4):
        probe.process(*pairs[k % 4])
        if k:
            for s, v in probe.stage_ms().items():
                stage[s] = stage.get(s, 0.0) + v / 3
    f = probe.get_disp()
    line = {"config": name, "h": h, "w": w, "d": d, "frames": n, "inflight": inflight, "frames_per_s": n / dt,
            "gdisp_px_per_s": n / dt * h * w * d / 1e9, "stage_ms": {k: round(v, 4) for k, v in stage.items()},
            "launches_per_frame": probe.launches, "device_mb_per_handle": round(probe.device_bytes / 1e6, 1),
            "valid_fraction": float((f <= d - 1).mean())}
    for e in engines + [probe]:
        e.close()
    return line


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=60)
    ap.add_argument("--only", default="", help="substring of the config names to run")
    args = ap.parse_args()
    for cfg in CONFIGS:
        if args.only not in cfg[0]:
            continue
        n = args.frames if cfg[2] * cfg[3] * cfg[4] < 3e8 else max(6, args.frames // 10)
        print(json.dumps(run(*cfg, frames=n)), flush=True)
