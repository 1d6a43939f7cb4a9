"""First-contact diagnostic (not a pytest): runs the CUDA path against the oracle stage by stage and prints
mismatch counts + stage timings, so one GPU call localises a bug.  python tests/diag_gpu.py [h w d scale]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import restate  # noqa: E402
from stereo_matching_b200 import capi, synth  # noqa: E402


def nd(a, b):
    a, b = np.asarray(a), np.asarray(b)
    if a.dtype.kind == "f":
        return int((a.view(np.uint32) != b.view(np.uint32)).sum())
    return int((a != b).sum())


def first_bad(a, b):
    a, b = np.asarray(a), np.asarray(b)
    m = (a.view(np.uint32) != b.view(np.uint32)) if a.dtype.kind == "f" else (a != b)
    idx = np.argwhere(m)
    if len(idx) == 0:
        return ""
    t = tuple(idx[0])
    return f" first@{t}: got {a[t]} want {b[t]}; rows {idx[:,0].min()}..{idx[:,0].max()} cols {idx[:,1].min()}..{idx[:,1].max()}"


def run(h, w, d, scale, noise=12, sky=False, algo="sgm", paths=8):
    L, R, _ = synth.random_dot_pair(h, w, seed=h + w + d, bg=5, fg=min(17, d // 2), noise=noise)
    sl, sr = synth.sky_masks(h // scale, w // scale) if sky else (None, None)
    t0 = time.time()
    want = restate.process(L, R, scale, d, algo, sl, sr, taps=restate.ALL_TAPS, paths=paths)
    t_cpu = time.time() - t0
    cls = dict(sgm=capi.SGM, bm=capi.BM, bm_sad=capi.BMSad)[algo]
    sv = cls(h, w, scale, d, flags=capi.FLAG_KEEP_VOLUMES | capi.FLAG_STAGE_TIMING, paths=paths)
    sv.process(L, R, sl, sr)
    sv.process(L, R, sl, sr)
    print(f"== {algo} {h}x{w} D={d} scale={scale} noise={noise} sky={sky} paths={paths}: oracle {t_cpu:.2f}s, gpu stages ms {sv.stage_ms()}, "
          f"launches {sv.launches}, device MB {sv.device_bytes/1e6:.0f}")
    views = (0,) if algo == "bm" else (0, 1)
    pairs = [("blur", "blur"), ("census", "census"), ("cost", "cost_hv" if algo != "bm_sad" else "cost_raw"), ("pathsum", "S"),
             ("disp", "disp"), ("subpixel", "fsub")]
    for v in views:
        for tap, key in pairs:
            if algo != "sgm" and tap == "pathsum":
                continue
            if algo == "bm_sad" and tap in ("blur", "census"):
                continue
            g = sv.tap(tap, v)
            n = nd(g, want[key][v])
            print(f"   {tap}[{v}]: {'ok' if n == 0 else str(n) + ' differ' + first_bad(g, want[key][v])}")
    for tap, key in (("lr", "f_lr"), ("median", "f_median"), ("final", "f_final")):
        g = sv.tap(tap)
        n = nd(g, want[key])
        print(f"   {tap}: {'ok' if n == 0 else str(n) + ' differ' + first_bad(g, want[key])}")
    sv.close()


if __name__ == "__main__":
    if len(sys.argv) >= 2:
        a = sys.argv[1:]
        run(int(a[0]), int(a[1]), int(a[2]), int(a[3]), noise=int(a[4]), sky=bool(int(a[5])), algo=a[6], paths=int(a[7]))
    else:
        import subprocess

        cases = ["96 160 32 1 12 0 sgm 8", "120 200 64 1 12 0 sgm 8", "120 200 64 1 40 0 sgm 8", "120 200 64 1 12 1 sgm 8",
                 "64 300 128 1 25 0 sgm 8", "200 320 64 2 10 1 sgm 8", "72 480 256 1 6 0 sgm 8", "80 144 64 1 12 0 sgm 4",
                 "96 160 32 1 12 0 bm 8", "90 250 64 1 30 0 bm_sad 8", "375 1242 128 1 8 0 sgm 8"]
        for c in cases:
            r = subprocess.run([sys.executable, __file__] + c.split(), capture_output=True, text=True)
            print(r.stdout, end="")
            if r.returncode != 0:
                print(f"== CASE {c} FAILED rc={r.returncode}: {r.stderr.strip().splitlines()[-1] if r.stderr.strip() else ''}")
            sys.stdout.flush()
