"""BASELINE.json configs[4]: a stream of 4096 1242x375 D=128 pairs over N GPUs from ONE C++ process (tests/cpp/host_shell_main.cpp
`stream` mode -> sgm_stream_*), first 8 results checked against the reference's golden vectors.  Dev / profile helper:
    python tests/stream_config4.py N_GPUS [FRAMES] [INFLIGHT]      -> one JSON line"""
import json
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from stereo_matching_b200 import synth  # noqa: E402

H, W, D, N_PAIRS = 375, 1242, 128, 8
n_gpus = int(sys.argv[1])
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
inflight = int(sys.argv[3]) if len(sys.argv) > 3 else 5
exe = os.path.join(ROOT, "stereo_matching_b200", "host", "_build", "host_shell_main")
with tempfile.TemporaryDirectory() as tmp:
    for k in range(N_PAIRS):
        L, R, _ = synth.random_dot_pair(H, W, seed=1234 + k, bg=12, fg=40, noise=4 if k else 0)
        L.tofile(os.path.join(tmp, f"l{k}.raw"))
        R.tofile(os.path.join(tmp, f"r{k}.raw"))
    p = subprocess.run([exe, "stream", str(H), str(W), "1", str(D), str(n_gpus), str(inflight), str(frames), tmp, str(N_PAIRS)],
                       capture_output=True, text=True, timeout=200)
last = [ln for ln in p.stderr.splitlines() if ln.startswith("{")]
if p.returncode != 0 or not last:
    raise SystemExit(f"stream run failed (rc {p.returncode}): {p.stderr[-400:]}")
j = json.loads(last[-1])
gold = {q["pair"]: q["f_final"] for q in json.load(open(os.path.join(ROOT, "tests", "golden", "bench_pairs.json")))["pairs"]}
j["parity_first_8"] = all(hh == gold[k] for k, hh in enumerate(j.pop("first_hashes")))
j["config"] = "configs[4]: stream of 1242x375 D=128 SGM pairs, one C++ process, sgm_stream over N GPUs, host buffers in, pinned views out"
print(json.dumps(j))
if not j["parity_first_8"]:
    raise SystemExit("first 8 results differ from tests/golden/bench_pairs.json")
