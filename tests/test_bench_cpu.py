"""Host-side logic of bench.py that the driver's contract depends on, checked without a GPU: both arms describe the same
workload, the reference arm's non-zero ranks leave without work, the roofline inputs come from where the JSON line says, and
the GPU arm refuses to run (instead of falling back to anything) when there is no device."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def test_both_arms_describe_the_same_workload():
    for n in (1, 2, 8):
        a, b = bench.config_dict(n), bench.config_dict(n)
        assert a == b and a["workload"].startswith("configs[2]")
        assert (a["width"], a["height"], a["max_disp"], a["paths"]) == (1242, 375, 128, 8)
        assert "model" not in a  # the contract's config names a workload, not a model
    assert bench.config_dict(1) != bench.config_dict(2)  # the sharding sentence names the GPU count


def test_golden_hashes_of_the_bench_pairs_are_committed():
    with open(os.path.join(ROOT, "tests", "golden", "bench_pairs.json")) as f:
        gold = json.load(f)
    assert (gold["h"], gold["w"], gold["d"]) == (bench.H, bench.W, bench.D)
    assert len(gold["pairs"]) == bench.N_PAIRS and all(len(p["f_final"]) == 16 for p in gold["pairs"])
    assert [p["seed"] for p in gold["pairs"]] == list(range(1234, 1234 + bench.N_PAIRS))  # the generator's seeds (config_dict)


def test_traffic_and_peak_sources():
    traffic, src = bench.load_traffic()
    with open(os.path.join(ROOT, "profiles", "aggregate_traffic.json")) as f:
        doc = json.load(f)
    assert traffic == float(doc["dram_bytes_per_frame"]) and "aggregate_traffic.json" in src
    algorithmic = 96.0 * bench.W * bench.H * bench.D
    assert 0.9 * algorithmic < traffic < 1.1 * algorithmic  # measured DRAM bytes ~ the 96 N algorithmic bytes
    peak, peak_src = bench.load_peaks()
    assert 3000.0 < peak < 9000.0 and peak_src


def test_reference_arm_world_size_2_rank_0_alone_prints(tmp_path):
    """`--impl reference` as the driver launches it for N = 2 (one process per rank, gloo rendezvous on 127.0.0.1): rank 0 runs
    the CPU reference and prints the line, rank 1 waits for it at the barrier and leaves with exit code 0 and no output."""
    from oracle import refbind

    if not refbind.available():
        pytest.skip("oracle/_ref is not built here")
    procs = []
    for rank in (0, 1):
        env = dict(os.environ, RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT="29713")
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                                       "--warmup", "1"], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = [p.communicate(timeout=400) for p in procs]
    assert [p.returncode for p in procs] == [0, 0], [o[1][-400:] for o in outs]
    assert outs[1][0].strip() == ""
    line = json.loads(outs[0][0].strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["n_gpus"] == 2 and line["config"] == bench.config_dict(2)
    assert line["value"] > 0 and line["cpu_baseline"]["kind"] == "reference" and line["e2e"]["h2d_bytes_per_step"] == 0


def test_gpu_arm_fails_loudly_without_a_device():
    import torch

    if torch.cuda.is_available():
        pytest.skip("this check is for the CPU-only container")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1", "--warmup", "0", "--no-configs", "--no-cpu-baseline"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode != 0  # no CPU fallback: the product path needs the CUDA engine
    assert '"metric"' not in r.stdout
