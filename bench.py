#!/usr/bin/env python
"""bench.py — headline benchmark of the stereo hot path (BASELINE.json: SGM frames/s and Gdisp.px/s at
1242x375 D=128, 8 paths, sub-pixel + LR-check + median + speckle).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One step = one batch of FRAMES_PER_STEP stereo pairs (distinct seeded synthetic random-dot pairs, cycled)
through the full SGM pipeline on this rank's GPU.  Ranks shard the frame stream (independent pairs, no
data-path collective: scaling = weak); torch.distributed is used only for the barrier and the max-over-ranks
of the device-timed region.  Prints ONE JSON line on rank 0.

  value     frames/s, whole job, inputs already resident in HBM (sgm_enqueue_device), results left on device
  e2e       frames/s through the reference-facing C ABI with HOST buffers (sgm_enqueue / sgm_wait:
            pinned staging + H2D + kernels + D2H inside the timed region)
  roofline  aggregation stage (horizontal pair kernel + down sweep + up sweep incl. the fused WTA epilogue): algorithmic
            96*W*H*D bytes per frame / its CUDA-event duration on the engine's stream, vs the measured
            HBM copy peak; `traffic` = DRAM bytes of those kernels from the committed ncu capture
  cpu_baseline  the reference's own CPU sources (oracle/_ref) on this box's host cores, bounded sample

--impl reference times the reference CPU implementation only (rank 0), same metric / unit / config.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

H, W, D, SCALE = 375, 1242, 128, 1
FRAMES_PER_STEP = 8
N_PAIRS = 8
WORKLOAD = "configs[2]: 1242x375 random-dot pairs, census 7x9 + recursive box + 8-path SGM D=128, sub-pixel, LR-check, median, speckle"
# --mode: the headline is "reference" (the reference's float32 arithmetic with its recursive cost filter).  "nofilter" switches
# the cost filter off (window 1x1, still float32); "integer" is the same computation on 16-bit integer volumes with DPX
# instructions (SGM_FLAG_INTEGER: identical results to "nofilter", tests/test_parity_gpu.py).  Secondary lines, not the metric.
MODES = {
    "reference": dict(over={}, flags=0, bytes_per_disp_px=96.0, dtype="f32", workload=WORKLOAD),
    "nofilter": dict(over=dict(box_w=1, box_h=1), flags=0, bytes_per_disp_px=96.0, dtype="f32",
                     workload=WORKLOAD.replace("recursive box", "no cost filter")),
    "integer": dict(over=dict(box_w=1, box_h=1), flags=8, bytes_per_disp_px=48.0, dtype="s16",
                    workload=WORKLOAD.replace("recursive box", "no cost filter") + "; INTEGER mode: s16x2 volumes, DPX min-plus"),
}


def config_dict(n_gpus, inflight, mode="reference"):
    return {
        "workload": MODES[mode]["workload"],
        "width": W, "height": H, "max_disp": D, "paths": 8, "scale": SCALE,
        "frames_per_step": FRAMES_PER_STEP, "distinct_pairs": N_PAIRS, "inflight_frames_per_gpu": inflight,
        "sharding": f"independent pairs round-robin over {n_gpus} GPU(s), no collective",
        "l2": "no explicit flush: each frame streams ~1.4 GB of fp32 volumes (>> 126 MB L2) and 8 distinct pairs are cycled",
    }


def make_pairs():
    from stereo_matching_b200 import synth

    pairs = []
    for k in range(N_PAIRS):
        L, R, _ = synth.random_dot_pair(H, W, seed=1234 + k, bg=12, fg=40, noise=4 if k else 0)
        pairs.append((L, R))
    return pairs


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons of one GPU while a timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])), mx.append(float(f[1])), pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "power_w_max": max(pw), "samples": len(sm),
                "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------
CPU_REF_TIMEOUT_S = 240  # one attempt of the CPU reference (a frame takes ~1.5 s on 16 threads); see cpu_reference_time


def cpu_reference_time(pairs, steps, warmup):
    """Times the reference's own CPU implementation (oracle/_ref = unmodified /root/reference sources, OpenMP on all
    host threads; the restatement if the compiled reference did not travel).  One step = ONE frame (pairs[0], seed 1234).

    Runs in a child process with a time limit: the reference's speckle filter races under OpenMP
    (src/Solver.cpp:525-553), so a run that does not come back is retried once and then reported as unavailable instead
    of taking the bench down.  Returns (kind, cores, times_ms) or (None, cores, reason)."""
    del pairs  # the child regenerates pair 0 (same generator and seed)
    try:
        cores = len(os.sched_getaffinity(0))  # torchrun exports OMP_NUM_THREADS=1, which must not throttle the baseline
    except AttributeError:
        cores = os.cpu_count() or 1
    cmd = [sys.executable, os.path.join(ROOT, "tests", "run_cpu_ref.py"), str(H), str(W), str(D), str(SCALE), str(warmup + steps), str(cores)]
    why = "not run"
    limit = max(CPU_REF_TIMEOUT_S, 30 * (warmup + steps))
    for attempt in range(2):
        try:
            r = subprocess.run(cmd, capture_output=True, text=True, timeout=limit)
        except subprocess.TimeoutExpired:
            why = f"no answer within {limit} s (attempt {attempt + 1})"
            continue
        last = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
        if r.returncode == 0 and last:
            j = json.loads(last[-1])
            return j["kind"], j["cores"], [float(x) for x in j["ms"][warmup:]]
        why = f"rc {r.returncode}: {r.stderr.strip()[-200:]}"
    return None, cores, why


def run_reference_arm(args, rank, world):
    if rank != 0:
        return  # rank 0 alone runs the CPU reference; the others exit 0 without work
    pairs = make_pairs()
    kind, cores, times = cpu_reference_time(pairs, args.steps, max(args.warmup, 1))
    if kind is None:
        print(json.dumps({"impl": "reference", "unavailable": f"CPU reference did not finish: {times}"}), flush=True)
        return
    total_s = sum(times) / 1e3
    fps = len(times) / total_s
    line = {
        "impl": "reference", "metric": "sgm_frames_per_s", "value": fps, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total_s / len(times), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config_dict(args.gpus, 0),
        "gdisp_px_per_s": fps * W * H * D / 1e9,
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": kind,
                         "sample": f"{len(times)} step(s) of ONE 1242x375 D=128 frame each through SGM::process (src/SGM.cpp:32-826)"},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------
def dist_setup(args):
    import torch

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist_mod

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        backend = "nccl" if (torch.cuda.is_available() and args.impl != "reference") else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
        dist_mod.init_process_group(backend=backend, rank=rank, world_size=world)
        dist = dist_mod
    return rank, world, local, dist


def barrier_sync(dist, torch, device):
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize(device)


def max_over_ranks(dist, torch, value, device):
    if dist is None:
        return value
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


# ---------------------------------------------------------------------------------------------
class Timed:
    """CUDA-event timing of a region that runs on several engine streams: `start` is recorded on an idle join
    stream that every engine stream then waits on; `end` is recorded after the join stream has waited on every
    engine stream."""

    def __init__(self, torch, device, engines):
        self.torch, self.device = torch, device
        self.streams = [torch.cuda.ExternalStream(sv.stream_ptr, device=device) for sv in engines]
        self.join = torch.cuda.Stream(device=device)
        self.start = torch.cuda.Event(enable_timing=True)
        self.end = torch.cuda.Event(enable_timing=True)

    def begin(self):
        self.start.record(self.join)
        for s in self.streams:
            s.wait_event(self.start)

    def finish(self):
        for s in self.streams:
            ev = self.torch.cuda.Event()
            ev.record(s)
            self.join.wait_event(ev)
        self.end.record(self.join)

    def ms(self):
        return self.start.elapsed_time(self.end)


def run_gpu_arm(args, rank, world, local, dist):
    import numpy as np
    import torch

    from stereo_matching_b200 import capi

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device visible; this engine has no CPU path (use --impl reference for the CPU baseline)")
    device = torch.device("cuda", local)
    torch.cuda.set_device(device)
    pairs = make_pairs()
    inflight = args.inflight
    mode = MODES[args.mode]
    engines = [capi.SGM(H, W, SCALE, D, device=local, flags=mode["flags"], **mode["over"]) for _ in range(inflight)]
    probe = capi.SGM(H, W, SCALE, D, device=local, flags=capi.FLAG_STAGE_TIMING | mode["flags"], **mode["over"])
    dev_pairs = [(torch.from_numpy(L).to(device), torch.from_numpy(R).to(device)) for L, R in pairs]
    out = np.empty((H, W), np.float32)
    torch.cuda.synchronize(device)
    n_frames = args.steps * FRAMES_PER_STEP
    n_warm = args.warmup * FRAMES_PER_STEP

    def resident(frames, first=0):
        launches = 0
        for f in range(first, first + frames):
            sv = engines[f % inflight]
            sv.wait()
            l, r = dev_pairs[f % N_PAIRS]
            sv.enqueue_device(l.data_ptr(), W, r.data_ptr(), W)
            launches += sv.launches
        return launches

    def host_e2e(frames, first=0):
        launches, checksum = 0, 0.0
        pending = [False] * inflight
        for f in range(first, first + frames):
            k = f % inflight
            sv = engines[k]
            if pending[k]:
                sv.wait()
                sv.get_disp_into(out)  # D2H'd result -> caller's buffer
                checksum += float(out[H // 2, W // 2])
            L, R = pairs[f % N_PAIRS]
            sv.enqueue(L, R)
            pending[k] = True
            launches += sv.launches
        for k in range(inflight):
            if pending[k]:
                engines[k].wait()
                engines[k].get_disp_into(out)
                checksum += float(out[H // 2, W // 2])
        return launches, checksum

    sampler = ClockSampler(local)
    sampler.start()

    # ---- value: inputs resident in HBM ----
    resident(n_warm)
    for sv in engines:
        sv.wait()
    t = Timed(torch, device, engines)
    barrier_sync(dist, torch, device)
    t.begin()
    launches = resident(n_frames, first=n_warm)
    t.finish()
    barrier_sync(dist, torch, device)
    for sv in engines:
        sv.wait()
    ms_value = max_over_ranks(dist, torch, t.ms(), device)

    # ---- e2e: host buffers through the C ABI ----
    host_e2e(n_warm)
    t2 = Timed(torch, device, engines)
    barrier_sync(dist, torch, device)
    t2.begin()
    host_e2e(n_frames, first=n_warm)
    t2.finish()
    barrier_sync(dist, torch, device)
    ms_e2e = max_over_ranks(dist, torch, t2.ms(), device)

    # ---- roofline: aggregation stage timed alone on its launching stream (serial frames, engine-side CUDA events) ----
    stage_sum = {}
    for f in range(n_warm + n_frames):
        l, r = dev_pairs[f % N_PAIRS]
        probe.enqueue_device(l.data_ptr(), W, r.data_ptr(), W)
        probe.wait()
        if f >= n_warm:
            for k, v in probe.stage_ms().items():
                stage_sum[k] = stage_sum.get(k, 0.0) + v
    clocks = sampler.stop()
    stage_ms = {k: v / n_frames for k, v in stage_sum.items()}
    return dict(ms_value=ms_value, ms_e2e=ms_e2e, launches=launches, stage_ms=stage_ms, clocks=clocks, n_frames=n_frames,
                inflight=inflight, device_bytes=engines[0].device_bytes)


def gpu_sgm_reference_time():
    """The reference's own CUDA class GPU_SGM (gpu_sgm/, compiled unmodified into oracle/_ref/libgpu_sgm_ref.so) timed on
    this GPU in a separate process: a reported baseline only (approximate algorithm, left view only, not bit-comparable)."""
    lib = os.path.join(ROOT, "oracle", "_ref", "libgpu_sgm_ref.so")
    if not os.path.exists(lib):
        return {"unavailable": "oracle/_ref/libgpu_sgm_ref.so not built"}
    try:
        r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "run_gpu_sgm_ref.py"), str(H), str(W), str(D), "5"],
                           capture_output=True, text=True, timeout=120)
        last = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
        if r.returncode != 0 or not last:
            return {"unavailable": f"run failed (rc {r.returncode})"}
        j = json.loads(last[-1])
        return {"value": j["frames_per_s_steady"], "unit": "frames/s", "ms_per_frame": j["ms_per_frame"], "kind": "reference gpu_sgm",
                "note": "GPU_SGM::process: approximate algorithm, left view only, no LR check; same 1242x375 D=128 pair"}
    except Exception as e:  # a baseline must never take the bench down
        return {"unavailable": f"{type(e).__name__}: {e}"}


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def load_traffic():
    """Per-frame DRAM bytes of the aggregation kernels from the committed ncu capture, if any."""
    path = os.path.join(ROOT, "profiles", "aggregate_traffic.json")
    try:
        with open(path) as f:
            return float(json.load(f)["dram_bytes_per_frame"])
    except Exception:
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--inflight", type=int, default=5, help="frames in flight per GPU (engine handles)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--mode", default="reference", choices=sorted(MODES), help="secondary lines: nofilter / integer (see MODES)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    # watchdog: a bench that does not finish must say where it is stuck and exit, not sit until somebody kills it
    import faulthandler
    faulthandler.dump_traceback_later(int(os.environ.get("SGM_B200_BENCH_WATCHDOG_S", "1200")), exit=True)

    rank, world, local, dist = dist_setup(args)
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return

    r = run_gpu_arm(args, rank, world, local, dist)
    if rank == 0:
        n_px = W * H * D
        total_frames = r["n_frames"] * world
        fps = total_frames / (r["ms_value"] / 1e3)
        fps_e2e = total_frames / (r["ms_e2e"] / 1e3)
        peak, peak_src = load_peaks()
        agg_ms = r["stage_ms"]["aggregate"]
        # 48 B per disp.px per pass, two passes per frame (float32 volumes); 24 B per pass with 16-bit volumes
        alg_bytes = MODES[args.mode]["bytes_per_disp_px"] * n_px
        achieved = alg_bytes / (agg_ms / 1e3) / 1e9
        line = {
            "metric": "sgm_frames_per_s", "value": fps, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": r["ms_value"] / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": MODES[args.mode]["dtype"], "data": "synthetic", "config": config_dict(world, r["inflight"], args.mode),
            "gdisp_px_per_s": fps * n_px / 1e9,
            "e2e": {"value": fps_e2e, "unit": "frames/s", "h2d_bytes_per_step": 2 * H * W * FRAMES_PER_STEP,
                    "d2h_bytes_per_step": 4 * H * W * FRAMES_PER_STEP, "gdisp_px_per_s": fps_e2e * n_px / 1e9},
            "gpu_launches": r["launches"],
            "roofline": {"kernel": "aggregation stage = sgm_hpair_async_kernel + sgm_sweep_kernel<DOWN> + sgm_sweep_kernel<UP> (3 launches per frame, both views per launch; the up sweep ends in the fused sum/WTA/uniqueness/sub-pixel epilogue)",
                         "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": load_traffic() if args.mode == "reference" else None, "peak_source": peak_src, "algorithmic_bytes_per_frame": alg_bytes,
                         "ms_per_frame": agg_ms, "timing": "engine-side CUDA events on the launching stream, frames run serially"},
            "stage_ms_per_frame": r["stage_ms"],
            "clocks": r["clocks"],
            "device_bytes_per_engine": r["device_bytes"],
        }
        if not args.no_cpu_baseline and world == 1:
            kind, cores, times = cpu_reference_time(None, 2, 1)
            if kind is None:
                line["cpu_baseline"] = {"value": None, "unit": "frames/s", "cores": cores, "kind": "reference", "sample": f"unavailable: {times}"}
            else:
                cfps = len(times) / (sum(times) / 1e3)
                line["cpu_baseline"] = {"value": cfps, "unit": "frames/s", "cores": cores, "kind": kind,
                                        "sample": f"{len(times)} frames of the same 1242x375 D=128 workload after 1 warm-up, SGM::process (src/SGM.cpp:32-826), OpenMP on all host threads"}
        else:
            line["cpu_baseline"] = None
        if not args.no_cpu_baseline and world == 1:
            line["gpu_sgm_reference"] = gpu_sgm_reference_time()
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
