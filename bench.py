#!/usr/bin/env python
"""bench.py — headline benchmark of the stereo hot path (BASELINE.json: SGM frames/s and Gdisp.px/s at
1242x375 D=128, 8 paths, sub-pixel + LR-check + median + speckle).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One step = FRAMES_PER_STEP stereo pairs (8 distinct seeded synthetic random-dot pairs, cycled) through the full SGM pipeline
on this rank's GPU.  Ranks shard the frame stream (independent pairs, no data-path collective: scaling = weak);
torch.distributed is used only for the barrier and the max-over-ranks of the device-timed region.  ONE JSON line on rank 0.

  value     frames/s, whole job, inputs already resident in HBM (sgm_enqueue_device), results left on the device
  e2e       frames/s through the C frame-stream runner (sgm_stream_submit / sgm_stream_next: the loop of the reference's
            caller, node.cpp:49,93,104) with HOST buffers: copy into pinned staging + H2D + kernels + D2H + copy out of
            pinned memory, all inside the timed region; the first 8 results are hashed against the reference's golden
            vectors (tests/golden/bench_pairs.json) outside it
  roofline  aggregation stage (horizontal pair kernel + down sweep + up sweep incl. the fused WTA epilogue): algorithmic
            96*W*H*D bytes per frame / its CUDA-event duration on the engine's stream, vs the measured
            HBM copy peak; `traffic` = DRAM bytes of those kernels from the committed ncu capture (`traffic_source`)
  configs   the other BASELINE.json configurations (and the integer mode), timed in the same run (N = 1 only)
  blocking  Solver::process() + get_disp() of the C++ shell, one frame at a time (what node.cpp as shipped would see)
  cpu_baseline  the reference's own CPU sources (oracle/_ref) on this box's host cores, bounded sample (N = 1 only)

--impl reference times the reference CPU implementation only (rank 0), same metric / unit / config.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

H, W, D, SCALE = 375, 1242, 128, 1
FRAMES_PER_STEP = 64   # GPU arm: 20 steps = 1280 frames per GPU, about 2 s of device time
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


def config_dict(n_gpus, mode="reference"):
    """Describes the WORKLOAD only, so that both arms (--impl b200 / reference) print the same dict: how many frames an arm
    puts into a step, or keeps in flight, is reported beside it (frames_per_step, inflight_frames_per_gpu)."""
    return {
        "workload": MODES[mode]["workload"],
        "width": W, "height": H, "max_disp": D, "paths": 8, "scale": SCALE, "distinct_pairs": N_PAIRS,
        "pair_generator": "std::mt19937 random dots, seeds 1234..1241, bg 12 / fg 40, uniform noise +-4 on pairs 1..7 (tests/golden/make_golden_bench.py)",
        "sharding": f"independent pairs round-robin over {n_gpus} GPU(s), no collective",
        "l2": "no explicit flush: each frame streams ~7 GB of fp32 volumes (>> 126 MB L2) and 8 distinct pairs are cycled",
    }


def make_pairs():
    from stereo_matching_b200 import synth

    return [synth.random_dot_pair(H, W, seed=1234 + k, bg=12, fg=40, noise=4 if k else 0)[:2] for k in range(N_PAIRS)]


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
REF_FRAME_DEADLINE_S = 30.0   # a 1242x375 D=128 frame takes ~1.4 s on 16 threads; construction + first frame ~6 s
REF_MAX_RESPAWNS = 3


def host_cores():
    try:
        return len(os.sched_getaffinity(0))  # torchrun exports OMP_NUM_THREADS=1, which must not throttle the baseline
    except AttributeError:
        return os.cpu_count() or 1


def cpu_reference_frames(n_frames, n_warm):
    """Times the reference's own CPU implementation (oracle/_ref = unmodified /root/reference sources, OpenMP on all host
    threads; the restatement if the compiled reference did not travel): frames 0 .. n_warm + n_frames - 1 of the bench's
    pair cycle through one solver, in a child process that prints a line per frame.

    Per-frame watchdog: the reference's speckle filter races under OpenMP (src/Solver.cpp:525-553) and a frame can hang;
    a frame that does not come back within REF_FRAME_DEADLINE_S gets the child killed (its exact PID) and a new child
    started AT THAT FRAME, keeping the frames already timed, so a hang costs ~30 s instead of the whole run.
    Returns (kind, cores, [ms of the timed frames], respawns) or (None, cores, reason, respawns)."""
    cores = host_cores()
    total = n_warm + n_frames
    times = {}
    kind, respawns, why = None, 0, "not run"
    while len(times) < total and respawns <= REF_MAX_RESPAWNS:
        first = len(times)
        cmd = [sys.executable, os.path.join(ROOT, "tests", "run_cpu_ref.py"), str(H), str(W), str(D), str(SCALE), str(first),
               str(total - first), str(cores)]
        child = subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        lines = []

        def pump(proc=child, sink=lines):
            for ln in proc.stdout:
                sink.append((time.monotonic(), ln))

        reader = threading.Thread(target=pump, daemon=True)
        reader.start()
        seen, last_progress = 0, time.monotonic()
        hung = False
        while True:
            while seen < len(lines):
                _, ln = lines[seen]
                seen += 1
                last_progress = time.monotonic()
                try:
                    j = json.loads(ln)
                except ValueError:
                    continue
                if "kind" in j:
                    kind, cores = j["kind"], j["cores"]
                elif "frame" in j:
                    times[int(j["frame"])] = float(j["ms"])
            if child.poll() is not None and seen >= len(lines):
                reader.join(timeout=1)
                if seen >= len(lines):
                    break
            # the first frame of a child also pays library load + a 2 GB allocation
            limit = REF_FRAME_DEADLINE_S * (2 if len(times) == first else 1)
            if time.monotonic() - last_progress > limit:
                hung = True
                child.kill()
                child.wait()
                break
            time.sleep(0.05)
        if len(times) >= total:
            break
        respawns += 1
        why = (f"frame {len(times)} gave no answer within {REF_FRAME_DEADLINE_S:.0f} s" if hung else
               f"child ended early (rc {child.returncode}): {(child.stderr.read() or '').strip()[-200:]}")
    if len(times) < total:
        return None, cores, why, respawns
    return kind, cores, [times[k] for k in range(n_warm, total)], respawns


def run_reference_arm(args, rank, world):
    if rank != 0:
        return  # rank 0 alone runs the CPU reference; the others exit 0 without work
    kind, cores, times, respawns = cpu_reference_frames(args.steps, max(args.warmup, 1))
    if kind is None:
        print(json.dumps({"impl": "reference", "unavailable": f"CPU reference did not finish: {times}"}), flush=True)
        return
    total_s = sum(times) / 1e3
    fps = len(times) / total_s
    sample = (f"{len(times)} step(s) of ONE 1242x375 D=128 frame each (pairs k mod 8 of the same cycle) through one SGM solver, "
              f"SGM::process (src/SGM.cpp:32-826), OpenMP on {cores} threads")
    line = {
        "impl": "reference", "metric": "sgm_frames_per_s", "value": fps, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total_s / len(times), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config_dict(args.gpus),
        "frames_per_step": 1, "inflight_frames_per_gpu": 0,
        "gdisp_px_per_s": fps * W * H * D / 1e9,
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": kind, "sample": sample, "child_respawns": respawns},
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


def stream_frames(st, pairs, out, frames, first=0, keep=None):
    """`frames` frames of the pair cycle through the C stream runner from ONE host thread: a result is collected whenever
    every slot is taken, the rest at the end.  keep: dict frame -> copy of the result for the frames it names."""
    done = 0
    for f in range(first, first + frames):
        if st.pending == st.capacity:
            k = st.next_into(out)
            if keep is not None and k in keep:
                keep[k] = out.copy()
            done += 1
        st.submit(*pairs[f % N_PAIRS])
    while st.pending:
        k = st.next_into(out)
        if keep is not None and k in keep:
            keep[k] = out.copy()
        done += 1
    return done


def check_first_results(results):
    """the first 8 results of the stream against the reference's golden vectors (outside the timed
    region).  Returns a dict for the JSON line; raises when a result differs (a fast wrong answer is not a benchmark)."""
    from stereo_matching_b200 import synth

    with open(os.path.join(ROOT, "tests", "golden", "bench_pairs.json")) as f:
        gold = {p["pair"]: p for p in json.load(f)["pairs"]}
    bad = [k for k, img in sorted(results.items()) if synth.fnv1a64(img) != gold[k % N_PAIRS]["f_final"]]
    if bad:
        raise SystemExit(f"bench.py: stream results of frames {bad} differ from the reference's golden vectors (tests/golden/bench_pairs.json)")
    return {"frames_checked": len(results), "against": "tests/golden/bench_pairs.json (reference CPU Solver, FNV-1a-64 of get_disp())", "ok": True}


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
    runner = capi.Stream(H, W, SCALE, D, devices=(local,), inflight=inflight, flags=mode["flags"], **mode["over"])
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

    # ---- e2e: host buffers through the C frame-stream runner (sgm_stream_*) ----
    # warm-up doubles as the in-stream parity check: its first 8 results are kept and hashed after the timed region
    first8 = {k: None for k in range(min(N_PAIRS, n_warm))}
    stream_frames(runner, pairs, out, n_warm, first=0, keep=first8)
    # every engine stream of the runner is idle before `start` and after the last sgm_stream_next, so two events on an
    # otherwise idle stream bracket exactly the device work of the region (the runner's streams are internal to it)
    side = torch.cuda.Stream(device=device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = runner.launches
    barrier_sync(dist, torch, device)
    e0.record(side)
    wall0 = time.perf_counter()
    stream_frames(runner, pairs, out, n_frames, first=n_warm)
    e1.record(side)
    barrier_sync(dist, torch, device)
    wall_e2e = (time.perf_counter() - wall0) * 1e3
    ms_e2e = max_over_ranks(dist, torch, e0.elapsed_time(e1), device)
    launches_e2e = runner.launches - launches0
    # the golden vectors are the reference's (float32, recursive cost filter): they apply to the headline mode only; the
    # secondary modes switch the cost filter off and have their own goldens in tests/ (kat_hashes_integer.json)
    parity = None
    if rank == 0:
        parity = check_first_results(first8) if args.mode == "reference" else {"frames_checked": 0, "ok": None, "note": "secondary mode: checked in tests/test_parity_gpu.py"}

    # ---- roofline: aggregation stage timed alone on its launching stream (serial frames, engine-side CUDA events) ----
    stage_sum = {}
    n_probe = min(n_frames, 160)
    for f in range(16 + n_probe):
        l, r = dev_pairs[f % N_PAIRS]
        probe.enqueue_device(l.data_ptr(), W, r.data_ptr(), W)
        probe.wait()
        if f >= 16:
            for k, v in probe.stage_ms().items():
                stage_sum[k] = stage_sum.get(k, 0.0) + v
    clocks = sampler.stop()
    stage_ms = {k: v / n_probe for k, v in stage_sum.items()}
    dev_bytes = probe.device_bytes
    runner.close()
    for sv in engines + [probe]:
        sv.close()
    return dict(ms_value=ms_value, ms_e2e=ms_e2e, wall_e2e=wall_e2e, launches=launches, launches_e2e=launches_e2e, stage_ms=stage_ms,
                clocks=clocks, n_frames=n_frames, inflight=inflight, device_bytes=dev_bytes, parity=parity)


# ---------------------------------------------------------------------------------------------
# The other BASELINE.json configurations, timed in the same run so that their numbers are driver-run (N = 1 only; the
# headline stays configs[2]).  Each: frames/s host to host through the stream runner, per-stage device times of frames run
# alone, and for the SGM ones the aggregation stage against the same roofline model.
OTHER_CONFIGS = [
    # name, algo, h, w, d, generator, integer mode, frames
    ("configs[0] 1242x375 census + 8-path SGM D=64", "sgm", 375, 1242, 64, dict(bg=12, fg=40), False, 240),
    ("configs[1] 1242x375 SAD 9x9 block matching D=64 + WTA + LR-check", "bm_sad", 375, 1242, 64, dict(bg=12, fg=40), False, 400),
    ("configs[1b] 1242x375 reference BM (census, left view) D=64", "bm", 375, 1242, 64, dict(bg=12, fg=40), False, 400),
    ("configs[3] 1920x1080 8-path SGM D=256", "sgm", 1080, 1920, 256, dict(bg=24, fg=200), False, 24),
    ("configs[2] integer / DPX mode (no cost filter, s16x2 volumes) D=128", "sgm", 375, 1242, 128, dict(bg=12, fg=40), True, 240),
    ("configs[0] integer / DPX mode D=64", "sgm", 375, 1242, 64, dict(bg=12, fg=40), True, 240),
]


def run_other_configs(local, peak):
    import numpy as np

    from stereo_matching_b200 import capi, synth

    algo_id = {"sgm": capi.ALGO_SGM, "bm": capi.ALGO_BM, "bm_sad": capi.ALGO_BM_SAD}
    lines = []
    for name, algo, h, w, d, gen, integer, frames in OTHER_CONFIGS:
        try:
            over = dict(box_w=1, box_h=1) if integer else {}
            flags = capi.FLAG_INTEGER if integer else 0
            pairs = [synth.random_dot_pair(h, w, seed=100 + k, noise=4 if k else 0, **gen)[:2] for k in range(4)]
            inflight = 5 if h * w * d < 3e8 else 3
            st = capi.Stream(h, w, 1, d, devices=(local,), inflight=inflight, algo=algo_id[algo], flags=flags, **over)
            out = np.empty((h, w), np.float32)

            def go(n, st=st, pairs=pairs, out=out):
                for f in range(n):
                    if st.pending == st.capacity:
                        st.next_into(out)
                    st.submit(*pairs[f % 4])
                while st.pending:
                    st.next_into(out)

            go(2 * inflight)
            t0 = time.perf_counter()
            go(frames)
            dt = time.perf_counter() - t0
            st.close()
            cls = {"sgm": capi.SGM, "bm": capi.BM, "bm_sad": capi.BMSad}[algo]
            probe = cls(h, w, 1, d, device=local, flags=flags | capi.FLAG_STAGE_TIMING, **over)
            stage, reps = {}, 6
            for k in range(reps + 2):
                probe.process(*pairs[k % 4])
                if k >= 2:
                    for sname, v in probe.stage_ms().items():
                        stage[sname] = stage.get(sname, 0.0) + v / reps
            line = {"config": name, "frames_per_s_e2e": frames / dt, "gdisp_px_per_s": frames / dt * h * w * d / 1e9, "frames": frames,
                    "inflight": inflight, "stage_ms_per_frame": {k: round(v, 4) for k, v in stage.items()},
                    "launches_per_frame": probe.launches, "device_mb_per_handle": round(probe.device_bytes / 1e6, 1)}
            if algo == "sgm":
                alg = (48.0 if integer else 96.0) * h * w * d
                ach = alg / (stage["aggregate"] / 1e3) / 1e9
                line["roofline"] = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                                    "algorithmic_bytes_per_frame": alg, "ms_per_frame": stage["aggregate"]}
            probe.close()
            lines.append(line)
        except Exception as e:  # a secondary line must never take the headline down
            lines.append({"config": name, "unavailable": f"{type(e).__name__}: {e}"})
    return lines


def blocking_solver_time(pairs):
    """Solver::process() + get_disp() of the C++ shell (stereo_matching_b200/host), one frame at a time, the way the
    reference's node.cpp:93,104 calls it: what a caller that does not pipeline frames sees.  Separate process, after the
    timed regions."""
    exe = os.path.join(ROOT, "stereo_matching_b200", "host", "_build", "host_shell_main")
    if not os.path.exists(exe):
        return {"unavailable": "stereo_matching_b200/host/_build/host_shell_main not built"}
    try:
        with tempfile.TemporaryDirectory() as tmp:
            l, r = os.path.join(tmp, "l.raw"), os.path.join(tmp, "r.raw")
            pairs[1][0].tofile(l)
            pairs[1][1].tofile(r)
            p = subprocess.run([exe, "bench", "sgm", str(H), str(W), str(SCALE), str(D), l, r, "200"], capture_output=True, text=True, timeout=180)
        last = [ln for ln in p.stderr.splitlines() if ln.startswith("{")]  # stdout carries the shells' per-frame prints
        if p.returncode != 0 or not last:
            return {"unavailable": f"rc {p.returncode}: {p.stderr.strip()[-160:]}"}
        j = json.loads(last[-1])
        j["what"] = "C++ shell: SolverPtr->process(l, r) + get_disp() per frame, blocking, one frame in flight (node.cpp:93,104)"
        return j
    except Exception as e:
        return {"unavailable": f"{type(e).__name__}: {e}"}


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
    """Per-frame DRAM bytes of the aggregation kernels: a constant read from the committed ncu --set full capture, not
    measured in this run (ncu counters cannot be read outside a profiler)."""
    path = os.path.join(ROOT, "profiles", "aggregate_traffic.json")
    try:
        with open(path) as f:
            j = json.load(f)
        return float(j["dram_bytes_per_frame"]), f"profiles/aggregate_traffic.json ({j.get('source', 'ncu --set full capture')}); constant, not measured in this run"
    except Exception:
        return None, "no committed ncu capture"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--inflight", type=int, default=6, help="frames in flight per GPU (engine handles); measured 3..8: profiles/r02_summary.md")
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the CPU baseline, the blocking-shell leg and the gpu_sgm baseline")
    ap.add_argument("--no-configs", action="store_true", help="skip the secondary `configs` lines")
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
        traffic, traffic_src = load_traffic() if args.mode == "reference" else (None, "no capture for this mode")
        line = {
            "metric": "sgm_frames_per_s", "value": fps, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": r["ms_value"] / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": MODES[args.mode]["dtype"], "data": "synthetic", "config": config_dict(world, args.mode),
            "frames_per_step": FRAMES_PER_STEP, "inflight_frames_per_gpu": r["inflight"],
            "gdisp_px_per_s": fps * n_px / 1e9,
            "e2e": {"value": fps_e2e, "unit": "frames/s", "h2d_bytes_per_step": 2 * H * W * FRAMES_PER_STEP,
                    "d2h_bytes_per_step": 4 * H * W * FRAMES_PER_STEP, "gdisp_px_per_s": fps_e2e * n_px / 1e9,
                    "api": "sgm_stream_submit / sgm_stream_next (C frame-stream runner, one host thread per rank), pageable host arrays in, host array out",
                    "timing": "CUDA events on an idle side stream around the region (every engine stream idle at both ends), max over ranks",
                    "wall_ms_rank0": r["wall_e2e"], "gpu_launches": r["launches_e2e"], "parity_first_frames": r["parity"]},
            "gpu_launches": r["launches"],
            "roofline": {"kernel": "aggregation stage = sgm_hpair_async_kernel + sgm_sweep_kernel<DOWN> + sgm_sweep_kernel<UP> (3 launches per frame, both views per launch; the up sweep ends in the fused sum/WTA/uniqueness/sub-pixel epilogue)",
                         "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "algorithmic_bytes_per_frame": alg_bytes,
                         "ms_per_frame": agg_ms, "timing": "engine-side CUDA events on the launching stream, frames run serially"},
            "stage_ms_per_frame": r["stage_ms"],
            "clocks": r["clocks"],
            "device_bytes_per_engine": r["device_bytes"],
        }
        extras = world == 1
        if extras and not args.no_configs:
            line["configs"] = run_other_configs(local, peak)
        if extras and not args.no_cpu_baseline:
            line["blocking"] = blocking_solver_time(make_pairs())
            kind, cores, times, respawns = cpu_reference_frames(2, 1)
            if kind is None:
                line["cpu_baseline"] = {"value": None, "unit": "frames/s", "cores": cores, "kind": "reference", "sample": f"unavailable: {times}"}
            else:
                cfps = len(times) / (sum(times) / 1e3)
                line["cpu_baseline"] = {"value": cfps, "unit": "frames/s", "cores": cores, "kind": kind, "child_respawns": respawns,
                                        "sample": f"{len(times)} frames of the same 1242x375 D=128 pair cycle after 1 warm-up, SGM::process (src/SGM.cpp:32-826), OpenMP on all host threads"}
            line["gpu_sgm_reference"] = gpu_sgm_reference_time()
        else:
            line["cpu_baseline"] = None
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
