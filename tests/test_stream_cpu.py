"""Host logic of the multi-GPU frame stream (stereo_matching_b200/stream.py), without a GPU: sharding arithmetic,
in-flight bookkeeping with a stand-in engine, and a world_size-2 gloo run that gathers and re-orders the ranks' results
(the N > 1 path of bench.py uses the same shard function; there is no data-path collective to test)."""
import os
import socket

import numpy as np
import pytest

from stereo_matching_b200 import stream

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class FakeEngine:
    """Stand-in for a capi.SGM handle: 'disparity' = a checksum image of the inputs, available only after wait()."""

    def __init__(self, log):
        self.busy, self.result, self.log = False, None, log

    def enqueue(self, L, R, *sky):
        assert not self.busy, "handle re-used before its frame was collected"
        self.busy = True
        self._next = (L.astype(np.float32) - R.astype(np.float32)) * 0.5
        self.log.append("enqueue")

    def wait(self):
        self.result, self.busy = self._next, False
        self.log.append("wait")

    def get_disp(self):
        assert not self.busy and self.result is not None
        return self.result.copy()


def _load(k):
    rng = np.random.default_rng(k)
    return rng.integers(0, 255, (4, 6), dtype=np.uint8), rng.integers(0, 255, (4, 6), dtype=np.uint8)


def _expected(k):
    L, R = _load(k)
    return (L.astype(np.float32) - R.astype(np.float32)) * 0.5


@pytest.mark.parametrize("n,world", [(0, 1), (1, 4), (7, 2), (4096, 8), (4097, 8), (13, 5)])
def test_shards_partition_the_stream(n, world):
    seen = sorted(k for r in range(world) for k in stream.shard_frames(n, r, world))
    assert seen == list(range(n))
    sizes = stream.shard_sizes(n, world)
    assert sum(sizes) == n and max(sizes) - min(sizes) <= 1


def test_bad_shard_arguments():
    for args in ((4, 2, 2), (4, -1, 2), (4, 0, 0), (-1, 0, 1)):
        with pytest.raises(ValueError):
            stream.shard_frames(*args)


@pytest.mark.parametrize("inflight", [1, 2, 3])
def test_stream_keeps_frames_in_flight_and_in_order(inflight):
    log = []
    engines = [FakeEngine(log) for _ in range(inflight)]
    out = list(stream.FrameStream(engines).run((k, _load(k)) for k in range(10)))
    assert [k for k, _ in out] == list(range(10))
    for k, d in out:
        assert np.array_equal(d, _expected(k))
    # the first `inflight` submissions happen before any wait: that is the overlap bench.py relies on
    assert log[:inflight] == ["enqueue"] * inflight and log.count("enqueue") == log.count("wait") == 10


def test_merge_detects_gaps():
    a = [(0, "x"), (2, "y")]
    with pytest.raises(RuntimeError):
        stream.merge_in_order([a, [(3, "z")]])
    assert [k for k, _ in stream.merge_in_order([a, [(1, "w"), (3, "z")]])] == [0, 1, 2, 3]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_frames, q):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    part = stream.run_sharded(n_frames, rank, world, _load, [FakeEngine([]) for _ in range(2)])
    gathered = [None] * world
    dist.all_gather_object(gathered, [(k, d.tolist()) for k, d in part])  # control plane only: results normally stay per rank
    dist.barrier()
    if rank == 0:
        q.put(gathered)
    dist.destroy_process_group()


def test_two_ranks_gloo_cover_the_stream_exactly_once():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port, n = _free_port(), 11
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    gathered = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    merged = stream.merge_in_order(gathered)
    assert len(merged) == n and [len(g) for g in gathered] == stream.shard_sizes(n, 2)
    for k, d in merged:
        assert np.array_equal(np.asarray(d, np.float32), _expected(k))


def test_c_stream_runner_host_logic_against_mock_engine(tmp_path):
    """The C frame-stream runner (stereo_matching_b200/csrc/stream_runner.cu: slots, back-pressure, in-order delivery, held views,
    device lanes, failing frames) compiled as plain C++ against a MOCK engine (tests/cpp/stream_runner_mock.cpp): the N>1 host
    logic without a GPU, with a producer and a consumer thread and with one thread alone."""
    import subprocess

    exe = str(tmp_path / "stream_mock")
    env = {k: v for k, v in os.environ.items() if k not in ("CXX", "CC")}
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-pthread", "-I/usr/local/cuda/include", "-I" + os.path.join(ROOT, "include"), "-x", "c++",
                           os.path.join(ROOT, "stereo_matching_b200", "csrc", "stream_runner.cu"),
                           os.path.join(ROOT, "tests", "cpp", "stream_runner_mock.cpp"), "-o", exe], env=env)
    out = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and "all checks passed" in out.stdout, out.stdout + out.stderr
