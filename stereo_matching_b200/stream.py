"""Frame-stream runner: shards a stream of independent stereo pairs over ranks (one process per GPU) and keeps several
frames in flight per GPU.

The reference processes one pair per `process()` call and keeps no state between frames, so
the only multi-GPU structure on this path is "frame k goes to rank k mod world": no data-path collective, no peer
traffic; results come back to the host over PCIe and are re-ordered by frame index.  This module holds that host
logic so it can be tested without a GPU (tests/test_stream_cpu.py, world_size-2 gloo) with a stand-in engine.
"""
from __future__ import annotations

from typing import Callable, Iterable, Iterator, List, Sequence, Tuple


def shard_frames(n_frames: int, rank: int, world: int) -> range:
    """Frame indices owned by `rank`: round-robin, so every rank gets floor or ceil of n_frames / world frames and a
    live stream (unknown length) can be dealt out frame by frame."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError(f"bad rank {rank} / world {world}")
    if n_frames < 0:
        raise ValueError("n_frames must be >= 0")
    return range(rank, n_frames, world)


def shard_sizes(n_frames: int, world: int) -> List[int]:
    return [len(shard_frames(n_frames, r, world)) for r in range(world)]


class FrameStream:
    """Runs frames through `engines` (objects with enqueue(L, R[, sky_l, sky_r]) / wait() / get_disp(), i.e. capi.SGM
    handles of ONE GPU) keeping len(engines) frames in flight; yields (frame_index, disparity) in submission order.

    The engine API is the C ABI's asynchronous pair (include/sgm_b200.h: sgm_enqueue / sgm_wait): enqueue copies the
    inputs into the handle's pinned staging and queues H2D + kernels + D2H on the handle's stream; wait blocks until
    that frame is done.  A handle is re-used only after its previous frame has been collected."""

    def __init__(self, engines: Sequence):
        if not engines:
            raise ValueError("need at least one engine handle")
        self.engines = list(engines)

    def run(self, frames: Iterable[Tuple[int, tuple]]) -> Iterator[Tuple[int, object]]:
        pending: List[Tuple[int, object]] = []  # (frame index, engine) in submission order
        free = list(self.engines)
        for index, inputs in frames:
            if not free:
                done_index, eng = pending.pop(0)
                eng.wait()
                yield done_index, eng.get_disp()
                free.append(eng)
            eng = free.pop(0)
            eng.enqueue(*inputs)
            pending.append((index, eng))
        for done_index, eng in pending:
            eng.wait()
            yield done_index, eng.get_disp()


def run_sharded(n_frames: int, rank: int, world: int, load: Callable[[int], tuple], engines: Sequence) -> List[Tuple[int, object]]:
    """This rank's share of an n_frames stream: [(frame index, disparity)], ascending frame index."""
    mine = shard_frames(n_frames, rank, world)
    return list(FrameStream(engines).run((k, load(k)) for k in mine))


def merge_in_order(per_rank: Sequence[Sequence[Tuple[int, object]]]) -> List[Tuple[int, object]]:
    """Re-orders the ranks' results by frame index (the host-side reorder queue of SURVEY.md 8(e)) and checks that
    every frame was produced exactly once."""
    merged = sorted((item for part in per_rank for item in part), key=lambda t: t[0])
    for want, (got, _) in enumerate(merged):
        if want != got:
            raise RuntimeError(f"frame stream has a gap or duplicate at frame {want} (saw {got})")
    return merged
