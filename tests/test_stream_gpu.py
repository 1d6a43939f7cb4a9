"""GPU tests of the C frame-stream runner (sgm_stream_*, stereo_matching_b200/csrc/stream_runner.cu): the loop of the
reference's caller (node.cpp:49,93,104: one Solver, process() per frame) over several frames in flight and several device
lanes.  Checker: the restatement (oracle) on the same pairs; results must come back in submission order, bit for bit.
A box with one GPU still exercises the round-robin over lanes: the same ordinal can be listed twice."""
import threading

import numpy as np
import pytest

from oracle import restate
from stereo_matching_b200 import capi, synth
from tests.common import bits_equal

pytestmark = pytest.mark.gpu

H_, W_, D_ = 96, 200, 64


@pytest.fixture(scope="module")
def pairs_and_truth():
    pairs = [synth.random_dot_pair(H_, W_, seed=40 + k, bg=4, fg=15, noise=10 + 10 * k)[:2] for k in range(3)]
    truth = [restate.process(L, R, 1, D_, "sgm")["f_final"] for L, R in pairs]
    return pairs, truth


@pytest.mark.parametrize("devices,inflight", [((0,), 1), ((0,), 3), ((0, 0), 2), ((0, 0, 0), 1)])
def test_stream_in_order_and_bit_exact(pairs_and_truth, devices, inflight):
    pairs, truth = pairs_and_truth
    st = capi.Stream(H_, W_, 1, D_, devices=devices, inflight=inflight)
    n, got = 23, []
    for k in range(n):                       # one thread: collect when every slot is taken
        if st.pending == st.capacity:
            got.append(st.next())
        assert st.submit(*pairs[k % 3]) == k
    while st.pending:
        got.append(st.next())
    assert [k for k, _ in got] == list(range(n))
    for k, f in got:
        assert bits_equal(f, truth[k % 3]), f"frame {k} differs from the oracle"
    assert st.launches >= n * 15
    with pytest.raises(capi.SgmError) as e:
        st.next()                            # nothing pending
    assert e.value.code == 5
    st.close()


def test_stream_producer_and_consumer_threads(pairs_and_truth):
    pairs, truth = pairs_and_truth
    st = capi.Stream(H_, W_, 1, D_, devices=(0, 0), inflight=2)
    n, bad = 40, []

    def consume():
        out = np.empty((H_, W_), np.float32)
        for want in range(n):
            while st.pending == 0:           # the producer may not have submitted yet
                pass
            k = st.next_into(out)
            if k != want or not bits_equal(out, truth[k % 3]):
                bad.append(k)

    t = threading.Thread(target=consume)
    t.start()
    for k in range(n):
        st.submit(*pairs[k % 3])             # blocks for back-pressure only
    t.join(timeout=120)
    assert not t.is_alive() and not bad, bad
    st.close()


def test_stream_zero_copy_view_and_masks():
    h, w, d = 120, 200, 64
    L, R, _ = synth.random_dot_pair(h, w, seed=9, bg=5, fg=17, noise=20)
    sl, sr = synth.sky_masks(h, w)
    want_sky = restate.process(L, R, 1, d, "sgm", sl, sr)["f_final"]
    want = restate.process(L, R, 1, d, "sgm")["f_final"]
    st = capi.Stream(h, w, 1, d, devices=(0,), inflight=2)
    st.submit(L, R, sl, sr)
    st.submit(L, R)                          # a frame without masks on the other handle
    k0, v0 = st.next_view()
    assert k0 == 0 and bits_equal(np.array(v0), want_sky)
    k1, v1 = st.next_view()                  # v0 is dead from here on
    assert k1 == 1 and bits_equal(np.array(v1), want)
    # a held view occupies its slot: it counts as pending (so a one-thread caller collects before it would wait for itself)
    # until it is released
    assert st.pending == 1
    st.submit(L, R)                          # frame 2 -> the other slot
    assert st.pending == st.capacity == 2
    st.release()
    assert st.pending == 1
    st.submit(L, R)                          # frame 3 -> the slot that was held
    assert st.next()[0] == 2 and st.next()[0] == 3 and st.pending == 0
    st.close()


def test_stream_other_algorithms_and_bad_arguments():
    h, w, d = 90, 250, 64
    L, R, _ = synth.random_dot_pair(h, w, seed=273 + d - 64 + 61, bg=4, fg=12, noise=30)
    want = restate.process(L, R, 1, d, "bm_sad")["f_final"]
    st = capi.Stream(h, w, 1, d, devices=(0,), inflight=3, algo=capi.ALGO_BM_SAD)
    for _ in range(3):                       # no more than `capacity` frames without collecting: submit would wait for a slot
        st.submit(L, R)
    assert st.pending == st.capacity == 3
    for k in range(3):
        kk, f = st.next()
        assert kk == k and bits_equal(f, want)
    st.close()
    with pytest.raises(capi.SgmError):
        capi.Stream(h, w, 1, d, devices=(0,), inflight=0)
    with pytest.raises(capi.SgmError):
        capi.Stream(h, w, 1, 48, devices=(0,), inflight=1)      # sgm_create's own validation comes through
    with pytest.raises(capi.SgmError):
        capi.Stream(h, w, 1, d, devices=(99,), inflight=1)


def test_cpp_stream_driver_matches_golden_and_uses_every_gpu():
    # BASELINE.json configs[4] in miniature: the C++ driver's `stream` mode (one process, producer + consumer threads,
    # sgm_stream over every GPU of the box) must return the reference's golden results for the bench pairs, in order
    import json
    import os
    import subprocess
    import sys

    torch = pytest.importorskip("torch")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    subprocess.check_call(["make", "-s", "-C", os.path.join(root, "stereo_matching_b200", "host"), "_build/host_shell_main"],
                          env={k: v for k, v in os.environ.items() if k not in ("CXX", "CC")})
    n = min(torch.cuda.device_count(), 8)
    out = subprocess.run([sys.executable, os.path.join(root, "tests", "stream_config4.py"), str(n), "96", "3"], capture_output=True, text=True, timeout=150)
    assert out.returncode == 0, out.stdout + out.stderr
    j = json.loads(out.stdout.strip().splitlines()[-1])
    assert j["parity_first_8"] and j["n_gpus"] == n and j["frames"] == 96 and j["launches"] > 96 * 15
