"""GPU test of sgm_speckle_filter_static (include/sgm_b200.h): the speckle stage with its label / area scratch in a static
__device__ array inside the library, for callers that must not allocate.  No counterpart in the reference;
the checker is the EXISTING entry point: sgm_process' own speckle stage, whose input and
output are the SGM_TAP_MEDIAN / SGM_TAP_FINAL taps (themselves bit-exact against the oracle in tests/test_parity_gpu.py).

Collected last on purpose (a failure here must not hide other tests).  Tolerance: bit-exact, like
every other tap.  torch only holds the caller-side device buffers here; the entry point itself allocates nothing.
"""
import numpy as np
import pytest

from stereo_matching_b200 import capi, synth
from tests.common import bits_equal

pytestmark = pytest.mark.gpu


def _frame(h, w, d, seed, noise):
    """(median tap = input of the speckle stage, final = its output) of one SGM frame through the existing entry point."""
    L, R, _ = synth.random_dot_pair(h, w, seed=seed, bg=5, fg=min(40, d // 2), noise=noise)
    sv = capi.SGM(h, w, 1, d)
    sv.process(L, R)
    med, fin = sv.tap("median"), sv.get_disp()
    sv.close()
    assert not bits_equal(med, fin), "the frame must have speckles to remove, or the check proves nothing"
    return med, fin


def test_static_scratch_speckle_matches_engine_on_consecutive_calls():
    torch = pytest.importorskip("torch")
    dev = torch.device("cuda", 0)
    # two frames of different size: the second call finds the first call's labels / areas in the shared array
    a_med, a_fin = _frame(375, 1242, 128, seed=77, noise=10)
    b_med, b_fin = _frame(200, 320, 64, seed=5, noise=40)
    a_in, b_in = torch.from_numpy(a_med).to(dev), torch.from_numpy(b_med).to(dev)
    a_out, b_out = torch.empty_like(a_in), torch.empty_like(b_in)
    torch.cuda.synchronize(dev)

    def run(t_in, t_out, d):
        h, w = t_in.shape
        capi.speckle_filter_static(t_in.data_ptr(), t_out.data_ptr(), h, w, float(d + 1), 1000, 2.0, device=0, stream=None)

    run(a_in, a_out, 128)                      # call 1 (also the per-device set-up)
    torch.cuda.synchronize(dev)
    free_before, _ = torch.cuda.mem_get_info(dev)
    run(b_in, b_out, 64)                       # call 2, straight after, other size
    run(a_in, a_out, 128)                      # call 3: the larger frame again, over the smaller one's leftovers
    torch.cuda.synchronize(dev)
    free_after, _ = torch.cuda.mem_get_info(dev)
    assert bits_equal(b_out.cpu().numpy(), b_fin), "call 2 differs from the engine's speckle stage"
    assert bits_equal(a_out.cpu().numpy(), a_fin), "call 3 differs from the engine's speckle stage"
    assert free_after == free_before, "the entry point must not allocate device memory"
    # in place (in == out) is allowed
    run(b_in, b_in, 64)
    torch.cuda.synchronize(dev)
    assert bits_equal(b_in.cpu().numpy(), b_fin)


def test_static_scratch_speckle_first_call_in_process_and_limits():
    torch = pytest.importorskip("torch")
    dev = torch.device("cuda", 0)
    med, fin = _frame(96, 192, 64, seed=3, noise=30)
    t_in = torch.from_numpy(med).to(dev)
    t_out = torch.empty_like(t_in)
    torch.cuda.synchronize(dev)
    for _ in range(2):  # two calls in a row, same buffers
        capi.speckle_filter_static(t_in.data_ptr(), t_out.data_ptr(), 96, 192, 65.0, 1000, 2.0)
        torch.cuda.synchronize(dev)
        assert bits_equal(t_out.cpu().numpy(), fin)
    # beyond the static array: a clear error, nothing launched
    with pytest.raises(capi.SgmError) as e:
        capi.speckle_filter_static(t_in.data_ptr(), t_out.data_ptr(), 2048, 1025, 65.0)
    assert e.value.code == 2 and "exceed the static scratch" in str(e.value)
    with pytest.raises(capi.SgmError) as e:
        capi.speckle_filter_static(None, t_out.data_ptr(), 96, 192, 65.0)
    assert e.value.code == 1
    # exactly at the limit is accepted (2048 x 1024 pixels): a flat image is one big component, nothing is removed
    big = torch.full((1024, 2048), 7.0, device=dev)
    out = torch.empty_like(big)
    torch.cuda.synchronize(dev)
    capi.speckle_filter_static(big.data_ptr(), out.data_ptr(), 1024, 2048, 65.0)
    torch.cuda.synchronize(dev)
    assert bool((out == 7.0).all())
    assert capi.SPECKLE_STATIC_MAX_PIXELS == 1024 * 2048
