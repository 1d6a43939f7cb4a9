"""CPU tests of the drop-in boundary (-m "not gpu"): the C-ABI library builds, loads and exports every
symbol include/sgm_b200.h declares; argument validation returns error codes; without a CUDA device the
product fails loudly (there is no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from stereo_matching_b200 import build as sgm_build
from stereo_matching_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    sgm_build.build()
    return capi.lib()


def _declared():
    src = open(os.path.join(ROOT, "include", "sgm_b200.h")).read()
    return sorted(set(re.findall(r"SGM_API[^;(]*?\b(sgm_\w+)\s*\(", src)))


def test_every_declared_symbol_is_exported(lib):
    names = _declared()
    assert len(names) >= 20 and sorted(capi.EXPORTS) == names
    for n in names:
        assert hasattr(lib, n), n
    assert lib.sgm_abi_version() == 1


def test_default_params_are_the_reference_constants(lib):
    p = capi.default_params(375, 1242, 1, 128)
    # inc/Solver.h:10-20, inc/SGM.h:7, src/SGM.cpp:27-28
    assert (p.census_h, p.census_w, p.box_h, p.box_w) == (7, 9, 3, 5)
    assert (p.P1, p.P2, p.paths) == (10, 100, 8)
    assert abs(p.unique_ratio - 0.7) < 1e-7 and p.lr_thresh == 1.0
    assert (p.median_h, p.median_w, p.speckle_size, p.speckle_dis) == (5, 5, 1000, 2)


def test_params_struct_layout_matches_header(lib):
    # 21 int32/float fields + flags
    assert C.sizeof(capi.Params) == 22 * 4


def test_validation_errors_before_touching_the_device(lib):
    for kw, code in ((dict(h=0, w=10, s=1, d=64), 1), (dict(h=100, w=100, s=3, d=64), 1), (dict(h=100, w=100, s=1, d=48), 2),
                     (dict(h=8, w=100, s=1, d=32), 1)):
        p = capi.default_params(kw["h"], kw["w"], kw["s"], kw["d"])
        handle = C.c_void_p()
        assert lib.sgm_create(C.byref(handle), C.byref(p)) == code
        assert handle.value is None and lib.sgm_last_error(None)
    assert lib.sgm_create(None, None) == 1


def test_integer_mode_parameter_checks_before_touching_the_device(lib):
    # SGM_FLAG_INTEGER is only accepted where every number on the path is an integer below 2^15:
    # a cost filter, penalties that overflow 16 bits, P1 > P2, D = 32 or a BM algorithm are refused with SGM_ERR_UNSUPPORTED
    bad = (dict(), dict(box_w=1, box_h=1, P2=5000), dict(box_w=1, box_h=1, P1=20, P2=10), dict(box_w=1, box_h=3),
           dict(box_w=1, box_h=1, flags=capi.FLAG_INTEGER | capi.FLAG_LINE_WALKERS))
    for over in bad:
        p = capi.default_params(64, 64, 1, 64, flags=capi.FLAG_INTEGER)
        for k, v in over.items():
            setattr(p, k, v)
        handle = C.c_void_p()
        assert lib.sgm_create(C.byref(handle), C.byref(p)) == 2, over
        assert handle.value is None and b"SGM_FLAG_INTEGER" in lib.sgm_last_error(None)
    p = capi.default_params(64, 64, 1, 32, flags=capi.FLAG_INTEGER, box_w=1, box_h=1)
    assert lib.sgm_create(C.byref(C.c_void_p()), C.byref(p)) == 2
    p = capi.default_params(64, 64, 1, 64, algo=capi.ALGO_BM, flags=capi.FLAG_INTEGER, box_w=1, box_h=1)
    assert lib.sgm_create(C.byref(C.c_void_p()), C.byref(p)) == 2


def test_no_cpu_fallback_without_a_device(lib):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(capi.SgmError) as e:
        capi.SGM(64, 64, 1, 32)
    assert e.value.code == 3 and "no CPU path" in str(e.value)
    # the frame-stream runner is built on the same handles: same answer, and its argument checks come before any device call
    with pytest.raises(capi.SgmError) as e:
        capi.Stream(64, 64, 1, 32, devices=(0,), inflight=2)
    assert e.value.code == 3 and "no CPU path" in str(e.value)
    with pytest.raises(capi.SgmError) as e:
        capi.Stream(64, 64, 1, 32, devices=(0,), inflight=0)
    assert e.value.code == 1
    # the allocation-free speckle entry point: size limit and argument checks need no device either
    with pytest.raises(capi.SgmError) as e:
        capi.speckle_filter_static(1, 1, 2048, 1025, 65.0)
    assert e.value.code == 2
    with pytest.raises(capi.SgmError) as e:
        capi.speckle_filter_static(1, 1, 64, 64, 65.0)
    assert e.value.code == 3


def test_too_wide_rows_are_rejected_before_touching_the_device(lib):
    # the Hamming kernels stage both census rows of an image row in shared memory: a row that does not fit is an unsupported
    # shape (error code), not a slower path
    with pytest.raises(capi.SgmError) as e:
        capi.SGM(64, 14000, 1, 128)
    assert e.value.code == 2 and "too wide" in str(e.value)
