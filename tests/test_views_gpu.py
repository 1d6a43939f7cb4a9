"""GPU parity of the result consumers: the device-rendered debug view against
Solver::show_disp (restated in oracle/restate.py and pinned against the compiled reference on the CPU side) and the
device point cloud against the restatement of node.cpp:119-143.  Bit-exact, including the double-precision coordinates
and the raster order of the points."""
import numpy as np
import pytest

from oracle import restate
from stereo_matching_b200 import capi, synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("cls,algo,h,w,s,d", [(capi.SGM, "sgm", 96, 192, 1, 64), (capi.SGM, "sgm", 120, 200, 2, 32),
                                              (capi.BM, "bm", 96, 160, 1, 32), (capi.BM, "bm", 120, 200, 2, 64),
                                              (capi.SGM, "sgm", 72, 300, 1, 256)])
def test_debug_view_matches_show_disp(cls, algo, h, w, s, d):
    L, R, _ = synth.random_dot_pair(h, w, seed=h + d, bg=4, fg=min(19, d // 2), noise=14)
    sv = cls(h, w, s, d)
    f = sv.process(L, R).get_disp()
    want = restate.debug_view(L, f, d, s, bm_rows=(algo == "bm"))
    got = sv.debug_view()
    assert got.shape == want.shape and np.array_equal(got, want)
    assert got[h // s:].any() and not got[-1].any()
    sv.close()


@pytest.mark.parametrize("h,w,s,d,cam", [(96, 192, 1, 64, (721.5, 721.5, 96.0, 48.0)), (120, 200, 2, 32, (350.25, 349.75, 99.5, 60.25)),
                                         (375, 1242, 1, 128, (718.856, 718.856, 607.1928, 185.2157))])
def test_point_cloud_matches_node_loop(h, w, s, d, cam):
    L, R, _ = synth.random_dot_pair(h, w, seed=w + d, bg=5, fg=min(21, d // 2), noise=10)
    sv = capi.SGM(h, w, s, d)
    f = sv.process(L, R).get_disp()
    # small max_range so that the range cut really drops points
    for max_range in (100.0, 25.0):
        want_xyz, want_i = restate.point_cloud(f, L, d, s, *cam, baseline=0.5, max_range=max_range)
        xyz, inten = sv.point_cloud(*cam, baseline=0.5, max_range=max_range)
        assert xyz.shape == want_xyz.shape and len(want_xyz) > 100
        assert np.array_equal(xyz.view(np.uint64), want_xyz.view(np.uint64)), "coordinates differ (double, bit-exact)"
        assert np.array_equal(inten, want_i)
    sv.close()


def test_point_cloud_small_capacity_reports_full_count():
    import ctypes as C

    h, w, d = 64, 128, 32
    L, R, _ = synth.random_dot_pair(h, w, seed=2, bg=4, fg=11)
    sv = capi.SGM(h, w, 1, d)
    sv.process(L, R)
    full, _ = sv.point_cloud(500.0, 500.0, 64.0, 32.0)
    cam = capi.Camera(500.0, 500.0, 64.0, 32.0, 0.5, 100.0)
    xyz = np.zeros((10, 3), np.float64)
    n = C.c_size_t()
    rc = capi.lib().sgm_point_cloud(sv._h, C.byref(cam), xyz.ctypes.data_as(C.POINTER(C.c_double)), None, 10, C.byref(n))
    assert rc == 0 and n.value == len(full) and np.array_equal(xyz, full[:10])
    sv.close()
