"""The C++ host shells (stereo_matching_b200/host: Solver / SGM / BM / cost.h with the reference's class surface).

CPU (-m "not gpu"):
  * the shells build against the stand-in OpenCV headers and export the classes node.cpp uses;
  * where /root/reference is mounted, the reference's UNMODIFIED node.cpp compiles against our inc/*.h (the reference
    tree is overlaid with symlinks exactly like a maintainer would drop the files in; ROS is stubbed);
  * the per-pixel probes of inc/cost.h agree with the oracle's dense volumes.
GPU (-m gpu): a node.cpp-style driver (tests/cpp/host_shell_main.cpp) runs frames through SolverPtr -> process ->
get_disp / show_disp and must reproduce the oracle bit for bit.
"""
import os
import shutil
import subprocess
import tempfile

import numpy as np
import pytest

from oracle import restate
from stereo_matching_b200 import synth
from tests.common import bits_equal

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "stereo_matching_b200", "host")
REF = "/root/reference"


def _env():
    env = dict(os.environ)
    env.pop("CXX", None)
    env.pop("CC", None)
    return env


def _make(target):
    subprocess.check_call(["make", "-s", "-C", HOST, target], env=_env())
    return os.path.join(HOST, target)


def test_shells_build_and_export_the_reference_classes():
    lib = _make("_build/libstereo_solver_b200.so")
    syms = subprocess.check_output(["nm", "-DC", "--defined-only", lib], text=True)
    for want in ("SGM::SGM(int, int, int, int)", "BM::BM(int, int, int, int)", "Solver::show_disp(cv::Mat&)",
                 "SGM::process(cv::Mat&, cv::Mat&, cv::Mat&, cv::Mat&)", "BM::process(cv::Mat&, cv::Mat&)",
                 "SAD(cv::Mat const&", "SSD(cv::Mat const&", "CT(cv::Mat const&", "CT_pts(cv::Mat const&", "hamming_cost("):
        assert want in syms, f"{want} not exported by the host shell library"
    # the shell must not carry a CPU matching path: everything goes through the C ABI of libsgm_b200.so
    undefined = subprocess.check_output(["nm", "-DC", "--undefined-only", lib], text=True)
    for want in ("sgm_create", "sgm_process", "sgm_get_disparity_f32", "sgm_destroy"):
        assert want in undefined


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "inc")), reason="reference tree not mounted")
def test_reference_node_cpp_compiles_unchanged_against_our_headers():
    with tempfile.TemporaryDirectory() as t:
        os.makedirs(os.path.join(t, "inc"))
        os.symlink(os.path.join(REF, "node.cpp"), os.path.join(t, "node.cpp"))
        for f in ("global.h", "utils.h", "roshelper.h"):          # the reference's own files, untouched
            os.symlink(os.path.join(REF, "inc", f), os.path.join(t, "inc", f))
        for f in ("Solver.h", "SGM.h", "BM.h", "cost.h"):          # ours, dropped in under the same names
            os.symlink(os.path.join(HOST, "inc", f), os.path.join(t, "inc", f))
        os.symlink(os.path.join(REF, "sky_detector"), os.path.join(t, "sky_detector"))
        cmd = ["g++", "-std=c++11", "-fsyntax-only", "-fopenmp", "-w", "-I", os.path.join(ROOT, "tests", "cpp", "ros_stubs"),
               "-I", os.path.join(ROOT, "oracle", "shim"), "-I", os.path.join(ROOT, "include"), "-I", t, os.path.join(t, "node.cpp")]
        r = subprocess.run(cmd, capture_output=True, text=True, env=_env())
        assert r.returncode == 0, r.stderr[-2000:]


def test_division_by_window_size_is_exact():
    # the cost filters divide by 5 / 3 (and 2 / 1 at scale 2) with a 4-instruction FMA sequence: must equal x / b
    exe = os.path.join(tempfile.gettempdir(), "verify_div_const")
    subprocess.check_call(["gcc", "-O2", "-fopenmp", "-ffp-contract=off", "-o", exe, os.path.join(ROOT, "tests", "cpp", "verify_div_const.c"),
                           "-lm"], env=_env())
    out = subprocess.run([exe, "61"], capture_output=True, text=True)   # every 61st float (70 M values), ~2 s; stride 1 = all
    assert out.returncode == 0, out.stdout


def _run_driver(tmp, algo, h, w, s, d, L, R, sky, env=None, want_probe=None):
    exe = _make("_build/host_shell_main")
    L.tofile(os.path.join(tmp, "l.raw"))
    R.tofile(os.path.join(tmp, "r.raw"))
    sl = sr = "-"
    if sky is not None:
        sl, sr = os.path.join(tmp, "sl.raw"), os.path.join(tmp, "sr.raw")
        sky[0].tofile(sl)
        sky[1].tofile(sr)
    prefix = os.path.join(tmp, "out")
    r = subprocess.run([exe, algo, str(h), str(w), str(s), str(d), os.path.join(tmp, "l.raw"), os.path.join(tmp, "r.raw"), sl, sr, prefix],
                       capture_output=True, text=True, env=dict(os.environ, **(env or {})))
    assert r.returncode == 0, r.stdout + r.stderr
    ih, iw = h // s, w // s
    disp = np.fromfile(prefix + "_disp.f32", np.float32).reshape(ih, iw)
    disp2 = np.fromfile(prefix + "_disp2.f32", np.float32).reshape(ih, iw)
    dbg = np.fromfile(prefix + "_debug.bgr", np.uint8).reshape(2 * ih, iw, 3)
    cloud = np.fromfile(prefix + "_cloud.f64", np.float64).reshape(-1, 3)
    if algo == "sgm" and want_probe is not None:  # the protected surface, dumped by a subclass (ProbeSGM in the driver)
        want_probe.update(
            ctl=np.fromfile(prefix + "_ctl.u64", np.uint64).reshape(ih, iw),
            cost_l=np.fromfile(prefix + "_cost_l.f32", np.float32).reshape(ih, iw, d),
            cost_r=np.fromfile(prefix + "_cost_r.f32", np.float32).reshape(ih, iw, d),
            sub_r=np.fromfile(prefix + "_sub_r.f32", np.float32).reshape(ih, iw),
            disp8=np.fromfile(prefix + "_disp8.u8", np.uint8).reshape(ih, iw),
            fbeta=np.fromfile(prefix + "_fbeta.f32", np.float32).reshape(ih, iw))
    return disp, disp2, dbg, cloud, r.stdout


def _colormap(f, d):
    """Restatement of Solver::colormap (src/Solver.cpp:652-707) for the debug-view check."""
    out = np.zeros(f.shape + (3,), np.uint8)
    g = np.float32(256 // d)
    for (i, j), v in np.ndenumerate(f):
        if v > d - 1:
            continue
        v = np.float32(v) * g
        if v <= 51:
            bgr = (255, int(np.float32(v * np.float32(5))), 0)
        elif v <= 102:
            bgr = (int(np.float32(255) - np.float32(v - np.float32(51)) * np.float32(5)), 255, 0)
        elif v <= 153:
            bgr = (0, 255, int(np.float32(v - np.float32(102)) * np.float32(5)))
        elif v <= 204:
            bgr = (0, 255 - int(128.0 * float(np.float32(v - np.float32(153))) / 51.0 + 0.5), 255)
        else:
            bgr = (0, 127 - int(127.0 * float(np.float32(v - np.float32(204))) / 51.0 + 0.5), 255)
        out[i, j] = [c & 255 for c in bgr]
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("algo,h,w,s,d,sky", [("sgm", 96, 192, 1, 64, False), ("sgm", 120, 200, 2, 64, True), ("bm", 96, 160, 1, 32, False),
                                              ("bm", 120, 200, 2, 64, False)])
def test_node_style_driver_matches_oracle(algo, h, w, s, d, sky):
    L, R, _ = synth.random_dot_pair(h, w, seed=h + d, bg=5, fg=min(17, d // 2), noise=15)
    masks = synth.sky_masks(h // s, w // s) if sky else None
    taps = restate.process(L, R, s, d, algo, *(masks or (None, None)), taps=restate.ALL_TAPS)
    want = taps["f_final"]
    probe = {}
    with tempfile.TemporaryDirectory() as tmp:
        disp, disp2, dbg, cloud, log = _run_driver(tmp, algo, h, w, s, d, L, R, masks, want_probe=probe)
    if algo == "sgm":
        # the protected members of inc/Solver.h:43-69 as a subclass sees them: census table, filtered cost volumes of both views
        # (build_dsi_from_table / _beta), compute_subpixel(disp_beta), the CV_8UC1 disp and filtered_disp_beta
        assert bits_equal(probe["ctl"], taps["census"][0])
        assert bits_equal(probe["cost_l"], taps["cost_hv"][0]) and bits_equal(probe["cost_r"], taps["cost_hv"][1])
        assert bits_equal(probe["sub_r"], taps["fsub"][1]) and bits_equal(probe["fbeta"], taps["fsub"][1])
        assert np.array_equal(probe["disp8"], np.minimum(taps["disp"][0], 255).astype(np.uint8))
    assert bits_equal(disp, want), "get_disp() of the C++ shell differs from the oracle"
    assert bits_equal(disp2, want), "second frame on the same solver differs"
    assert "time cost" in log
    ih, iw = h // s, w // s
    # debug view: decimated left image on top, colour map below starting at row ih - 1 (src/Solver.cpp:85-92)
    row_step = 1 if (algo == "bm" or s == 1) else s
    gray = L[::row_step][:ih, ::s][:, :iw]
    assert np.array_equal(dbg[: ih - 1, :, 0], gray[: ih - 1]) and np.array_equal(dbg[: ih - 1, :, 2], gray[: ih - 1])
    assert np.array_equal(dbg[ih - 1: 2 * ih - 1], _colormap(want, d)) and np.array_equal(dbg, restate.debug_view(L, want, d, s, bm_rows=(algo == "bm")))
    assert not dbg[2 * ih - 1].any()
    # get_point_cloud(): node.cpp:119-143 on the device
    want_xyz, _ = restate.point_cloud(want, L, d, s, 718.856, 718.856, 607.1928, 185.2157, 0.5, 100.0)
    assert cloud.shape == want_xyz.shape and np.array_equal(cloud.view(np.uint64), want_xyz.view(np.uint64))


@pytest.mark.gpu
@pytest.mark.parametrize("h,w,s,d,sky", [(96, 192, 1, 64, False), (120, 200, 2, 64, False), (120, 200, 1, 64, True)])
def test_node_style_driver_integer_mode(h, w, s, d, sky):
    # SGM_B200_INTEGER=1: the C++ shell opens the engine in its integer / DPX mode (SGM without the cost filter);
    # oracle = the restatement with the cost filter window at 1x1.  node.cpp:93 always passes sky masks: they are accepted.
    L, R, _ = synth.random_dot_pair(h, w, seed=h + d, bg=5, fg=min(17, d // 2), noise=15)
    masks = synth.sky_masks(h // s, w // s) if sky else None
    want = restate.process(L, R, s, d, "sgm", *(masks or (None, None)), box_w=s, box_h=s)["f_final"]
    with tempfile.TemporaryDirectory() as tmp:
        disp, disp2, _, _, _ = _run_driver(tmp, "sgm", h, w, s, d, L, R, masks, env={"SGM_B200_INTEGER": "1"})
    assert bits_equal(disp, want) and bits_equal(disp2, want)


def test_cost_h_probes_match_oracle_dense_volumes():
    # SAD()/CT_pts()/hamming_cost() are host utilities with the reference's semantics: compare them, through a tiny C++
    # driver, with the oracle's dense SAD volume and census tables
    _make("_build/libstereo_solver_b200.so")
    src = os.path.join(ROOT, "tests", "cpp", "cost_probe_main.cpp")
    exe = os.path.join(HOST, "_build", "cost_probe_main")
    subprocess.check_call(["g++", "-std=c++11", "-O2", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "tests", "cpp", "stubs", "inc"),
                           "-I", os.path.join(ROOT, "oracle", "shim"), "-I", HOST, "-o", exe, src, "-L", os.path.join(HOST, "_build"),
                           "-lstereo_solver_b200", "-Wl,-rpath," + os.path.join(HOST, "_build"), "-L", os.path.join(ROOT, "stereo_matching_b200"),
                           "-lsgm_b200", "-Wl,-rpath," + os.path.join(ROOT, "stereo_matching_b200")], env=_env())
    h, w, d = 24, 40, 16
    L, R, _ = synth.random_dot_pair(h, w, seed=5, bg=2, fg=6, noise=20)
    with tempfile.TemporaryDirectory() as tmp:
        L.tofile(os.path.join(tmp, "l.raw"))
        R.tofile(os.path.join(tmp, "r.raw"))
        subprocess.check_call([exe, str(h), str(w), str(d), os.path.join(tmp, "l.raw"), os.path.join(tmp, "r.raw"), os.path.join(tmp, "out")])
        sad = np.fromfile(os.path.join(tmp, "out_sad.f32"), np.float32).reshape(h, w, d)
        ctl = np.fromfile(os.path.join(tmp, "out_ctl.u64"), np.uint64).reshape(h, w)
        ctr = np.fromfile(os.path.join(tmp, "out_ctr.u64"), np.uint64).reshape(h, w)
        ham = np.fromfile(os.path.join(tmp, "out_ham.i32"), np.int32).reshape(h, w)
    # oracle: census of the UNBLURRED images is not a tap, so compare SAD with the dense restatement and the census /
    # hamming probes with a direct numpy restatement
    want_sad = restate.process(L, R, 1, 32, "bm_sad", taps=("cost_raw",), sad_h=9, sad_w=9)["cost_raw"][0][:, :, :d]
    assert bits_equal(sad, np.ascontiguousarray(want_sad))

    def census(img):
        out = np.zeros(img.shape, np.uint64)
        pad = np.pad(img, ((3, 3), (4, 4)), mode="edge")
        for a in range(7):
            for b in range(9):
                if a == 3 and b == 4:
                    continue
                out = (out << np.uint64(1)) | (pad[a:a + h, b:b + w] > img).astype(np.uint64)
        return out

    assert np.array_equal(ctl, census(L)) and np.array_equal(ctr, census(R))
    x = ctl ^ ctr
    assert np.array_equal(ham, np.array([bin(int(v)).count("1") for v in x.ravel()], np.int32).reshape(h, w))
