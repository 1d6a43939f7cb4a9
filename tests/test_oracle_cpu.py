"""CPU tests of the oracle (-m "not gpu"): the restatement (oracle/sgm_oracle.c) is pinned against
(1) the committed golden vectors produced by the reference itself (tests/golden/, made by
make_golden.py from oracle/_ref = the unmodified /root/reference sources) and (2), where the compiled
reference is present, live runs of it on fresh random inputs; the blur stand-in is pinned against cv2."""
import os

import numpy as np
import pytest

from oracle import refbind, restate
from stereo_matching_b200 import synth
from tests.common import bits_equal, case_inputs, n_diff

GOLD = os.path.join(os.path.dirname(__file__), "golden")
H = restate.fnv1a64


def test_generator_matches_std_mt19937():
    # first outputs of std::mt19937(1234); the KAT hashes of SURVEY.md 8(c) depend on it
    L, R, gt = synth.random_dot_pair(4, 8, seed=1234)
    assert R.ravel()[:3].tolist() == [822569775 & 255, 2137449171 & 255, 2671936806 & 255]
    assert synth.fnv1a64(R) == H(R)


def test_blur_restatement_matches_cv2():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.RandomState(0)
    for shp in [(16, 16), (375, 1242), (37, 53), (1, 9), (9, 1), (2, 2)]:
        img = rng.randint(0, 256, size=shp).astype(np.uint8)
        want = cv2.GaussianBlur(img, (3, 3), 2, sigmaY=1)
        got = restate.process(img, img, 1, 32, "sgm", taps=("blur",))["blur"][0] if min(shp) >= 8 else None
        if got is not None:
            assert bits_equal(got, want), shp
        if refbind.available():
            assert bits_equal(refbind.gaussian_blur(img), want), shp


def test_survey_known_answers(golden):
    # SURVEY.md 8(c) quotes these for the reference on the mt19937(1234) pair
    h = golden["kitti_sgm_d128_clean"]["hashes"]
    assert (h["disp"], h["disp_beta"], h["f_final"]) == ("13c4a67b7e022475", "6d8bb4348527823e", "233489a1f55f2e3f")
    h = golden["kitti_sgm_d64_clean"]["hashes"]
    assert (h["disp"], h["disp_beta"], h["f_final"]) == ("89ff0b11223194b5", "6b8a0eb1f5555b26", "6c9b06a2424d5188")
    assert golden["kitti_bm_d64_clean"]["hashes"]["disp"] == "06a66f0562967f35"


FULL = ["kitti_sgm_d128_clean", "kitti_sgm_d128_noisy_sky", "kitti_sgm_d64_noisy", "kitti_sgm_d64_scale2",
        "kitti_bm_d64_noisy", "odd_sgm_d32", "tall_sgm_d32"]


@pytest.mark.parametrize("name", FULL)
def test_restatement_against_golden(golden, name):
    c = golden[name]
    L, R, (sl, sr) = case_inputs(c)
    want = c["hashes"]
    assert (H(L), H(R)) == (want["left"], want["right"])
    out = restate.process(L, R, c["scale"], c["d"], c["algo"], sl, sr, taps=restate.ALL_TAPS)
    got = dict(census_l=H(out["census"][0]), census_r=H(out["census"][1]), cost_raw_l=H(out["cost_raw"][0]),
               cost_h_l=H(out["cost_h"][0]), cost_hv_l=H(out["cost_hv"][0]), disp=H(out["disp"][0].astype(np.uint8)))
    if c["algo"] == "sgm":
        got.update(cost_raw_r=H(out["cost_raw"][1]), cost_hv_r=H(out["cost_hv"][1]), S_r=H(out["S"][1]),
                   disp_beta=H(out["disp"][1].astype(np.uint8)), fsub_r=H(out["fsub"][1]), f_final=H(out["f_final"]))
    for k, v in got.items():
        assert v == want[k], f"{name}: stage {k} differs from the reference"


def test_small_fixture_arrays():
    z = np.load(os.path.join(GOLD, "small_sgm.npz"))
    d = int(z["d"])
    out = restate.process(z["left"], z["right"], 1, d, "sgm", taps=restate.ALL_TAPS)
    assert bits_equal(out["census"][0], z["census_l"]) and bits_equal(out["census"][1], z["census_r"])
    assert bits_equal(out["cost_hv"][0], z["cost_hv_l"]) and bits_equal(out["cost_hv"][1], z["cost_hv_r"])
    assert bits_equal(out["S"][1], z["S_r"])
    assert bits_equal(out["disp"][0].astype(np.uint8), z["disp"])
    assert bits_equal(out["disp"][1].astype(np.uint8), z["disp_beta"])
    assert bits_equal(out["fsub"][1], z["fsub_r"]) and bits_equal(out["f_final"], z["f_final"])
    b = np.load(os.path.join(GOLD, "small_bm.npz"))
    ob = restate.process(b["left"], b["right"], 1, d, "bm")
    assert bits_equal(ob["disp"][0].astype(np.uint8), b["disp"])


@pytest.mark.skipif(not refbind.available(), reason="compiled reference (oracle/_ref) not present")
@pytest.mark.parametrize("h,w,d,scale,noise,sky", [(96, 160, 32, 1, 6, False), (120, 200, 64, 1, 40, True),
                                                   (200, 320, 64, 2, 10, True), (64, 300, 128, 1, 25, False)])
def test_restatement_against_live_reference(h, w, d, scale, noise, sky):
    L, R, _ = synth.random_dot_pair(h, w, seed=h * 7 + d, bg=5, fg=min(17, d // 2), noise=noise)
    sl, sr = synth.sky_masks(h // scale, w // scale) if sky else (None, None)
    refbind.set_threads(1)
    ref = refbind.process("sgm", L, R, scale, d, sl, sr, want_cost=True)
    out = restate.process(L, R, scale, d, "sgm", sl, sr, taps=restate.ALL_TAPS)
    for beta, m in ((False, sl), (True, sr)):
        st = refbind.stages("sgm", L, R, scale, d, beta=beta, sky=m)
        v = int(beta)
        assert bits_equal(st["cost_raw"], out["cost_raw"][v]) and bits_equal(st["cost_h"], out["cost_h"][v])
        assert bits_equal(st["cost_hv"], out["cost_hv"][v])
    assert bits_equal(ref["cost_last"], out["S"][1])
    assert n_diff(ref["disp"], out["disp"][0]) == 0 and n_diff(ref["disp_beta"], out["disp"][1]) == 0
    assert bits_equal(ref["fdisp_beta"], out["fsub"][1]) and bits_equal(ref["fdisp"], out["f_final"])
    # reference BM (census + box + WTA with |dd| > 2)
    rb = refbind.process("bm", L, R, scale, d, sl, sr)
    ob = restate.process(L, R, scale, d, "bm", sl, sr)
    assert n_diff(rb["disp"], ob["disp"][0]) == 0


@pytest.mark.skipif(not refbind.available(), reason="compiled reference (oracle/_ref) not present")
def test_sad_volume_against_reference_function():
    # oracle's SAD volume vs the reference's own SAD() (src/cost.cpp:4-30) evaluated per pixel
    L, R, _ = synth.random_dot_pair(24, 40, seed=2, bg=3, fg=7, noise=12)
    want = refbind.cost_volume("sad", L, R, 16, 9, 9)
    got = restate.process(L, R, 1, 16, "bm_sad", taps=("cost_raw",))["cost_raw"][0]
    assert bits_equal(want, got)


@pytest.mark.parametrize("h,w,d,scale,paths,noise", [(48, 96, 64, 1, 8, 20), (40, 150, 128, 1, 8, 50), (33, 70, 64, 1, 4, 30),
                                                      (60, 120, 64, 2, 8, 15)])
def test_integer_restatement_equals_float_restatement_without_cost_filter(h, w, d, scale, paths, noise):
    # the engine's integer / DPX mode (SGM_FLAG_INTEGER): with the cost filter window at 1x1 every number on the
    # reference's float path is a small integer, so a 16-bit integer restatement of the same loops must reproduce the
    # float restatement (which is pinned against the reference) exactly - that is the oracle of the integer mode
    L, R, _ = synth.random_dot_pair(h, w, seed=h + w + d, bg=3, fg=min(15, d // 3), noise=noise)
    out = restate.process(L, R, scale, d, "sgm", taps=restate.ALL_TAPS, box_w=scale, box_h=scale, paths=paths)
    for v in (0, 1):
        raw, hv, S = out["cost_raw"][v], out["cost_hv"][v], out["S"][v]
        assert bits_equal(raw, hv)                                    # window 1: the filters are the identity
        assert raw.min() >= 0 and raw.max() <= 62 and np.array_equal(raw, np.floor(raw))
        assert S.max() <= paths * (62 + 100) and np.array_equal(S, np.floor(S))
        Si = restate.aggregate_u16(raw.astype(np.uint16), 10, 100, paths)
        assert np.array_equal(Si.astype(np.float32), S)


@pytest.mark.skipif(not refbind.available("nofilter"), reason="oracle/_ref/libsgm_ref_nofilter.so not built")
@pytest.mark.parametrize("h,w,d,noise", [(96, 160, 64, 6), (64, 300, 128, 25), (97, 161, 32, 60)])
def test_restatement_without_cost_filter_against_reference_nofilter(h, w, d, noise):
    # the DIRECT pin of the integer / DPX mode's oracle: the reference's own sources compiled with COST_WIN_H = COST_WIN_W = 1
    # (make -C oracle ref_nofilter) against the restatement with box_w = box_h = 1, tap by tap, and its 16-bit integer form
    L, R, _ = synth.random_dot_pair(h, w, seed=h * 5 + d, bg=3, fg=max(4, min(17, d // 2)), noise=noise)
    mine = restate.process(L, R, 1, d, "sgm", taps=restate.ALL_TAPS, box_w=1, box_h=1)
    with refbind.use("nofilter"):
        refbind.set_threads(1)
        ref = refbind.process("sgm", L, R, 1, d, want_cost=True)
        st = [refbind.stages("sgm", L, R, 1, d, beta=b) for b in (False, True)]
    assert bits_equal(st[0]["cost_hv"], st[0]["cost_raw"])                      # the filter really is off in that build
    assert bits_equal(mine["cost_hv"][0], st[0]["cost_hv"]) and bits_equal(mine["cost_hv"][1], st[1]["cost_hv"])
    assert bits_equal(mine["S"][1], ref["cost_last"])
    assert n_diff(mine["disp"][0], ref["disp"]) == 0 and n_diff(mine["disp"][1], ref["disp_beta"]) == 0
    assert bits_equal(mine["fsub"][1], ref["fdisp_beta"]) and bits_equal(mine["f_final"], ref["fdisp"])
    Si = restate.aggregate_u16(st[1]["cost_hv"].astype(np.uint16), 10, 100, 8)  # 16-bit form on the REFERENCE's volume
    assert np.array_equal(Si.astype(np.float32), ref["cost_last"])


@pytest.mark.skipif(not refbind.available("nofilter"), reason="oracle/_ref/libsgm_ref_nofilter.so not built")
def test_integer_mode_golden_equals_reference_nofilter(golden_integer):
    # tests/golden/kat_hashes_integer.json at KITTI size against the reference built without its cost filter
    c = golden_integer["kitti_int_d64_clean"]
    L, R, _ = synth.random_dot_pair(c["h"], c["w"], seed=c["seed"], bg=c["bg"], fg=c["fg"], noise=c["noise"])
    with refbind.use("nofilter"):
        refbind.set_threads(1)
        ref = refbind.process("sgm", L, R, 1, c["d"], want_cost=True)
    want = c["hashes"]
    assert H(ref["cost_last"]) == want["S_r"] and H(ref["disp"].astype(np.uint16)) == want["disp"]
    assert H(ref["disp_beta"].astype(np.uint16)) == want["disp_beta"] and H(ref["fdisp_beta"]) == want["fsub_r"]
    assert H(ref["fdisp"]) == want["f_final"]


def test_integer_mode_golden_is_reproducible(golden_integer):
    # tests/golden/kat_hashes_integer.json must be what the restatement (window 1x1) produces today
    c = golden_integer["kitti_int_d64_scale2"]
    L, R, _ = synth.random_dot_pair(c["h"], c["w"], seed=c["seed"], bg=c["bg"], fg=c["fg"], noise=c["noise"])
    s = c["scale"]
    out = restate.process(L, R, s, c["d"], "sgm", taps=("S", "disp", "f_final"), box_w=s, box_h=s)
    want = c["hashes"]
    assert (H(L), H(R)) == (want["left"], want["right"])
    assert H(out["S"][0]) == want["S_l"] and H(out["disp"][1]) == want["disp_beta"] and H(out["f_final"]) == want["f_final"]
    assert int(out["S"][0].max()) == c["stats"]["S_max"] <= 8 * (62 + 100)


@pytest.mark.parametrize("name", ["kitti_bmsad_d64_noisy", "hd_sgm_d256_clean"])
def test_fullsize_golden_is_reproducible(golden_fullsize, name):
    # tests/golden/kat_hashes_fullsize.json (the configurations the reference cannot run, at BASELINE.json's full sizes)
    # must be what the widened restatement produces today; the 1080p case takes about a minute
    c = golden_fullsize[name]
    L, R, _ = synth.random_dot_pair(c["h"], c["w"], seed=c["seed"], bg=c["bg"], fg=c["fg"], noise=c["noise"])
    out = restate.process(L, R, c["scale"], c["d"], c["algo"], taps=("disp", "fsub", "f_lr", "f_median", "f_final"))
    want = c["hashes"]
    assert (H(L), H(R)) == (want["left"], want["right"])
    assert H(out["disp"][0]) == want["disp_l"] and H(out["disp"][1]) == want["disp_r"] and H(out["fsub"][1]) == want["fsub_r"]
    assert H(out["f_lr"]) == want["f_lr"] and H(out["f_median"]) == want["f_median"] and H(out["f_final"]) == want["f_final"]
    assert int((out["f_final"] <= c["d"] - 1).sum()) == c["stats"]["valid"]


@pytest.mark.skipif(not refbind.available(), reason="compiled reference (oracle/_ref) not built")
@pytest.mark.parametrize("kind,wh,ww,scale", [("ssd", 7, 9, 1), ("ssd", 9, 9, 2), ("ssd", 3, 5, 1), ("ct", 7, 9, 1), ("ct", 7, 9, 2), ("ct", 3, 3, 1)])
def test_window_cost_restatement_against_reference_functions(kind, wh, ww, scale):
    # SSD() / CT() of the reference (src/cost.cpp:33-96), called for every (pixel, d) by the harness, against the numpy
    # restatement the GPU tests use as their checker
    L, R, _ = synth.random_dot_pair(24, 45, seed=5 + wh, bg=2, fg=6, noise=20)
    want = refbind.cost_volume(kind, L, R, 20, wh, ww, scale)
    assert bits_equal(restate.window_cost_volume(kind, L, R, 20, wh, ww, scale), want)


def test_widened_restatement_runs_d256():
    # D=256 cannot run through the reference (src/Solver.cpp:10 asserts); the restatement is the oracle there
    L, R, _ = synth.random_dot_pair(48, 400, seed=4, bg=24, fg=200, noise=4)
    out = restate.process(L, R, 1, 256, "sgm")
    f = out["f_final"]
    assert out["disp"][0].max() <= 257 and (f <= 255).sum() > 0


@pytest.mark.skipif(not refbind.available(), reason="compiled reference (oracle/_ref) not built")
@pytest.mark.parametrize("algo,h,w,scale,d", [("sgm", 64, 160, 1, 64), ("sgm", 96, 200, 2, 32), ("bm", 64, 160, 1, 32)])
def test_debug_view_restatement_vs_reference_show_disp(algo, h, w, scale, d):
    # pins oracle/restate.py's colormap / debug_view (used to check the device rendering) against Solver::show_disp
    L, R, _ = synth.random_dot_pair(h, w, seed=h + d, bg=3, fg=min(20, d // 2), noise=12)
    refbind.set_threads(1)
    view, f = refbind.show_disp(algo, L, R, scale, d)
    if algo == "bm":
        # the reference's BM never writes filtered_disp (src/BM.cpp:52-96): colour-map whatever buffer it returned
        want = restate.debug_view(L, f, d, scale, bm_rows=True)
    else:
        want = restate.debug_view(L, f, d, scale)
    assert np.array_equal(view, want)
    # every branch of the colour ramp, including the double-precision segments, on a synthetic ramp
    ramp = np.linspace(0, d + 1, 4 * 257, dtype=np.float32).reshape(4, 257)
    cm = restate.colormap(ramp, d)
    assert cm[ramp > d - 1].max() == 0 and (cm[ramp <= d - 1].max(axis=-1) == 255).all()


def test_point_cloud_restatement_basics():
    # node.cpp:119-143 restated: invalid pixels and points beyond max_range are dropped, order is raster order
    d = 64
    f = np.full((4, 6), d + 1, np.float32)
    f[1, 2], f[1, 4], f[3, 0] = 10.0, 0.001, 32.5     # 0.001 px -> Z far beyond 100 m -> dropped
    left = np.arange(24, dtype=np.uint8).reshape(4, 6)
    xyz, inten = restate.point_cloud(f, left, d, 1, 700.0, 720.0, 3.0, 2.0)
    assert xyz.shape == (2, 3) and list(inten) == [left[1, 2], left[3, 0]]
    z0 = (np.float64(np.float32(700.0) + np.float32(720.0)) / 2.0) * 0.5 / (np.float64(np.float32(10.0)) + 1e-6)
    assert xyz[0, 2] == z0 and xyz[0, 0] == np.float64(np.float32(2 - 3.0)) * z0 / np.float64(np.float32(700.0))


NODE_REF = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "node_ref")


@pytest.mark.skipif(not os.path.exists(NODE_REF), reason="oracle/_ref/node_ref (the reference's node.cpp, linked) not built")
@pytest.mark.parametrize("h,w,d,scale,sky_rows,noise", [(120, 200, 64, 1, 0, 15), (150, 260, 64, 2, 20, 10), (96, 192, 128, 1, 24, 30)])
def test_point_cloud_restatement_against_reference_node_cpp(tmp_path, h, w, d, scale, sky_rows, noise):
    # N3 pin: the disparity -> point cloud loop exists only inside main() of the reference's node.cpp (:119-143).  The
    # UNMODIFIED node.cpp is linked with the reference's own solver and I/O stubs (oracle/node_harness.cpp), fed a synthetic
    # pair, and its stereo_pts / stereo_pixel vectors are compared with restate.point_cloud — the checker of the GPU tests
    # (tests/test_host_shell.py, tests/test_views_gpu.py) — bit for bit.
    import subprocess

    L, R, _ = synth.random_dot_pair(h, w, seed=h + d, bg=5, fg=min(17, d // 2), noise=noise)
    L.tofile(tmp_path / "l.raw")
    R.tofile(tmp_path / "r.raw")
    env = dict(os.environ, NODE_LEFT_RAW=str(tmp_path / "l.raw"), NODE_RIGHT_RAW=str(tmp_path / "r.raw"), NODE_OUT=str(tmp_path / "out"),
               NODE_SKY_ROWS=str(sky_rows), ROS_PARAM__sgm_node_img_h=str(h), ROS_PARAM__sgm_node_img_w=str(w),
               ROS_PARAM__sgm_node_scale=str(scale), ROS_PARAM__sgm_node_max_disp=str(d), OMP_NUM_THREADS="1")
    r = subprocess.run([NODE_REF], env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-500:] + r.stderr[-500:]
    ih, iw = h // scale, w // scale
    disp = np.fromfile(tmp_path / "out_disp.f32", np.float32).reshape(ih, iw)
    xyz = np.fromfile(tmp_path / "out_xyz.f64", np.float64).reshape(-1, 3)
    inten = np.fromfile(tmp_path / "out_int.u8", np.uint8)
    # the disparity node.cpp published is the reference solver's (same sources as oracle/_ref/libsgm_ref.so) ...
    sl = sr = None
    if sky_rows:
        sl = np.zeros((ih, iw), np.uint8)
        sl[:sky_rows] = 255
        sr = sl.copy()
    refbind.set_threads(1)
    assert bits_equal(disp, refbind.process("sgm", L, R, scale, d, sl, sr)["fdisp"])
    # ... and its point cloud is what the restatement computes from that disparity
    want_xyz, want_int = restate.point_cloud(disp, L, d, scale, 718.856, 718.856, 607.1928, 185.2157, 0.5, 100.0)
    assert f"pointcloud size: {len(want_int)}, {len(want_int)}" in r.stdout
    assert xyz.shape == want_xyz.shape and np.array_equal(xyz.view(np.uint64), want_xyz.view(np.uint64))
    assert np.array_equal(inten, want_int) and len(inten) > 50   # (scale 2 halves the disparities: most points pass the 100 m range cut)
