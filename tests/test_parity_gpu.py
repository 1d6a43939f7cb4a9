"""GPU parity tests (-m gpu): the CUDA path, called through the C ABI (stereo_matching_b200.capi ->
libsgm_b200.so), against the oracle on the same seeded inputs.

  * small / mid sizes: every stage tap bit-for-bit against the restatement (oracle/sgm_oracle.c) and,
    where the compiled reference travelled (oracle/_ref), against the reference itself;
  * BASELINE.json's KITTI-size configs: against the golden hashes the reference produced
    (tests/golden/kat_hashes.json) — no oracle needed at run time;
  * sizes the reference cannot run (D=256, 1080p): against the widened restatement at a reduced size and
    through size-independent properties at full size.
Tolerances: integer disparities, the invalid mask and every float tap are compared BIT-EXACT (stricter
than north_star's 1e-3 px for sub-pixel values).
"""
import numpy as np
import pytest

from oracle import refbind, restate
from stereo_matching_b200 import capi, synth
from tests.common import bits_equal, case_inputs, n_diff

pytestmark = pytest.mark.gpu
H = restate.fnv1a64


def _check_sgm_against_restatement(L, R, scale, d, sl=None, sr=None, paths=8, flags=0, **over):
    want = restate.process(L, R, scale, d, "sgm", sl, sr, taps=restate.ALL_TAPS, paths=paths, **over)
    sv = capi.SGM(L.shape[0], L.shape[1], scale, d, flags=capi.FLAG_KEEP_VOLUMES | flags, paths=paths, **over)
    sv.process(L, R, sl, sr)
    bad = {}
    for v in (0, 1):
        for tap, key in (("blur", "blur"), ("census", "census"), ("cost", "cost_hv"), ("pathsum", "S"), ("disp", "disp"),
                         ("subpixel", "fsub")):
            nd = n_diff(sv.tap(tap, v), want[key][v])
            if nd:
                bad[f"{tap}[{v}]"] = nd
    for tap, key in (("lr", "f_lr"), ("median", "f_median"), ("final", "f_final")):
        nd = n_diff(sv.tap(tap), want[key])
        if nd:
            bad[tap] = nd
    nd = n_diff(sv.get_disp(), want["f_final"])
    if nd:
        bad["get_disp"] = nd
    sv.close()
    assert not bad, f"stages differing from the oracle (count of differing elements): {bad}"
    return want


@pytest.mark.parametrize("h,w,d,scale,noise,sky", [
    (96, 160, 32, 1, 6, False),
    (120, 200, 64, 1, 40, True),
    (200, 320, 64, 2, 10, True),
    (64, 300, 128, 1, 25, False),
    (97, 161, 64, 1, 60, False),   # odd sizes, heavy noise: many holes for the median / speckle stages
    (300, 200, 32, 1, 12, True),   # W < H
])
def test_sgm_stages_bit_exact_vs_oracle(h, w, d, scale, noise, sky):
    L, R, _ = synth.random_dot_pair(h, w, seed=h * 7 + d, bg=5, fg=min(17, d // 2), noise=noise)
    sl, sr = synth.sky_masks(h // scale, w // scale) if sky else (None, None)
    want = _check_sgm_against_restatement(L, R, scale, d, sl, sr)
    assert (want["f_final"] <= d - 1).any()


@pytest.mark.parametrize("h,w,d", [(16, 16, 32), (17, 33, 32), (16, 161, 128), (40, 21, 128), (19, 75, 64), (24, 48, 256)])
def test_sgm_small_and_ragged_sizes_vs_oracle(h, w, d):
    # smallest accepted images, widths that leave a single-column last block in the sweeps (161 = 16*10 + 1 at D=128,
    # 33 = 4*8 + 1 at D=32), fewer rows than rows kept in flight, W < D
    L, R, _ = synth.random_dot_pair(h, w, seed=h * w + d, bg=2, fg=min(9, w // 3), noise=25)
    _check_sgm_against_restatement(L, R, 1, d)


@pytest.mark.parametrize("h,w,d,integer", [
    (18, 1250, 128, False), (17, 1255, 64, False), (16, 1258, 128, False),  # 17-column blocks: last block of 9 / 14 / 17 columns
    (20, 760, 128, False), (18, 1243, 32, False),                          # ... of 12 / 2 columns
    (18, 1259, 128, False), (16, 1290, 64, False),                         # too wide for 17-column blocks: 19 columns, last block 5 / 17
    (18, 1250, 128, True), (16, 1259, 128, True), (17, 1255, 64, True),
    (16, 1920, 256, False), (16, 1500, 256, False),                        # 13-column blocks, one view per launch: last block 9 / 5
    (16, 1910, 256, True), (18, 1500, 128, False),
])
def test_sgm_wide_rows_one_block_per_sm_shapes_vs_oracle(h, w, d, integer):
    # wide images switch the sweeps to one wide block per SM (sweep_kernels.cu: wide_shape picks 13, 17 or 19 columns); the
    # hardware warp -> column permutation of those blocks depends on how many columns the ragged last block of a row has
    L, R, _ = synth.random_dot_pair(h, w, seed=h * w + d, bg=3, fg=min(17, d // 2), noise=20)
    if integer:
        _check_sgm_against_restatement(L, R, 1, d, flags=capi.FLAG_INTEGER, box_w=1, box_h=1)
    else:
        _check_sgm_against_restatement(L, R, 1, d)


def test_sgm_halo_tag_sequence_wraps(monkeypatch):
    # the cross-block exchange of the sweeps tags every vector with (launch sequence, row); the sequence wraps after 2^20
    # sweep launches (~5e5 frames), where the ring is wiped and the numbering restarts: run frames across that point
    monkeypatch.setenv("SGM_B200_HALO_SEQ_START", str((1 << 20) - 4))
    h, w, d = 48, 170, 128
    pairs = [synth.random_dot_pair(h, w, seed=90 + k, bg=3, fg=11, noise=15)[:2] for k in range(2)]
    want = [restate.process(L, R, 1, d, "sgm")["f_final"] for L, R in pairs]
    sv = capi.SGM(h, w, 1, d)
    for k in range(6):  # 2 sweep launches per frame: the wrap happens inside this loop
        sv.process(*pairs[k % 2])
        assert bits_equal(sv.get_disp(), want[k % 2]), f"frame {k} after the tag wrap differs from the oracle"
    sv.close()


@pytest.mark.parametrize("flags", [0, capi.FLAG_LINE_WALKERS])
def test_sgm_four_paths_vs_oracle(flags):
    L, R, _ = synth.random_dot_pair(80, 144, seed=21, bg=4, fg=13, noise=10)
    _check_sgm_against_restatement(L, R, 1, 64, paths=4, flags=flags)


# ---- integer / DPX mode (SGM_FLAG_INTEGER) ----------------------------------------------------------
# Oracle: the restatement with the cost filter window at 1x1 (pinned against the reference for the default window, and
# against its own 16-bit integer form in tests/test_oracle_cpu.py).  Every tap is compared bit for bit, like the float mode.
@pytest.mark.parametrize("h,w,d,scale,noise", [
    (96, 160, 64, 1, 6),
    (64, 300, 128, 1, 25),
    (97, 161, 128, 1, 60),    # odd sizes, heavy noise
    (300, 200, 64, 1, 12),    # W < H
    (72, 250, 256, 1, 10),    # 128 words per pixel
    (200, 320, 64, 2, 10),    # scale 2 (box window 2 / 2 = 1)
    (16, 16, 64, 1, 20), (16, 161, 128, 1, 25), (40, 21, 128, 1, 25), (19, 75, 64, 1, 25), (24, 48, 256, 1, 25),
])
def test_integer_mode_stages_bit_exact_vs_oracle(h, w, d, scale, noise):
    L, R, _ = synth.random_dot_pair(h, w, seed=h * 5 + d, bg=3, fg=max(4, min(17, d // 2, w // 3)), noise=noise)
    _check_sgm_against_restatement(L, R, scale, d, flags=capi.FLAG_INTEGER, box_w=scale, box_h=scale)


def test_integer_mode_four_paths_and_other_penalties():
    L, R, _ = synth.random_dot_pair(80, 144, seed=21, bg=4, fg=13, noise=10)
    _check_sgm_against_restatement(L, R, 1, 64, paths=4, flags=capi.FLAG_INTEGER, box_w=1, box_h=1)
    _check_sgm_against_restatement(L, R, 1, 128, flags=capi.FLAG_INTEGER, box_w=1, box_h=1, P1=7, P2=300)
    _check_sgm_against_restatement(L, R, 1, 64, flags=capi.FLAG_INTEGER, box_w=1, box_h=1, P1=0, P2=0)


def test_integer_mode_equals_float_mode_at_kitti_size():
    # BASELINE.json config 2 shape (1242x375, D=128): the float engine with the cost filter off is the reference's
    # arithmetic; the integer engine must return the same bits for every result
    h, w, d = 375, 1242, 128
    L, R, gt = synth.random_dot_pair(h, w, seed=1234, bg=12, fg=40, noise=8)
    a = capi.SGM(h, w, 1, d, box_w=1, box_h=1)
    b = capi.SGM(h, w, 1, d, box_w=1, box_h=1, flags=capi.FLAG_INTEGER)
    a.process(L, R), b.process(L, R)
    for v in (0, 1):
        assert bits_equal(a.tap("disp", v), b.tap("disp", v)) and bits_equal(a.tap("subpixel", v), b.tap("subpixel", v))
    assert bits_equal(a.tap("lr"), b.tap("lr")) and bits_equal(a.get_disp(), b.get_disp())
    f = b.get_disp()
    ok = (f <= d - 1) & (np.arange(w)[None, :] >= d)
    assert ok.mean() > 0.5 and (np.abs(f[ok] - gt[ok]) <= 1).mean() > 0.95
    assert b.device_bytes < 0.6 * a.device_bytes     # 16-bit volumes
    for _ in range(2):                                # deterministic, handle reusable
        assert bits_equal(b.process(L, R).get_disp(), f)
    a.close(), b.close()


@pytest.mark.parametrize("name", ["kitti_int_d128_noisy", "kitti_int_d64_clean", "kitti_int_d64_scale2"])
def test_integer_mode_kitti_size_against_golden(golden_integer, name):
    # full BASELINE.json size against the committed vectors of the oracle (restatement with the cost filter at 1x1,
    # cross-checked with its 16-bit form when they were generated): every tap by hash, no oracle run on the GPU box
    c = golden_integer[name]
    L, R, _ = synth.random_dot_pair(c["h"], c["w"], seed=c["seed"], bg=c["bg"], fg=c["fg"], noise=c["noise"])
    want, s = c["hashes"], c["scale"]
    assert (H(L), H(R)) == (want["left"], want["right"])
    sv = capi.SGM(c["h"], c["w"], s, c["d"], flags=capi.FLAG_INTEGER | capi.FLAG_KEEP_VOLUMES, box_w=s, box_h=s)
    sv.process(L, R)
    got = dict(cost_l=H(sv.tap("cost", 0)), cost_r=H(sv.tap("cost", 1)), S_l=H(sv.tap("pathsum", 0)), S_r=H(sv.tap("pathsum", 1)),
               disp=H(sv.tap("disp", 0)), disp_beta=H(sv.tap("disp", 1)), fsub_l=H(sv.tap("subpixel", 0)),
               fsub_r=H(sv.tap("subpixel", 1)), f_lr=H(sv.tap("lr")), f_median=H(sv.tap("median")), f_final=H(sv.get_disp()))
    bad = sorted(k for k, v in got.items() if v != want[k])
    assert not bad, f"{name}: taps whose hash differs from the oracle's golden vectors: {bad}"
    assert int((sv.get_disp() <= c["d"] - 1).sum()) == c["stats"]["valid"]
    sv.close()


def test_integer_mode_rejects_what_it_cannot_represent():
    for over in (dict(), dict(box_w=1, box_h=1, P2=5000), dict(box_w=1, box_h=1, P1=20, P2=10)):
        with pytest.raises(capi.SgmError) as e:
            capi.SGM(64, 64, 1, 64, flags=capi.FLAG_INTEGER, **over)   # default box 5x3 / too large penalties / P1 > P2
        assert e.value.code == 2
    with pytest.raises(capi.SgmError) as e:
        capi.SGM(64, 64, 1, 32, flags=capi.FLAG_INTEGER, box_w=1, box_h=1)
    assert e.value.code == 2
    sv = capi.SGM(64, 64, 1, 64, flags=capi.FLAG_INTEGER, box_w=1, box_h=1, P2=2000)
    z = np.zeros((64, 64), np.uint8)
    sv.process(z, z)                                                   # fine without masks (P2 <= 4033) ...
    with pytest.raises(capi.SgmError) as e:
        sv.process(z, z, z, z)                                         # ... but the masked pixels' stand-in cost 2 * P2 needs P2 <= 1365
    assert e.value.code == 2
    sv.close()


@pytest.mark.parametrize("h,w,d,scale,P1,P2,frac", [(120, 200, 64, 1, 10, 100, 0.25), (97, 161, 128, 1, 7, 300, 0.4), (200, 320, 64, 2, 10, 100, 0.25),
                                                    (64, 300, 128, 1, 10, 1365, 0.5)])
def test_integer_mode_with_sky_masks_vs_oracle(h, w, d, scale, P1, P2, frac):
    # the reference fills masked pixels with 999999 (src/Solver.cpp:165-179); the integer mode uses 2 * P2 and must still return
    # the reference's results everywhere: checker = the float restatement (cost filter 1x1) with the same masks
    L, R, _ = synth.random_dot_pair(h, w, seed=h * 3 + d, bg=3, fg=max(4, min(17, d // 2, w // 3)), noise=20)
    sl, sr = synth.sky_masks(h // scale, w // scale, frac)
    sr[:, : w // scale // 3] = 0                                       # the two masks differ
    want = restate.process(L, R, scale, d, "sgm", sl, sr, taps=restate.ALL_TAPS, box_w=scale, box_h=scale, P1=P1, P2=P2)
    sv = capi.SGM(h, w, scale, d, flags=capi.FLAG_INTEGER | capi.FLAG_KEEP_VOLUMES, box_w=scale, box_h=scale, P1=P1, P2=P2)
    sv.process(L, R, sl, sr)
    for v, m in ((0, sl), (1, sr)):
        assert n_diff(sv.tap("disp", v), want["disp"][v]) == 0 and bits_equal(sv.tap("subpixel", v), want["fsub"][v])
        free = m != 255                                                # the volumes agree wherever no mask applies ...
        assert bits_equal(sv.tap("cost", v)[free], want["cost_hv"][v][free]) and bits_equal(sv.tap("pathsum", v)[free], want["S"][v][free])
        c = sv.tap("cost", v)[~free]                                   # ... and masked pixels carry {0, 2 * P2, 2 * P2, ...}
        assert (c[:, 0] == 0).all() and (c[:, 1:] == 2 * P2).all() and (want["cost_hv"][v][~free][:, 1:] == 999999).all()
        assert (sv.tap("disp", v)[~free] == 0).all()                   # a masked pixel's winner is d = 0 and stays valid
    assert bits_equal(sv.tap("lr"), want["f_lr"]) and bits_equal(sv.tap("median"), want["f_median"])
    assert bits_equal(sv.get_disp(), want["f_final"])
    sv.close()


@pytest.mark.parametrize("h,w,d,scale,sky", [(120, 200, 64, 1, True), (64, 300, 128, 1, False), (72, 250, 256, 1, False)])
def test_sgm_line_walker_schedule_vs_oracle(h, w, d, scale, sky):
    # the fall-back aggregation schedule (independent line walkers, used when an image is too wide for the resident
    # fused sweeps) must give the same bits as the default fused schedule
    L, R, _ = synth.random_dot_pair(h, w, seed=h + d, bg=5, fg=min(17, d // 2), noise=20)
    sl, sr = synth.sky_masks(h // scale, w // scale) if sky else (None, None)
    _check_sgm_against_restatement(L, R, scale, d, sl, sr, flags=capi.FLAG_LINE_WALKERS)


def test_sgm_subpixel_slanted_plane():
    L, R, gt = synth.slanted_pair(128, 256, d0=6.0, d1=40.0)
    want = _check_sgm_against_restatement(L, R, 1, 64)
    f = want["f_final"]
    ok = (f <= 63) & (np.arange(256)[None, :] >= 64)
    assert ok.sum() > 1000 and np.abs(f[ok] - gt[ok]).mean() < 1.0
    assert (f[ok] != np.floor(f[ok])).mean() > 0.3  # the parabola fit really produces fractions


def test_sgm_widened_d256_vs_restatement():
    # D=256 cannot run through the reference (src/Solver.cpp:10); oracle = widened restatement
    L, R, _ = synth.random_dot_pair(72, 480, seed=4, bg=24, fg=200, noise=6)
    _check_sgm_against_restatement(L, R, 1, 256)


@pytest.mark.skipif(not refbind.available(), reason="compiled reference (oracle/_ref) did not travel")
def test_sgm_against_compiled_reference_live():
    h, w, d = 150, 400, 64
    L, R, _ = synth.random_dot_pair(h, w, seed=77, bg=9, fg=30, noise=20)
    sl, sr = synth.sky_masks(h, w, 0.2)
    refbind.set_threads(1)
    ref = refbind.process("sgm", L, R, 1, d, sl, sr, want_cost=True)
    sv = capi.SGM(h, w, 1, d, flags=capi.FLAG_KEEP_VOLUMES)
    sv.process(L, R, sl, sr)
    assert bits_equal(sv.get_disp_u8(0), ref["disp"])
    assert bits_equal(sv.get_disp_u8(1), ref["disp_beta"])
    assert bits_equal(sv.tap("pathsum", 1), ref["cost_last"])
    assert bits_equal(sv.tap("subpixel", 1), ref["fdisp_beta"])
    got, want = sv.get_disp(), ref["fdisp"]
    assert np.array_equal(got > d - 1, want > d - 1)          # invalid mask exact
    assert np.abs(got - want).max() <= 1e-3                      # north_star tolerance ...
    assert bits_equal(got, want)                                 # ... and in fact bit-equal
    for beta, m in ((False, sl), (True, sr)):
        st = refbind.stages("sgm", L, R, 1, d, beta=beta, sky=m)
        assert bits_equal(sv.tap("cost", int(beta)), st["cost_hv"])
    assert bits_equal(sv.tap("census", 0), st["ct_l"]) and bits_equal(sv.tap("census", 1), st["ct_r"])
    bm = capi.BM(h, w, 1, d)
    bm.process(L, R, sl, sr)
    assert bits_equal(bm.get_disp_u8(0), refbind.process("bm", L, R, 1, d, sl, sr)["disp"])


KITTI_SGM = ["kitti_sgm_d128_clean", "kitti_sgm_d64_clean", "kitti_sgm_d128_noisy_sky", "kitti_sgm_d64_noisy",
             "kitti_sgm_d64_scale2", "odd_sgm_d32", "tall_sgm_d32"]


@pytest.mark.parametrize("name", KITTI_SGM)
def test_sgm_kitti_size_against_reference_golden(golden, name):
    c = golden[name]
    L, R, (sl, sr) = case_inputs(c)
    want = c["hashes"]
    assert (H(L), H(R)) == (want["left"], want["right"])
    sv = capi.SGM(c["h"], c["w"], c["scale"], c["d"], flags=capi.FLAG_KEEP_VOLUMES)
    sv.process(L, R, sl, sr)
    got = dict(census_l=H(sv.tap("census", 0)), census_r=H(sv.tap("census", 1)), cost_hv_l=H(sv.tap("cost", 0)),
               cost_hv_r=H(sv.tap("cost", 1)), S_r=H(sv.tap("pathsum", 1)), disp=H(sv.get_disp_u8(0)),
               disp_beta=H(sv.get_disp_u8(1)), fsub_r=H(sv.tap("subpixel", 1)), f_final=H(sv.get_disp()))
    bad = {k: (v, want[k]) for k, v in got.items() if v != want[k]}
    assert not bad, f"{name}: taps whose hash differs from the reference: {sorted(bad)}"
    f = sv.get_disp()
    assert int((f <= c["d"] - 1).sum()) == c["stats"]["valid"]
    sv.close()


@pytest.mark.parametrize("name", ["kitti_bm_d64_clean", "kitti_bm_d64_noisy"])
def test_bm_kitti_size_against_reference_golden(golden, name):
    c = golden[name]
    L, R, _ = case_inputs(c)
    bm = capi.BM(c["h"], c["w"], c["scale"], c["d"])
    bm.process(L, R)
    want = c["hashes"]
    assert H(bm.tap("census", 0)) == want["census_l"] and H(bm.tap("cost", 0)) == want["cost_hv_l"]
    assert H(bm.get_disp_u8(0)) == want["disp"]
    bm.close()


@pytest.mark.parametrize("h,w,d,scale,noise", [(96, 160, 32, 1, 10), (150, 260, 64, 2, 30), (80, 200, 128, 1, 50)])
def test_bm_census_vs_oracle(h, w, d, scale, noise):
    L, R, _ = synth.random_dot_pair(h, w, seed=h + d, bg=4, fg=12, noise=noise)
    want = restate.process(L, R, scale, d, "bm", taps=restate.ALL_TAPS)
    bm = capi.BM(h, w, scale, d)
    bm.process(L, R)
    assert bits_equal(bm.tap("cost", 0), want["cost_hv"][0])
    assert n_diff(bm.tap("disp", 0), want["disp"][0]) == 0
    assert bits_equal(bm.tap("subpixel", 0), want["fsub"][0])
    assert bits_equal(bm.tap("median"), want["f_median"]) and bits_equal(bm.get_disp(), want["f_final"])
    bm.close()


@pytest.mark.parametrize("h,w,d,scale,noise", [(64, 128, 32, 1, 8), (90, 250, 64, 1, 30), (120, 200, 64, 2, 12)])
def test_bm_sad_vs_oracle(h, w, d, scale, noise):
    # BASELINE.json config 2 (SAD 9x9 + WTA + LR-check): oracle = restatement built on the reference's SAD()
    L, R, _ = synth.random_dot_pair(h, w, seed=h * 3 + d, bg=4, fg=12, noise=noise)
    want = restate.process(L, R, scale, d, "bm_sad", taps=restate.ALL_TAPS)
    bm = capi.BMSad(h, w, scale, d)
    bm.process(L, R)
    for v in (0, 1):
        assert bits_equal(bm.tap("cost", v), want["cost_raw"][v]), f"SAD volume view {v}"
        assert n_diff(bm.tap("disp", v), want["disp"][v]) == 0
    assert bits_equal(bm.tap("lr"), want["f_lr"])
    assert bits_equal(bm.tap("median"), want["f_median"]) and bits_equal(bm.get_disp(), want["f_final"])
    bm.close()


def test_dense_sad_entry_point_vs_oracle():
    L, R, _ = synth.random_dot_pair(40, 70, seed=2, bg=3, fg=7, noise=12)
    for wh, ww in ((9, 9), (7, 9), (3, 5)):
        want = restate.process(L, R, 1, 32, "bm_sad", taps=("cost_raw",), sad_h=wh, sad_w=ww)["cost_raw"][0]
        assert bits_equal(capi.cost_volume_sad(L, R, 32, wh, ww), want), (wh, ww)


def test_median_3x3_window_vs_oracle():
    # sgm_create accepts MEDIAN_FILTER_H/W = 3 (inc/Solver.h:17-18 are compile-time constants of the reference; the
    # restatement takes them as parameters): the other instantiation of the median kernels
    L, R, _ = synth.random_dot_pair(97, 161, seed=15, bg=5, fg=17, noise=60)
    _check_sgm_against_restatement(L, R, 1, 64, median_h=3, median_w=3)
    L, R, _ = synth.random_dot_pair(120, 200, seed=16, bg=4, fg=12, noise=30)
    want = restate.process(L, R, 1, 64, "bm_sad", median_h=3, median_w=3)
    bm = capi.BMSad(120, 200, 1, 64, median_h=3, median_w=3)
    assert bits_equal(bm.process(L, R).tap("median"), want["f_median"]) and bits_equal(bm.get_disp(), want["f_final"])
    bm.close()


@pytest.mark.parametrize("algo,noise", [("sgm", 200), ("bm_sad", 120), ("bm", 120)])
def test_median_many_candidates_spans_several_chunks(algo, noise):
    # very heavy noise leaves far more fillable holes (8e3 for SGM, 5e4 - 7e4 for block matching: a third of the frame is
    # invalid) than one shared-memory chunk of the median solver holds (3072): chunks are resolved in raster order and later
    # chunks read the earlier ones' fills; block matching also has dependency chains hundreds of holes deep
    h, w, d = 400, 1000, 64
    L, R, _ = synth.random_dot_pair(h, w, seed=23, bg=6, fg=25, noise=noise)
    want = restate.process(L, R, 1, d, algo)
    filled = int((want["f_median"].view(np.uint32) != want["f_lr"].view(np.uint32)).sum())
    assert filled > 2 * 3072, f"only {filled} holes were filled: the case does not span several chunks"
    sv = {"sgm": capi.SGM, "bm_sad": capi.BMSad, "bm": capi.BM}[algo](h, w, 1, d)
    sv.process(L, R)
    assert bits_equal(sv.tap("lr"), want["f_lr"]) and bits_equal(sv.tap("median"), want["f_median"])
    assert bits_equal(sv.get_disp(), want["f_final"])
    sv.close()


def test_dense_sad_generic_kernel_vs_oracle():
    # disparity counts that are not a power of two (and very wide windows) take the per-row SAD kernel instead of the banded one
    L, R, _ = synth.random_dot_pair(40, 70, seed=2, bg=3, fg=7, noise=12)
    for d, wh, ww in ((24, 9, 9), (40, 5, 7), (32, 3, 27)):
        want = restate.process(L, R, 1, 64, "bm_sad", taps=("cost_raw",), sad_h=wh, sad_w=ww)["cost_raw"][0][:, :, :d]
        assert bits_equal(capi.cost_volume_sad(L, R, d, wh, ww), np.ascontiguousarray(want)), (d, wh, ww)


@pytest.mark.parametrize("kind,wh,ww,scale,d", [("ssd", 7, 9, 1, 32), ("ssd", 9, 9, 2, 24), ("ssd", 3, 5, 1, 64), ("ssd", 15, 17, 1, 16),
                                                ("ct", 7, 9, 1, 32), ("ct", 7, 9, 2, 24), ("ct", 3, 3, 1, 64), ("ct", 5, 11, 1, 40)])
def test_dense_ssd_and_ct_volumes_vs_oracle(kind, wh, ww, scale, d):
    # SSD / CT of inc/cost.h:8-9 as dense device volumes (row A12): checker = the numpy restatement (pinned against the
    # reference's own functions on the CPU) and, where the compiled reference travelled, those functions themselves
    L, R, _ = synth.random_dot_pair(41, 77, seed=7 + wh + d, bg=3, fg=9, noise=25)
    got = capi.cost_volume(kind, L, R, d, wh, ww, scale)
    assert bits_equal(got, restate.window_cost_volume(kind, L, R, d, wh, ww, scale)), (kind, wh, ww, scale, d)
    if refbind.available():
        assert bits_equal(got, refbind.cost_volume(kind, L, R, d, wh, ww, scale))


def test_dense_ssd_and_ct_reject_oversized_windows():
    z = np.zeros((32, 32), np.uint8)
    with pytest.raises(capi.SgmError) as e:
        capi.cost_volume("ssd", z, z, 16, 17, 17)   # 289 pixels * 255^2 >= 2^24: float accumulation would round
    assert e.value.code == 2
    with pytest.raises(capi.SgmError) as e:
        capi.cost_volume("ct", z, z, 16, 9, 9)      # 80 taps do not fit the reference's 64-bit strings
    assert e.value.code == 2


def test_device_resident_pitched_sky_masks():
    # sgm_enqueue_device with sky masks whose row pitch is larger than the width (densified on the device), and with the
    # caller's left image overwritten right after sgm_wait (the engine keeps its own copy for sgm_point_cloud)
    torch = pytest.importorskip("torch")
    h, w, d = 120, 200, 64
    L, R, _ = synth.random_dot_pair(h, w, seed=31, bg=5, fg=17, noise=15)
    sl, sr = synth.sky_masks(h, w, 0.3)
    want = restate.process(L, R, 1, d, "sgm", sl, sr)["f_final"]
    tl, tr = torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()
    pitch = w + 56
    tsl, tsr = torch.zeros((h, pitch), dtype=torch.uint8, device="cuda"), torch.zeros((h, pitch), dtype=torch.uint8, device="cuda")
    tsl[:, :w], tsr[:, :w] = torch.from_numpy(sl).cuda(), torch.from_numpy(sr).cuda()
    tsl[:, w:], tsr[:, w:] = 255, 255            # padding bytes that must not be read as mask
    torch.cuda.synchronize()
    sv = capi.SGM(h, w, 1, d)
    sv.enqueue_device(tl.data_ptr(), w, tr.data_ptr(), w, tsl.data_ptr(), pitch, tsr.data_ptr(), pitch)
    sv.wait()
    assert bits_equal(sv.get_disp(), want)
    xyz0, int0 = sv.point_cloud(718.856, 718.856, 607.1928, 185.2157)
    tl.zero_()                                    # the caller reuses its buffer
    torch.cuda.synchronize()
    xyz1, int1 = sv.point_cloud(718.856, 718.856, 607.1928, 185.2157)
    assert np.array_equal(int0, int1) and np.array_equal(xyz0, xyz1) and int0.any()
    sv.close()


def test_strided_inputs_repeat_and_two_handles():
    h, w, d = 64, 160, 32
    L, R, _ = synth.random_dot_pair(h, w, seed=9, bg=4, fg=11, noise=10)
    want = restate.process(L, R, 1, d, "sgm")["f_final"]
    big_l, big_r = np.zeros((h, w + 37), np.uint8), np.zeros((h, w + 5), np.uint8)
    big_l[:, :w], big_r[:, :w] = L, R
    a, b = capi.SGM(h, w, 1, d), capi.SGM(h, w, 1, d)
    a.enqueue(big_l[:, :w], big_r[:, :w])
    b.enqueue(L, R)
    a.wait(), b.wait()
    assert bits_equal(a.get_disp(), want) and bits_equal(b.get_disp(), want)
    for _ in range(3):  # the handle is reusable and deterministic
        a.process(L, R)
        assert bits_equal(a.get_disp(), want)
    assert a.launches > 0
    a.close(), b.close()


def test_device_resident_inputs_match_host_path():
    torch = pytest.importorskip("torch")
    h, w, d = 96, 192, 64
    L, R, _ = synth.random_dot_pair(h, w, seed=10, bg=6, fg=20, noise=10)
    sv = capi.SGM(h, w, 1, d)
    want = sv.process(L, R).get_disp()
    tl, tr = torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()
    torch.cuda.synchronize()
    sv.enqueue_device(tl.data_ptr(), w, tr.data_ptr(), w)
    sv.wait()
    assert bits_equal(sv.get_disp(), want)
    sv.close()


def test_error_codes_instead_of_asserts():
    with pytest.raises(capi.SgmError) as e:
        capi.SGM(100, 100, 3, 64)          # src/Solver.cpp:8 asserts scale in {1,2}
    assert e.value.code == 1
    with pytest.raises(capi.SgmError) as e:
        capi.SGM(100, 100, 1, 48)          # src/Solver.cpp:10 asserts d in {32,64,128}
    assert e.value.code == 2
    with pytest.raises(capi.SgmError):
        capi.SGM(0, 100, 1, 64)
    sv = capi.SGM(64, 64, 1, 32)
    with pytest.raises(capi.SgmError):
        sv.get_disp()                       # nothing processed yet
    with pytest.raises(capi.SgmError):
        sv.process(np.zeros((64, 64), np.float32), np.zeros((64, 64), np.uint8))   # src/SGM.cpp:38
    with pytest.raises(capi.SgmError):
        sv.process(np.zeros((64, 65), np.uint8), np.zeros((64, 65), np.uint8))     # src/SGM.cpp:34-35
    with pytest.raises(capi.SgmError):
        sv.get_disp_u8(2)
    sv.close()


def _hd_taps(sv):
    got = {}
    for v, sfx in ((0, "_l"), (1, "_r")):
        got["census" + sfx] = H(sv.tap("census", v))
        got["cost_hv" + sfx] = H(sv.tap("cost", v))
        got["S" + sfx] = H(sv.tap("pathsum", v))
        got["disp" + sfx] = H(sv.tap("disp", v))
        got["fsub" + sfx] = H(sv.tap("subpixel", v))
    got.update(f_lr=H(sv.tap("lr")), f_median=H(sv.tap("median")), f_final=H(sv.get_disp()))
    return got


@pytest.mark.parametrize("name", ["hd_sgm_d256_clean", "hd_sgm_d256_noisy", "hd_int_d256_noisy"])
def test_full_size_1080p_d256_against_golden(golden_fullsize, name):
    # BASELINE.json configs[3] (1920x1080, D=256, 8 paths) at its full size, every tap by hash against the widened
    # restatement (tests/golden/make_golden_fullsize.py; the reference asserts d <= 128, src/Solver.cpp:10).  This size takes
    # code paths no smaller test reaches: one view per sweep launch (the two views do not fit a resident cooperative grid),
    # 8 words per lane, the tall-image median.
    c = golden_fullsize[name]
    h, w, d = c["h"], c["w"], c["d"]
    L, R, gt = synth.random_dot_pair(h, w, seed=c["seed"], bg=c["bg"], fg=c["fg"], noise=c["noise"])
    want = c["hashes"]
    assert (H(L), H(R)) == (want["left"], want["right"])
    over = dict(box_w=1, box_h=1, flags=capi.FLAG_KEEP_VOLUMES | capi.FLAG_INTEGER) if c["integer"] else dict(flags=capi.FLAG_KEEP_VOLUMES)
    sv = capi.SGM(h, w, 1, d, **over)
    sv.process(L, R)
    got = _hd_taps(sv)
    bad = sorted(k for k, v in got.items() if v != want[k])
    assert not bad, f"{name}: taps whose hash differs from the widened restatement: {bad}"
    f1 = sv.get_disp()
    assert int((f1 <= d - 1).sum()) == c["stats"]["valid"]
    assert int((sv.tap("median").view(np.uint32) != sv.tap("lr").view(np.uint32)).sum()) == c["stats"]["holes_filled"]
    # size-independent properties on top of the hashes: the handle is reusable and deterministic, the result is close to
    # the ground truth, invalid pixels carry exactly D+1
    assert bits_equal(sv.process(L, R).get_disp(), f1)
    valid = f1 <= d - 1
    cols = np.arange(w)[None, :] >= d
    assert ((valid & cols & (np.abs(f1 - gt) <= 1)).sum() > 0.9 * cols.sum() * h) and set(np.unique(f1[~valid])) == {float(d + 1)}
    assert sv.get_disp_u16(0).max() <= d + 1
    sv.close()


def test_full_size_1080p_flat_frame_has_no_texture():
    # a frame that is constant along x has no texture: identical left / right => disparity 0 or invalid everywhere
    h, w, d = 1080, 1920, 256
    flat = np.repeat(np.arange(h, dtype=np.uint8)[:, None] * 3, w, axis=1)
    sv = capi.SGM(h, w, 1, d)
    assert set(np.unique(sv.process(flat, flat).get_disp())) <= {0.0, float(d + 1)}
    sv.close()


@pytest.mark.parametrize("name", ["kitti_bmsad_d64_clean", "kitti_bmsad_d64_noisy"])
def test_bm_sad_kitti_size_against_golden(golden_fullsize, name):
    # BASELINE.json configs[1]: SAD 9x9 block matching + WTA + LR-check at 1242x375, D=64, every tap by hash against the
    # restatement built on the reference's own SAD() (src/cost.cpp:4-30; pinned in tests/test_oracle_cpu.py)
    c = golden_fullsize[name]
    L, R, _ = synth.random_dot_pair(c["h"], c["w"], seed=c["seed"], bg=c["bg"], fg=c["fg"], noise=c["noise"])
    want = c["hashes"]
    assert (H(L), H(R)) == (want["left"], want["right"])
    bm = capi.BMSad(c["h"], c["w"], c["scale"], c["d"])
    bm.process(L, R)
    got = dict(f_lr=H(bm.tap("lr")), f_median=H(bm.tap("median")), f_final=H(bm.get_disp()))
    for v, sfx in ((0, "_l"), (1, "_r")):
        got["cost_raw" + sfx] = H(bm.tap("cost", v))
        got["disp" + sfx] = H(bm.tap("disp", v))
        got["fsub" + sfx] = H(bm.tap("subpixel", v))
    bad = sorted(k for k, v in got.items() if v != want[k])
    assert not bad, f"{name}: taps whose hash differs from the restatement: {bad}"
    assert int((bm.get_disp() <= c["d"] - 1).sum()) == c["stats"]["valid"]
    bm.close()


def _fuzz_cases(n, seed):
    rng = np.random.RandomState(seed)
    cases = []
    for k in range(n):
        algo = ("sgm", "sgm", "sgm_int", "bm", "bm_sad")[k % 5]
        scale = int(rng.choice([1, 1, 2]))
        d = int(rng.choice([64, 128] if algo == "sgm_int" else [32, 64, 128]))
        h = int(rng.randint(16 * scale, 90 * scale))
        w = int(rng.randint(16 * scale, 260 * scale))
        cases.append(dict(algo=algo, h=h, w=w, d=d, scale=scale, noise=int(rng.choice([0, 5, 20, 60, 120])), sky=bool(rng.rand() < 0.4),
                          paths=int(rng.choice([8, 8, 4])) if algo.startswith("sgm") else 8, median=int(rng.choice([5, 5, 3])),
                          bg=int(rng.randint(0, 8)), fg=int(rng.randint(8, 30)), seed=int(rng.randint(1, 10 ** 6))))
    return cases


@pytest.mark.parametrize("c", _fuzz_cases(100, seed=20261105), ids=lambda c: f"{c['algo']}-{c['h']}x{c['w']}-d{c['d']}-s{c['scale']}")
def test_fuzz_random_shapes_and_modes_vs_oracle(c):
    # deterministic fuzzing over shapes (down to the 16x16 minimum, widths below D, odd sizes), scales, noise levels (none ..
    # a third of the frame invalid), sky masks, 4 / 8 paths, both median windows and every pipeline, final and intermediate
    # results against the oracle bit for bit
    h, w, d, s = c["h"], c["w"], c["d"], c["scale"]
    L, R, _ = synth.random_dot_pair(h, w, seed=c["seed"], bg=c["bg"], fg=min(c["fg"], max(1, w // s // 3)), noise=c["noise"])
    sl, sr = synth.sky_masks(h // s, w // s, 0.3) if c["sky"] else (None, None)
    over = dict(median_h=c["median"], median_w=c["median"])
    if c["algo"] in ("sgm", "sgm_int"):
        integer = c["algo"] == "sgm_int"
        if integer:
            over.update(box_w=s, box_h=s)
        want = restate.process(L, R, s, d, "sgm", sl, sr, taps=restate.ALL_TAPS, paths=c["paths"], **over)
        sv = capi.SGM(h, w, s, d, flags=(capi.FLAG_INTEGER if integer else 0), paths=c["paths"], **over)
    else:
        want = restate.process(L, R, s, d, c["algo"], sl, sr, taps=restate.ALL_TAPS, **over)
        sv = (capi.BM if c["algo"] == "bm" else capi.BMSad)(h, w, s, d, **over)
    sv.process(L, R, sl, sr)
    bad = []
    views = (0,) if c["algo"] == "bm" else (0, 1)
    for v in views:
        if n_diff(sv.tap("disp", v), want["disp"][v]):
            bad.append(f"disp[{v}]")
        if not bits_equal(sv.tap("subpixel", v), want["fsub"][v]):
            bad.append(f"subpixel[{v}]")
    for tap, key in (("lr", "f_lr"), ("median", "f_median")):
        if not bits_equal(sv.tap(tap), want[key]):
            bad.append(tap)
    if not bits_equal(sv.get_disp(), want["f_final"]):
        bad.append("final")
    sv.close()
    assert not bad, f"{c}: {bad}"
