"""ctypes bindings of oracle/_build/libsgm_oracle.so — our plain-C restatement of the reference's
SGM / BM CPU path (oracle/sgm_oracle.c).  TEST INFRASTRUCTURE: the checker, never the product.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_build", "libsgm_oracle.so")
ALGO = {"sgm": 0, "bm": 1, "bm_sad": 2}


class Params(C.Structure):
    _fields_ = [(k, C.c_int) for k in ("height", "width", "scale", "max_disp", "algo", "paths", "P1", "P2", "census_h",
                                        "census_w", "box_w", "box_h")] + [("unique_ratio", C.c_float),
                                                                           ("lr_thresh", C.c_float)] + [
                                                                               (k, C.c_int)
                                                                               for k in ("median_h", "median_w",
                                                                                         "speckle_size", "speckle_dis",
                                                                                         "sad_h", "sad_w")
                                                                           ]


class Taps(C.Structure):
    _fields_ = [(k, C.c_void_p * 2) for k in ("img", "blur", "census", "cost_raw", "cost_h", "cost_hv", "S", "disp",
                                              "fsub")] + [(k, C.c_void_p) for k in ("f_lr", "f_median", "f_final")]


_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "restate"])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        L.orc_default_params.argtypes = [C.POINTER(Params)] + [C.c_int] * 5
        L.orc_process.restype = C.c_int
        L.orc_process.argtypes = [C.POINTER(Params)] + [C.c_void_p] * 4 + [C.POINTER(Taps)]
        L.orc_fnv1a64.restype = C.c_uint64
        L.orc_fnv1a64.argtypes = [C.c_void_p, C.c_size_t]
        _lib = L
    return _lib


def default_params(h, w, scale, d, algo="sgm", **over) -> Params:
    p = Params()
    lib().orc_default_params(C.byref(p), h, w, scale, d, ALGO[algo])
    for k, v in over.items():
        setattr(p, k, v)
    return p


_TAP_SPECS = {  # name -> (dtype, per-view?, has D axis?)
    "img": (np.uint8, True, False),
    "blur": (np.uint8, True, False),
    "census": (np.uint64, True, False),
    "cost_raw": (np.float32, True, True),
    "cost_h": (np.float32, True, True),
    "cost_hv": (np.float32, True, True),
    "S": (np.float32, True, True),
    "disp": (np.uint16, True, False),
    "fsub": (np.float32, True, False),
    "f_lr": (np.float32, False, False),
    "f_median": (np.float32, False, False),
    "f_final": (np.float32, False, False),
}
LIGHT_TAPS = ("census", "disp", "fsub", "f_lr", "f_median", "f_final")
ALL_TAPS = tuple(_TAP_SPECS)


def process(left, right, scale, d, algo="sgm", sky_l=None, sky_r=None, taps=LIGHT_TAPS, **over):
    """Runs the restatement; returns {tap: array or [left, beta] arrays}."""
    left = np.ascontiguousarray(left, np.uint8)
    right = np.ascontiguousarray(right, np.uint8)
    h, w = left.shape
    ih, iw = h // scale, w // scale
    p = default_params(h, w, scale, d, algo, **over)
    t = Taps()
    out = {}
    for name in taps:
        dt, per_view, has_d = _TAP_SPECS[name]
        shape = (ih, iw, d) if has_d else (ih, iw)
        if per_view:
            arrs = [np.zeros(shape, dt), np.zeros(shape, dt)]
            getattr(t, name)[0] = arrs[0].ctypes.data
            getattr(t, name)[1] = arrs[1].ctypes.data
            out[name] = arrs
        else:
            a = np.zeros(shape, dt)
            setattr(t, name, a.ctypes.data)
            out[name] = a
    sl = None if sky_l is None else np.ascontiguousarray(sky_l, np.uint8)
    sr = None if sky_r is None else np.ascontiguousarray(sky_r, np.uint8)
    rc = lib().orc_process(C.byref(p), left.ctypes.data, right.ctypes.data, None if sl is None else sl.ctypes.data,
                           None if sr is None else sr.ctypes.data, C.byref(t))
    if rc != 0:
        raise RuntimeError("orc_process failed")
    return out


def aggregate_u16(cost, P1=10, P2=100, paths=8) -> np.ndarray:
    """orc_aggregate_u16: the eight SGM paths and their sum in 16-bit integers on an (h, w, d) uint16 cost volume."""
    cost = np.ascontiguousarray(cost, np.uint16)
    h, w, d = cost.shape
    S = np.zeros_like(cost)
    f = lib().orc_aggregate_u16
    f.restype = None
    f.argtypes = [C.c_void_p] + [C.c_int] * 6 + [C.c_void_p]
    f(cost.ctypes.data, h, w, d, P1, P2, paths, S.ctypes.data)
    return S


def fnv1a64(buf) -> str:
    a = np.ascontiguousarray(buf)
    return f"{lib().orc_fnv1a64(a.ctypes.data, a.nbytes):016x}"


# ---------------------------------------------------------------------------------------------
# Consumers of the final disparity (numpy restatements; TEST INFRASTRUCTURE like the rest of oracle/)
def colormap(f, d) -> np.ndarray:
    """Solver::colormap (src/Solver.cpp:652-707): float32 disparity -> (h, w, 3) BGR, with the reference's float / double
    arithmetic and float -> uchar truncations.  Pinned against the compiled reference's show_disp in tests/test_oracle_cpu.py."""
    f = np.asarray(f, np.float32)
    out = np.zeros(f.shape + (3,), np.uint8)
    v = f * np.float32(256 // d)
    valid = ~(f > np.float32(d - 1))
    s1 = valid & (v <= 51)
    s2 = valid & ~s1 & (v <= 102)
    s3 = valid & ~s1 & ~s2 & (v <= 153)
    s4 = valid & ~s1 & ~s2 & ~s3 & (v <= 204)
    s5 = valid & ~s1 & ~s2 & ~s3 & ~s4
    five = np.float32(5)
    out[s1, 0] = 255
    out[s1, 1] = (v[s1] * five).astype(np.int32).astype(np.uint8)
    t = v[s2] - np.float32(51)
    out[s2, 0] = (np.float32(255) - t * five).astype(np.int32).astype(np.uint8)
    out[s2, 1] = 255
    t = v[s3] - np.float32(102)
    out[s3, 1] = 255
    out[s3, 2] = (t * five).astype(np.int32).astype(np.uint8)
    t = (v[s4] - np.float32(153)).astype(np.float64)
    out[s4, 1] = (255 - (128.0 * t / 51.0 + 0.5).astype(np.int32).astype(np.uint8).astype(np.int32)).astype(np.uint8)
    out[s4, 2] = 255
    t = (v[s5] - np.float32(204)).astype(np.float64)
    out[s5, 1] = (127 - (127.0 * t / 51.0 + 0.5).astype(np.int32).astype(np.uint8).astype(np.int32)).astype(np.uint8)
    out[s5, 2] = 255
    return out


def debug_view(left, f, d, scale=1, bm_rows=False) -> np.ndarray:
    """Solver::show_disp (src/Solver.cpp:55-93): working-resolution left image as BGR on top, colour map copied over
    rows h-1 .. 2h-2, last row zero.  bm_rows: BM decimates with source row i instead of i*scale (src/BM.cpp:23-24)."""
    ih, iw = f.shape
    gray = np.asarray(left)[:: (1 if (bm_rows or scale == 1) else scale)][:ih, ::scale][:, :iw]
    out = np.zeros((2 * ih, iw, 3), np.uint8)
    out[:ih] = gray[:, :, None]
    out[ih - 1: 2 * ih - 1] = colormap(f, d)
    return out


def point_cloud(f, left, d, scale, fx, fy, cx, cy, baseline=0.5, max_range=100.0):
    """node.cpp:119-143: raster-order list of camera-frame points and grey values.  (fx + fy) is a float add, the rest of
    Zc is double; (j*scale - cx) is a float subtraction; the grey value is left(i, j) with WORKING-resolution indices."""
    f = np.asarray(f, np.float32)
    ih, iw = f.shape
    fx, fy, cx, cy = (np.float32(x) for x in (fx, fy, cx, cy))
    jj, ii = np.meshgrid(np.arange(iw, dtype=np.int32), np.arange(ih, dtype=np.int32))
    with np.errstate(divide="ignore"):
        Z = np.float64(fx + fy) / 2.0 * np.float64(baseline) / (f.astype(np.float64) + 1e-6)
    keep = (f != np.float32(d + 1)) & ~(Z > np.float64(np.float32(max_range)))
    X = ((jj * scale).astype(np.float32) - cx).astype(np.float64) * Z / np.float64(fx)
    Y = ((ii * scale).astype(np.float32) - cy).astype(np.float64) * Z / np.float64(fy)
    xyz = np.stack([X[keep], Y[keep], Z[keep]], axis=1)
    inten = np.asarray(left)[:ih, :iw][keep]
    return xyz, inten


def window_cost_volume(kind, left, right, d, win_h, win_w, scale=1) -> np.ndarray:
    """Dense numpy restatement of SSD (src/cost.cpp:33-59) / CT (:62-96) as Solver::build_dsi (src/Solver.cpp:96-117) would call
    them for every (pixel, disparity) of the left view -> (h, w, d) float32.  Pinned against the reference's own functions
    (refbind.cost_volume) in tests/test_oracle_cpu.py.
      SSD: sum over the window of (L(y, x_l) - R(y, max(x_l - k/scale, 0)))^2 with replicate-clamped (y, x_l), / win_h / win_w.
      CT : number of window taps (centre skipped, taps with x_r = x_l - k/scale < 0 dropped) whose comparisons against the
           centre pixels L(y, x) and R(y, max(x - k, 0)) — k NOT divided by scale — disagree."""
    L = np.ascontiguousarray(left, np.uint8).astype(np.int32)
    R = np.ascontiguousarray(right, np.uint8).astype(np.int32)
    h, w = L.shape
    rh, rw = win_h // 2, win_w // 2
    ys, xs = np.arange(h), np.arange(w)
    out = np.zeros((h, w, d), np.float32)
    for k in range(d):
        ks = k // scale
        acc = np.zeros((h, w), np.int64)
        if kind == "ct":
            cl = L
            cr = R[:, np.maximum(xs - k, 0)]
        for a in range(-rh, rh + 1):
            y = np.clip(ys + a, 0, h - 1)
            for b in range(-rw, rw + 1):
                xl = np.clip(xs + b, 0, w - 1)
                if kind == "ssd":
                    xr = np.maximum(xl - ks, 0)
                    df = L[y][:, xl] - R[y][:, xr]
                    acc += df * df
                else:
                    if a == 0 and b == 0:
                        continue
                    xr = xl - ks
                    keep = xr >= 0
                    bl = L[y][:, xl] > cl
                    br = R[y][:, np.maximum(xr, 0)] > cr
                    acc += (bl != br) & keep[None, :]
        if kind == "ssd":
            out[:, :, k] = (acc.astype(np.float32) / np.float32(win_h)) / np.float32(win_w)
        else:
            out[:, :, k] = acc.astype(np.float32)
    return out
