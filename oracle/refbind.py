"""ctypes bindings of oracle/_ref/libsgm_ref.so — the reference's own CPU sources
(/root/reference/src/{Solver,SGM,BM,cost}.cpp) compiled unmodified by oracle/Makefile.

TEST INFRASTRUCTURE: the checker, never the thing measured as the product.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libsgm_ref.so")
# the same unmodified sources compiled against a copy of inc/Solver.h with COST_WIN_H = COST_WIN_W = 1 (make -C oracle
# ref_nofilter): the reference with its cost filter off, the direct pin of the integer / DPX mode's oracle
LIB_PATH_NOFILTER = os.path.join(_HERE, "_ref", "libsgm_ref_nofilter.so")

_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
_libs = {}
_variant = "default"


def _path(variant):
    return LIB_PATH_NOFILTER if variant == "nofilter" else LIB_PATH


def available(variant="default") -> bool:
    return os.path.exists(_path(variant))


class use:
    """`with refbind.use("nofilter"): ...` routes every call of this module to libsgm_ref_nofilter.so."""

    def __init__(self, variant):
        self.variant = variant

    def __enter__(self):
        global _variant
        self.prev, _variant = _variant, self.variant
        return self

    def __exit__(self, *exc):
        global _variant
        _variant = self.prev


def lib():
    if _variant not in _libs:
        if not available(_variant):
            raise RuntimeError(f"{_path(_variant)} missing: run `make -C oracle ref ref_nofilter` where /root/reference exists")
        L = C.CDLL(_path(_variant))
        L.ref_process.restype = C.c_double
        L.ref_process.argtypes = [C.c_int] * 5 + [C.c_void_p] * 9
        L.ref_time_process.restype = C.c_int
        L.ref_time_process.argtypes = [C.c_int] * 5 + [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.ref_stages.restype = C.c_int
        L.ref_stages.argtypes = [C.c_int] * 5 + [C.c_void_p] * 3 + [C.c_int] + [C.c_void_p] * 5
        L.ref_cost_volume.restype = C.c_int
        L.ref_cost_volume.argtypes = [C.c_int] * 7 + [C.c_void_p] * 3
        L.ref_hamming_cost.restype = C.c_int
        L.ref_hamming_cost.argtypes = [C.c_uint64, C.c_uint64]
        L.ref_gaussian_blur.restype = C.c_int
        L.ref_gaussian_blur.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_double, C.c_double]
        if hasattr(L, "ref_show_disp"):
            L.ref_show_disp.restype = C.c_int
            L.ref_show_disp.argtypes = [C.c_int] * 5 + [C.c_void_p] * 4
        if hasattr(L, "ref_open"):
            L.ref_open.restype = C.c_void_p
            L.ref_open.argtypes = [C.c_int] * 5
            L.ref_run.restype = C.c_double
            L.ref_run.argtypes = [C.c_void_p] * 4
            L.ref_close.argtypes = [C.c_void_p]
        L.ref_set_threads.argtypes = [C.c_int]
        L.ref_set_quiet.argtypes = [C.c_int]
        L.ref_max_threads.restype = C.c_int
        _libs[_variant] = L
    return _libs[_variant]


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _u8(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.uint8)


def set_threads(n: int):
    lib().ref_set_threads(int(n))


def max_threads() -> int:
    return lib().ref_max_threads()


ALGO = {"sgm": 0, "bm": 1}


def process(algo, left, right, scale, d, sky_l=None, sky_r=None, want_cost=False):
    """Runs SGM::process / BM::process of the reference.  Returns a dict of numpy arrays + 'ms'."""
    left, right = _u8(left), _u8(right)
    h, w = left.shape
    ih, iw = h // scale, w // scale
    sky_l, sky_r = _u8(sky_l), _u8(sky_r)
    out = dict(
        disp=np.zeros((ih, iw), np.uint8),
        disp_beta=np.zeros((ih, iw), np.uint8),
        fdisp=np.zeros((ih, iw), np.float32),
        fdisp_beta=np.zeros((ih, iw), np.float32),
    )
    cost = np.zeros((ih, iw, d), np.float32) if want_cost else None
    ms = lib().ref_process(ALGO[algo], h, w, scale, d, _p(left), _p(right), _p(sky_l), _p(sky_r), _p(out["disp"]),
                           _p(out["disp_beta"]), _p(out["fdisp"]), _p(out["fdisp_beta"]), _p(cost))
    if ms < 0:
        raise RuntimeError("ref_process failed")
    out["ms"] = ms
    if want_cost:
        out["cost_last"] = cost
    return out


def show_disp(algo, left, right, scale, d):
    """process() then Solver::show_disp: (debug view uint8 [2*ih, iw, 3], final float disparity)."""
    left, right = _u8(left), _u8(right)
    h, w = left.shape
    ih, iw = h // scale, w // scale
    view = np.zeros((2 * ih, iw, 3), np.uint8)
    f = np.zeros((ih, iw), np.float32)
    if lib().ref_show_disp(ALGO[algo], h, w, scale, d, _p(left), _p(right), _p(view), _p(f)) != 0:
        raise RuntimeError("ref_show_disp failed")
    return view, f


def time_process(algo, left, right, scale, d, iters):
    left, right = _u8(left), _u8(right)
    h, w = left.shape
    ms = np.zeros(iters, np.float64)
    rc = lib().ref_time_process(ALGO[algo], h, w, scale, d, _p(left), _p(right), iters, _p(ms))
    if rc != 0:
        raise RuntimeError("ref_time_process failed")
    return ms


class Session:
    """One reference solver kept across frames (node.cpp:49,93): run(left, right) -> wall ms of process()."""

    def __init__(self, algo, h, w, scale, d):
        self.shape = (h // scale, w // scale)
        self._s = lib().ref_open(ALGO[algo], h, w, scale, d)
        if not self._s:
            raise RuntimeError("ref_open failed")

    def run(self, left, right, want_disp=False):
        left, right = _u8(left), _u8(right)
        out = np.zeros(self.shape, np.float32) if want_disp else None
        ms = lib().ref_run(self._s, _p(left), _p(right), _p(out))
        if ms < 0:
            raise RuntimeError("ref_run failed")
        return (ms, out) if want_disp else ms

    def close(self):
        if self._s:
            lib().ref_close(self._s)
            self._s = None


def stages(algo, left, right, scale, d, beta=False, sky=None):
    """Intermediate taps: census tables, raw / h-filtered / hv-filtered cost volume of one view."""
    left, right = _u8(left), _u8(right)
    h, w = left.shape
    ih, iw = h // scale, w // scale
    sky = _u8(sky)
    out = dict(
        ct_l=np.zeros((ih, iw), np.uint64),
        ct_r=np.zeros((ih, iw), np.uint64),
        cost_raw=np.zeros((ih, iw, d), np.float32),
        cost_h=np.zeros((ih, iw, d), np.float32),
        cost_hv=np.zeros((ih, iw, d), np.float32),
    )
    rc = lib().ref_stages(ALGO[algo], h, w, scale, d, _p(left), _p(right), _p(sky), int(beta), _p(out["ct_l"]),
                          _p(out["ct_r"]), _p(out["cost_raw"]), _p(out["cost_h"]), _p(out["cost_hv"]))
    if rc != 0:
        raise RuntimeError("ref_stages failed")
    return out


def cost_volume(kind, left, right, d, win_h, win_w, scale=1):
    left, right = _u8(left), _u8(right)
    h, w = left.shape
    out = np.zeros((h, w, d), np.float32)
    rc = lib().ref_cost_volume({"sad": 0, "ssd": 1, "ct": 2}[kind], h, w, d, win_h, win_w, scale, _p(left), _p(right),
                               _p(out))
    if rc != 0:
        raise RuntimeError("ref_cost_volume failed")
    return out


def gaussian_blur(img, sx=2.0, sy=1.0):
    img = _u8(img)
    out = np.zeros_like(img)
    lib().ref_gaussian_blur(img.shape[0], img.shape[1], _p(img), _p(out), float(sx), float(sy))
    return out
