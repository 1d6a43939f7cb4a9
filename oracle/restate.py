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


def fnv1a64(buf) -> str:
    a = np.ascontiguousarray(buf)
    return f"{lib().orc_fnv1a64(a.ctypes.data, a.nbytes):016x}"
