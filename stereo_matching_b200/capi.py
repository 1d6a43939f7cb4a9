"""ctypes binding of the C ABI (include/sgm_b200.h -> stereo_matching_b200/libsgm_b200.so) and a thin
Python mirror of the reference's class surface (inc/Solver.h:23-70: Solver(h, w, s, d), process(l, r[, sky,
sky_beta]), get_disp()), used by the tests and bench.py.  There is no CPU path: if the library is missing
or no CUDA device is visible, construction raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# SGM_B200_LIBRARY (read here, in the Python test harness, never by the C library): load another build of the library, e.g. the
# tuning build libsgm_b200_tuning.so with its experiment knobs
LIB_PATH = os.environ.get("SGM_B200_LIBRARY") or os.path.join(HERE, "libsgm_b200.so")

ALGO_SGM, ALGO_BM, ALGO_BM_SAD = 0, 1, 2
FLAG_KEEP_VOLUMES, FLAG_STAGE_TIMING, FLAG_LINE_WALKERS, FLAG_INTEGER, FLAG_HOST_TAPS = 1, 2, 4, 8, 16
TAP = dict(census=0, cost=1, pathsum=2, disp=3, subpixel=4, lr=5, median=6, final=7, blur=8)
_TAP_DTYPE = dict(census=np.uint64, cost=np.float32, pathsum=np.float32, disp=np.uint16, subpixel=np.float32,
                  lr=np.float32, median=np.float32, final=np.float32, blur=np.uint8)
STAGES = ("upload", "cost", "aggregate", "post", "download")

EXPORTS = [
    "sgm_abi_version", "sgm_default_params", "sgm_create", "sgm_destroy", "sgm_last_error", "sgm_process", "sgm_enqueue",
    "sgm_wait", "sgm_enqueue_device", "sgm_get_disparity_f32", "sgm_get_disparity_u8", "sgm_get_disparity_u16",
    "sgm_disparity_device", "sgm_tap_bytes", "sgm_get_tap", "sgm_get_size", "sgm_get_launch_count", "sgm_get_stage_ms",
    "sgm_get_stream", "sgm_device_bytes", "sgm_cost_volume_sad", "sgm_render_debug_view", "sgm_point_cloud",
    "sgm_get_subpixel_beta", "sgm_speckle_filter_static", "sgm_stream_create", "sgm_stream_destroy", "sgm_stream_submit",
    "sgm_stream_next", "sgm_stream_release", "sgm_stream_pending", "sgm_stream_launches", "sgm_stream_last_error", "sgm_cost_volume_ssd", "sgm_cost_volume_ct",
]


class Params(C.Structure):
    _fields_ = ([(k, C.c_int32) for k in ("height", "width", "scale", "max_disp", "algo", "paths", "P1", "P2", "census_h",
                                          "census_w", "box_w", "box_h")] + [("unique_ratio", C.c_float), ("lr_thresh", C.c_float)] +
                [(k, C.c_int32) for k in ("median_h", "median_w", "speckle_size", "speckle_dis", "sad_h", "sad_w", "device")] +
                [("flags", C.c_uint32)])


class Camera(C.Structure):
    """sgm_camera: CamIntrinsics (inc/utils.h:14-20) + baseline / max_range of node.cpp:117-118."""
    _fields_ = [("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float), ("baseline", C.c_double),
                ("max_range", C.c_float)]


class SgmError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"sgm_b200 error {code}: {msg}")
        self.code = code


_lib = None


def lib():
    """Loads libsgm_b200.so; fails loudly when it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: run `python -m stereo_matching_b200.build` (needs nvcc)")
        L = C.CDLL(LIB_PATH)
        vp, sz, i = C.c_void_p, C.c_size_t, C.c_int
        L.sgm_abi_version.restype = i
        L.sgm_default_params.argtypes = [C.POINTER(Params), i, i, i, i, i]
        L.sgm_default_params.restype = None
        L.sgm_create.argtypes = [C.POINTER(vp), C.POINTER(Params)]
        L.sgm_destroy.argtypes = [vp]
        L.sgm_destroy.restype = None
        L.sgm_last_error.argtypes = [vp]
        L.sgm_last_error.restype = C.c_char_p
        for name in ("sgm_process", "sgm_enqueue", "sgm_enqueue_device"):
            getattr(L, name).argtypes = [vp, vp, sz, vp, sz, vp, sz, vp, sz]
        L.sgm_wait.argtypes = [vp]
        L.sgm_get_disparity_f32.argtypes = [vp, vp, sz]
        L.sgm_get_disparity_u8.argtypes = [vp, i, vp, sz]
        L.sgm_get_disparity_u16.argtypes = [vp, i, vp, sz]
        L.sgm_get_subpixel_beta.argtypes = [vp, vp, sz]
        L.sgm_disparity_device.argtypes = [vp]
        L.sgm_disparity_device.restype = vp
        L.sgm_tap_bytes.argtypes = [vp, i]
        L.sgm_tap_bytes.restype = sz
        L.sgm_get_tap.argtypes = [vp, i, i, vp, sz]
        L.sgm_get_size.argtypes = [vp, C.POINTER(i), C.POINTER(i), C.POINTER(i)]
        L.sgm_get_launch_count.argtypes = [vp]
        L.sgm_get_stage_ms.argtypes = [vp, C.POINTER(C.c_float), i]
        L.sgm_get_stream.argtypes = [vp]
        L.sgm_get_stream.restype = vp
        L.sgm_device_bytes.argtypes = [vp]
        L.sgm_device_bytes.restype = sz
        for name in ("sgm_cost_volume_sad", "sgm_cost_volume_ssd", "sgm_cost_volume_ct"):
            getattr(L, name).argtypes = [i, vp, vp, i, i, i, i, i, i, vp]
        L.sgm_speckle_filter_static.argtypes = [i, vp, vp, i, i, C.c_float, i, C.c_float, vp]
        L.sgm_stream_create.argtypes = [C.POINTER(vp), C.POINTER(Params), C.POINTER(i), i, i]
        L.sgm_stream_destroy.argtypes = [vp]
        L.sgm_stream_destroy.restype = None
        L.sgm_stream_submit.argtypes = [vp, vp, sz, vp, sz, vp, sz, vp, sz, C.POINTER(C.c_uint64)]
        L.sgm_stream_next.argtypes = [vp, vp, sz, C.POINTER(vp), C.POINTER(C.c_uint64)]
        L.sgm_stream_release.argtypes = [vp]
        L.sgm_stream_pending.argtypes = [vp]
        L.sgm_stream_launches.argtypes = [vp]
        L.sgm_stream_launches.restype = C.c_uint64
        L.sgm_stream_last_error.argtypes = [vp]
        L.sgm_stream_last_error.restype = C.c_char_p
        L.sgm_render_debug_view.argtypes = [vp, vp, sz]
        L.sgm_point_cloud.argtypes = [vp, C.POINTER(Camera), C.POINTER(C.c_double), vp, sz, C.POINTER(sz)]
        _lib = L
    return _lib


def default_params(h, w, scale, d, algo=ALGO_SGM, **over) -> Params:
    p = Params()
    lib().sgm_default_params(C.byref(p), h, w, scale, d, algo)
    for k, v in over.items():
        if not hasattr(p, k):
            raise AttributeError(k)
        setattr(p, k, v)
    return p


def _img(a):
    a = np.asarray(a)
    if a.dtype != np.uint8 or a.ndim != 2:
        raise SgmError(1, "images must be 2-D uint8 arrays (the reference asserts CV_8UC1, src/SGM.cpp:38)")
    if a.strides[1] != 1:
        a = np.ascontiguousarray(a)
    return a


class Solver:
    """Python mirror of the reference's Solver(h, w, s, d) surface over the C ABI."""

    algo = ALGO_SGM

    def __init__(self, h, w, s, d, device=0, flags=0, **over):
        self._h = None
        self.params = default_params(h, w, s, d, self.algo, device=device, flags=flags, **over)
        handle = C.c_void_p()
        rc = lib().sgm_create(C.byref(handle), C.byref(self.params))
        if rc != 0:
            raise SgmError(rc, lib().sgm_last_error(None).decode())
        self._h = handle
        hh, ww, dd = C.c_int(), C.c_int(), C.c_int()
        lib().sgm_get_size(self._h, C.byref(hh), C.byref(ww), C.byref(dd))
        self.img_h, self.img_w, self.max_disp = hh.value, ww.value, dd.value
        self.invalid_disp = d + 1
        self._keep = []

    def _check(self, rc):
        if rc != 0:
            raise SgmError(rc, lib().sgm_last_error(self._h).decode())

    def close(self):
        if self._h is not None:
            lib().sgm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _args(self, img_l, img_r, sky_mask, sky_mask_beta):
        l, r = _img(img_l), _img(img_r)
        if l.shape != r.shape or l.shape != (self.params.height, self.params.width):
            raise SgmError(1, f"image shape {l.shape}/{r.shape} does not match the solver ({self.params.height}, {self.params.width})")
        sl = None if sky_mask is None else _img(sky_mask)
        sr = None if sky_mask_beta is None else _img(sky_mask_beta)
        for m in (sl, sr):
            if m is not None and m.shape != (self.img_h, self.img_w):
                raise SgmError(1, "sky mask shape must be (h/scale, w/scale)")
        self._keep = [l, r, sl, sr]
        p = lambda a: None if a is None else a.ctypes.data
        st = lambda a: 0 if a is None else a.strides[0]
        return self._h, p(l), st(l), p(r), st(r), p(sl), st(sl), p(sr), st(sr)

    def process(self, img_l, img_r, sky_mask=None, sky_mask_beta=None):
        self._check(lib().sgm_process(*self._args(img_l, img_r, sky_mask, sky_mask_beta)))
        return self

    def enqueue(self, img_l, img_r, sky_mask=None, sky_mask_beta=None):
        self._check(lib().sgm_enqueue(*self._args(img_l, img_r, sky_mask, sky_mask_beta)))

    def enqueue_device(self, left_ptr, left_stride, right_ptr, right_stride, sky_l_ptr=None, sky_l_stride=0, sky_r_ptr=None,
                       sky_r_stride=0):
        self._check(lib().sgm_enqueue_device(self._h, left_ptr, left_stride, right_ptr, right_stride, sky_l_ptr, sky_l_stride,
                                             sky_r_ptr, sky_r_stride))

    def wait(self):
        self._check(lib().sgm_wait(self._h))

    def get_disp(self) -> np.ndarray:
        out = np.empty((self.img_h, self.img_w), np.float32)
        self._check(lib().sgm_get_disparity_f32(self._h, out.ctypes.data, out.strides[0]))
        return out

    def get_disp_into(self, out: np.ndarray) -> None:
        """get_disp() into a caller-owned float32 (img_h, img_w) array (no allocation)."""
        self._check(lib().sgm_get_disparity_f32(self._h, out.ctypes.data, out.strides[0]))

    def get_disp_u8(self, view=0) -> np.ndarray:
        out = np.empty((self.img_h, self.img_w), np.uint8)
        self._check(lib().sgm_get_disparity_u8(self._h, view, out.ctypes.data, out.strides[0]))
        return out

    def get_disp_u16(self, view=0) -> np.ndarray:
        out = np.empty((self.img_h, self.img_w), np.uint16)
        self._check(lib().sgm_get_disparity_u16(self._h, view, out.ctypes.data, out.strides[0]))
        return out

    def get_subpixel_beta(self) -> np.ndarray:
        """Solver::filtered_disp_beta (src/SGM.cpp:801): the right view's sub-pixel disparity."""
        out = np.empty((self.img_h, self.img_w), np.float32)
        self._check(lib().sgm_get_subpixel_beta(self._h, out.ctypes.data, out.strides[0]))
        return out

    def tap(self, name, view=0) -> np.ndarray:
        dt = _TAP_DTYPE[name]
        shape = (self.img_h, self.img_w, self.max_disp) if name in ("cost", "pathsum") else (self.img_h, self.img_w)
        out = np.empty(shape, dt)
        self._check(lib().sgm_get_tap(self._h, TAP[name], view, out.ctypes.data, out.nbytes))
        return out

    def debug_view(self) -> np.ndarray:
        """Solver::show_disp(debug_view) (src/Solver.cpp:55-93): (2*img_h, img_w, 3) BGR."""
        out = np.empty((2 * self.img_h, self.img_w, 3), np.uint8)
        self._check(lib().sgm_render_debug_view(self._h, out.ctypes.data, out.strides[0]))
        return out

    def point_cloud(self, fx, fy, cx, cy, baseline=0.5, max_range=100.0):
        """node.cpp:119-143: (xyz float64 [n, 3], intensity uint8 [n]) in raster order."""
        cam = Camera(fx, fy, cx, cy, baseline, max_range)
        cap = self.img_h * self.img_w
        xyz = np.empty((cap, 3), np.float64)
        inten = np.empty(cap, np.uint8)
        n = C.c_size_t()
        self._check(lib().sgm_point_cloud(self._h, C.byref(cam), xyz.ctypes.data_as(C.POINTER(C.c_double)), inten.ctypes.data, cap, C.byref(n)))
        return xyz[: n.value].copy(), inten[: n.value].copy()

    @property
    def launches(self) -> int:
        return lib().sgm_get_launch_count(self._h)

    @property
    def device_bytes(self) -> int:
        return lib().sgm_device_bytes(self._h)

    def stage_ms(self) -> dict:
        ms = (C.c_float * len(STAGES))()
        self._check(lib().sgm_get_stage_ms(self._h, ms, len(STAGES)))
        return dict(zip(STAGES, list(ms)))

    @property
    def stream_ptr(self) -> int:
        """cudaStream_t the engine launches on (for CUDA-event timing on the launching stream)."""
        return lib().sgm_get_stream(self._h)

    def disparity_device_ptr(self) -> int:
        return lib().sgm_disparity_device(self._h)


class SGM(Solver):
    """inc/SGM.h:10-27"""
    algo = ALGO_SGM


class BM(Solver):
    """inc/BM.h:6-18"""
    algo = ALGO_BM


class BMSad(Solver):
    """SAD block matching (BASELINE.json config 2), built from the reference's SAD() (src/cost.cpp:4-30)."""
    algo = ALGO_BM_SAD


class Stream:
    """sgm_stream: frames of one size over `devices` (list of CUDA ordinals), `inflight` frames in flight per device, results
    in submission order.  submit() blocks only for back-pressure; next() blocks until the oldest pending frame is done."""

    def __init__(self, h, w, s, d, devices=(0,), inflight=4, algo=ALGO_SGM, flags=0, **over):
        self._h = None
        self.params = default_params(h, w, s, d, algo, flags=flags, **over)
        self.img_h, self.img_w = h // s, w // s
        self.capacity = len(devices) * inflight
        devs = (C.c_int * len(devices))(*devices)
        handle = C.c_void_p()
        rc = lib().sgm_stream_create(C.byref(handle), C.byref(self.params), devs, len(devices), inflight)
        if rc != 0:
            raise SgmError(rc, lib().sgm_stream_last_error(None).decode())
        self._h = handle

    def _check(self, rc):
        if rc != 0:
            raise SgmError(rc, lib().sgm_stream_last_error(self._h).decode())

    def close(self):
        if self._h is not None:
            lib().sgm_stream_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def submit(self, img_l, img_r, sky_l=None, sky_r=None) -> int:
        l, r = _img(img_l), _img(img_r)
        sl = None if sky_l is None else _img(sky_l)
        sr = None if sky_r is None else _img(sky_r)
        p = lambda a: None if a is None else a.ctypes.data
        st = lambda a: 0 if a is None else a.strides[0]
        k = C.c_uint64()
        self._check(lib().sgm_stream_submit(self._h, p(l), st(l), p(r), st(r), p(sl), st(sl), p(sr), st(sr), C.byref(k)))
        return k.value  # the inputs were copied into pinned staging: the arrays may be reused at once

    def next_into(self, out: np.ndarray) -> int:
        k = C.c_uint64()
        self._check(lib().sgm_stream_next(self._h, out.ctypes.data, out.strides[0], None, C.byref(k)))
        return k.value

    def next(self):
        out = np.empty((self.img_h, self.img_w), np.float32)
        return self.next_into(out), out

    def next_view(self):
        """(frame index, zero-copy float32 view of the pinned result); the view keeps its slot until release() or the next
        next*() call."""
        k, ptr = C.c_uint64(), C.c_void_p()
        self._check(lib().sgm_stream_next(self._h, None, 0, C.byref(ptr), C.byref(k)))
        buf = (C.c_float * (self.img_h * self.img_w)).from_address(ptr.value)
        return k.value, np.frombuffer(buf, np.float32).reshape(self.img_h, self.img_w)

    def release(self) -> None:
        self._check(lib().sgm_stream_release(self._h))

    @property
    def pending(self) -> int:
        return lib().sgm_stream_pending(self._h)

    @property
    def launches(self) -> int:
        return lib().sgm_stream_launches(self._h)


SPECKLE_STATIC_MAX_PIXELS = 1 << 21


def speckle_filter_static(in_ptr, out_ptr, h, w, invalid_value, max_size=1000, max_dis=2.0, device=0, stream=None) -> None:
    """sgm_speckle_filter_static: the speckle stage on caller-owned DEVICE buffers (raw pointers), no allocation; queued on
    `stream` (None = default stream), not synchronised."""
    rc = lib().sgm_speckle_filter_static(device, in_ptr, out_ptr, h, w, invalid_value, max_size, max_dis, stream)
    if rc != 0:
        raise SgmError(rc, lib().sgm_last_error(None).decode())


def cost_volume(kind, left, right, max_disp, win_h, win_w, scale=1, device=0) -> np.ndarray:
    """Dense forms of SAD / SSD / CT (inc/cost.h:7-10) as Solver::build_dsi would fill them: (h, w, max_disp) float32."""
    fn = getattr(lib(), {"sad": "sgm_cost_volume_sad", "ssd": "sgm_cost_volume_ssd", "ct": "sgm_cost_volume_ct"}[kind])
    l, r = np.ascontiguousarray(_img(left)), np.ascontiguousarray(_img(right))
    h, w = l.shape
    out = np.empty((h, w, max_disp), np.float32)
    rc = fn(device, l.ctypes.data, r.ctypes.data, h, w, max_disp, win_h, win_w, scale, out.ctypes.data)
    if rc != 0:
        raise SgmError(rc, lib().sgm_last_error(None).decode())
    return out


def cost_volume_sad(left, right, max_disp, win_h, win_w, scale=1, device=0) -> np.ndarray:
    return cost_volume("sad", left, right, max_disp, win_h, win_w, scale, device)
