"""ctypes binding of libvo_b200.so (include/vo_b200.h).

There is no CPU fallback here: if the library is missing it is built (nvcc, sm_100a); if it
cannot be loaded, or no CUDA device is usable, an exception is raised.
"""
import ctypes
import os

import numpy as np

from . import _build

c_void_p = ctypes.c_void_p
c_int = ctypes.c_int
c_i64 = ctypes.c_int64
c_u64 = ctypes.c_uint64
c_double = ctypes.c_double


class VoError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libvo_b200 error %d: %s" % (code, msg))
        self.code = code


class KltParams(ctypes.Structure):
    _fields_ = [("win", c_int), ("max_level", c_int), ("max_iters", c_int), ("epsilon", c_double),
                ("min_eig_thr", c_double), ("fb_check", c_int), ("fb_thresh", c_double)]


class PnpParams(ctypes.Structure):
    _fields_ = [("max_iters", c_int), ("confidence", c_double), ("reproj_thresh", c_double),
                ("refine_iters", c_int), ("min_inliers", c_int)]


# name -> (restype, argtypes); every symbol include/vo_b200.h declares must be listed here
# (tests/test_abi.py cross-checks this table against the header).
SIGNATURES = {
    "vo_klt_params_default": (None, [ctypes.POINTER(KltParams)]),
    "vo_pnp_params_default": (None, [ctypes.POINTER(PnpParams)]),
    "vo_abi_version": (c_int, []),
    "vo_device_count": (c_int, []),
    "vo_create": (c_int, [ctypes.POINTER(c_void_p), c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "vo_destroy": (c_int, [c_void_p]),
    "vo_last_error": (ctypes.c_char_p, [c_void_p]),
    "vo_synchronize": (c_int, [c_void_p]),
    "vo_launch_count": (c_i64, [c_void_p]),
    "vo_alloc_pinned": (c_int, [ctypes.POINTER(c_void_p), c_u64]),
    "vo_free_pinned": (c_int, [c_void_p]),
    "vo_upload_frame": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_i64]),
    "vo_has_frame": (c_int, [c_void_p, c_i64]),
    "vo_drop_frames": (c_int, [c_void_p]),
    "vo_get_pyramid_level": (c_int, [c_void_p, c_i64, c_int, c_void_p, c_void_p,
                                     ctypes.POINTER(c_int), ctypes.POINTER(c_int)]),
    "vo_calc_optical_flow_pyr_lk": (c_int, [c_void_p, c_i64, c_i64, c_void_p, c_int, c_void_p, c_void_p,
                                            c_void_p, c_void_p, ctypes.POINTER(KltParams)]),
}

_lib = None


def lib_path():
    return _build.LIB


def load(build_if_missing=True):
    """Load (building first if needed) libvo_b200.so and set the prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    if build_if_missing and _build.needs_build():
        _build.build()
    if not os.path.exists(_build.LIB):
        raise ImportError("libvo_b200.so is missing (%s) and there is no CPU fallback" % _build.LIB)
    lib = ctypes.CDLL(_build.LIB)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export it
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def _ptr(a):
    return None if a is None else c_void_p(a.ctypes.data)


def klt_defaults():
    p = KltParams()
    load().vo_klt_params_default(ctypes.byref(p))
    return p


def pnp_defaults():
    p = PnpParams()
    load().vo_pnp_params_default(ctypes.byref(p))
    return p


class PinnedArray:
    """A numpy view over page-locked host memory owned by the library."""

    def __init__(self, shape, dtype=np.uint8):
        self.lib = load()
        nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = c_void_p()
        r = self.lib.vo_alloc_pinned(ctypes.byref(p), max(nbytes, 1))
        if r != 0:
            raise VoError(r, self.lib.vo_last_error(None).decode())
        self.ptr = p
        buf = (ctypes.c_uint8 * max(nbytes, 1)).from_address(p.value)
        self.array = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    def free(self):
        if self.ptr is not None:
            self.array = None
            self.lib.vo_free_pinned(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Context:
    """One vo_ctx: one CUDA stream on one GPU, owner of all device state of one VO stream."""

    def __init__(self, device=0, max_w=2048, max_h=1024, track_capacity=4096, hyp_capacity=4096,
                 stream=None):
        self.lib = load()
        n = self.lib.vo_device_count()
        if n <= 0:
            raise VoError(n, "no usable CUDA device (vo_device_count=%d); this package has no CPU path" % n)
        h = c_void_p()
        r = self.lib.vo_create(ctypes.byref(h), device, max_w, max_h, track_capacity, hyp_capacity,
                               c_void_p(stream) if stream else None)
        if r != 0:
            raise VoError(r, self.lib.vo_last_error(None).decode())
        self.h = h
        self.max_w, self.max_h = max_w, max_h

    def close(self):
        if getattr(self, "h", None):
            self.lib.vo_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, r):
        if r != 0:
            raise VoError(r, self.lib.vo_last_error(self.h).decode())

    def synchronize(self):
        self._check(self.lib.vo_synchronize(self.h))

    def launch_count(self):
        return int(self.lib.vo_launch_count(self.h))

    # ---- frames ------------------------------------------------------------------------------
    def upload_frame(self, img, key):
        if img.dtype != np.uint8 or img.ndim != 2 or img.strides[1] != 1:
            raise ValueError("frames must be 2-D uint8 arrays with contiguous rows")
        h, w = img.shape
        self._check(self.lib.vo_upload_frame(self.h, c_void_p(img.ctypes.data), w, h, img.strides[0], key))

    def has_frame(self, key):
        return bool(self.lib.vo_has_frame(self.h, key))

    def drop_frames(self):
        self._check(self.lib.vo_drop_frames(self.h))

    def pyramid_level(self, key, level, want_deriv=True):
        w, h = c_int(), c_int()
        self._check(self.lib.vo_get_pyramid_level(self.h, key, level, None, None, ctypes.byref(w), ctypes.byref(h)))
        img = np.empty((h.value, w.value), np.uint8)
        der = np.empty((h.value, w.value, 2), np.int16) if want_deriv else None
        self._check(self.lib.vo_get_pyramid_level(self.h, key, level, _ptr(img), _ptr(der), None, None))
        return img, der

    # ---- unit-level: cv2.calcOpticalFlowPyrLK ---------------------------------------------------
    def calc_optical_flow_pyr_lk(self, prev_key, next_key, pts, params=None, want_iters=False):
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 2)
        n = len(pts)
        out = np.empty((n, 2), np.float32)
        status = np.empty(n, np.uint8)
        err = np.empty(n, np.float32)
        iters = np.empty(n, np.int32) if want_iters else None
        self._check(self.lib.vo_calc_optical_flow_pyr_lk(
            self.h, prev_key, next_key, _ptr(pts), n, _ptr(out), _ptr(status), _ptr(err), _ptr(iters),
            ctypes.byref(params) if params is not None else None))
        if want_iters:
            return out, status, err, iters
        return out, status, err
