"""CPU ORACLE - test infrastructure only.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.  The product path (the package next to it and libvo_b200.so) never does.

What is here
  lk.c, pnp.c, dlt.c   plain-C restatements of the OpenCV routines the reference calls on its hot
                       path (cv2.calcOpticalFlowPyrLK, cv2.solvePnPRansac, cv2.triangulatePoints);
                       the arithmetic lives in OpenCV (third party, not vendored by the reference;
                       pinned opencv-python==4.8.1.78, installed here 4.13.0.92).
  ref_pipeline.py      a port of the reference's three per-frame methods over cv2 (CPU baseline).
Parity pin: tests/golden/*.npz hold outputs of cv2 4.13.0 and of the reference's own classes
(imported from /root/reference in the build container by tests/golden/make_golden.py).
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_build", "liboracle.so")
_lib = None


def build(force=False):
    srcs = [os.path.join(HERE, f) for f in os.listdir(HERE) if f.endswith(".c")]
    if (not force and os.path.exists(LIB)
            and os.path.getmtime(LIB) >= max(os.path.getmtime(s) for s in srcs)):
        return LIB
    r = subprocess.run(["make", "-C", HERE, "-B", "_build/liboracle.so"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + r.stdout + r.stderr)
    return LIB


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(LIB)
    return _lib


def _p(a):
    return None if a is None else ctypes.c_void_p(a.ctypes.data)


def pyr_down(img):
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    out = np.empty(((h + 1) // 2, (w + 1) // 2), np.uint8)
    lib().vo_oracle_pyr_down(_p(img), w, h, _p(out))
    return out


def calc_optical_flow_pyr_lk(prev, nxt, pts, win=21, max_level=3, max_iters=30, eps=0.01,
                             min_eig_thr=1e-4, want_iters=False):
    """cv2.calcOpticalFlowPyrLK(prev, next, pts, None) restated (oracle/lk.c)."""
    prev = np.ascontiguousarray(prev, np.uint8)
    nxt = np.ascontiguousarray(nxt, np.uint8)
    pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 2)
    h, w = prev.shape
    n = len(pts)
    out = np.zeros((n, 2), np.float32)
    st = np.zeros(n, np.uint8)
    err = np.zeros(n, np.float32)
    it = np.zeros(n, np.int32)
    f = lib().vo_oracle_calc_optical_flow_pyr_lk
    f.restype = ctypes.c_int
    f.argtypes = ([ctypes.c_void_p] * 2 + [ctypes.c_int] * 2 + [ctypes.c_void_p, ctypes.c_int]
                  + [ctypes.c_void_p] * 3 + [ctypes.c_int] * 3 + [ctypes.c_double] * 2 + [ctypes.c_void_p])
    f(_p(prev), _p(nxt), w, h, _p(pts), n, _p(out), _p(st), _p(err), win, max_level, max_iters,
      eps, min_eig_thr, _p(it))
    if want_iters:
        return out, st, err, it
    return out, st, err


def scharr(img):
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    out = np.empty((h, w, 2), np.int16)
    lib().vo_oracle_scharr(_p(img), w, h, _p(out))
    return out
