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


# ---- PnP-RANSAC (oracle/pnp.c) -------------------------------------------------------------------
def _f(name, restype, argtypes):
    fn = getattr(lib(), name)
    fn.restype = restype
    fn.argtypes = argtypes
    return fn


_VP, _I, _D = ctypes.c_void_p, ctypes.c_int, ctypes.c_double


def ransac_sets(n, m, H):
    """The H x m minimal-set table cv2's RANSAC draws for n points (cv::RNG(2^64-1))."""
    out = np.empty((H, m), np.int32)
    _f("vo_oracle_ransac_sets", None, [_I, _I, _I, _VP])(n, m, H, _p(out))
    return out


def update_num_iters(p, ep, m, max_iters):
    return _f("vo_oracle_update_num_iters", _I, [_D, _D, _I, _I])(p, ep, m, max_iters)


def p3p_all(X3, p3, K):
    X3 = np.ascontiguousarray(X3, np.float64).reshape(3, 3)
    p3 = np.ascontiguousarray(p3, np.float64).reshape(3, 2)
    K = np.ascontiguousarray(K, np.float64)
    Rs = np.zeros((4, 3, 3)); ts = np.zeros((4, 3))
    ns = _f("vo_oracle_p3p_all", _I, [_VP] * 5)(_p(X3), _p(p3), _p(K), _p(Rs), _p(ts))
    return Rs[:ns], ts[:ns]


def p3p4(X4, p4, K):
    """cv2.solvePnP(flags=SOLVEPNP_P3P) on 4 correspondences -> (ok, R, t)."""
    X4 = np.ascontiguousarray(X4, np.float64).reshape(4, 3)
    p4 = np.ascontiguousarray(p4, np.float64).reshape(4, 2)
    K = np.ascontiguousarray(K, np.float64)
    R = np.zeros((3, 3)); t = np.zeros(3)
    ok = _f("vo_oracle_p3p4", _I, [_VP] * 5)(_p(X4), _p(p4), _p(K), _p(R), _p(t))
    return bool(ok), R, t


def score(X, P, K, R, t, thr=2.0):
    X = np.ascontiguousarray(np.asarray(X, np.float32), np.float64)
    P = np.ascontiguousarray(P, np.float32)
    n = len(X)
    mask = np.zeros(n, np.uint8)
    _f("vo_oracle_score", _I, [_VP, _VP, _I, _VP, _VP, _VP, _D, _VP])(
        _p(X), _p(P), n, _p(np.ascontiguousarray(K, np.float64)), _p(np.ascontiguousarray(R, np.float64)),
        _p(np.ascontiguousarray(t, np.float64)), thr, _p(mask))
    return mask


def refine_lm(X, P, mask, K, R, t, max_iters=50):
    X = np.ascontiguousarray(np.asarray(X, np.float32), np.float64)
    P = np.ascontiguousarray(P, np.float32)
    R = np.array(R, np.float64).reshape(3, 3).copy(); t = np.array(t, np.float64).reshape(3).copy()
    m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
    it = _f("vo_oracle_refine_lm", _I, [_VP, _VP, _VP, _I, _VP, _VP, _VP, _I])(
        _p(X), _p(P), _p(m), len(X), _p(np.ascontiguousarray(K, np.float64)), _p(R), _p(t), max_iters)
    return R, t, it


def rotation_to_rvec(R):
    R = np.ascontiguousarray(R, np.float64)
    r = np.zeros(3)
    _f("vo_oracle_rotation_to_rvec", None, [_VP, _VP])(_p(R), _p(r))
    return r


def solve_pnp_ransac(X, P, K, max_iters=100, confidence=0.99999, thr=2.0, refine_iters=50, sets=None):
    """cv2.solvePnPRansac(X, P, K, 0, flags=SOLVEPNP_P3P-hypotheses, ...) + ITERATIVE refit restated.
    Returns dict(ok, R, t, Rh, th, mask, iters, best_iter, n_inliers, lm_iters)."""
    X = np.ascontiguousarray(X, np.float64).reshape(-1, 3)
    P = np.ascontiguousarray(P, np.float32).reshape(-1, 2)
    K = np.ascontiguousarray(K, np.float64)
    n = len(X)
    R = np.zeros((3, 3)); t = np.zeros(3); Rh = np.zeros((3, 3)); th = np.zeros(3)
    mask = np.zeros(n, np.uint8); info = np.zeros(4, np.int32)
    s = None if sets is None else np.ascontiguousarray(sets, np.int32)
    if s is not None:
        assert s.shape[0] >= max_iters and s.shape[1] == 4
    ok = _f("vo_oracle_solve_pnp_ransac", _I, [_VP, _VP, _I, _VP, _I, _D, _D, _I] + [_VP] * 7)(
        _p(X), _p(P), n, _p(K), max_iters, confidence, thr, refine_iters, _p(s), _p(R), _p(t), _p(Rh),
        _p(th), _p(mask), _p(info))
    return dict(ok=bool(ok), R=R, t=t, Rh=Rh, th=th, mask=mask, iters=int(info[0]), best_iter=int(info[1]),
                n_inliers=int(info[2]), lm_iters=int(info[3]))


# ---- DLT triangulation + bearing test (oracle/dlt.c) ---------------------------------------------
def triangulate_points(M1, M2, p1, p2):
    """cv2.triangulatePoints(M1, M2, p1.T, p2.T).T restated -> n x 4 homogeneous (unit norm)."""
    M1 = np.ascontiguousarray(M1, np.float64); M2 = np.ascontiguousarray(M2, np.float64)
    p1 = np.ascontiguousarray(p1, np.float64).reshape(-1, 2); p2 = np.ascontiguousarray(p2, np.float64).reshape(-1, 2)
    out = np.zeros((len(p1), 4))
    _f("vo_oracle_triangulate_points", None, [_VP] * 4 + [_I, _VP])(_p(M1), _p(M2), _p(p1), _p(p2), len(p1), _p(out))
    return out


def bearing_angle_deg(c, f):
    c = np.ascontiguousarray(c, np.float64); f = np.ascontiguousarray(f, np.float64)
    return _f("vo_oracle_bearing_angle_deg", _D, [_VP, _VP])(_p(c), _p(f))


def bearing_triangulate(C, F, T, Tnew, K, alpha_deg):
    """Promotion step of LandmarkTriangulator.triangulate_landmark -> (promote mask, X n x 3, alpha)."""
    C = np.ascontiguousarray(C, np.float64).reshape(-1, 2); F = np.ascontiguousarray(F, np.float64).reshape(-1, 2)
    n = len(C)
    T = np.ascontiguousarray(T, np.float64).reshape(n, 16)
    promote = np.zeros(n, np.uint8); X = np.zeros((n, 3)); alpha = np.zeros(n)
    _f("vo_oracle_bearing_triangulate", _I, [_VP, _VP, _VP, _I, _VP, _VP, _D, _VP, _VP, _VP])(
        _p(C), _p(F), _p(T), n, _p(np.ascontiguousarray(Tnew, np.float64)), _p(np.ascontiguousarray(K, np.float64)),
        alpha_deg, _p(promote), _p(X), _p(alpha))
    return promote, X, alpha


def corner_min_eigen_val(img, block=4, vec_cols=-1):
    """cv2.cornerMinEigenVal(img, block, ksize=3) restated (oracle/gftt.c)."""
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    eig = np.empty((h, w), np.float32)
    f = lib().vo_oracle_corner_min_eigen_val
    f.restype = None
    f.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    f(_p(img), w, h, block, vec_cols, _p(eig))
    return eig


def sobel_scaled(img, block=4, vec_cols=-1):
    """cv2.Sobel(img, CV_32F, 1|0, 0|1, ksize=3, scale=1/(4*block*255)) restated (oracle/gftt.c)."""
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    dx, dy = np.empty((h, w), np.float32), np.empty((h, w), np.float32)
    f = lib().vo_oracle_sobel_scaled
    f.restype = None
    f.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    f(_p(img), w, h, block, vec_cols, _p(dx), _p(dy))
    return dx, dy


def good_features_to_track(img, max_corners=500, quality=0.03, min_distance=10.0, mask=None, block=4,
                           vec_cols=-1, want_eig=False):
    """cv2.goodFeaturesToTrack(img, max_corners, quality, min_distance, mask=mask, blockSize=block,
    gradientSize=3, useHarrisDetector=False) restated (oracle/gftt.c; reference vo_pipeline.py:381).
    Returns (n, 2) float32 corners (x, y) in cv2's order."""
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    if mask is not None:
        mask = np.ascontiguousarray(mask, np.uint8)
    cap = max_corners if max_corners > 0 else w * h
    corners = np.zeros((cap, 2), np.float32)
    eig = np.empty((h, w), np.float32) if want_eig else None
    nc = ctypes.c_int()
    f = lib().vo_oracle_good_features_to_track
    f.restype = ctypes.c_int
    f.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_double,
                  ctypes.c_double, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                  ctypes.POINTER(ctypes.c_int)]
    n = f(_p(img), w, h, _p(mask), max_corners, quality, min_distance, block, vec_cols, _p(corners), cap, _p(eig),
          ctypes.byref(nc))
    if want_eig:
        return corners[:n].copy(), eig, nc.value
    return corners[:n].copy()


# ---- bootstrap: five-point essential-matrix RANSAC + recoverPose (oracle/fivept.c) -----------------------------------
def five_point(q1, q2):
    """Essential matrices (m x 3 x 3, m <= 10, unit Frobenius norm) through 5 normalised correspondences."""
    q1 = np.ascontiguousarray(q1, np.float64).reshape(5, 2)
    q2 = np.ascontiguousarray(q2, np.float64).reshape(5, 2)
    Es = np.zeros((10, 9))
    m = _f("vo_oracle_five_point", _I, [_VP, _VP, _VP])(_p(q1), _p(q2), _p(Es))
    return Es[:m].reshape(m, 3, 3)


def find_essential_mat(p1, p2, K, prob=0.999, thr=1.0, max_iters=1000):
    """cv2.findEssentialMat(p1, p2, K, cv2.RANSAC, prob, thr) restated -> dict(ok, E, mask, iters, best_iter, best_model, n_inliers)."""
    p1 = np.ascontiguousarray(p1, np.float64).reshape(-1, 2)
    p2 = np.ascontiguousarray(p2, np.float64).reshape(-1, 2)
    K = np.ascontiguousarray(K, np.float64)
    n = len(p1)
    E = np.zeros((3, 3)); mask = np.zeros(max(n, 1), np.uint8); info = np.zeros(4, np.int32)
    ok = _f("vo_oracle_find_essential_mat", _I, [_VP, _VP, _I, _VP, _D, _D, _I, _VP, _VP, _VP])(
        _p(p1), _p(p2), n, _p(K), prob, thr, max_iters, _p(E), _p(mask), _p(info))
    return dict(ok=bool(ok), E=E, mask=mask[:n], iters=int(info[0]), best_iter=int(info[1]), best_model=int(info[2]), n_inliers=int(info[3]))


def recover_pose(E, p1, p2, K, mask=None):
    """cv2.recoverPose(E, p1, p2, K, mask=mask) restated -> (count, R, t, mask ANDed with the winner's cheirality test, which)."""
    E = np.ascontiguousarray(E, np.float64)
    p1 = np.ascontiguousarray(p1, np.float64).reshape(-1, 2)
    p2 = np.ascontiguousarray(p2, np.float64).reshape(-1, 2)
    K = np.ascontiguousarray(K, np.float64)
    n = len(p1)
    m = np.ones(max(n, 1), np.uint8) if mask is None else np.ascontiguousarray(np.asarray(mask).ravel() != 0, np.uint8).copy()
    R = np.zeros((3, 3)); t = np.zeros(3); which = np.zeros(1, np.int32)
    g = _f("vo_oracle_recover_pose", _I, [_VP, _VP, _VP, _I, _VP, _VP, _VP, _VP, _VP])(
        _p(E), _p(p1), _p(p2), n, _p(K), _p(m), _p(R), _p(t), _p(which))
    return g, R, t, m[:n], int(which[0])
