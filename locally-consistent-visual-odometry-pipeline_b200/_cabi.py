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


class GfttParams(ctypes.Structure):
    _fields_ = [("max_corners", c_int), ("quality_level", c_double), ("min_distance", c_double),
                ("block_size", c_int), ("vec_cols", c_int), ("use_cluster", c_int)]


class BootParams(ctypes.Structure):
    _fields_ = [("prob", c_double), ("threshold", c_double), ("max_iters", c_int)]


class Counters(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int32) for n in (
        "n_landmarks", "n_candidates", "n_tracked", "n_lost", "n_inliers", "pnp_ok", "ransac_iters",
        "best_hypothesis", "lm_iters", "n_promoted", "n_candidates_lost", "cf_alias", "klt_iters",
        "klt_iters_candidates")] + [
        ("reproj_cost", c_double), ("n_appended", ctypes.c_int32), ("reserved_", ctypes.c_int32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


_PK = ctypes.POINTER(KltParams)
_PP = ctypes.POINTER(PnpParams)
_PI = ctypes.POINTER(c_int)

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
    "vo_prefetch_frame": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_i64]),
    "vo_prefetch_frame_deferred": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_i64]),
    "vo_has_frame": (c_int, [c_void_p, c_i64]),
    "vo_drop_frames": (c_int, [c_void_p]),
    "vo_get_pyramid_level": (c_int, [c_void_p, c_i64, c_int, c_void_p, c_void_p,
                                     ctypes.POINTER(c_int), ctypes.POINTER(c_int)]),
    "vo_calc_optical_flow_pyr_lk": (c_int, [c_void_p, c_i64, c_i64, c_void_p, c_int, c_void_p, c_void_p,
                                            c_void_p, c_void_p, ctypes.POINTER(KltParams)]),
    "vo_set_state": (c_int, [c_void_p, c_void_p, c_void_p, c_int]),
    "vo_get_state": (c_int, [c_void_p, c_void_p, c_void_p, c_int, _PI]),
    "vo_snapshot_state": (c_int, [c_void_p]),
    "vo_get_snapshot": (c_int, [c_void_p, c_void_p, c_void_p, c_int, _PI]),
    "vo_set_candidates": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int]),
    "vo_get_candidates": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, _PI]),
    "vo_get_lost": (c_int, [c_void_p, c_void_p, c_int, _PI]),
    "vo_get_counters": (c_int, [c_void_p, ctypes.POINTER(Counters)]),
    "vo_peek_counters": (c_int, [c_void_p, ctypes.POINTER(Counters)]),
    "vo_klt_track_landmarks": (c_int, [c_void_p, c_i64, c_i64, _PK, _PI]),
    "vo_pnp_ransac": (c_int, [c_void_p, c_void_p, _PP, c_void_p, _PI, _PI]),
    "vo_klt_track_landmarks_add": (c_int, [c_void_p, c_i64, c_i64, _PK, c_void_p, c_int, c_void_p, _PI]),
    "vo_klt_track_landmarks_ex": (c_int, [c_void_p, c_i64, c_i64, _PK, c_void_p, c_int, c_void_p, c_double, _PI]),
    "vo_track_rollback": (c_int, [c_void_p]),
    "vo_debug_pnp_stamps": (c_int, [c_void_p, c_void_p]),
    "vo_map_reset": (c_int, [c_void_p]),
    "vo_map_accumulate": (c_int, [c_void_p, c_void_p, c_double, c_double, c_int, _PI]),
    "vo_map_get": (c_int, [c_void_p, c_void_p, c_int, _PI, c_void_p, c_int, _PI, c_void_p]),
    "vo_map_render": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "vo_pnp_ransac_promote": (c_int, [c_void_p, c_void_p, _PP, c_double, c_void_p, _PI, _PI, _PI]),
    "vo_promote_commit": (c_int, [c_void_p]),
    "vo_promote_rollback": (c_int, [c_void_p]),
    "vo_promote_pending": (c_int, [c_void_p]),
    "vo_ahead_adopt": (c_int, [c_void_p]),
    "vo_candidates_tracked": (c_int, [c_void_p, c_i64, c_i64]),
    "vo_get_prev_state": (c_int, [c_void_p, c_void_p, c_void_p, c_int, _PI]),
    "vo_get_prev2_state": (c_int, [c_void_p, c_void_p, c_void_p, c_int, _PI]),
    "vo_track_triangulate_promote": (c_int, [c_void_p, c_i64, c_i64, c_void_p, c_void_p, c_double, _PK, _PI]),
    "vo_add_candidates": (c_int, [c_void_p, c_void_p, c_int, c_void_p]),
    "vo_boot_params_default": (None, [ctypes.POINTER(BootParams)]),
    "vo_bootstrap_pose": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, ctypes.POINTER(BootParams), c_void_p, c_void_p,
                                  c_void_p, c_void_p]),
    "vo_append_landmarks": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int]),
    "vo_queue_landmarks": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int]),
    "vo_track_candidates": (c_int, [c_void_p, c_i64, c_i64, _PK]),
    "vo_get_candidate_tracks": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, _PI]),
    "vo_process_frame": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_i64, c_i64, c_void_p, _PK, _PP,
                                 c_double, c_void_p, c_int, c_void_p, ctypes.POINTER(Counters)]),
    "vo_process_frame_pipelined": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_i64, c_i64, c_void_p, _PK, _PP,
                                 c_double, c_void_p, c_int, c_void_p, ctypes.POINTER(Counters)]),
    "vo_set_profiling": (c_int, [c_void_p, c_int]),
    "vo_get_stage_times": (c_int, [c_void_p, c_void_p, _PI]),
    "vo_solve_pnp_ransac": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, _PP, c_void_p, c_void_p,
                                    c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, _PI]),
    "vo_solve_pnp_ransac_noalloc": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, _PP, c_void_p, c_void_p,
                                            c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, _PI]),
    "vo_gftt_params_default": (None, [ctypes.POINTER(GfttParams)]),
    "vo_good_features_to_track": (c_int, [c_void_p, c_i64, c_void_p, c_void_p, c_int, ctypes.POINTER(GfttParams),
                                          c_void_p, c_int, _PI, c_void_p, c_void_p]),
    "vo_detect_add_candidates_shi": (c_int, [c_void_p, c_i64, ctypes.POINTER(GfttParams), c_void_p]),
    "vo_triangulate_points": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p]),
}

_lib = None


def lib_path():
    return _build.LIB


NOALLOC_MAX_POINTS = 4096   # include/vo_b200.h: VO_NOALLOC_MAX_POINTS / VO_NOALLOC_MAX_ITERS
NOALLOC_MAX_ITERS = 1024


def load(build_if_missing=True):
    """Load (building first if needed) libvo_b200.so and set the prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    if build_if_missing and _build.needs_build():
        # N ranks of one job may get here together on a fresh checkout: one of them builds, the others wait
        import fcntl
        os.makedirs(_build.OBJ, exist_ok=True)
        with open(os.path.join(_build.OBJ, ".build.lock"), "w") as lock:
            fcntl.flock(lock, fcntl.LOCK_EX)
            try:
                if _build.needs_build():
                    _build.build()
            finally:
                fcntl.flock(lock, fcntl.LOCK_UN)
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


def gftt_defaults():
    p = GfttParams()
    load().vo_gftt_params_default(ctypes.byref(p))
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

    def prefetch_frame(self, img, key, deferred=False):
        """vo_prefetch_frame: upload + pyramid on the ingest stream (overlaps the stages enqueued after it).
        deferred=True: enqueued by the next pose call, under the GPU's PnP (vo_prefetch_frame_deferred);
        the caller keeps `img` alive until then."""
        if img.dtype != np.uint8 or img.ndim != 2 or img.strides[1] != 1:
            raise ValueError("frames must be 2-D uint8 arrays with contiguous rows")
        h, w = img.shape
        fn = self.lib.vo_prefetch_frame_deferred if deferred else self.lib.vo_prefetch_frame
        self._check(fn(self.h, c_void_p(img.ctypes.data), w, h, img.strides[0], key))

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

    # ---- Markov state ------------------------------------------------------------------------------
    def set_state(self, P, X):
        P = np.ascontiguousarray(P, np.float32).reshape(-1, 2)
        X = np.ascontiguousarray(X, np.float64).reshape(-1, 3)
        if len(P) != len(X):
            raise ValueError("P and X differ in length")
        self._check(self.lib.vo_set_state(self.h, _ptr(P), _ptr(X), len(P)))

    def n_state(self):
        n = c_int()
        self._check(self.lib.vo_get_state(self.h, None, None, 0, ctypes.byref(n)))
        return n.value

    def get_state(self):
        n = self.n_state()
        P = np.empty((n, 2), np.float32)
        X = np.empty((n, 3), np.float64)
        m = c_int()
        self._check(self.lib.vo_get_state(self.h, _ptr(P), _ptr(X), n, ctypes.byref(m)))
        return P, X

    def snapshot_state(self):
        self._check(self.lib.vo_snapshot_state(self.h))

    def get_snapshot(self):
        n = c_int()
        self._check(self.lib.vo_get_snapshot(self.h, None, None, 0, ctypes.byref(n)))
        P = np.empty((n.value, 2), np.float32)
        X = np.empty((n.value, 3), np.float64)
        self._check(self.lib.vo_get_snapshot(self.h, _ptr(P), _ptr(X), n.value, ctypes.byref(n)))
        return P, X

    def set_candidates(self, C, F, T, cf_alias=False):
        C = np.ascontiguousarray(C, np.float32).reshape(-1, 2)
        F = np.ascontiguousarray(F, np.float64).reshape(-1, 2)
        T = np.ascontiguousarray(T, np.float64).reshape(-1, 16)
        if not (len(C) == len(F) == len(T)):
            raise ValueError("C, F, T differ in length")
        self._check(self.lib.vo_set_candidates(self.h, _ptr(C), _ptr(F), _ptr(T), len(C), int(cf_alias)))

    def get_candidates(self):
        n = c_int()
        self._check(self.lib.vo_get_candidates(self.h, None, None, None, 0, ctypes.byref(n)))
        C = np.empty((n.value, 2), np.float32)
        F = np.empty((n.value, 2), np.float64)
        T = np.empty((n.value, 4, 4), np.float64)
        self._check(self.lib.vo_get_candidates(self.h, _ptr(C), _ptr(F), _ptr(T), n.value, ctypes.byref(n)))
        return C, F, T

    def get_lost(self):
        n = c_int()
        self._check(self.lib.vo_get_lost(self.h, None, 0, ctypes.byref(n)))
        lost = np.empty((n.value, 2), np.float32)
        self._check(self.lib.vo_get_lost(self.h, _ptr(lost), n.value, ctypes.byref(n)))
        return lost

    def counters(self, sync=True):
        cnt = Counters()
        fn = self.lib.vo_get_counters if sync else self.lib.vo_peek_counters
        self._check(fn(self.h, ctypes.byref(cnt)))
        return cnt.as_dict()

    # ---- per-frame stages ------------------------------------------------------------------------
    def klt_track_landmarks(self, prev_key, next_key, params=None, sync=False, add=None, early_alpha=None):
        """a1.  add=(pts n x 2 float64, T 4x4): the previous frame's new candidates, appended on the side
        stream in front of the candidates' KLT (vo_klt_track_landmarks_add).  early_alpha (degrees): the
        pose-independent half of the coming promotion step runs behind that KLT (vo_klt_track_landmarks_ex)."""
        n = c_int()
        pk = ctypes.byref(params) if params is not None else None
        pn = ctypes.byref(n) if sync else None
        have_add = add is not None and len(add[0]) > 0
        if have_add or early_alpha is not None:
            pts = T = None
            if have_add:
                pts = np.ascontiguousarray(add[0], np.float64).reshape(-1, 2)
                T = np.ascontiguousarray(add[1], np.float64)
            self._check(self.lib.vo_klt_track_landmarks_ex(self.h, prev_key, next_key, pk, _ptr(pts), 0 if pts is None else len(pts),
                                                           _ptr(T), -1.0 if early_alpha is None else float(early_alpha), pn))
        else:
            self._check(self.lib.vo_klt_track_landmarks(self.h, prev_key, next_key, pk, pn))
        return n.value if sync else None

    def pnp_ransac(self, K, params=None):
        K = np.ascontiguousarray(K, np.float64)
        T = np.empty((4, 4), np.float64)
        ninl, ok = c_int(), c_int()
        self._check(self.lib.vo_pnp_ransac(self.h, _ptr(K), ctypes.byref(params) if params is not None else None,
                                           _ptr(T), ctypes.byref(ninl), ctypes.byref(ok)))
        return bool(ok.value), T, ninl.value

    def pnp_ransac_promote(self, K, params, alpha_deg, early=False):
        """estimatePose + speculative promotion with the device pose, one sync (vo_pnp_ransac_promote).
        early=True: wait for the pose only; the promotion step is still running when this returns
        (n_promoted is then reported as None)."""
        K = np.ascontiguousarray(K, np.float64)
        T = np.empty((4, 4), np.float64)
        ninl, ok, nprom = c_int(), c_int(), c_int()
        self._check(self.lib.vo_pnp_ransac_promote(self.h, _ptr(K), ctypes.byref(params) if params is not None else None,
                                                   float(alpha_deg), _ptr(T), ctypes.byref(ninl), ctypes.byref(ok),
                                                   None if early else ctypes.byref(nprom)))
        return bool(ok.value), T, ninl.value, (None if early else nprom.value)

    def pnp_stamps(self):
        """ns between the stamps of the last RANSAC run: P3P, scoring, bookkeeping, mask, LM, pruning."""
        t = np.zeros(8, np.uint64)
        self._check(self.lib.vo_debug_pnp_stamps(self.h, _ptr(t)))
        return np.diff(t[:7].astype(np.int64))

    def track_rollback(self):
        self._check(self.lib.vo_track_rollback(self.h))

    def promote_commit(self):
        self._check(self.lib.vo_promote_commit(self.h))

    def promote_rollback(self):
        self._check(self.lib.vo_promote_rollback(self.h))

    def candidates_tracked(self, prev_key, next_key):
        return bool(self.lib.vo_candidates_tracked(self.h, prev_key, next_key))

    def ahead_adopt(self):
        self._check(self.lib.vo_ahead_adopt(self.h))

    def promote_pending(self):
        return bool(self.lib.vo_promote_pending(self.h))

    def get_prev2_state(self):
        """The landmarks as they were two tracking steps back (see vo_get_prev2_state)."""
        n = c_int()
        self._check(self.lib.vo_get_prev2_state(self.h, None, None, 0, ctypes.byref(n)))
        P = np.empty((n.value, 2), np.float32)
        X = np.empty((n.value, 3), np.float64)
        if n.value:
            self._check(self.lib.vo_get_prev2_state(self.h, _ptr(P), _ptr(X), n.value, ctypes.byref(n)))
        return P, X

    def get_prev_state(self):
        n = c_int()
        self._check(self.lib.vo_get_prev_state(self.h, None, None, 0, ctypes.byref(n)))
        P = np.empty((n.value, 2), np.float32)
        X = np.empty((n.value, 3), np.float64)
        self._check(self.lib.vo_get_prev_state(self.h, _ptr(P), _ptr(X), n.value, ctypes.byref(n)))
        return P, X

    def track_triangulate_promote(self, prev_key, next_key, K, T44, alpha_deg, params=None, sync=False):
        K = np.ascontiguousarray(K, np.float64)
        T44 = np.ascontiguousarray(T44, np.float64)
        n = c_int()
        self._check(self.lib.vo_track_triangulate_promote(
            self.h, prev_key, next_key, _ptr(K), _ptr(T44), float(alpha_deg),
            ctypes.byref(params) if params is not None else None, ctypes.byref(n) if sync else None))
        return n.value if sync else None

    def track_candidates(self, prev_key, next_key, params=None):
        self._check(self.lib.vo_track_candidates(self.h, prev_key, next_key,
                                                 ctypes.byref(params) if params is not None else None))

    def candidate_tracks(self):
        n = c_int()
        self._check(self.lib.vo_get_candidate_tracks(self.h, None, None, None, 0, ctypes.byref(n)))
        a = np.empty((n.value, 2), np.float32)
        b = np.empty((n.value, 2), np.float32)
        st = np.empty(n.value, np.uint8)
        self._check(self.lib.vo_get_candidate_tracks(self.h, _ptr(a), _ptr(b), _ptr(st), n.value, ctypes.byref(n)))
        return a, b, st

    def add_candidates(self, pts, T44):
        pts = np.ascontiguousarray(pts, np.float64).reshape(-1, 2)
        T44 = np.ascontiguousarray(T44, np.float64)
        self._check(self.lib.vo_add_candidates(self.h, _ptr(pts), len(pts), _ptr(T44)))

    def append_landmarks(self, P, X, target_total=0, queue=False):
        """Append landmarks behind P/X (reference vo_pipeline.py:318-320); target_total > 0: top-up to that
        many rows, decided on the device.  queue=True: with the next process_frame call.  Host arrays are
        copied before the call returns (pageable) - keep pinned / device rows alive yourself."""
        P = np.ascontiguousarray(P, np.float32).reshape(-1, 2)
        X = np.ascontiguousarray(X, np.float64).reshape(-1, 3)
        if len(P) != len(X):
            raise ValueError("P and X must have the same number of rows")
        if queue:
            self._queued_rows = getattr(self, "_queued_rows", [])[-6:] + [(P, X)]   # keep the arrays alive
        fn = self.lib.vo_queue_landmarks if queue else self.lib.vo_append_landmarks
        self._check(fn(self.h, _ptr(P), _ptr(X), len(P), int(target_total)))

    def process_frame_tracked(self, prev_key, next_key, K, alpha_deg, feed, klt, pnp, T_out, cnt):
        """vo_process_frame_pipelined on a tracking step that is already on the stream (the class surface's hot path):
        no conversions - K float64 C-contiguous 3x3, feed float64 C-contiguous n x 2 or None, T_out 4x4 float64, cnt a
        Counters instance, both filled in place."""
        r = self.lib.vo_process_frame_pipelined(
            self.h, None, 0, 0, 0, prev_key, next_key, K.ctypes.data, ctypes.byref(klt), ctypes.byref(pnp), alpha_deg,
            feed.ctypes.data if feed is not None else None, 0 if feed is None else len(feed), T_out.ctypes.data, ctypes.byref(cnt))
        if r != 0:
            self._check(r)

    def queue_landmarks_raw(self, P, X, target_total):
        """vo_queue_landmarks without conversions (P float32 n x 2, X float64 n x 3, both C-contiguous; the caller keeps
        them alive until the call after the next process_frame returns)."""
        r = self.lib.vo_queue_landmarks(self.h, P.ctypes.data, X.ctypes.data, len(P), target_total)
        if r != 0:
            self._check(r)

    def bootstrap_pose(self, kps1, kps2, K, prob=0.999, threshold=1.0, max_iters=1000):
        """cv2.findEssentialMat(kps1, kps2, K, cv2.RANSAC, prob, threshold) + cv2.recoverPose(E, kps1, kps2, K, mask=mask)
        on the device (reference vo_pipeline.py:139-162) -> dict(ok, T34, mask, E, iters, best_iter, best_model, n_inliers, n_good)."""
        kps1 = np.ascontiguousarray(kps1, np.float64).reshape(-1, 2)
        kps2 = np.ascontiguousarray(kps2, np.float64).reshape(-1, 2)
        K = np.ascontiguousarray(K, np.float64)
        n = len(kps1)
        prm = BootParams(float(prob), float(threshold), int(max_iters))
        T = np.zeros((3, 4)); E = np.zeros((3, 3)); mask = np.zeros(max(n, 1), np.uint8); info = np.zeros(6, np.int32)
        self._check(self.lib.vo_bootstrap_pose(self.h, _ptr(kps1), _ptr(kps2), n, _ptr(K), ctypes.byref(prm), _ptr(T), _ptr(mask),
                                               _ptr(E), _ptr(info)))
        return dict(ok=bool(info[0]), T34=T, mask=mask[:n], E=E, iters=int(info[1]), best_iter=int(info[2]), best_model=int(info[3]),
                    n_inliers=int(info[4]), n_good=int(info[5]))

    def process_frame(self, img, prev_key, next_key, K, alpha_deg, new_candidates=None, klt=None, pnp=None,
                      want_counters=True, pipelined=False):
        K = np.ascontiguousarray(K, np.float64)
        T = np.empty((4, 4), np.float64)
        cnt = Counters() if want_counters else None
        if new_candidates is not None:
            new_candidates = np.ascontiguousarray(new_candidates, np.float64).reshape(-1, 2)
        nn = 0 if new_candidates is None else len(new_candidates)
        if img is not None:
            h, w = img.shape
            ip, pitch = c_void_p(img.ctypes.data), img.strides[0]
        else:
            h = w = pitch = 0
            ip = None
        fn = self.lib.vo_process_frame_pipelined if pipelined else self.lib.vo_process_frame
        self._check(fn(
            self.h, ip, w, h, pitch, prev_key, next_key, _ptr(K), ctypes.byref(klt) if klt is not None else None,
            ctypes.byref(pnp) if pnp is not None else None, float(alpha_deg),
            _ptr(new_candidates) if nn else None, nn, _ptr(T), ctypes.byref(cnt) if cnt is not None else None))
        return T, (cnt.as_dict() if cnt is not None else None)

    # ---- trajectory / map bookkeeping (main.py:239-255, :262-271) ---------------------
    def map_reset(self):
        self._check(self.lib.vo_map_reset(self.h))

    def map_accumulate(self, T=None, r_min=3.0, r_max=50.0, use_prev_state=True, sync=False):
        """Append the camera position of pose T (None: the device pose of the last PnP) to the path and the
        landmarks with r_min < |X - pos| < r_max to the map; returns the rows added when sync=True."""
        n = c_int()
        Tc = None if T is None else np.ascontiguousarray(T, np.float64)
        self._check(self.lib.vo_map_accumulate(self.h, None if Tc is None else _ptr(Tc), float(r_min), float(r_max),
                                               int(bool(use_prev_state)), ctypes.byref(n) if sync else None))
        return n.value if sync else None

    def map_get(self):
        """(path n x 3, landmarks m x 3, last camera position)"""
        nt, npts = c_int(), c_int()
        pos = np.zeros(3)
        self._check(self.lib.vo_map_get(self.h, None, 0, ctypes.byref(nt), None, 0, ctypes.byref(npts), _ptr(pos)))
        traj = np.empty((nt.value, 3), np.float64)
        pts = np.empty((npts.value, 3), np.float64)
        self._check(self.lib.vo_map_get(self.h, _ptr(traj), nt.value, ctypes.byref(nt), _ptr(pts), npts.value,
                                        ctypes.byref(npts), _ptr(pos)))
        return traj, pts, pos

    def map_render(self, w=1200, h=600, margin=40, radius=3):
        """Top-down X-Z plot as an (h, w, 3) BGR uint8 image (cv2.imwrite-ready) and its axis limits."""
        img = np.empty((h, w, 3), np.uint8)
        bounds = np.zeros(4)
        self._check(self.lib.vo_map_render(self.h, _ptr(img), w, h, margin, radius, _ptr(bounds)))
        return img, bounds

    def set_profiling(self, enable):
        self._check(self.lib.vo_set_profiling(self.h, int(enable)))

    def stage_times(self):
        ms = np.zeros(6)
        n = c_int()
        self._check(self.lib.vo_get_stage_times(self.h, _ptr(ms), ctypes.byref(n)))
        names = ("upload_pyramid", "klt_landmarks", "compact", "pnp", "promote", "klt_candidates")
        return {k: float(v) / max(n.value, 1) for k, v in zip(names, ms)}, n.value

    # ---- unit-level: cv2.solvePnPRansac / cv2.triangulatePoints ------------------------------------
    def solve_pnp_ransac(self, X, P, K, params=None, sets=None, noalloc=False):
        """noalloc=True: vo_solve_pnp_ransac_noalloc - same results out of the library's static device scratch
        (at most NOALLOC_MAX_POINTS correspondences and NOALLOC_MAX_ITERS iterations; no allocation in the call)."""
        X = np.ascontiguousarray(X, np.float64).reshape(-1, 3)
        P = np.ascontiguousarray(P, np.float32).reshape(-1, 2)
        K = np.ascontiguousarray(K, np.float64)
        n = len(X)
        if params is None:
            params = pnp_defaults()
        if sets is not None:
            sets = np.ascontiguousarray(sets, np.int32)
            if sets.shape[0] < params.max_iters or sets.shape[1] != 4:
                raise ValueError("minimal-set table must be max_iters x 4")
        T = np.zeros((4, 4)); Th = np.zeros(12); rvec = np.zeros(3); tvec = np.zeros(3)
        mask = np.zeros(n, np.uint8); info = np.zeros(4, np.int32); ok = c_int()
        fn = self.lib.vo_solve_pnp_ransac_noalloc if noalloc else self.lib.vo_solve_pnp_ransac
        self._check(fn(self.h, _ptr(X), _ptr(P), n, _ptr(K), ctypes.byref(params), _ptr(sets),
                       _ptr(T), _ptr(Th), _ptr(rvec), _ptr(tvec), _ptr(mask), _ptr(info), ctypes.byref(ok)))
        return dict(ok=bool(ok.value), T=T, R=T[:3, :3].copy(), t=T[:3, 3].copy(), Rh=Th[:9].reshape(3, 3),
                    th=Th[9:], rvec=rvec, tvec=tvec, mask=mask, iters=int(info[0]), best_iter=int(info[1]),
                    n_inliers=int(info[2]), lm_iters=int(info[3]))

    def triangulate_points(self, M1, M2, p1, p2):
        M1 = np.ascontiguousarray(M1, np.float64)
        M2 = np.ascontiguousarray(M2, np.float64).reshape(3, 4)
        p1 = np.ascontiguousarray(p1, np.float64).reshape(-1, 2)
        p2 = np.ascontiguousarray(p2, np.float64).reshape(-1, 2)
        n = len(p1)
        per_point = int(M1.size == 12 * n and n > 1 or M1.ndim == 3)
        out = np.zeros((n, 4))
        self._check(self.lib.vo_triangulate_points(self.h, _ptr(M1), _ptr(M2), per_point, _ptr(p1), _ptr(p2), n, _ptr(out)))
        return out

    # ---- unit-level: cv2.goodFeaturesToTrack (reference vo_pipeline.py:381) ---------------------------
    def good_features_to_track(self, frame_key, mask=None, exclude_pts=None, params=None, want_eig=False, shape=None):
        """Shi-Tomasi corners of a resident frame in cv2's order: (n, 2) float32 (x, y).  `mask` is a
        (h, w) uint8 image or None; `exclude_pts` (n, 2) builds the reference's single-pixel mask."""
        prm = params if params is not None else gftt_defaults()
        if mask is not None:
            mask = np.ascontiguousarray(mask, np.uint8)
            shape = mask.shape
        ex = None
        if exclude_pts is not None and len(exclude_pts):
            ex = np.ascontiguousarray(exclude_pts, np.float32).reshape(-1, 2)
        if want_eig and shape is None:
            raise ValueError("want_eig needs the frame shape")
        cap = prm.max_corners if prm.max_corners > 0 else int(shape[0] * shape[1])
        out = np.zeros((max(cap, 1), 2), np.float32)
        eig = np.empty(shape, np.float32) if want_eig else None
        n = c_int()
        info = np.zeros(4, np.int32)
        self._check(self.lib.vo_good_features_to_track(self.h, frame_key, _ptr(mask), _ptr(ex), 0 if ex is None else len(ex),
                                                       ctypes.byref(prm), _ptr(out), cap, ctypes.byref(n), _ptr(eig), _ptr(info)))
        if want_eig:
            return out[:n.value].copy(), eig, int(info[0])
        return out[:n.value].copy()

    def detect_add_candidates_shi(self, frame_key, params=None, T=None):
        """find_new_candidates_shi + the candidate append on the device, no synchronisation."""
        prm = params if params is not None else gftt_defaults()
        T = None if T is None else np.ascontiguousarray(T, np.float64).reshape(4, 4)
        self._check(self.lib.vo_detect_add_candidates_shi(self.h, frame_key, ctypes.byref(prm), _ptr(T)))
