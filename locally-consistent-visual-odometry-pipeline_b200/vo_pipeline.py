"""Drop-in replacement for the reference's src/vo_pipeline.py (same module name, same class surface).

Put this directory ahead of /root/reference/src on sys.path and the reference's src/main.py runs
unchanged (`from vo_pipeline import *`, main.py:7): the three per-frame stages

    KeypointsToLandmarksAssociator.associateKeypointsToLandmarks   (reference vo_pipeline.py:188-227)
    PoseEstimator.estimatePose                                     (reference vo_pipeline.py:245-335)
    LandmarkTriangulator.triangulate_landmark                      (reference vo_pipeline.py:470-750)

run as hand-written sm_100a CUDA kernels behind the C-ABI of libvo_b200.so (include/vo_b200.h),
with the Markov state P/X/C/F/T resident on the GPU.  The dicts handed back to main.py are lazy
host mirrors: `state['X']` is copied from the device the first time it is read in a frame
(main.py:248 reads it every frame).  Bootstrap (VOInitializer) and the candidate detectors
(SIFT / Shi-Tomasi) are out of the hot path and stay on the CPU through cv2, as in the reference.

There is no CPU fallback for the three stages: without a CUDA device the first call raises.

Differences a caller can observe (DESIGN.md lists them with reasons):
  - hypotheses of the PnP-RANSAC come from P3P (BASELINE.json north_star), not EPnP-5;
  - row matching is by tracking status, not by the reference's element-wise np.isin (App. B-4);
  - the per-frame debug prints of the reference are omitted unless config['debug'] is set.
"""
import os
import sys
import time
import zlib
from typing import Dict

import numpy as np
import yaml

try:   # cv2 is needed only for bootstrap, detectors and the re-bootstrap fallback
    import cv2
except Exception:   # pragma: no cover - the hot path does not need it
    cv2 = None
try:   # the reference leaks `plt` through its star import; main.py imports its own
    import matplotlib.pyplot as plt
except Exception:   # pragma: no cover
    plt = None

_HERE = os.path.dirname(os.path.abspath(__file__))
if __package__:
    from . import _cabi
else:   # imported as top-level `vo_pipeline` (the drop-in case): load the sibling binding by path
    import importlib.util as _ilu
    _pkg_name = "_vo_b200_pkg"
    if _pkg_name not in sys.modules:
        _spec = _ilu.spec_from_file_location(_pkg_name, os.path.join(_HERE, "__init__.py"),
                                             submodule_search_locations=[_HERE])
        _mod = _ilu.module_from_spec(_spec)
        sys.modules[_pkg_name] = _mod
        _spec.loader.exec_module(_mod)
    _cabi = sys.modules[_pkg_name]._cabi

# ---- configuration: same file, same keys, loaded at import from the CWD (reference vo_pipeline.py:11-17)
config_path = 'config.yaml'


def load_config(config_path):
    with open(config_path, 'r') as file:
        return yaml.safe_load(file)


config = load_config('config.yaml')

# optional keys of this implementation; an unmodified config.yaml reproduces the reference's settings
_DEFAULTS = {
    'pnp_max_iters': 100,        # cv2.solvePnPRansac default iterationsCount (never passed, vo:258-264)
    'pnp_refine_iters': 20,
    'klt_fb_check': False,       # the reference has no forward-backward check
    'klt_fb_thresh': 1.0,
    'speculative_promotion': True,   # enqueue the promotion step ahead of estimatePose's sync
    'speculative_tracking': True,    # track into a frame handed over early right behind the promotion step
    # 'feed' mode only: estimatePose enqueues the WHOLE rest of the frame (promotion, candidate feed, landmark batches) and
    # the next frame's tracking step before it waits for the pose, so the GPU never waits for the host between frames.
    # triangulate_landmark must then be called as main.py:280-284 calls it (the state, frames and pose just returned);
    # anything else raises instead of silently diverging.
    'frame_ahead': False,
    'emulate_cf_alias': True,    # reference quirk App. B-5 (first candidate batch: F follows C)
    'device': 0,
    'gpu_detector': True,        # find_new_candidates_shi on the device (csrc/det.cu); False: cv2 on the CPU
    'gpu_bootstrap': True,       # five-point RANSAC + recoverPose on the device (csrc/boot.cu); False: cv2 on the CPU
}


def _cfg(key):
    return config.get(key, _DEFAULTS[key]) if isinstance(config, dict) else _DEFAULTS[key]


# ---- runtime: one vo_ctx per process ---------------------------------------------------------------
class _Runtime:
    def __init__(self, w, h):
        self.ctx = _cabi.Context(device=int(_cfg('device')), max_w=max(w, 2048), max_h=max(h, 1024))
        self.klt = _cabi.klt_defaults()
        self.klt.fb_check = int(bool(_cfg('klt_fb_check')))
        self.klt.fb_thresh = float(_cfg('klt_fb_thresh'))
        self.frames = {}          # id(array) -> (key, strong ref, fingerprint)
        self.hint = None          # frame passed to prefetch_frame, not yet on the ingest stream
        self.prefetched = {}      # data pointer -> (key, strong ref, fingerprint) of frames handed over early
        self.next_key = 1
        self.state_version = 0    # bumped whenever the device landmarks change
        self.ext_version = 0      # bumped whenever the device candidates change
        self.timers = {}
        self.prev_token = 0       # bumped at every tracking step (whose input state stays readable)
        self.pair = None          # (old_frame, new_frame, k_old, k_new) of the last associate call
        self.pending_add = None   # (points, pose) of a deferred vo_add_candidates
        self.spec = None          # assumptions of a pending speculative promotion
        self.spec_track = None    # assumptions of a tracking step enqueued ahead of its associate call
        self.last_prefetch = None # (key, frame, fingerprint) of the newest frame on the ingest stream
        self.ahead = None         # frame_ahead: what estimatePose assumed about the coming triangulate_landmark call
        self.tracks_ahead = 0     # tracking steps the device has taken beyond the last associate call (0 or 1)
        self.need_adopt = False   # an adopted step ahead the C side has not been told about yet
        self.cnt = _cabi.Counters()   # filled in place by the frame_ahead pose call
        self.keep_alive = []      # landmark rows queued for the device, kept until it has read them

    # frames arrive as host uint8 arrays; the same object is passed to several stages and then
    # becomes the "old" frame of the next iteration (main.py:293), so upload + pyramid happen once
    def frame_key(self, img):
        if img.dtype != np.uint8 or img.ndim != 2:
            raise ValueError("frames must be 2-D uint8 (cv2.IMREAD_GRAYSCALE) arrays")
        if img.strides[1] != 1:
            img = np.ascontiguousarray(img)
        # cheap content fingerprint: catches a buffer that was refilled in place
        fp = (img.shape, img.__array_interface__['data'][0], zlib.adler32(img[img.shape[0] >> 1]))
        ent = self.frames.get(id(img))
        if ent is not None and ent[2] == fp and self.ctx.has_frame(ent[0]):
            return ent[0]
        if self.hint is not None and self.hint.__array_interface__['data'][0] == fp[1]:
            self.hint = None      # needed right now: nothing left to hide the ingest under
        pre = self.prefetched.pop(fp[1], None) if self.prefetched else None
        if pre is not None and self.last_prefetch is not None and pre[0] == self.last_prefetch[0]:
            self.last_prefetch = None
        if pre is not None and pre[2] == fp and self.ctx.has_frame(pre[0]):
            key = pre[0]          # handed over early by prefetch_frame: already on its way / resident
        else:
            key = self.next_key
            self.next_key += 1
            self.ctx.upload_frame(img, key)
        self.frames[id(img)] = (key, img, fp)
        if len(self.frames) > 3:   # the ctx keeps 4 pyramids; drop the oldest host references
            for k in sorted(self.frames, key=lambda i: self.frames[i][0])[:-3]:
                del self.frames[k]
        return key

    def prefetch(self, img):
        """Frame ingest: remember a FUTURE frame; its upload + pyramid go to the ingest
        stream as soon as the current frame's tracking is on the main stream (issue_hint), so both the
        host-side enqueue and the device-side work hide under the current frame's stages."""
        if img.dtype != np.uint8 or img.ndim != 2 or img.strides[1] != 1:
            raise ValueError("frames must be 2-D uint8 (cv2.IMREAD_GRAYSCALE) arrays with contiguous rows")
        self.hint = img

    def issue_hint(self, deferred=False):
        img = self.hint
        if img is None:
            return
        self.hint = None
        fp = (img.shape, img.__array_interface__['data'][0], zlib.adler32(img[img.shape[0] >> 1]))
        key = self.next_key
        self.next_key += 1
        self.ctx.prefetch_frame(img, key, deferred=deferred)   # (self.prefetched keeps img alive)
        if len(self.prefetched) >= 2:    # hints that were never used
            self.prefetched.pop(next(iter(self.prefetched)))
        self.prefetched[fp[1]] = (key, img, fp)
        self.last_prefetch = (key, img, fp)

    def drop_spec_track(self):
        """Take back a tracking step that was enqueued ahead of an associate call that never came."""
        if self.spec_track is not None:
            self.spec_track = None
            self.tracks_ahead = 0
            self.ctx.track_rollback()

    def pair_keys(self, old_frame, new_frame):
        """Keys of a frame pair; the pair the last associate call saw is recognised by identity."""
        pr = self.pair
        if pr is not None and pr[0] is old_frame and pr[1] is new_frame and self.ctx.has_frame(pr[2]) \
                and self.ctx.has_frame(pr[3]):
            return pr[2], pr[3]
        return self.frame_key(old_frame), self.frame_key(new_frame)

    # the append of new candidates is deferred until the next frame's upload is on the stream (the
    # copy engine and the pyramid kernels then start ~20 us earlier); anything that looks at the
    # candidates flushes it first
    def flush_pending_add(self):
        if self.pending_add is not None:
            pts, pose = self.pending_add
            self.pending_add = None
            self.ctx.add_candidates(pts, pose)

    def drop_speculation(self):
        """Undo a speculative promotion whose assumptions were not confirmed."""
        if self.spec is not None:
            self.spec = None
            if self.ctx.promote_pending():
                self.ctx.promote_rollback()


_rt = None


def _runtime(img=None):
    global _rt
    if _rt is None:
        h, w = (img.shape if img is not None else (1024, 2048))
        _rt = _Runtime(w, h)
    return _rt


def _reset_runtime():
    """Drop the process-wide context (tests)."""
    global _rt
    if _rt is not None:
        _rt.ctx.close()
    _rt = None


class DeviceState(dict):
    """state {'P','X'} whose arrays live on the GPU; reading an item mirrors both to the host once."""

    def __init__(self, rt, version):
        super().__init__()
        self._rt, self._version, self._dirty = rt, version, False

    def _current(self):
        return not self._dirty and self._rt is _rt and self._version == self._rt.state_version

    def _load(self):
        if not dict.__contains__(self, 'P'):
            rt = self._rt
            st = rt.spec_track
            ah = rt.ahead
            if ah is not None and ah['state'] is self and self._current():
                # frame_ahead, between estimatePose and triangulate_landmark: the inlier rows of this frame; the device may
                # already have tracked them into the next frame (they are then the head of the buffer it tracked FROM)
                P, X = rt.ctx.get_prev_state() if rt.tracks_ahead else rt.ctx.get_state()
                P, X = P[:ah['n_inl']], X[:ah['n_inl']]
            elif st is not None and st['state'] is self and self._current():
                P, X = self._rt.ctx.get_prev_state()   # the device is one tracking step ahead of this dict
            elif self._current():
                if rt.need_adopt:     # the step ahead is this frame's tracking step now: do not let the query take it back
                    rt.ctx.ahead_adopt()
                    rt.need_adopt = False
                P, X = self._rt.ctx.get_state()
                lim = getattr(self, '_limit', None)
                if lim is not None and self._rt.spec is not None:   # speculatively promoted rows are not visible yet
                    P, X = P[:lim], X[:lim]
            elif self._rt is _rt and getattr(self, '_prev', None) == self._rt.prev_token:
                # previous frame's state (main.py:248); two tracking steps back when the pose call already enqueued the next one
                P, X = self._rt.ctx.get_prev2_state() if self._rt.tracks_ahead else self._rt.ctx.get_prev_state()
            else:
                raise RuntimeError("stale DeviceState: the device landmarks have moved on to a later frame")
            dict.__setitem__(self, 'P', P)
            dict.__setitem__(self, 'X', X)

    def __getitem__(self, k):
        if k in ('P', 'X'):
            self._load()
        return dict.__getitem__(self, k)

    def __setitem__(self, k, v):
        if k in ('P', 'X'):
            self._load()
            self._dirty = True     # the host copy is now the authority; the next stage re-uploads
        dict.__setitem__(self, k, v)

    def __contains__(self, k):
        return k in ('P', 'X') or dict.__contains__(self, k)

    def __len__(self):
        return max(2, dict.__len__(self))

    def __iter__(self):
        self._load()
        return dict.__iter__(self)

    def keys(self):
        self._load()
        return dict.keys(self)

    def items(self):
        self._load()
        return dict.items(self)

    def values(self):
        self._load()
        return dict.values(self)

    def get(self, k, d=None):
        return self[k] if k in self else d

    def _rebind(self, version):
        dict.clear(self)
        self._version, self._dirty = version, False
        self._limit = None


class DeviceExtendedState(dict):
    """extended_state {'C','F','T'} resident on the GPU, mirrored lazily like DeviceState."""

    def __init__(self, rt, version):
        super().__init__()
        self._rt, self._version, self._dirty = rt, version, False

    def _current(self):
        return not self._dirty and self._rt is _rt and self._version == self._rt.ext_version

    def _load(self):
        if not dict.__contains__(self, 'C'):
            if not self._current():
                raise RuntimeError("stale DeviceExtendedState: the device candidates have moved on")
            self._rt.drop_speculation()
            st = self._rt.spec_track
            if st is not None and st.get('ahead'):   # (the query would take the step ahead back behind our back)
                self._rt.drop_spec_track()
            self._rt.flush_pending_add()
            C, F, T = self._rt.ctx.get_candidates()
            dict.__setitem__(self, 'C', C.astype(np.float64))
            dict.__setitem__(self, 'F', F)
            dict.__setitem__(self, 'T', T)

    def __getitem__(self, k):
        if k in ('C', 'F', 'T'):
            self._load()
        return dict.__getitem__(self, k)

    def __setitem__(self, k, v):
        if k in ('C', 'F', 'T'):
            self._load()
            self._dirty = True
        dict.__setitem__(self, k, v)

    def __contains__(self, k):
        return k in ('C', 'F', 'T') or dict.__contains__(self, k)

    def __len__(self):
        return max(3, dict.__len__(self))

    def __iter__(self):
        self._load()
        return dict.__iter__(self)

    def keys(self):
        self._load()
        return dict.keys(self)

    def items(self):
        self._load()
        return dict.items(self)

    def _rebind(self, version):
        dict.clear(self)
        self._version, self._dirty = version, False


def _sync_state_to_device(rt, state):
    """Make the device landmarks equal to `state` (upload unless it already mirrors the device)."""
    if isinstance(state, DeviceState) and state._current():
        return
    P = np.asarray(state['P'], np.float32).reshape(-1, 2)      # reference vo_pipeline.py:214
    X = np.asarray(state['X'], np.float64).reshape(-1, 3)
    rt.ctx.set_state(P, X)
    rt.state_version += 1
    if isinstance(state, DeviceState):
        state._version, state._dirty = rt.state_version, False


def _sync_ext_to_device(rt, ext):
    if isinstance(ext, DeviceExtendedState) and ext._current():
        return
    rt.pending_add = None     # superseded: the caller's arrays are the authority now
    C = np.asarray(ext['C'], np.float64).reshape(-1, 2)
    F = np.asarray(ext['F'], np.float64).reshape(-1, 2)
    T = np.asarray(ext['T'], np.float64).reshape(-1, 4, 4)
    alias = bool(_cfg('emulate_cf_alias')) and len(C) > 0 and (ext['C'] is ext['F'])
    rt.ctx.set_candidates(C.astype(np.float32), F, T, cf_alias=alias)
    rt.ext_version += 1
    if isinstance(ext, DeviceExtendedState):
        ext._version, ext._dirty = rt.ext_version, False


# ---- the reference's class surface --------------------------------------------------------------------
class BestVision():
    """State holder (reference vo_pipeline.py:20-38): K, previous_image, state{P,X}, extended_state{C,F,T}."""

    def __init__(self, K: np.ndarray):
        self.K = K
        self.previous_image = np.ndarray
        self.state = {'P': np.ndarray, 'X': np.ndarray}
        self.extended_state = {'C': np.ndarray, 'F': np.ndarray, 'T': np.ndarray}


class VOInitializer():
    """Bootstrap from two frames (reference vo_pipeline.py:41-177).  Runs once per sequence (and in
    the rare re-bootstrap of estimatePose), so it stays on the CPU through cv2 like the reference."""

    def __init__(self, K):
        self.K = K

    @staticmethod
    def _need_cv2():
        if cv2 is None:
            raise ImportError("VOInitializer / the candidate detectors need cv2 (CPU, out of the hot path)")

    def getKeypointMatches(self, frame1: np.ndarray, frame2: np.ndarray):
        """Detector from config['init_detector'] ('sift' | 'shi'), SIFT descriptors, brute-force 2-NN,
        0.8 ratio test.  Returns matched pixel positions in frame1 and frame2."""
        self._need_cv2()
        sift = cv2.SIFT.create()
        detector = config['init_detector']
        if detector == 'shi':
            shi = dict(maxCorners=600, qualityLevel=0.03, minDistance=10, mask=None, blockSize=3,
                       gradientSize=3, useHarrisDetector=False, k=0.04)
            kps = [[cv2.KeyPoint(x=float(c[0][0]), y=float(c[0][1]), size=20)
                    for c in cv2.goodFeaturesToTrack(f, **shi)] for f in (frame1, frame2)]
        elif detector == 'sift':
            kps = [sift.detect(f, None) for f in (frame1, frame2)]
        else:
            raise ValueError("init_detector must be 'sift' or 'shi'")
        if config['init_descriptor'] != 'sift':
            raise ValueError("init_descriptor must be 'sift'")
        (kps1, des1), (kps2, des2) = (sift.compute(f, k) for f, k in zip((frame1, frame2), kps))
        pairs = cv2.BFMatcher.create(cv2.NORM_L2, crossCheck=False).knnMatch(des1, des2, k=2)
        good = [m for m, n in pairs if m.distance < 0.8 * n.distance]
        return (np.array([kps1[m.queryIdx].pt for m in good]), np.array([kps2[m.trainIdx].pt for m in good]))

    def getPoseEstimate(self, kps_f1, kps_f2):
        """5-point essential-matrix RANSAC + cheirality: 3x4 [R|t] of frame2 w.r.t. frame1, inlier mask (n x 1 uint8, as
        cv2 returns it and as recoverPose leaves it: RANSAC inliers that pass the cheirality test).  On the device
        (csrc/boot.cu); config gpu_bootstrap: False keeps the cv2 calls of the reference."""
        if _cfg('gpu_bootstrap'):
            try:
                r = _runtime().ctx.bootstrap_pose(kps_f1, kps_f2, self.K, 0.999, 1.0)
            except _cabi.VoError:
                r = None
            if r is not None:
                if not r['ok']:    # cv2 hands back E = None here and recoverPose raises on it: same loud failure
                    raise ValueError("bootstrap: no essential matrix found for %d matches" % len(kps_f1))
                return r['T34'], r['mask'].reshape(-1, 1)
        self._need_cv2()
        E, mask = cv2.findEssentialMat(kps_f1, kps_f2, self.K, cv2.RANSAC, 0.999, 1.0, None)
        _, R, t, _ = cv2.recoverPose(E, kps_f1, kps_f2, self.K, mask=mask)
        return np.hstack((R, t)), mask

    def get_2D_3D_landmarks_association(self, kps_1: np.ndarray, kps_2: np.ndarray, T_img1_img2: np.ndarray) -> Dict:
        """Triangulate the matches; state['P'] is taken from frame 1 as in the reference (App. B-1).
        The batched DLT runs on the GPU (same kernel as the per-frame promotion) when a device is there."""
        m1 = self.K @ np.eye(3, 4)
        m2 = self.K @ T_img1_img2
        try:
            XH = _runtime().ctx.triangulate_points(m1, m2, kps_1, kps_2)
        except _cabi.VoError:
            self._need_cv2()
            XH = cv2.triangulatePoints(m1, m2, kps_1.T, kps_2.T).T
        return {'X': XH[:, 0:3] / XH[:, 3:4], 'P': kps_1}


class KeypointsToLandmarksAssociator():
    def __init__(self, K, current_pose):
        self.K = K
        self.current_pose = current_pose     # stored and never used, as in the reference (App. B-11)

    def associateKeypointsToLandmarks(self, old_frame: np.ndarray, new_frame: np.ndarray, state: dict):
        """KLT of state['P'] from old_frame to new_frame; returns (new_state, points_tracking_lost)
        with new_state = rows of P (new positions, float32) and X whose status is 1."""
        rt = _runtime(new_frame)
        rt.drop_speculation()      # estimatePose without a matching triangulate_landmark
        if rt.ahead is not None:   # frame_ahead: that frame's stage is on the device already; its inputs are consumed
            rt.ahead = None
            if _candidate_feed:
                _candidate_feed.pop(0)
            if _landmark_feed:
                _landmark_feed.pop(0)
            rt.state_version += 1
            rt.ext_version += 1
            if rt.tracks_ahead:
                rt.tracks_ahead = 0
                rt.ctx.track_rollback()
        st = rt.spec_track
        if st is not None:
            # triangulate_landmark already enqueued this very tracking step behind its promotion (the frame
            # had been handed over early): if the arguments are the ones it assumed, there is nothing to do
            rt.spec_track = None
            fp = (new_frame.shape, new_frame.__array_interface__['data'][0], zlib.adler32(new_frame[new_frame.shape[0] >> 1])) \
                if new_frame.ndim == 2 and new_frame.dtype == np.uint8 and new_frame.strides[1] == 1 else None
            if (st['old'] is old_frame and fp == st['fp'] and state is st['state'] and state._current()
                    and rt.ctx.has_frame(st['k_new'])):
                rt.prefetched.pop(fp[1], None)
                rt.last_prefetch = None
                rt.frames[id(new_frame)] = (st['k_new'], new_frame, fp)
                if len(rt.frames) > 3:
                    fr = rt.frames
                    del fr[min(fr, key=lambda i: fr[i][0])]
                rt.pair = (old_frame, new_frame, st['k_old'], st['k_new'])
                rt.tracks_ahead = 0
                rt.need_adopt = bool(st.get('ahead'))
                rt.prev_token += 1
                if not dict.__contains__(state, 'P'):
                    state._prev = rt.prev_token
                # The next frame's ingest.  A step that was enqueued one frame ahead (frame_ahead) leaves the GPU busy
                # until the coming pose call, which tracks INTO that frame right behind this frame's tail: the ingest
                # (H2D + pyramid, ~35 us) must be on its way now, not halfway through that call - KLT(k+2) would wait for
                # it (15 us of idle GPU per frame).  Otherwise it goes out under the coming PnP.
                rt.issue_hint(deferred=not st.get('ahead'))
                rt.state_version += 1
                return (DeviceState(rt, rt.state_version), _LazyLost(rt, rt.state_version))
            rt.ctx.track_rollback()
            rt.tracks_ahead = 0
        pr = rt.pair
        if pr is not None and pr[1] is old_frame and rt.ctx.has_frame(pr[3]):
            k_old = pr[3]          # main.py:293 (img1 = img2): the frame tracked INTO last time, as uploaded then
        else:
            k_old = rt.frame_key(old_frame)
        k_new = rt.frame_key(new_frame)   # upload + pyramid first (a no-op for a frame handed over early)
        rt.pair = (old_frame, new_frame, k_old, k_new)
        # last frame's new candidates must be tracked from this frame on: they ride on the side stream in
        # front of the candidates' KLT, behind the landmark launch
        add, rt.pending_add = rt.pending_add, None
        _sync_state_to_device(rt, state)
        rt.prev_token += 1
        if not isinstance(state, DeviceState):
            state['P'] = np.asarray(state['P']).astype(np.float32)   # reference side effect (vo:214)
        elif not dict.__contains__(state, 'P'):
            # the caller may still read this (now previous) state after the device has moved on:
            # the landmark rotation keeps it until the next tracking step
            state._prev = rt.prev_token
        rt.ctx.klt_track_landmarks(k_old, k_new, rt.klt, add=add, early_alpha=_early_alpha())
        rt.issue_hint(deferred=True)   # next frame's ingest: enqueued by the pose call, under the GPU's PnP
        rt.state_version += 1
        return (DeviceState(rt, rt.state_version), _LazyLost(rt, rt.state_version))


class _LazyLost:
    """points_tracking_lost (only printed by main.py in debug mode): fetched on first use."""

    def __init__(self, rt, version):
        self._rt, self._version, self._a = rt, version, None

    def _arr(self):
        if self._a is None:
            if self._rt.state_version != self._version:
                raise RuntimeError("lost-track positions of an earlier frame are gone")
            self._a = self._rt.ctx.get_lost()
        return self._a

    def __array__(self, dtype=None, copy=None):
        return self._arr() if dtype is None else self._arr().astype(dtype)

    def __len__(self):
        return len(self._arr())

    def __getitem__(self, i):
        return self._arr()[i]

    def __repr__(self):
        return repr(self._arr())


Associator = KeypointsToLandmarksAssociator
KeypointsToLandmarksAssociation = KeypointsToLandmarksAssociator   # name used by the README / north_star


class PoseEstimator():
    def __init__(self, K):
        self.K = K
        self.REPOJ_THRESH = 2
        self.CONFIDENCE = 0.99999
        if config['find_new_candidates_method'] == 'sift-sift':
            self.CONFIDENCE = config['sift_ransac_confidence']
        if config['find_new_candidates_method'] == 'shi-mask':
            self.CONFIDENCE = config['shi_ransac_confidence']
        self.last = {}

    def _params(self):
        key = (int(_cfg('pnp_max_iters')), float(self.CONFIDENCE), float(self.REPOJ_THRESH), int(_cfg('pnp_refine_iters')))
        if getattr(self, '_pcache', (None,))[0] == key:
            return self._pcache[1]
        p = self._build_params()
        self._pcache = (key, p)
        return p

    def _build_params(self):
        p = _cabi.pnp_defaults()
        p.max_iters = int(_cfg('pnp_max_iters'))
        p.confidence = float(self.CONFIDENCE)
        p.reproj_thresh = float(self.REPOJ_THRESH)
        p.refine_iters = int(_cfg('pnp_refine_iters'))
        return p

    def estimatePose(self, associations, img_idx) -> np.ndarray:
        """PnP-RANSAC + LM refit over associations{'P','X'}; prunes associations to the inliers in
        place and returns the 4x4 world->camera pose (host ndarray: main.py:239 inverts it)."""
        rt = _runtime()
        rt.drop_speculation()
        rt.drop_spec_track()
        rt.issue_hint(deferred=True)
        uploaded = not (isinstance(associations, DeviceState) and associations._current())
        _sync_state_to_device(rt, associations)
        # The promotion step of triangulate_landmark does not need the host: enqueue it ahead of the
        # sync this method ends in, with the pose taken from the device, and let triangulate_landmark
        # confirm it when its arguments are the ones assumed (main.py always passes them).
        method = config['find_new_candidates_method']
        if (bool(_cfg('frame_ahead')) and method == 'feed' and bool(_cfg('speculative_promotion')) and not uploaded
                and rt.pair is not None and rt.ctx.candidates_tracked(rt.pair[2], rt.pair[3])):
            return self._estimate_frame_ahead(rt, associations, img_idx)
        speculate = (bool(_cfg('speculative_promotion')) and not uploaded
                     and (method in ('feed', 'sift-sift') or (method == 'shi-mask' and _device_shi()))
                     and rt.pair is not None and rt.ctx.candidates_tracked(rt.pair[2], rt.pair[3]))
        if speculate:
            alpha = float(_alpha_threshold())
            ok, T, n_inl, _ = rt.ctx.pnp_ransac_promote(self.K, self._params(), alpha, early=True)
        else:
            ok, T, n_inl = rt.ctx.pnp_ransac(self.K, self._params())
        self.last = {'ok': ok, 'n_inliers': n_inl}
        if config.get('debug'):
            print('len inliers', n_inl)
        if not ok or n_inl < 4:
            rt.drop_speculation()      # (no model: already rolled back inside the library)
            return self._rebootstrap(associations, img_idx)
        rt.state_version += 1
        if isinstance(associations, DeviceState):
            associations._rebind(rt.state_version)
        else:   # a plain dict was passed: prune it on the host too
            P, X = rt.ctx.get_state()
            associations['P'], associations['X'] = P[:n_inl], X[:n_inl]
        if speculate:
            rt.spec = {'T': T.tobytes(), 'alpha': alpha, 'K': np.asarray(self.K, np.float64).tobytes(), 'pair': rt.pair,
                       'state_version': rt.state_version, 'ext_version': rt.ext_version}
            if isinstance(associations, DeviceState):
                associations._limit = n_inl
        return T

    def _estimate_frame_ahead(self, rt, associations, img_idx):
        """config frame_ahead: PnP, then - before the pose is waited for - the promotion step, the append of this
        frame's candidate feed and landmark batches and the tracking step into the frame handed over early
        (vo_process_frame_pipelined on the tracking step associate made).  triangulate_landmark only confirms.
        This is the host's critical path between two frames: no conversions that are not needed."""
        alpha = float(_alpha_threshold())
        feed = _candidate_feed[0] if _candidate_feed else None
        batches = _landmark_feed[0] if _landmark_feed else None
        keep = rt.keep_alive            # rows the device reads after this call returns
        if batches:
            for bP, bX, target in batches:
                if len(bP):
                    if not (isinstance(bP, np.ndarray) and bP.dtype == np.float32 and bP.flags.c_contiguous):
                        bP = np.ascontiguousarray(bP, np.float32).reshape(-1, 2)
                    if not (isinstance(bX, np.ndarray) and bX.dtype == np.float64 and bX.flags.c_contiguous):
                        bX = np.ascontiguousarray(bX, np.float64).reshape(-1, 3)
                    keep.append((bP, bX))
                    rt.ctx.queue_landmarks_raw(bP, bX, int(target))
        del keep[:-8]
        fd = None
        if feed is not None and len(feed):
            fd = feed if (isinstance(feed, np.ndarray) and feed.dtype == np.float64 and feed.flags.c_contiguous) \
                else np.ascontiguousarray(feed, np.float64).reshape(-1, 2)
        Kc = self.K if (isinstance(self.K, np.ndarray) and self.K.dtype == np.float64 and self.K.flags.c_contiguous) \
            else np.ascontiguousarray(self.K, np.float64)
        T = np.empty((4, 4), np.float64)
        cnt = rt.cnt
        rt.ctx.process_frame_tracked(rt.pair[2], rt.pair[3], Kc, alpha, fd, rt.klt, self._params(), T, cnt)
        ok, n_inl = bool(cnt.pnp_ok), cnt.n_inliers
        self.last = {'ok': ok, 'n_inliers': n_inl}
        rt.need_adopt = False
        rt.tracks_ahead = cnt.reserved_ & 1
        if not ok or n_inl < 4:
            rt.tracks_ahead = 0
            if _candidate_feed:
                _candidate_feed.pop(0)
            if _landmark_feed:
                _landmark_feed.pop(0)
            rt.state_version += 1
            rt.ext_version += 1
            return self._rebootstrap(associations, img_idx)
        rt.state_version += 1
        associations._rebind(rt.state_version)
        rt.ahead = {'T': T, 'alpha': alpha, 'K': Kc, 'pair': rt.pair, 'state': associations, 'n_inl': n_inl, 'feed': feed,
                    'batches': batches, 'ext_version': rt.ext_version, 'pre': rt.last_prefetch if rt.tracks_ahead else None}
        return T

    def _rebootstrap(self, associations, img_idx):
        """Fewer than 4 inliers: bootstrap again from frames img_idx and img_idx+4 read from disk,
        append the new landmarks and retry (reference vo_pipeline.py:274-324).  CPU, rare."""
        VOInitializer._need_cv2()
        print('less than 4 inliers (or None)! re-bootstrapping at frame', img_idx)
        init = VOInitializer(self.K)
        ds = config['dataset']
        if ds == 'kitti':
            names = ['kitti-dataset/05/image_0/%06d.png' % i for i in (img_idx, img_idx + 4)]
        elif ds == 'malaga':
            d = 'malaga-urban-dataset-extract-07/malaga-urban-dataset-extract-07_rectified_800x600_Images'
            left = sorted(f for f in os.listdir(d) if f.endswith('left.jpg'))
            names = [os.path.join(d, left[i]) for i in (img_idx, img_idx + 4)]
        elif ds == 'parking':
            names = ['parking/images/img_%05d.png' % i for i in (img_idx, img_idx + 4)]
        else:
            raise ValueError("Invalid dataset selection")
        cur, nxt = (cv2.imread(n, cv2.IMREAD_GRAYSCALE) for n in names)
        k1, k2 = init.getKeypointMatches(cur, nxt)
        T12, mask = init.getPoseEstimate(k1, k2)
        mask = np.hstack(mask).astype(np.bool_)
        fresh = init.get_2D_3D_landmarks_association(k1[mask, :], k2[mask, :], T12)
        P = np.vstack((np.asarray(associations['P'], np.float64).reshape(-1, 2), fresh['P']))
        X = np.vstack((np.asarray(associations['X'], np.float64).reshape(-1, 3), fresh['X']))
        associations['P'], associations['X'] = P, X
        return self.estimatePose(associations, img_idx)


class LandmarkTriangulator():
    """Detects new candidate keypoints, tracks the candidates from frame to frame and promotes them
    to landmarks once the bearing angle between first and current observation exceeds a threshold
    (reference vo_pipeline.py:337-750).  Constructed every frame by main.py:280: O(1), no allocation."""

    def __init__(self, K, old_des):
        self.K = K
        self.old_des = old_des

    # ---- candidate detectors: CPU / cv2, outside the timed hot path ----------
    @staticmethod
    def _pixel_mask(frame, keypoints_well_tracked):
        mask = np.full(frame.shape[:2], 255, np.uint8)
        pts = np.asarray(keypoints_well_tracked, np.float64).reshape(-1, 2)
        if len(pts):
            x = pts[:, 0].astype(np.int64)      # int() truncation toward zero, as the reference
            y = pts[:, 1].astype(np.int64)
            ok = (y >= 0) & (y < mask.shape[0]) & (x >= 0) & (x < mask.shape[1])
            mask[y[ok], x[ok]] = 0
        return mask

    def find_new_candidates_shi(self, new_frame: np.ndarray, state: dict, keypoints_well_tracked):
        """Shi-Tomasi corners (500, quality 0.03, min distance 10, block 4) outside the single pixels
        of the well-tracked candidates (reference vo_pipeline.py:357-390)."""
        if _cfg('gpu_detector'):
            # same corners in the same order as cv2.goodFeaturesToTrack with the reference's mask
            # (tests/test_gftt_gpu.py); the mask is built on the device from the point list
            rt = _runtime(new_frame)
            pts = np.asarray(keypoints_well_tracked, np.float64).reshape(-1, 2)
            return rt.ctx.good_features_to_track(rt.frame_key(new_frame), exclude_pts=pts).astype(np.float64)
        VOInitializer._need_cv2()
        mask = self._pixel_mask(new_frame, keypoints_well_tracked)
        corners = cv2.goodFeaturesToTrack(new_frame, maxCorners=500, qualityLevel=0.03, minDistance=10, mask=mask,
                                          blockSize=4, gradientSize=3, useHarrisDetector=False, k=0.04)
        if corners is None:
            return np.zeros((0, 2))
        return corners.reshape(-1, 2).astype(np.float64)

    def find_new_candidates_sift_mask(self, new_frame: np.ndarray, state: dict, keypoints_well_tracked):
        """SIFT keypoints outside the masked pixels (reference vo_pipeline.py:392-425)."""
        VOInitializer._need_cv2()
        mask = self._pixel_mask(new_frame, keypoints_well_tracked)
        kps = cv2.SIFT.create().detect(new_frame, mask=mask)
        return np.array([kp.pt for kp in kps]).reshape(-1, 2)

    def find_new_candidates_sift_sift(self, new_frame: np.ndarray, state: dict, extended_state: dict):
        """SIFT keypoints of the new frame minus those whose descriptor matches self.old_des (ratio
        0.8); returns (new_keypoints, descriptors) (reference vo_pipeline.py:427-467)."""
        VOInitializer._need_cv2()
        keypoints, cur_des = cv2.SIFT.create().detectAndCompute(new_frame, None)
        pairs = cv2.BFMatcher.create(cv2.NORM_L2, crossCheck=False).knnMatch(self.old_des, cur_des, k=2)
        good = [m for m, n in pairs if m.distance < 0.8 * n.distance]
        old_keypoints = np.array([keypoints[m.trainIdx].pt for m in good])
        keypoints = np.array([kp.pt for kp in keypoints])
        keep = np.isin(keypoints, old_keypoints, invert=True).all(axis=1)   # element-wise, as the reference
        return (keypoints[keep], cur_des)

    def _confirm_frame_ahead(self, rt, old_frame, new_frame, state, extended_state, new_pose):
        """config frame_ahead: estimatePose has enqueued this whole stage already; check that it is being asked for
        what it assumed (main.py:280-284 always does) and hand the state objects on."""
        ah, rt.ahead = rt.ahead, None
        if isinstance(extended_state, DeviceExtendedState):
            ext_ok = extended_state._current()
        else:
            ext_ok = rt.ext_version == 0 and len(np.asarray(extended_state['C'])) == 0
        feed = _candidate_feed.pop(0) if _candidate_feed else None
        batches = _landmark_feed.pop(0) if _landmark_feed else None
        same_pose = new_pose is ah['T'] or (isinstance(new_pose, np.ndarray) and new_pose.shape == (4, 4) and np.array_equal(new_pose, ah['T']))
        same_K = self.K is ah['K'] or np.array_equal(np.asarray(self.K, np.float64), ah['K'])
        if not (ext_ok and ah['ext_version'] == rt.ext_version and state is ah['state'] and state._current()
                and ah['pair'] is rt.pair and rt.pair[0] is old_frame and rt.pair[1] is new_frame
                and float(_alpha_threshold()) == ah['alpha'] and same_pose and same_K
                and feed is ah['feed'] and batches is ah['batches']):
            raise RuntimeError("frame_ahead: triangulate_landmark was not called with the frames, state and pose of the "
                               "estimatePose call before it (set frame_ahead: False for arbitrary call sequences)")
        rt.state_version += 1
        rt.ext_version += 1
        dict.clear(state)
        state._version, state._dirty, state._limit = rt.state_version, False, None
        pre = ah['pre']
        if pre is not None:    # the tracking step into the frame handed over early is on the stream already
            rt.spec_track = {'old': new_frame, 'fp': pre[2], 'k_old': rt.pair[3], 'k_new': pre[0], 'state': state, 'ahead': True}
        if isinstance(extended_state, DeviceExtendedState):
            extended_state._rebind(rt.ext_version)
        else:
            extended_state = DeviceExtendedState(rt, rt.ext_version)
        return (state, extended_state, [])

    # ---- the per-frame stage ---------------------------------------------------------------------------
    def triangulate_landmark(self, old_frame: np.ndarray, new_frame: np.ndarray, state: dict, extended_state: dict,
                             new_pose):
        """Track C old->new, drop lost rows of C/F/T, add the new candidates (F := C, T := new_pose),
        promote every candidate whose bearing angle alpha(C, F) exceeds the threshold: triangulate it
        with (T, new_pose), append it to state P/X and remove it from C/F/T.
        Returns (state, extended_state, cur_des)."""
        rt = _runtime(new_frame)
        if rt.ahead is not None:
            return self._confirm_frame_ahead(rt, old_frame, new_frame, state, extended_state, new_pose)
        rt.drop_spec_track()
        method = config['find_new_candidates_method']
        threshold = _alpha_threshold()
        new_pose = np.ascontiguousarray(new_pose, np.float64)
        sp = rt.spec
        if isinstance(extended_state, DeviceExtendedState):
            ext_ok = extended_state._current()
        else:   # first frame: main.py passes empty numpy arrays (main.py:185-188); the ctx is fresh and empty too
            ext_ok = rt.ext_version == 0 and len(np.asarray(extended_state['C'])) == 0
        confirmed = (sp is not None and ext_ok and sp['ext_version'] == rt.ext_version
                     and isinstance(state, DeviceState) and state._current()
                     and state._version == sp['state_version'] and sp['pair'] is rt.pair
                     and rt.pair[0] is old_frame and rt.pair[1] is new_frame
                     and float(threshold) == sp['alpha'] and new_pose.tobytes() == sp['T']
                     and np.asarray(self.K, np.float64).tobytes() == sp['K'])
        committed = True
        if confirmed:
            # the promotion ran ahead of estimatePose's sync with exactly these arguments
            rt.spec = None
            committed = False     # (left to the tracking call below when there is one: one ctypes call less)
        else:
            rt.drop_speculation()
            k_old, k_new = rt.pair_keys(old_frame, new_frame)
            rt.flush_pending_add()
            _sync_state_to_device(rt, state)
            if not isinstance(extended_state, DeviceExtendedState) and len(np.asarray(extended_state['C'])) == 0:
                extended_state = {'C': np.zeros((0, 2)), 'F': np.zeros((0, 2)), 'T': np.zeros((0, 4, 4))}   # main.py:185-188
            _sync_ext_to_device(rt, extended_state)

            # 1) + 3) on the device: candidate KLT (already done with the landmarks' launch when the same
            # frame pair went through associateKeypointsToLandmarks), removal, bearing test, DLT, promotion
            rt.ctx.track_candidates(k_old, k_new, rt.klt)
        well_tracked_prev = []
        # shi-mask with the stock detector method: detection, mask and append stay on the device
        device_shi = (method == 'shi-mask' and _device_shi()
                      and type(self).find_new_candidates_shi is LandmarkTriangulator.find_new_candidates_shi)
        if method in ('shi-mask', 'sift-mask') and not device_shi:
            prev, _, st = rt.ctx.candidate_tracks()          # the detectors mask these pixels
            well_tracked_prev = prev[st == 1]
        if not confirmed:
            rt.ctx.track_triangulate_promote(k_old, k_new, self.K, new_pose, float(threshold), rt.klt)

        # 2) new candidates: CPU detectors run while the GPU works on the promotion
        new_candidates_list, cur_des = [], []
        if device_shi:
            # vo_pipeline.py:381 + :574-591 in one enqueue: corners masked at the previous positions of
            # the candidates KLT kept, appended with T = new_pose; nothing returns to the host
            rt.ctx.detect_add_candidates_shi(rt.pair[3] if confirmed else k_new, T=new_pose.copy())
        elif method == 'shi-mask':
            new_candidates_list = self.find_new_candidates_shi(new_frame, state, well_tracked_prev)
        elif method == 'sift-mask':
            new_candidates_list = self.find_new_candidates_sift_mask(new_frame, state, well_tracked_prev)
        elif method == 'sift-sift':
            new_candidates_list, cur_des = self.find_new_candidates_sift_sift(new_frame, state, extended_state)
        elif method == 'feed':   # pre-generated candidate feed (benchmarks / tests)
            new_candidates_list = _candidate_feed.pop(0) if _candidate_feed else []
        if len(new_candidates_list) > 0:
            # appended on the device with the next tracking step (or when someone looks); the rows are
            # copied when that happens, so the caller's feed array is snapshotted here
            rt.pending_add = (np.array(new_candidates_list, np.float64), new_pose.copy())   # T = eye(4) @ new_pose (vo:582)

        # landmark batches from outside the tracking loop (set_landmark_feed): behind the promoted rows, in
        # front of the next tracking step
        if _landmark_feed:
            batches = _landmark_feed.pop(0)
            for bP, bX, target in batches:
                if len(bP):
                    rt.ctx.append_landmarks(bP, bX, target)
                    committed = True      # (an append behind a speculative promotion confirms it)
        rt.state_version += 1
        rt.ext_version += 1
        if isinstance(state, DeviceState):
            dict.clear(state)            # (_rebind, inlined: this sits between the pose and the next launch)
            state._version, state._dirty, state._limit = rt.state_version, False, None
        else:
            state = DeviceState(rt, rt.state_version)
        # The next frame is already on the ingest stream (prefetch_frame / FrameIngest): enqueue its
        # tracking now, right behind the promotion step, instead of leaving the GPU idle until main.py
        # comes round to associateKeypointsToLandmarks - which then finds the work done, or takes it back
        # (vo_track_rollback) if it is called with anything else than (new_frame, that frame, this state).
        pre = rt.last_prefetch
        if (confirmed and pre is not None and bool(_cfg('speculative_tracking'))
                and (method == 'feed' or device_shi)):
            add = rt.pending_add
            try:    # (confirms the promotion as well; fails only if the frame ahead has left the slot cache)
                rt.ctx.klt_track_landmarks(rt.pair[3], pre[0], rt.klt, add=add, early_alpha=_early_alpha())
                rt.pending_add = None
                committed = True
                rt.spec_track = {'old': new_frame, 'fp': pre[2], 'k_old': rt.pair[3], 'k_new': pre[0], 'state': state}
            except _cabi.VoError as e:
                if e.code != -4:      # anything but VO_ERR_NOFRAME is a real failure
                    raise
                rt.last_prefetch = None
        if not committed:
            rt.ctx.promote_commit()
        if isinstance(extended_state, DeviceExtendedState):
            extended_state._rebind(rt.ext_version)
        else:
            extended_state = DeviceExtendedState(rt, rt.ext_version)
        return (state, extended_state, cur_des)


def _early_alpha():
    """Threshold for classifying the candidates on the side stream during the tracking step (the pose-independent
    half of the promotion): only when estimatePose will enqueue the promotion itself, with this very alpha."""
    method = config['find_new_candidates_method']
    if bool(_cfg('speculative_promotion')) and (method in ('feed', 'sift-sift') or (method == 'shi-mask' and _device_shi())):
        return float(_alpha_threshold())
    return None


def _device_shi():
    return bool(_cfg('gpu_detector'))


def _alpha_threshold():
    """Bearing-angle threshold in degrees (reference vo_pipeline.py:608-612)."""
    method = config['find_new_candidates_method']
    if method == 'sift-sift':
        return config['sift_alpha']
    if method == 'shi-mask':
        return config['shi_alpha']
    return 1


# candidate feed for find_new_candidates_method == 'feed': a list of (n x 2) arrays, one per frame
_candidate_feed = []


def set_candidate_feed(feed):
    """Per-frame candidate positions used when config['find_new_candidates_method'] == 'feed'."""
    _candidate_feed[:] = list(feed)


# landmark feed: per frame a list of (P n x 2, X n x 3, target_total) batches appended behind state P/X at the
# end of triangulate_landmark - what the reference does with the fresh landmarks of its re-bootstrap
# (vo_pipeline.py:318-320), offered here as a hook for map-based relocalisation and for benchmarks
_landmark_feed = []


def set_landmark_feed(feed):
    """feed: one entry per coming frame, each a list of (P, X, target_total) batches; target_total > 0 appends
    only as many rows as bring |P| up to it (decided on the device), 0 appends all rows."""
    _landmark_feed[:] = list(feed)


def append_landmarks(state, P, X, target_total=0):
    """state['P'], state['X'] := vstack with (P, X) (reference vo_pipeline.py:318-320) without a round trip
    through the host when `state` mirrors the device; returns the state to use from here on."""
    rt = _runtime()
    if isinstance(state, DeviceState) and state._current():
        rt.drop_speculation()
        rt.drop_spec_track()
        rt.ctx.append_landmarks(P, X, target_total)
        rt.state_version += 1
        state._rebind(rt.state_version)
        return state
    P0 = np.asarray(state['P'], np.float64).reshape(-1, 2)
    X0 = np.asarray(state['X'], np.float64).reshape(-1, 3)
    n = len(P) if target_total <= 0 else int(np.clip(target_total - len(P0), 0, len(P)))
    return {'P': np.vstack((P0, np.asarray(P, np.float64).reshape(-1, 2)[:n])),
            'X': np.vstack((X0, np.asarray(X, np.float64).reshape(-1, 3)[:n]))}


# ---- frame ingest (the reference reads frame k+1 from disk only after frame k is done,
# main.py:216-223) ---------------------------------------------------------------------------------------
def prefetch_frame(img):
    """Optional hint: `img` (2-D uint8, ideally pinned - see FrameIngest) will be passed as new_frame
    soon.  Its H2D copy and pyramid are enqueued on the ingest stream right after the tracking of the
    current frame (inside the next associateKeypointsToLandmarks / estimatePose call), so they run under
    the stages of the current frame; the stage that first uses the frame waits for it on the device.
    Results do not change."""
    _runtime(img).prefetch(img)


class FrameIngest:
    """Pre-decoded frames in a ring of pinned host buffers, decoded ahead of their use by a pool of worker
    threads (cv2.imread releases the GIL; main.py:216-223 decodes one frame per iteration, 3-4 ms each) and handed
    to the device one frame ahead:

        for img2 in FrameIngest(paths, workers=8):          # instead of cv2.imread(...) per iteration
            new_state, lost = associate.associateKeypointsToLandmarks(img1, img2, vision.state)

    `source` is a sequence of file paths (read with cv2.IMREAD_GRAYSCALE, as main.py:217-223 does) or of
    2-D uint8 arrays.  Frames come out in order whatever order the workers finish in.  Yields numpy views over
    pinned memory; a view stays valid for two more iterations (main.py keeps the previous frame as img1).
    depth: ring slots (default workers + 3: previous, current and next frame are live, the rest is decode-ahead)."""

    def __init__(self, source, depth=None, prefetch=True, workers=1):
        workers = max(int(workers), 1)
        if depth is None:
            depth = workers + 3
        if depth < 3:
            raise ValueError("FrameIngest needs depth >= 3 (previous, current and next frame are live)")
        self.source, self.depth, self.do_prefetch, self.workers = source, depth, prefetch, workers
        self.ring, self.shape = [], None

    def _decode(self, item):
        if isinstance(item, (str, os.PathLike)):
            VOInitializer._need_cv2()
            img = cv2.imread(os.fspath(item), cv2.IMREAD_GRAYSCALE)
            if img is None:
                raise FileNotFoundError(item)
            return img
        return np.asarray(item)

    def _slot(self, i, img):
        if img.dtype != np.uint8 or img.ndim != 2:
            raise ValueError("frames must be 2-D uint8")
        if self.shape != img.shape:   # (re)allocate the ring for this frame size
            self.ring = [_cabi.PinnedArray(img.shape, np.uint8) for _ in range(self.depth)]
            self.shape = img.shape
        dst = self.ring[i % self.depth].array
        dst[...] = img
        return dst

    def __len__(self):
        return len(self.source)

    def __iter__(self):
        import threading
        n = len(self.source)
        # frame i may be written into its ring slot once frame i - depth is dead.  `consumed` = the last frame whose
        # iteration the caller has finished: it still holds that one (img1), the generator holds the next two, every
        # older frame is dead - so frame i is free to decode while i <= consumed + depth - 1
        cond = threading.Condition()
        done = {}                    # index -> pinned view, or the exception its decode raised
        state = {"next": 0, "consumed": -1, "stop": False}

        def worker():
            while True:
                with cond:
                    while not state["stop"] and (state["next"] >= n or state["next"] > state["consumed"] + self.depth - 1):
                        if state["next"] >= n:
                            return
                        cond.wait()
                    if state["stop"]:
                        return
                    i = state["next"]
                    state["next"] += 1
                try:
                    img = self._decode(self.source[i])
                    with cond:       # (ring allocation and the copy into the slot; slots of different frames differ)
                        if self.shape != img.shape:
                            self._slot(i, img)
                            res = self.ring[i % self.depth].array
                        else:
                            res = None
                    if res is None:
                        res = self.ring[i % self.depth].array
                        res[...] = img
                except BaseException as e:   # surface decode errors in the consumer, at the frame's position
                    res = e
                with cond:
                    done[i] = res
                    cond.notify_all()

        threads = [threading.Thread(target=worker, daemon=True) for _ in range(min(self.workers, max(n, 1)))]
        for t in threads:
            t.start()

        def take(i):
            if i >= n:
                return None
            with cond:
                while i not in done:
                    cond.wait()
                r = done.pop(i)
            if isinstance(r, BaseException):
                with cond:
                    state["stop"] = True
                    cond.notify_all()
                raise r
            return r

        try:
            cur = take(0)
            j = 0
            while cur is not None:
                nxt = take(j + 1)
                if nxt is not None and self.do_prefetch:
                    prefetch_frame(nxt)
                yield cur
                with cond:
                    state["consumed"] = j
                    cond.notify_all()
                cur = nxt
                j += 1
        finally:
            with cond:
                state["stop"] = True
                cond.notify_all()


# ---- trajectory / map bookkeeping on the device (main.py:239-255, :262-271, :296-305) ----
class TrajectoryMap:
    """poses_plotting / X_plotting of main.py kept on the device.  Instead of

        pos = (np.linalg.inv(new_pose) @ pos_h)[0:3]; X = vision.state['X']      # D2H of all landmarks
        X_plotting = np.vstack((X_plotting, X[(norm(X - pos) < 50) & (norm(X - pos) > 3)]))

    every frame, call `tmap.update(new_pose)` right after estimatePose: one small kernel appends the camera
    position and the filtered landmarks (the ones the frame started from, as main.py:248 reads them);
    nothing comes back to the host until `arrays()` / `savefig()`."""

    def __init__(self, r_min=3.0, r_max=50.0):
        self.r_min, self.r_max = float(r_min), float(r_max)
        _runtime().ctx.map_reset()

    def update(self, new_pose=None):
        """new_pose None: the pose the last estimatePose left on the device."""
        _runtime().ctx.map_accumulate(new_pose, self.r_min, self.r_max, use_prev_state=True)

    def arrays(self):
        """(poses_plotting n x 3, X_plotting m x 3) as host arrays."""
        traj, pts, _ = _runtime().ctx.map_get()
        return traj, pts

    def render(self, w=1200, h=600, margin=40, radius=3):
        """The X-Z plot of main.py:262-271 as an (h, w, 3) BGR image: landmarks blue, travelled path red."""
        return _runtime().ctx.map_render(w, h, margin, radius)[0]

    def savefig(self, path, **kw):
        VOInitializer._need_cv2()
        cv2.imwrite(path, self.render(**kw))
