"""ORACLE / CPU BASELINE (test infrastructure only - never imported by the product path).

A port of the reference's three per-frame methods over cv2, for machines where /root/reference is
not present (the GPU box): same cv2 calls, same Python loops, same numpy glue, so that its timing
stands for "the reference's own CPU implementation of the path" (kind = "port") and its outputs
for the reference's.  Validated against the real classes imported from /root/reference/src in the
build container (tests/golden/make_golden.py `seq`, tests/test_pipeline_oracle.py).

  associate()    KeypointsToLandmarksAssociator.associateKeypointsToLandmarks  vo_pipeline.py:188-227
  estimate()     PoseEstimator.estimatePose (without the disk re-bootstrap)     vo_pipeline.py:245-335
  triangulate()  LandmarkTriangulator.triangulate_landmark                      vo_pipeline.py:470-750
Candidate detection is replaced by a pre-generated feed (detectors are excluded from
the timed region).  pnp_flags selects the literal reference call (cv2.SOLVEPNP_ITERATIVE = EPnP-5
hypotheses) or the P3P-hypothesis variant this project gates parity on: RANSAC mask
from flags=SOLVEPNP_P3P, pose from cv2.solvePnP(ITERATIVE) on the inliers.
"""
import numpy as np

try:
    import cv2
except Exception:   # pragma: no cover
    cv2 = None


class RefPipeline:
    def __init__(self, K, confidence=0.99999, reproj=2.0, alpha=1.0, pnp="p3p", max_iters=100, row_match="isin",
                 detector=None):
        # detector="shi": config find_new_candidates_method 'shi-mask' (vo_pipeline.py:357-390, :519-520):
        # cv2.goodFeaturesToTrack with the reference's single-pixel mask instead of a candidate feed
        self.detector = detector
        self.K = np.asarray(K, np.float64)
        self.confidence, self.reproj, self.alpha, self.pnp, self.max_iters = confidence, reproj, alpha, pnp, max_iters
        self.row_match = row_match
        self.info = {}

    # vo_pipeline.py:188-227
    def associate(self, old_frame, new_frame, state):
        state['P'] = state['P'].astype(np.float32)
        next_points, status, err = cv2.calcOpticalFlowPyrLK(old_frame, new_frame, state['P'], None)
        ok = np.hstack(status).astype(np.bool_)
        return ({'P': next_points[ok], 'X': state['X'][ok]}, next_points[np.logical_not(ok)])

    # vo_pipeline.py:245-335 (the <4 inliers branch raises instead of reading frames from disk)
    def estimate(self, associations):
        P, X = associations['P'], associations['X']
        if self.pnp == "literal":
            ok, rvec, t, inliers = cv2.solvePnPRansac(X, P, self.K, np.zeros(4), flags=cv2.SOLVEPNP_ITERATIVE,
                                                      confidence=self.confidence, reprojectionError=self.reproj)
        else:
            ok, rvec, t, inliers = cv2.solvePnPRansac(X, P, self.K, np.zeros(4), flags=cv2.SOLVEPNP_P3P,
                                                      confidence=self.confidence, reprojectionError=self.reproj,
                                                      iterationsCount=self.max_iters)
            if inliers is not None and len(inliers) >= 4:
                idx = np.hstack(inliers)
                Xi = np.asarray(X, np.float64)[idx].astype(np.float32).astype(np.float64)
                _, rvec, t = cv2.solvePnP(Xi, np.asarray(P, np.float64)[idx], self.K, np.zeros(4), flags=cv2.SOLVEPNP_ITERATIVE)
        R, _ = cv2.Rodrigues(rvec)
        if inliers is None or len(inliers) < 4:
            raise RuntimeError("fewer than 4 PnP inliers: the reference would re-bootstrap from disk here")
        inliers = np.hstack(inliers)
        self.info['n_inliers'] = len(inliers)
        associations['P'] = associations['P'][inliers, :]
        associations['X'] = associations['X'][inliers, :]
        return np.concatenate([np.concatenate([R, t], axis=-1), np.array([[0, 0, 0, 1]])], axis=0)

    # vo_pipeline.py:470-750 with the detector replaced by `new_candidates`
    # vo_pipeline.py:357-390
    def find_new_candidates_shi(self, new_frame, keypoints_well_tracked):
        mask = np.full(new_frame.shape[:2], 255, np.uint8)
        for i in range(len(keypoints_well_tracked)):
            y, x = int(keypoints_well_tracked[i][1]), int(keypoints_well_tracked[i][0])
            if 0 <= y < mask.shape[0] and 0 <= x < mask.shape[1]:
                mask[y, x] = 0
        kp = cv2.goodFeaturesToTrack(new_frame, maxCorners=500, qualityLevel=0.03, minDistance=10, mask=mask, blockSize=4,
                                     gradientSize=3, useHarrisDetector=False, k=0.04)
        return np.array([cv2.KeyPoint(x=pt[0][0], y=pt[0][1], size=20).pt for pt in kp])

    def triangulate(self, old_frame, new_frame, state, ext, new_pose, new_candidates):
        c1 = ext['C'].astype(np.float32)
        c1_ok = []
        if len(c1) > 0:
            c2, st, err = cv2.calcOpticalFlowPyrLK(old_frame, new_frame, c1, None)
            st = np.hstack(st).astype(np.bool_)
            c2_ok, c1_ok, c1_lost = c2[st == 1], c1[st == 1], c1[st == 0]
            if self.row_match == "isin":      # the reference's element-wise "row" matching (App. B-4)
                if self.detector == "shi":
                    mask = np.isin(c2, c2_ok).all(axis=1)        # vo_pipeline.py:519-520
                else:
                    mask = np.isin(ext['C'], c1_ok).all(axis=1)  # vo_pipeline.py:521-522
                ext['C'][mask] = c2_ok
                mask = np.isin(ext['C'], c1_lost).all(axis=1)
            else:                              # what it means: status-mask selection
                ext['C'][st] = c2_ok
                mask = ~st
            rm = np.where(mask)[0]
            if len(rm) > 0:
                ext['C'] = np.delete(ext['C'], rm, axis=0)
                ext['F'] = np.delete(ext['F'], rm, axis=0)
                ext['T'] = np.delete(ext['T'], rm, axis=0)
        if self.detector == "shi" and new_candidates is None:
            new_candidates = self.find_new_candidates_shi(new_frame, c1_ok)
        if len(new_candidates) > 0:
            newT = np.stack([(np.eye(4) @ new_pose) for _ in range(len(new_candidates))])
            if len(ext['C']) == 0:
                ext['C'] = new_candidates
                ext['F'] = new_candidates          # same ndarray: the C/F aliasing quirk (App. B-5)
                ext['T'] = newT
            else:
                ext['C'] = np.concatenate((ext['C'], new_candidates), axis=0)
                ext['F'] = np.concatenate((ext['F'], new_candidates), axis=0)
                ext['T'] = np.concatenate((ext['T'], newT), axis=0)
        validate = []
        for idx in range(len(ext['C'])):
            uc, vc = ext['C'][idx]
            wc = np.sqrt(uc**2 + vc**2 + 1)
            a = np.array([uc / wc, vc / wc, 1 / wc])
            uf, vf = ext['F'][idx]
            wf = np.sqrt(uf**2 + vf**2 + 1)
            b = np.array([uf / wf, vf / wf, 1 / wf])
            cos = np.clip(np.dot(a, b) / (np.linalg.norm(a) * np.linalg.norm(b)), -1.0, 1.0)
            if np.degrees(np.arccos(cos)) > self.alpha:
                validate.append(idx)
        for idx in validate:
            Xh = cv2.triangulatePoints(self.K @ ext['T'][idx][0:3, :], self.K @ new_pose[0:3, :], ext['F'][idx].T, ext['C'][idx].T)
            X3 = (Xh[0:3].T / Xh[3].T).T
            state['P'] = np.concatenate((state['P'], ext['C'][idx].reshape(1, 2)), axis=0)
            state['X'] = np.concatenate((state['X'], X3.reshape(1, 3)), axis=0)
        if len(validate) > 0:
            ext['C'] = np.delete(ext['C'], validate, axis=0)
            ext['F'] = np.delete(ext['F'], validate, axis=0)
            ext['T'] = np.delete(ext['T'], validate, axis=0)
        self.info['n_promoted'] = len(validate)
        return state, ext

    def frame(self, old_frame, new_frame, state, ext, new_candidates):
        new_state, _ = self.associate(old_frame, new_frame, state)
        T = self.estimate(new_state)
        new_state, ext = self.triangulate(old_frame, new_frame, new_state, ext, T, new_candidates)
        return new_state, ext, T


def append_landmarks(state, P, X, target_total=0):
    """np.vstack of externally supplied landmarks onto state P/X (vo_pipeline.py:318-320); target_total > 0:
    only as many rows as bring |P| up to it - the rule of vo_append_landmarks, restated."""
    P0 = np.asarray(state['P'], np.float64).reshape(-1, 2)
    X0 = np.asarray(state['X'], np.float64).reshape(-1, 3)
    n = len(P) if target_total <= 0 else int(np.clip(target_total - len(P0), 0, len(P)))
    return {'P': np.vstack((P0, np.asarray(P, np.float64).reshape(-1, 2)[:n])),
            'X': np.vstack((X0, np.asarray(X, np.float64).reshape(-1, 3)[:n]))}


def empty_ext():
    return {'C': np.zeros((0, 2)), 'F': np.zeros((0, 2)), 'T': np.zeros((0, 4, 4))}
