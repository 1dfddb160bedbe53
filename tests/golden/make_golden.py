"""Generates tests/golden/*.npz in the BUILD container, where cv2 and /root/reference exist.

  python tests/golden/make_golden.py

The goldens pin the oracle (and through it the CUDA path) to outputs of the real thing:
cv2 (version recorded in each file) called exactly as the reference calls it, and the reference's
own classes imported from /root/reference/src (matplotlib stubbed; it is not installed).
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import cv2  # noqa: E402
import synth  # noqa: E402


def klt():
    from test_klt_oracle import KLT_CASES
    for name, (h, w, seed, n) in KLT_CASES.items():
        prev, nxt, pts = synth.klt_pair(h, w, seed, n)
        # exactly the reference's call: vo_pipeline.py:215 / :501
        next_pts, status, err = cv2.calcOpticalFlowPyrLK(prev, nxt, pts, None)
        np.savez_compressed(os.path.join(HERE, "klt_%s.npz" % name), pts=pts, next_pts=next_pts,
                            status=status, err=err, prev_sum=prev.astype(np.int64).sum(),
                            next_sum=nxt.astype(np.int64).sum(), cv2_version=cv2.__version__)
    img = synth.texture(75, 101, 11)
    lv, out = img, {}
    for l in range(1, 4):
        lv = cv2.pyrDown(lv)
        out["l%d" % l] = lv
    np.savez_compressed(os.path.join(HERE, "pyrdown.npz"), cv2_version=cv2.__version__, **out)


PNP_CASES = {  # name: (n, outlier_frac, H, seed) - SURVEY.md 8 d1
    "n500_o20_h100": (500, 0.2, 100, 1),
    "n2000_o20_h100": (2000, 0.2, 100, 2),
    "n2000_o20_h2000": (2000, 0.2, 2000, 3),
    "n2000_o50_h2000": (2000, 0.5, 2000, 4),
    "n10000_o50_h4096": (10000, 0.5, 4096, 5),
    "n300_o60_h2000": (300, 0.6, 2000, 6),
    "n50_o30_h100": (50, 0.3, 100, 7),
}


def pnp():
    """Primary PnP gate: masks of cv2.solvePnPRansac(flags=SOLVEPNP_P3P, iterationsCount=H)
    - its RANSAC loop with cv::RNG(2^64-1) minimal sets - and the pose of cv2.solvePnP(ITERATIVE)
    on that inlier set (= the refit the reference's literal call performs, vo_pipeline.py:258-264).
    Report-only: the literal reference call (EPnP-5 hypotheses, <=100 iterations)."""
    out = {"cv2_version": cv2.__version__}
    for name, (n, of, H, seed) in PNP_CASES.items():
        X, P, R, t, _ = synth.pnp_scene(n, of, seed)
        ok, rv, tv, inl = cv2.solvePnPRansac(X, P, synth.K_KITTI, np.zeros(4), flags=cv2.SOLVEPNP_P3P,
                                             confidence=0.99999, reprojectionError=2, iterationsCount=H)
        assert ok
        mask = np.zeros(n, np.uint8)
        mask[inl.ravel()] = 1
        Xi = X[mask == 1].astype(np.float32).astype(np.float64)
        ok2, rv2, tv2 = cv2.solvePnP(Xi, P[mask == 1].astype(np.float64), synth.K_KITTI, np.zeros(4),
                                     flags=cv2.SOLVEPNP_ITERATIVE)
        assert ok2
        okl, rvl, tvl, inll = cv2.solvePnPRansac(X, P, synth.K_KITTI, np.zeros(4), flags=cv2.SOLVEPNP_ITERATIVE,
                                                 confidence=0.99999, reprojectionError=2)
        ml = np.zeros(n, np.uint8)
        ml[inll.ravel()] = 1
        out[name + "_mask"] = np.packbits(mask)
        out[name + "_rvec"] = rv2.ravel()
        out[name + "_tvec"] = tv2.ravel()
        out[name + "_xsum"] = X.sum()
        out[name + "_literal_mask"] = np.packbits(ml)
        out[name + "_literal_rvec"] = rvl.ravel()
        out[name + "_literal_tvec"] = tvl.ravel()
    np.savez_compressed(os.path.join(HERE, "pnp.npz"), **out)


TRI_CASES = {"n500": (500, 1), "n2000": (2000, 2), "n3000": (3000, 3)}


def tri():
    """cv2.triangulatePoints on two-view scenes (dehomogenised: the sign of the homogeneous vector is
    arbitrary), called batched as at vo_pipeline.py:172 and one point per call as at :666; plus the
    reference's bearing-angle loop (vo_pipeline.py:621-642) evaluated with numpy exactly as written."""
    out = {"cv2_version": cv2.__version__}
    K = synth.K_KITTI
    for name, (n, seed) in TRI_CASES.items():
        X, T1, T2, p1, p2 = synth.two_view_scene(n, seed)
        Xh = cv2.triangulatePoints(K @ T1[:3], K @ T2[:3], p1.T.copy(), p2.T.copy()).T
        out[name + "_X"] = Xh[:, :3] / Xh[:, 3:]
        one = np.stack([cv2.triangulatePoints(K @ T1[0:3, :], K @ T2[0:3, :], p1[i].T, p2[i].T) for i in range(16)])
        assert np.allclose(one[:, :3, 0] / one[:, 3:, 0], out[name + "_X"][:16], rtol=1e-9)
    C, F = synth.bearing_pairs(4000, 9)
    al = []
    for idx in range(len(C)):
        uc, vc = C[idx]
        wc = np.sqrt(uc**2 + vc**2 + 1)
        vector_a = np.array([uc / wc, vc / wc, 1 / wc])
        uf, vf = F[idx]
        wf = np.sqrt(uf**2 + vf**2 + 1)
        vector_b = np.array([uf / wf, vf / wf, 1 / wf])
        cos = np.dot(vector_a, vector_b) / (np.linalg.norm(vector_a) * np.linalg.norm(vector_b))
        al.append(np.degrees(np.arccos(np.clip(cos, -1.0, 1.0))))
    out["bearing_alpha"] = np.array(al)
    np.savez_compressed(os.path.join(HERE, "tri.npz"), **out)


from make_golden_cases import GFTT_CASES, gftt_image, reference_mask  # noqa: E402


def gftt():
    """cv2.goodFeaturesToTrack exactly as the reference calls it (vo_pipeline.py:381), with the
    reference's mask, plus cv2.cornerMinEigenVal (the map behind it) as a checksum + samples."""
    import hashlib
    out = {"cv2_version": cv2.__version__}
    for name in GFTT_CASES:
        img, pts = gftt_image(name)
        mask = reference_mask(img, pts)
        c = cv2.goodFeaturesToTrack(img, maxCorners=500, qualityLevel=0.03, minDistance=10, mask=mask, blockSize=4,
                                    gradientSize=3, useHarrisDetector=False, k=0.04)
        out[name + "_corners"] = np.zeros((0, 2), np.float32) if c is None else c.reshape(-1, 2)
        c = cv2.goodFeaturesToTrack(img, maxCorners=0, qualityLevel=0.01, minDistance=5, blockSize=4)
        out[name + "_all_corners"] = c.reshape(-1, 2).astype(np.uint16)
        eig = cv2.cornerMinEigenVal(img, 4, ksize=3)
        out[name + "_eig_sha1"] = np.frombuffer(hashlib.sha1(eig.tobytes()).digest(), np.uint8)
        out[name + "_eig_rows"] = eig[:: max(1, img.shape[0] // 6)]
        out[name + "_img_sum"] = img.astype(np.int64).sum()
    np.savez_compressed(os.path.join(HERE, "gftt.npz"), **out)


SEQ_CASES = {  # name: (shape, scale, n_frames, n_tracks, n_feed, seed)
    "kitti_half": ("kitti", 0.5, 36, 600, 80, 3),
    "parking_half": ("parking", 0.5, 24, 300, 60, 4),
    "malaga_half": ("malaga", 0.5, 20, 900, 100, 5),
    # full resolution, the BASELINE configs' sizes
    "kitti_full": ("kitti", 1.0, 61, 2000, 150, 7),
    "malaga_full": ("malaga", 1.0, 31, 3000, 200, 8),
}


SHI_SEQ_CASES = {  # name: (shape, scale, n_frames, n_tracks, seed)
    "kitti_half_shi": ("kitti", 0.5, 14, 300, 6),
}


def _import_reference(workdir, method="sift-sift"):
    """Import the reference's own src/vo_pipeline.py (config.yaml in CWD, matplotlib stubbed)."""
    import importlib
    for name in ("matplotlib", "matplotlib.pyplot"):
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.__getattr__ = lambda k: (lambda *a, **kw: None)
            sys.modules[name] = m
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    os.makedirs(workdir, exist_ok=True)
    with open(os.path.join(workdir, "config.yaml"), "w") as f:
        f.write("dataset: 'kitti'\nfind_new_candidates_method: '%s'\ninit_detector: 'sift'\n"
                "init_descriptor: 'sift'\ndebug: False\nvisualization: False\nanimation: False\n"
                "sift_alpha: 1\nsift_ransac_confidence: 0.99999\nshi_alpha: 1.5\nshi_ransac_confidence: 0.999993\n" % method)
    cwd = os.getcwd()
    os.chdir(workdir)
    sys.path.insert(0, "/root/reference/src")
    try:
        sys.modules.pop("vo_pipeline", None)
        ref = importlib.import_module("vo_pipeline")
    finally:
        sys.path.remove("/root/reference/src")
        os.chdir(cwd)
    sys.modules.pop("vo_pipeline", None)
    return ref


def seq():
    """The reference's OWN classes (imported from /root/reference/src) driven in main.py's order
    (main.py:229-284) over synthetic sequences; the SIFT detector is replaced by the pre-generated
    candidate feed (detectors are outside the hot path).  Two PnP variants are recorded:
      literal - the reference as is (EPnP-5 hypotheses, <=100 iterations): the trajectory/ATE gate;
      p3p     - cv2.solvePnPRansac patched to flags=SOLVEPNP_P3P + ITERATIVE refit on the inliers:
                the primary parity oracle of SURVEY.md 8c (frame-by-frame state sizes and poses)."""
    import contextlib
    import importlib
    import io
    scene_mod = importlib.import_module("locally-consistent-visual-odometry-pipeline_b200.scene")
    ref = _import_reference("/tmp/ref_golden_cwd")
    real_ransac = cv2.solvePnPRansac

    def p3p_ransac(X, P, K, dist, flags=None, confidence=0.99, reprojectionError=8.0, **kw):
        ok, rv, tv, inl = real_ransac(X, P, K, dist, flags=cv2.SOLVEPNP_P3P, confidence=confidence,
                                      reprojectionError=reprojectionError, iterationsCount=100)
        if inl is not None and len(inl) >= 4:
            idx = np.hstack(inl)
            Xi = np.asarray(X, np.float64)[idx].astype(np.float32).astype(np.float64)
            _, rv, tv = cv2.solvePnP(Xi, np.asarray(P, np.float64)[idx], K, dist, flags=cv2.SOLVEPNP_ITERATIVE)
        return ok, rv, tv, inl

    out = {"cv2_version": cv2.__version__}
    for name, (shape, scale, nf, nt, nfeed, seed) in SEQ_CASES.items():
        sq = scene_mod.Sequence(shape, n_frames=nf, n_tracks=nt, n_feed=nfeed, seed=seed, scale=scale)
        out[name + "_checksum"] = sq.checksum()
        for variant in ("literal", "p3p"):
            ref.cv2.solvePnPRansac = p3p_ransac if variant == "p3p" else real_ransac
            feeds = [f.copy() for f in sq.feeds]
            ref.LandmarkTriangulator.find_new_candidates_sift_sift = lambda self, fr, st, ex: (feeds.pop(0), [])
            state = {"P": sq.P0.astype(np.float64), "X": sq.X0.copy()}
            ext = {"C": np.array([]), "F": np.array([]), "T": np.array([])}
            assoc = ref.KeypointsToLandmarksAssociator(sq.K, np.eye(4))
            pe = ref.PoseEstimator(sq.K)
            feeds.pop(0)
            Ts, nP, nC, nInl, nTrk, margin = [], [], [], [], [], []
            with contextlib.redirect_stdout(io.StringIO()):
                for k in range(1, nf):
                    new_state, _ = assoc.associateKeypointsToLandmarks(sq.frames[k - 1], sq.frames[k], state)
                    nTrk.append(len(new_state["P"]))
                    if variant == "p3p":
                        # how close does a reprojection error of the winning hypothesis come to the 2 px threshold?
                        # (the C restatement gives the hypothesis; cv2 only returns the refit) - what decides whether
                        # another implementation's 1e-4 px differences can flip an inlier decision on this frame
                        import oracle
                        Pi, Xi = np.asarray(new_state["P"], np.float32), np.asarray(new_state["X"], np.float64)
                        o = oracle.solve_pnp_ransac(Xi, Pi, sq.K, max_iters=100)
                        Xc = Xi.astype(np.float32).astype(np.float64) @ o["Rh"].T + o["th"]
                        uv = np.stack([sq.K[0, 0] * Xc[:, 0] / Xc[:, 2] + sq.K[0, 2], sq.K[1, 1] * Xc[:, 1] / Xc[:, 2] + sq.K[1, 2]], 1)
                        e = np.sqrt(((uv - Pi) ** 2).sum(1))
                        margin.append((float(np.abs(e - 2.0).min()), int(o["n_inliers"])))
                    T = pe.estimatePose(new_state, k)
                    nInl.append(len(new_state["P"]))
                    if variant == "p3p":
                        assert margin[-1][1] == nInl[-1], "C restatement and cv2 disagree on the inlier count"

                    tri = ref.LandmarkTriangulator(sq.K, None)
                    state, ext, _ = tri.triangulate_landmark(sq.frames[k - 1], sq.frames[k], new_state, ext, T)
                    Ts.append(T); nP.append(len(state["P"])); nC.append(len(ext["C"]))
            out["%s_%s_T" % (name, variant)] = np.array(Ts)
            out["%s_%s_nP" % (name, variant)] = np.array(nP)
            out["%s_%s_nC" % (name, variant)] = np.array(nC)
            out["%s_%s_nInl" % (name, variant)] = np.array(nInl)
            out["%s_%s_nTrk" % (name, variant)] = np.array(nTrk)
            if variant == "p3p":
                out[name + "_p3p_margin"] = np.array([m[0] for m in margin])
            if variant == "p3p":
                out[name + "_p3p_P_last"] = np.asarray(state["P"], np.float64)
                out[name + "_p3p_X_last"] = np.asarray(state["X"], np.float64)
                out[name + "_p3p_C_last"] = np.asarray(ext["C"], np.float64)
            gt = sq.poses[1:]
            print(name, variant, "ATE %.4f m over %.1f m" % (scene_mod.ate(Ts, gt), sq.scene.step * (nf - 1)),
                  "nP", nP[0], nP[-1], "nC", nC[0], nC[-1])
        ref.cv2.solvePnPRansac = real_ransac
    # config find_new_candidates_method 'shi-mask': the reference's own Shi-Tomasi detector, mask,
    # alpha (shi_alpha) and RANSAC confidence (shi_ransac_confidence) - SURVEY.md 8f-1
    ref.config['find_new_candidates_method'] = 'shi-mask'
    for name, (shape, scale, nf, nt, seed) in SHI_SEQ_CASES.items():
        sq = scene_mod.Sequence(shape, n_frames=nf, n_tracks=nt, n_feed=1, seed=seed, scale=scale)
        out[name + "_checksum"] = sq.checksum()
        for variant in ("literal", "p3p"):
            ref.cv2.solvePnPRansac = p3p_ransac if variant == "p3p" else real_ransac
            state = {"P": sq.P0.astype(np.float64), "X": sq.X0.copy()}
            ext = {"C": np.array([]), "F": np.array([]), "T": np.array([])}
            assoc = ref.KeypointsToLandmarksAssociator(sq.K, np.eye(4))
            pe = ref.PoseEstimator(sq.K)
            Ts, nP, nC = [], [], []
            with contextlib.redirect_stdout(io.StringIO()):
                for k in range(1, nf):
                    new_state, _ = assoc.associateKeypointsToLandmarks(sq.frames[k - 1], sq.frames[k], state)
                    T = pe.estimatePose(new_state, k)
                    tri = ref.LandmarkTriangulator(sq.K, None)
                    state, ext, _ = tri.triangulate_landmark(sq.frames[k - 1], sq.frames[k], new_state, ext, T)
                    Ts.append(T); nP.append(len(state["P"])); nC.append(len(ext["C"]))
            out["%s_%s_T" % (name, variant)] = np.array(Ts)
            out["%s_%s_nP" % (name, variant)] = np.array(nP)
            out["%s_%s_nC" % (name, variant)] = np.array(nC)
            if variant == "p3p":
                out[name + "_p3p_C_last"] = np.asarray(ext["C"], np.float64)
            print(name, variant, "ATE %.4f m over %.1f m" % (scene_mod.ate(Ts, sq.poses[1:]), sq.scene.step * (nf - 1)),
                  "nP", nP[0], nP[-1], "nC", nC[0], nC[-1])
        ref.cv2.solvePnPRansac = real_ransac
    ref.config['find_new_candidates_method'] = 'sift-sift'
    np.savez_compressed(os.path.join(HERE, "seq.npz"), **out)


BOOT_CASES = {  # name: (n, outlier_frac, seed, K name) - SURVEY.md 8f-3
    "n300_o20": (300, 0.2, 2, "kitti"), "n800_o30": (800, 0.3, 3, "kitti"), "n1500_o45": (1500, 0.45, 4, "kitti"),
    "n200_o0": (200, 0.0, 5, "kitti"), "n60_o10": (60, 0.1, 6, "parking"), "n2000_o20": (2000, 0.2, 7, "malaga"),
    "n500_o50": (500, 0.5, 8, "kitti"), "n5_exact": (5, 0.0, 9, "kitti"),
}


def bootstrap():
    """The bootstrap's two cv2 calls exactly as the reference makes them (vo_pipeline.py:139-162):
    findEssentialMat(k1, k2, K, RANSAC, 0.999, 1.0, None) and recoverPose(E, k1, k2, K, mask=mask), whose
    mask argument is ANDed with the cheirality test in place."""
    Ks = {"kitti": synth.K_KITTI, "parking": synth.K_PARKING, "malaga": synth.K_MALAGA}
    out = {"cv2_version": cv2.__version__}
    for name, (n, of, seed, kn) in BOOT_CASES.items():
        K = Ks[kn]
        p1, p2, R, t, _ = synth.bootstrap_scene(n, of, seed, K=K, noise=0.0 if n == 5 else 0.3)
        E, mask = cv2.findEssentialMat(p1, p2, K, cv2.RANSAC, 0.999, 1.0, None)
        out[name + "_psum"] = float(p1.sum() + p2.sum())
        out[name + "_E_all"] = E                          # n == 5: every model cv2 found, stacked
        E0 = E[:3]
        out[name + "_E"] = E0
        out[name + "_mask_E"] = np.packbits(mask.ravel() != 0)
        m2 = mask.copy()
        good, Rc, tc, m2o = cv2.recoverPose(E0, p1, p2, K, mask=m2)
        assert np.array_equal(m2, m2o)                    # modified in place, as the reference relies on
        out[name + "_good"] = good
        out[name + "_R"] = Rc
        out[name + "_t"] = tc.ravel()
        out[name + "_mask_pose"] = np.packbits(m2o.ravel() != 0)
        print(name, "inliers", int((mask != 0).sum()), "after recoverPose", good)
    np.savez_compressed(os.path.join(HERE, "bootstrap.npz"), **out)


if __name__ == "__main__":
    which = sys.argv[1:] or ["klt", "pnp", "tri", "gftt", "seq", "bootstrap"]
    for w in which:
        globals()[w]()
    print("golden vectors written with cv2", cv2.__version__)
