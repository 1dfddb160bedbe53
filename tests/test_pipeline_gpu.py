"""GPU tests (B200 box) of the whole per-frame path: the drop-in vo_pipeline classes driven in
main.py's order (main.py:229-294), against golden runs of the reference's own classes
(tests/golden/seq.npz) and, when cv2 is importable, the CPU port run live on the same frames.
Gates: frame-by-frame poses close to the P3P-variant reference run, trajectory ATE not worse than the
literal reference's by more than 1 % (north_star), state sizes within a few rows (threshold-straddling
tracks), fused per-frame call bit-identical to the staged calls."""
import os

import numpy as np
import pytest

from test_pipeline_oracle import GOLD, SEQ_CASES, SHI_SEQ_CASES, get_sequence, get_shi_sequence, run_port, run_port_shi
from conftest import parity_report
from test_pnp_oracle import rot_angle

pytestmark = pytest.mark.gpu


def drive_dropin(vo, sq, read_prev_state=True, want_counts=False):
    """main.py's loop body with the candidate feed in place of the detector."""
    vo.set_candidate_feed([f.copy() for f in sq.feeds[1:]])
    vision = vo.BestVision(sq.K)
    vision.state = {"P": sq.P0.astype(np.float64), "X": sq.X0.copy()}
    vision.extended_state = {"C": np.array([]), "F": np.array([]), "T": np.array([])}
    associate = vo.KeypointsToLandmarksAssociator(sq.K, np.eye(4))
    pose_estimator = vo.PoseEstimator(sq.K)
    Ts, nP, nC, nTrk, nInl = [], [], [], [], []
    img1 = sq.frames[0]
    for k in range(1, len(sq.frames)):
        img2 = sq.frames[k]
        new_state, lost = associate.associateKeypointsToLandmarks(img1, img2, vision.state)
        new_pose = pose_estimator.estimatePose(new_state, k)
        c = vo._runtime().ctx.counters(sync=False)                  # as of the pose read-back
        nTrk.append(c["n_tracked"]); nInl.append(c["n_inliers"])
        pos = np.linalg.inv(new_pose)[0:3, 3]                       # main.py:239
        if read_prev_state:
            X = vision.state["X"]                                   # main.py:248: the PREVIOUS state
            assert X.ndim == 2 and X.shape[1] == 3
            _ = X[(np.linalg.norm(X - pos, axis=1) > 3)]
        tri = vo.LandmarkTriangulator(sq.K, None)                   # main.py:280
        new_state, ext, cur_des = tri.triangulate_landmark(img1, img2, new_state, vision.extended_state, new_pose)
        vision.state, vision.extended_state = new_state, ext
        img1 = img2
        Ts.append(new_pose)
        vo._runtime().flush_pending_add()      # the append is deferred to the next frame's upload
        cnt = vo._runtime().ctx.counters()
        nP.append(cnt["n_landmarks"]); nC.append(cnt["n_candidates"])
    if want_counts:
        return np.array(Ts), np.array(nP), np.array(nC), vision, np.array(nTrk), np.array(nInl)
    return np.array(Ts), np.array(nP), np.array(nC), vision


@pytest.mark.parametrize("name", list(SEQ_CASES))
def test_dropin_sequence_vs_reference_classes(dropin, scene_mod, name):
    g = np.load(os.path.join(GOLD, "seq.npz"))
    sq = get_sequence(scene_mod, name)
    same_frames = sq.checksum() == int(g[name + "_checksum"])
    Ts, nP, nC, vision, nTrk, nInl = drive_dropin(dropin, sq, want_counts=True)
    gt = sq.poses[1:]
    ate_ours = scene_mod.ate(Ts, gt)
    margin = None
    if same_frames:
        Tg, nPg, nCg = g[name + "_p3p_T"], g[name + "_p3p_nP"], g[name + "_p3p_nC"]
        nTrkg, nInlg, margin = g[name + "_p3p_nTrk"], g[name + "_p3p_nInl"], g[name + "_p3p_margin"]
        Tl = g[name + "_literal_T"]
    else:   # frames rendered differently here: compare with the port run live instead
        pytest.importorskip("cv2")
        Tg, nPg, nCg, _, _ = run_port(sq, "p3p")
        Tl = run_port(sq, "literal")[0]
        nTrkg = nInlg = None
    ate_p3p, ate_lit = scene_mod.ate(Tg, gt), scene_mod.ate(Tl, gt)
    path_len = sq.scene.step * (len(sq.frames) - 1)
    dR = max(rot_angle(a[:3, :3], b[:3, :3]) for a, b in zip(Ts, Tg))
    dt = max(np.linalg.norm(a[:3, 3] - b[:3, 3]) for a, b in zip(Ts, Tg))
    print("%s: ATE ours %.4f, reference P3P-variant %.4f, literal %.4f (m over %.1f m); max dR %.2e rad, max dt %.2e m; "
          "max |dnP| %d, max |dnC| %d" % (name, ate_ours, ate_p3p, ate_lit, path_len, dR, dt,
                                          np.abs(nP - nPg).max(), np.abs(nC - nCg).max()))
    # State sizes are identical until the first track that sits on a decision threshold flips (the minimal sets
    # are drawn as rng % n, so one row more or less re-deals every later RANSAC).  The split must have such a cause:
    same = (nP == nPg) & (nC == nCg)
    if nTrkg is not None:
        same &= (nTrk == nTrkg) & (nInl == nInlg)
    first_diff = int(np.argmin(same)) if not same.all() else len(same)
    cause = None
    if first_diff < len(same) and nTrkg is not None:
        f = first_diff
        if nTrk[f] != nTrkg[f]:       # a KLT status flag (north_star allows 0.5 % of them to differ from cv2)
            cause = "KLT status flag: %d vs %d tracked" % (nTrk[f], nTrkg[f])
            assert abs(int(nTrk[f]) - int(nTrkg[f])) <= max(1, int(0.005 * nTrkg[f]))
        elif nInl[f] != nInlg[f]:     # an inlier decision: some reprojection error must sit at the 2 px threshold
            cause = "inlier decision, closest reprojection error %.2e px from the threshold" % margin[f]
            # (the margin is that of the GOLDEN's winning hypothesis; ours comes from minimal-set points that differ by
            # up to 1e-4 px, which an ill-conditioned P3P sample amplifies - hence 0.02 px, not 1e-4)
            assert margin[f] <= 2e-2, "inlier counts split at frame %d without an error near the threshold (%.3g px)" % (f, margin[f])
            assert abs(int(nInl[f]) - int(nInlg[f])) <= 3
        else:                         # the candidate side: a candidate's status flag or a bearing angle at alpha
            cause = "candidate status flag / bearing angle at the threshold: |dnC| %d, |dnP| %d" % (
                abs(int(nC[f]) - int(nCg[f])), abs(int(nP[f]) - int(nPg[f])))
            assert abs(int(nC[f]) - int(nCg[f])) <= max(2, int(0.005 * nCg[f])) and abs(int(nP[f]) - int(nPg[f])) <= 2
    print("   state sizes identical for the first %d of %d frames; split caused by: %s" % (first_diff, len(same), cause))
    assert np.abs(nP - nPg).max() <= 0.15 * nPg.max()
    assert np.abs(nC - nCg).max() <= 0.05 * nCg.max()
    # while the two runs see identical inputs their poses agree to the north_star PnP tolerance
    # (1e-4 rad / 1e-4 relative translation); afterwards they are two RANSAC realisations of the same
    # trajectory and only the ATE gate applies
    m = max(first_diff, 1)
    dR0 = max(rot_angle(a[:3, :3], b[:3, :3]) for a, b in zip(Ts[:m], Tg[:m]))
    dt0 = max(np.linalg.norm(a[:3, 3] - b[:3, 3]) / np.linalg.norm(b[:3, 3]) for a, b in zip(Ts[:m], Tg[:m]))
    print("   over the identical-input prefix: max dR %.2e rad, max dt_rel %.2e" % (dR0, dt0))
    assert dR0 <= 1e-4 and dt0 <= 1e-4
    # north_star: trajectory ATE within 1 % of the reference's.  Against the same-solver oracle as long as the run IS
    # that oracle's run (identical masks throughout, or over the identical prefix); once a threshold flip has re-dealt
    # the minimal sets the run is another RANSAC realisation, and the reference's two realisations (P3P-variant and
    # literal, e.g. 0.095 m vs 0.246 m on malaga_full) bound what 1 % can mean
    if m >= 3:
        ate_pre_ours, ate_pre_ref = scene_mod.ate(Ts[:m], gt[:m]), scene_mod.ate(Tg[:m], gt[:m])
        assert ate_pre_ours <= 1.01 * ate_pre_ref + 1e-4
    if first_diff == len(same):
        assert ate_ours <= 1.01 * ate_p3p + 1e-4
    else:
        assert ate_ours <= 1.01 * max(ate_lit, ate_p3p) + 1e-3
    parity_report("sequence", name, dict(frames=len(same), ate_ours_m=ate_ours, ate_reference_p3p_m=ate_p3p, ate_reference_literal_m=ate_lit,
                                         path_m=path_len, identical_prefix_frames=first_diff, split_cause=cause,
                                         prefix_max_dR_rad=dR0, prefix_max_dt_rel=dt0, all_max_dR_rad=dR, all_max_dt_m=dt,
                                         max_abs_dnP=int(np.abs(nP - nPg).max()), max_abs_dnC=int(np.abs(nC - nCg).max()),
                                         golden_cv2=str(g["cv2_version"]), same_frames_as_golden=bool(same_frames)))
    # the lazily mirrored state is consistent
    P, X = vision.state["P"], vision.state["X"]
    assert len(P) == len(X) == nP[-1] and P.dtype == np.float32 and X.dtype == np.float64
    C, F, T = (vision.extended_state[k] for k in "CFT")
    assert len(C) == len(F) == len(T) == nC[-1] and T.shape[1:] == (4, 4)


@pytest.mark.parametrize("name", list(SHI_SEQ_CASES))
@pytest.mark.parametrize("gpu_detector", [True, False])
def test_dropin_shi_mask_sequence_vs_reference_classes(dropin_shi, scene_mod, name, gpu_detector):
    """config 'shi-mask' end to end: candidates come from the Shi-Tomasi detector every frame -
    on the device (detect + mask + append without a host round trip, csrc/det.cu) or, with
    gpu_detector False, from cv2 on the CPU as before - against the reference's own classes run in
    that mode (golden).  The detector is bit-exact, so both must give the same candidate sets."""
    vo = dropin_shi
    vo.config["gpu_detector"] = gpu_detector
    if not gpu_detector:
        pytest.importorskip("cv2")
    g = np.load(os.path.join(GOLD, "seq.npz"))
    sq = get_shi_sequence(scene_mod, name)
    Ts, nP, nC, vision = drive_dropin(vo, sq)
    if sq.checksum() == int(g[name + "_checksum"]):
        Tg, nPg, nCg, Cg = g[name + "_p3p_T"], g[name + "_p3p_nP"], g[name + "_p3p_nC"], g[name + "_p3p_C_last"]
        Tl = g[name + "_literal_T"]
    else:
        pytest.importorskip("cv2")
        Tg, nPg, nCg, _, e = run_port_shi(sq, "p3p")
        Cg = np.asarray(e["C"], np.float64)
        Tl = run_port_shi(sq, "literal")[0]
    gt = sq.poses[1:]
    same = (nP == nPg) & (nC == nCg)
    first_diff = int(np.argmin(same)) if not same.all() else len(same)
    print("%s gpu_detector=%s: state sizes identical for the first %d of %d frames; nC %s vs %s" % (
        name, gpu_detector, first_diff, len(same), nC[-3:], nCg[-3:]))
    assert first_diff >= len(same) // 2
    assert np.abs(nP - nPg).max() <= 0.15 * nPg.max() and np.abs(nC - nCg).max() <= 0.05 * nCg.max()
    m = max(first_diff, 1)
    dR0 = max(rot_angle(a[:3, :3], b[:3, :3]) for a, b in zip(Ts[:m], Tg[:m]))
    dt0 = max(np.linalg.norm(a[:3, 3] - b[:3, 3]) / np.linalg.norm(b[:3, 3]) for a, b in zip(Ts[:m], Tg[:m]))
    assert dR0 <= 1e-4 and dt0 <= 1e-4
    assert scene_mod.ate(Ts, gt) <= 1.01 * max(scene_mod.ate(Tl, gt), scene_mod.ate(Tg, gt)) + 1e-3
    if first_diff == len(same):      # identical inputs all the way: the candidate sets are the same rows
        C = np.asarray(vision.extended_state["C"], np.float64)
        assert C.shape == Cg.shape and np.abs(C - Cg).max() <= 0.05


def test_fused_frame_equals_staged_calls(pkg, scene_mod):
    """vo_process_frame (one sync per frame) == upload + klt_track_landmarks + pnp_ransac +
    track_triangulate_promote + add_candidates, bit for bit."""
    sq = get_sequence(scene_mod, "kitti_half")
    K = sq.K
    pnp = pkg.pnp_defaults()
    a = pkg.Context(device=0, max_w=2048, max_h=1024)
    b = pkg.Context(device=0, max_w=2048, max_h=1024)
    try:
        for c in (a, b):
            c.upload_frame(sq.frames[0], 1)
            c.set_state(sq.P0, sq.X0)
            c.set_candidates(np.zeros((0, 2)), np.zeros((0, 2)), np.zeros((0, 16)))
        for k in range(1, 12):
            Ta, cnt = a.process_frame(sq.frames[k], k, k + 1, K, 1.0, sq.feeds[k], pnp=pnp)
            b.upload_frame(sq.frames[k], k + 1)
            b.klt_track_landmarks(k, k + 1)
            ok, Tb, ninl = b.pnp_ransac(K, pnp)
            b.track_triangulate_promote(k, k + 1, K, Tb, 1.0)
            b.add_candidates(sq.feeds[k], Tb)
            assert ok and cnt["pnp_ok"] == 1 and ninl == cnt["n_inliers"]
            assert np.array_equal(Ta, Tb)
            cb = b.counters()
            for key in ("n_landmarks", "n_candidates", "n_promoted", "n_candidates_lost", "cf_alias"):
                assert cnt[key] == cb[key], key
        Pa, Xa = a.get_state(); Pb, Xb = b.get_state()
        assert np.array_equal(Pa, Pb) and np.array_equal(Xa, Xb)
        for u, v in zip(a.get_candidates(), b.get_candidates()):
            assert np.array_equal(u, v)
    finally:
        a.close(); b.close()


@pytest.mark.parametrize("early", [False, True])
def test_speculative_promotion_commit_and_rollback(pkg, scene_mod, early):
    """early: the host waits for the pose only (n_promoted == NULL), the promotion is still running when
    the call returns and the peeked counters are those before it.
    vo_pnp_ransac_promote (+commit) == vo_pnp_ransac + vo_track_triangulate_promote, bit for bit; a
    rollback restores the state exactly as vo_pnp_ransac leaves it, and the staged promotion can then
    still run on the pending candidate tracking."""
    sq = get_sequence(scene_mod, "kitti_half")
    K = sq.K
    pnp = pkg.pnp_defaults()
    a = pkg.Context(device=0, max_w=2048, max_h=1024)   # staged
    b = pkg.Context(device=0, max_w=2048, max_h=1024)   # speculative, alternating commit / rollback
    try:
        for c in (a, b):
            c.upload_frame(sq.frames[0], 1)
            c.set_state(sq.P0, sq.X0)
            c.set_candidates(np.zeros((0, 2)), np.zeros((0, 2)), np.zeros((0, 16)))
        for k in range(1, 14):
            for c in (a, b):
                c.upload_frame(sq.frames[k], k + 1)
                c.klt_track_landmarks(k, k + 1)
            ok, Ta, ninl = a.pnp_ransac(K, pnp)
            Pa0, Xa0 = a.get_state()                     # pruned to the inliers, nothing promoted yet
            Ca0 = a.get_candidates()
            assert b.candidates_tracked(k, k + 1)
            okb, Tb, ninlb, nprom = b.pnp_ransac_promote(K, pnp, 1.0, early=early)
            assert ok and okb and ninl == ninlb and np.array_equal(Ta, Tb) and b.promote_pending()
            if early:
                assert nprom is None
                pk = b.counters(sync=False)              # as of the pose read-back: before the promotion
                # landmarks pruned, promoted rows not appended yet; candidates untouched (plain tracking call) or,
                # had the tracking call classified them early, final counts
                assert pk["n_landmarks"] == ninl
                assert pk["n_candidates"] + pk["n_promoted"] + pk["n_candidates_lost"] == len(Ca0[0])
                nprom = None
            if k % 2 == 0:
                b.promote_rollback()
                assert not b.promote_pending()
                Pb0, Xb0 = b.get_state()
                assert np.array_equal(Pa0, Pb0) and np.array_equal(Xa0, Xb0)
                for u, v in zip(Ca0, b.get_candidates()):
                    assert np.array_equal(u, v)
                b.track_triangulate_promote(k, k + 1, K, Tb, 1.0)
            else:
                Pb, Xb = b.get_state()                   # promoted rows sit behind the inlier rows
                assert np.array_equal(Pb[:ninl], Pa0) and np.array_equal(Xb[:ninl], Xa0)
                assert len(Pb) == ninl + (b.counters()["n_promoted"] if nprom is None else nprom)
                b.promote_commit()
            a.track_triangulate_promote(k, k + 1, K, Ta, 1.0)
            for c in (a, b):
                c.add_candidates(sq.feeds[k], Ta)
            ca, cb = a.counters(), b.counters()
            for key in ("n_landmarks", "n_candidates", "n_promoted", "n_candidates_lost", "cf_alias"):
                assert ca[key] == cb[key], (k, key)
            Pa, Xa = a.get_state(); Pb, Xb = b.get_state()
            assert np.array_equal(Pa, Pb) and np.array_equal(Xa, Xb)
            for u, v in zip(a.get_candidates(), b.get_candidates()):
                assert np.array_equal(u, v)
            # the state this frame started from is still readable (main.py:248)
            Pp, Xp = b.get_prev_state()
            assert len(Pp) == prev_rows if k > 1 else len(Pp) == len(sq.P0)
            assert np.array_equal(Xp, prev_X)  if k > 1 else np.array_equal(Xp, sq.X0)
            prev_rows, prev_X = len(Pb), Xb
    finally:
        a.close(); b.close()


def test_dropin_speculation_is_transparent(dropin, scene_mod):
    """The drop-in classes give the same poses and states with the speculative promotion on and off,
    and the previous frame's state (main.py:248) is served with the right content either way."""
    sq = get_sequence(scene_mod, "kitti_half")
    runs = []
    for spec in (True, False):
        dropin.config['speculative_promotion'] = spec
        dropin._reset_runtime()
        dropin.set_candidate_feed([f.copy() for f in sq.feeds[1:]])
        vision = dropin.BestVision(sq.K)
        vision.state = {"P": sq.P0.astype(np.float64), "X": sq.X0.copy()}
        vision.extended_state = {"C": np.array([]), "F": np.array([]), "T": np.array([])}
        associate = dropin.KeypointsToLandmarksAssociator(sq.K, np.eye(4))
        pose_estimator = dropin.PoseEstimator(sq.K)
        Ts, prevX, endX = [], [], []
        img1 = sq.frames[0]
        for k in range(1, 16):
            img2 = sq.frames[k]
            new_state, _ = associate.associateKeypointsToLandmarks(img1, img2, vision.state)
            new_pose = pose_estimator.estimatePose(new_state, k)
            prevX.append(np.array(vision.state["X"]))            # previous state, read late
            if k % 5 == 0:
                mid = np.array(new_state["X"])                       # inlier rows only, even while a speculation is pending
                assert len(mid) == pose_estimator.last['n_inliers']
            tri = dropin.LandmarkTriangulator(sq.K, None)
            vision.state, vision.extended_state, _ = tri.triangulate_landmark(img1, img2, new_state, vision.extended_state, new_pose)
            if k % 3 == 0:
                endX.append(np.array(vision.state["X"]))         # current state, read early
            img1 = img2
            Ts.append(new_pose)
        C, F, T = (np.array(vision.extended_state[key]) for key in "CFT")
        runs.append((np.array(Ts), prevX, endX, np.array(vision.state["P"]), C, F, T))
    dropin.config.pop('speculative_promotion')
    on, off = runs
    assert np.array_equal(on[0], off[0])
    for u, v in zip(on[1] + on[2], off[1] + off[2]):
        assert np.array_equal(u, v)
    for i in range(3, 7):
        assert np.array_equal(on[i], off[i])
    # content of the late read: frame k's previous state == frame k-1's end state
    assert np.array_equal(on[1][3], on[2][0]) and np.array_equal(on[1][6], on[2][1])
    assert np.array_equal(on[1][0], sq.X0)


def test_state_roundtrip_and_edge_cases(ctx):
    rng = np.random.default_rng(0)
    P = rng.uniform(0, 500, (777, 2)).astype(np.float32)
    X = rng.normal(0, 10, (777, 3))
    ctx.set_state(P, X)
    P2, X2 = ctx.get_state()
    assert np.array_equal(P, P2) and np.array_equal(X, X2)
    ctx.set_state(P[:0], X[:0])                               # empty state
    assert ctx.n_state() == 0
    ok, T, ninl = ctx.pnp_ransac(np.eye(3))                   # nothing to estimate from: ok = False
    assert not ok and ninl == 0
    C = rng.uniform(0, 300, (5, 2)).astype(np.float32)
    F = C.astype(np.float64) + 1.0
    T5 = np.repeat(np.eye(4)[None], 5, 0)
    ctx.set_candidates(C, F, T5)
    C2, F2, T2 = ctx.get_candidates()
    assert np.array_equal(C, C2) and np.array_equal(F, F2) and np.array_equal(T2, T5)
    # capacity growth: more candidates than the initial capacity, appended in several batches
    big = rng.uniform(0, 1000, (9000, 2))
    ctx.set_candidates(C[:0], F[:0], T5[:0])
    for i in range(3):
        ctx.add_candidates(big[3000 * i:3000 * (i + 1)], np.eye(4) * (i + 1))
    Cn, Fn, Tn = ctx.get_candidates()
    assert len(Cn) == 9000 and np.array_equal(Fn, big) and np.array_equal(Cn, big.astype(np.float32))
    assert Tn[0, 0, 0] == 1 and Tn[3000, 0, 0] == 2 and Tn[8999, 0, 0] == 3


def test_cf_alias_quirk(ctx, scene_mod):
    """App. B-5: a candidate batch that arrives while C is empty has F bound to the same array, so
    one frame later F equals the tracked position; later batches keep their first observation."""
    sq = get_sequence(scene_mod, "kitti_half")
    K = sq.K
    ctx.upload_frame(sq.frames[0], 7000)
    ctx.upload_frame(sq.frames[1], 7001)
    ctx.upload_frame(sq.frames[2], 7002)
    ctx.set_state(sq.P0, sq.X0)
    ctx.set_candidates(np.zeros((0, 2)), np.zeros((0, 2)), np.zeros((0, 16)))
    ctx.track_triangulate_promote(7000, 7001, K, sq.poses[1], 1.0)      # nothing to track yet
    ctx.add_candidates(sq.feeds[1], sq.poses[1])                         # first batch: C is F
    assert ctx.counters()["cf_alias"] == 1
    ctx.track_triangulate_promote(7001, 7002, K, sq.poses[2], 1e9)       # huge alpha: no promotion
    C, F, T = ctx.get_candidates()
    assert len(C) > 0 and np.array_equal(F, C.astype(np.float64))        # F followed C
    assert not np.array_equal(F, sq.feeds[1][:len(F)])
    ctx.add_candidates(sq.feeds[2], sq.poses[2])                         # second batch: concatenate, no alias
    assert ctx.counters()["cf_alias"] == 0


def test_frame_ingest_prefetch_is_bit_identical(pkg, scene_mod):
    """frames handed over one step ahead on the ingest stream (vo_prefetch_frame) give the
    same poses, landmarks and candidates, bit for bit, as in-step uploads - from pinned, pageable and
    strided sources, across slot recycling, and when a prefetched frame is never used."""
    sq = get_sequence(scene_mod, "kitti_half")
    K = sq.K
    pnp = pkg.pnp_defaults()
    h, w = sq.frames[0].shape
    a = pkg.Context(device=0, max_w=2048, max_h=1024)   # in-step uploads
    b = pkg.Context(device=0, max_w=2048, max_h=1024)   # ingest one frame ahead
    n = 14
    pinned = [pkg.PinnedArray((h, w), np.uint8) for _ in range(n + 2)]
    wide = np.zeros((h, w + 24), np.uint8)               # strided pageable source
    try:
        for p, f in zip(pinned, sq.frames):
            p.array[...] = f
        for c in (a, b):
            c.upload_frame(sq.frames[0], 1)
            c.set_state(sq.P0, sq.X0)
            c.set_candidates(np.zeros((0, 2)), np.zeros((0, 2)), np.zeros((0, 16)))
        b.prefetch_frame(pinned[1].array, 2)
        for k in range(1, n):
            # next frame to the ingest stream BEFORE this frame's stages: pinned / pageable / strided in turn
            if k % 3 == 0:
                src = pinned[k + 1].array
            elif k % 3 == 1:
                src = sq.frames[k + 1]
            else:
                wide[:, :w] = sq.frames[k + 1]
                src = wide[:, :w]
            b.prefetch_frame(src, k + 2)
            if k == 5:
                b.prefetch_frame(sq.frames[0], 1000)     # a hint that is never consumed
            assert b.has_frame(k + 1)
            Ta, ca = a.process_frame(sq.frames[k], k, k + 1, K, 1.0, sq.feeds[k], pnp=pnp)
            Tb, cb = b.process_frame(None, k, k + 1, K, 1.0, sq.feeds[k], pnp=pnp)
            assert ca["pnp_ok"] == 1 and np.array_equal(Ta, Tb), k
            for key in ("n_landmarks", "n_candidates", "n_promoted", "n_inliers", "klt_iters"):
                assert ca[key] == cb[key], (k, key)
        Pa, Xa = a.get_state(); Pb, Xb = b.get_state()
        assert np.array_equal(Pa, Pb) and np.array_equal(Xa, Xb)
        for u, v in zip(a.get_candidates(), b.get_candidates()):
            assert np.array_equal(u, v)
        # the pyramid an ingest leaves behind is the one an upload builds
        a.upload_frame(sq.frames[3], 77)
        b.prefetch_frame(sq.frames[3], 77)
        for lv in range(4):
            ia, da = a.pyramid_level(77, lv)
            ib, db = b.pyramid_level(77, lv)
            assert np.array_equal(ia, ib) and np.array_equal(da, db)
    finally:
        a.close(); b.close()
        for p in pinned:
            p.free()


def test_dropin_prefetch_hint_and_frame_ingest(dropin, scene_mod):
    """vo_pipeline.prefetch_frame / FrameIngest: the drop-in classes driven with frames handed over one
    iteration early return exactly the poses of the plain drive."""
    sq = get_sequence(scene_mod, "kitti_half")
    Ts0, nP0, nC0, _ = drive_dropin(dropin, sq, read_prev_state=False)
    dropin._reset_runtime()

    vo = dropin
    vo.set_candidate_feed([f.copy() for f in sq.feeds[1:]])
    vision = vo.BestVision(sq.K)
    vision.state = {"P": sq.P0.astype(np.float64), "X": sq.X0.copy()}
    vision.extended_state = {"C": np.array([]), "F": np.array([]), "T": np.array([])}
    associate = vo.KeypointsToLandmarksAssociator(sq.K, np.eye(4))
    pose_estimator = vo.PoseEstimator(sq.K)
    Ts = []
    img1 = None
    ingest = vo.FrameIngest(sq.frames, depth=4)
    assert len(ingest) == len(sq.frames)
    for k, img2 in enumerate(ingest):
        assert np.array_equal(img2, sq.frames[k])
        if img1 is None:
            img1 = img2
            continue
        new_state, lost = associate.associateKeypointsToLandmarks(img1, img2, vision.state)
        new_pose = pose_estimator.estimatePose(new_state, k)
        tri = vo.LandmarkTriangulator(sq.K, None)
        vision.state, vision.extended_state, _ = tri.triangulate_landmark(img1, img2, new_state, vision.extended_state, new_pose)
        img1 = img2
        Ts.append(new_pose)
    assert np.array_equal(np.array(Ts), Ts0)
    # every frame but the first two came in through the ingest stream: no key was uploaded twice
    assert vo._runtime().next_key == len(sq.frames) + 1


def test_track_with_folded_append_equals_add_then_track(pkg, scene_mod):
    """vo_klt_track_landmarks_add (the previous frame's new candidates appended on the side stream, in
    front of the candidates' KLT) == vo_add_candidates followed by vo_klt_track_landmarks, bit for bit,
    including the C/F aliasing quirk of the first batch and state reads while the append is in flight."""
    sq = get_sequence(scene_mod, "kitti_half")
    K = sq.K
    pnp = pkg.pnp_defaults()
    a = pkg.Context(device=0, max_w=2048, max_h=1024)
    b = pkg.Context(device=0, max_w=2048, max_h=1024)
    try:
        for c in (a, b):
            c.upload_frame(sq.frames[0], 1)
            c.set_state(sq.P0, sq.X0)
            c.set_candidates(np.zeros((0, 2)), np.zeros((0, 2)), np.zeros((0, 16)))
        pend = None
        for k in range(1, 12):
            for c in (a, b):
                c.upload_frame(sq.frames[k], k + 1)
            if pend is not None:
                a.add_candidates(*pend)
            a.klt_track_landmarks(k, k + 1)
            b.klt_track_landmarks(k, k + 1, add=pend)
            if k % 4 == 0:      # a synchronising read while the side-stream append may still be running
                assert a.counters()["n_candidates"] == b.counters()["n_candidates"]
            oka, Ta, na = a.pnp_ransac(K, pnp)
            if k % 2:
                okb, Tb, nb, _ = b.pnp_ransac_promote(K, pnp, 1.0, early=True)
                b.promote_commit()
            else:
                okb, Tb, nb = b.pnp_ransac(K, pnp)
                b.track_triangulate_promote(k, k + 1, K, Tb, 1.0)
            a.track_triangulate_promote(k, k + 1, K, Ta, 1.0)
            assert oka and okb and na == nb and np.array_equal(Ta, Tb), k
            pend = (sq.feeds[k].copy(), Ta.copy()) if len(sq.feeds[k]) else None
            ca, cb = a.counters(), b.counters()
            for key in ("n_landmarks", "n_candidates", "n_promoted", "n_candidates_lost", "cf_alias"):
                assert ca[key] == cb[key], (k, key)
        Pa, Xa = a.get_state(); Pb, Xb = b.get_state()
        assert np.array_equal(Pa, Pb) and np.array_equal(Xa, Xb)
        for u, v in zip(a.get_candidates(), b.get_candidates()):
            assert np.array_equal(u, v)
    finally:
        a.close(); b.close()


def test_track_rollback_restores_the_tracking_input(pkg, scene_mod):
    """vo_track_rollback: after tracking into a wrong frame and taking it back, the run continues bit for
    bit like one that never did (landmarks, candidates incl. a folded append, poses)."""
    sq = get_sequence(scene_mod, "kitti_half")
    K = sq.K
    pnp = pkg.pnp_defaults()
    a = pkg.Context(device=0, max_w=2048, max_h=1024)
    b = pkg.Context(device=0, max_w=2048, max_h=1024)
    try:
        for c in (a, b):
            c.upload_frame(sq.frames[0], 1)
            c.set_state(sq.P0, sq.X0)
            c.set_candidates(np.zeros((0, 2)), np.zeros((0, 2)), np.zeros((0, 16)))
        pend = None
        for k in range(1, 10):
            for c in (a, b):
                c.upload_frame(sq.frames[k], k + 1)
            a.klt_track_landmarks(k, k + 1, add=pend)
            if k % 2:
                b.upload_frame(sq.frames[k + 2], 500 + k)
                b.klt_track_landmarks(k, 500 + k, add=pend)       # wrong target frame
                Pw, Xw = b.get_state()                            # (synchronises: the wrong step really ran)
                b.track_rollback()
                P0, X0 = b.get_state()
                assert len(P0) >= len(Pw)
                b.klt_track_landmarks(k, k + 1)                   # the append stays: not repeated
            else:
                b.klt_track_landmarks(k, k + 1, add=pend)
            oka, Ta, na = a.pnp_ransac(K, pnp)
            okb, Tb, nb = b.pnp_ransac(K, pnp)
            assert oka and okb and na == nb and np.array_equal(Ta, Tb), k
            for c in (a, b):
                c.track_triangulate_promote(k, k + 1, K, Ta, 1.0)
            pend = (sq.feeds[k].copy(), Ta.copy()) if len(sq.feeds[k]) else None
            Pa, Xa = a.get_state(); Pb, Xb = b.get_state()
            assert np.array_equal(Pa, Pb) and np.array_equal(Xa, Xb)
            for u, v in zip(a.get_candidates(), b.get_candidates()):
                assert np.array_equal(u, v)
        with pytest.raises(pkg.VoError):
            b.track_rollback()                                    # nothing to undo after a PnP
    finally:
        a.close(); b.close()


def test_dropin_speculative_tracking_hit_and_miss(dropin, scene_mod):
    """Frames hinted one iteration ahead: triangulate_landmark enqueues the next tracking step itself.
    Right hints (consumed) and wrong hints (rolled back) both reproduce the plain drive's poses exactly,
    and the state dict handed back in between still reads as the pre-tracking landmarks."""
    sq = get_sequence(scene_mod, "kitti_half")
    Ts0, nP0, nC0, _ = drive_dropin(dropin, sq, read_prev_state=False)
    for wrong_every in (0, 3):
        dropin._reset_runtime()
        vo = dropin
        vo.set_candidate_feed([f.copy() for f in sq.feeds[1:]])
        vision = vo.BestVision(sq.K)
        vision.state = {"P": sq.P0.astype(np.float64), "X": sq.X0.copy()}
        vision.extended_state = {"C": np.array([]), "F": np.array([]), "T": np.array([])}
        associate = vo.KeypointsToLandmarksAssociator(sq.K, np.eye(4))
        pose_estimator = vo.PoseEstimator(sq.K)
        Ts, hits = [], 0
        img1 = sq.frames[0]
        n = len(sq.frames)
        for k in range(1, n):
            img2 = sq.frames[k]
            nxt = k + 1 if not (wrong_every and k % wrong_every == 0) else k + 2
            if nxt < n:
                vo.prefetch_frame(sq.frames[nxt])
            hits += vo._runtime().spec_track is not None
            new_state, lost = associate.associateKeypointsToLandmarks(img1, img2, vision.state)
            new_pose = pose_estimator.estimatePose(new_state, k)
            tri = vo.LandmarkTriangulator(sq.K, None)
            vision.state, vision.extended_state, _ = tri.triangulate_landmark(img1, img2, new_state, vision.extended_state, new_pose)
            if k == 4:      # read the returned state while the next tracking step is already enqueued
                assert vo._runtime().spec_track is not None
                n_now = len(vision.state["X"])
                assert n_now == len(vision.state["P"]) and n_now >= vo._runtime().ctx.counters()["n_landmarks"]
            img1 = img2
            Ts.append(new_pose)
        assert np.array_equal(np.array(Ts), Ts0), wrong_every
        assert hits >= n - 4


def test_pipelined_frames_equal_the_synchronous_call(pkg, scene_mod):
    """vo_process_frame_pipelined (pose-only wait, next frame's tracking enqueued behind the promotion when
    that frame was handed over early) gives the poses and the final state of vo_process_frame bit for
    bit - with hits (next call matches), a miss (taken back) and steps without a frame ahead."""
    sq = get_sequence(scene_mod, "kitti_half")
    K = sq.K
    pnp = pkg.pnp_defaults()
    a = pkg.Context(device=0, max_w=2048, max_h=1024)
    b = pkg.Context(device=0, max_w=2048, max_h=1024)
    n = 14
    try:
        for c in (a, b):
            c.upload_frame(sq.frames[0], 1)
            c.set_state(sq.P0, sq.X0)
            c.set_candidates(np.zeros((0, 2)), np.zeros((0, 2)), np.zeros((0, 16)))
        for k in range(1, n):
            Ta, ca = a.process_frame(sq.frames[k], k, k + 1, K, 1.0, sq.feeds[k], pnp=pnp)
            if k % 5 == 3:
                b.prefetch_frame(sq.frames[k + 2], 900 + k)        # a frame that is NOT the next one: miss
            elif k % 5 != 4 and k + 1 < len(sq.frames):
                b.prefetch_frame(sq.frames[k + 1], k + 2)          # the next frame, one step ahead
            Tb, cb = b.process_frame(sq.frames[k], k, k + 1, K, 1.0, sq.feeds[k], pnp=pnp, pipelined=True)
            assert ca["pnp_ok"] == 1 and cb["pnp_ok"] == 1 and np.array_equal(Ta, Tb), k
            assert ca["n_inliers"] == cb["n_inliers"]
            assert cb["n_landmarks"] == cb["n_inliers"]            # as of the pose: pruned, promoted rows not appended yet
            if k > 1:   # the candidates were classified on the side stream during the tracking: final counts
                assert cb["n_promoted"] == ca["n_promoted"] and cb["n_candidates_lost"] == ca["n_candidates_lost"]
                assert cb["n_candidates"] + len(sq.feeds[k]) == ca["n_candidates"]
        # leave the pipelined mode: one synchronous call on both, then compare everything
        Ta, ca = a.process_frame(sq.frames[n], n, n + 1, K, 1.0, sq.feeds[n], pnp=pnp)
        Tb, cb = b.process_frame(sq.frames[n], n, n + 1, K, 1.0, sq.feeds[n], pnp=pnp)
        assert np.array_equal(Ta, Tb)
        for key in ("n_landmarks", "n_candidates", "n_promoted", "n_candidates_lost", "n_inliers", "cf_alias"):
            assert ca[key] == cb[key], key
        Pa, Xa = a.get_state(); Pb, Xb = b.get_state()
        assert np.array_equal(Pa, Pb) and np.array_equal(Xa, Xb)
        for u, v in zip(a.get_candidates(), b.get_candidates()):
            assert np.array_equal(u, v)
    finally:
        a.close(); b.close()


def test_frame_ingest_from_files(dropin, scene_mod, tmp_path):
    """FrameIngest over file paths (cv2.imread on the worker thread, as main.py:217-223 reads them): frames
    arrive bit-identical, in order, in pinned buffers, and a missing file surfaces in the consumer."""
    cv2 = pytest.importorskip("cv2")
    sq = get_sequence(scene_mod, "kitti_half")
    paths = []
    for k in range(6):
        p = str(tmp_path / ("%06d.png" % k))
        cv2.imwrite(p, sq.frames[k])
        paths.append(p)
    got = []
    for img in dropin.FrameIngest(paths, depth=3):
        assert img.dtype == np.uint8 and img.flags["C_CONTIGUOUS"]
        got.append(img.copy())
    assert len(got) == 6 and all(np.array_equal(a, b) for a, b in zip(got, sq.frames))
    with pytest.raises(FileNotFoundError):
        for _ in dropin.FrameIngest(paths[:2] + [str(tmp_path / "missing.png")]):
            pass
    with pytest.raises(ValueError):
        dropin.FrameIngest(paths, depth=2)


def test_early_classification_paths(pkg, scene_mod):
    """vo_klt_track_landmarks_ex: the candidate side of the promotion classified on the side stream during
    the tracking.  Finished by a promotion with the same alpha, redone with another alpha, taken back by a
    look at the candidates, and without a pose (fewer than 4 landmarks): always the state the plain
    path reaches, bit for bit."""
    sq = get_sequence(scene_mod, "kitti_half")
    K = sq.K
    pnp = pkg.pnp_defaults()
    a = pkg.Context(device=0, max_w=2048, max_h=1024)   # plain
    b = pkg.Context(device=0, max_w=2048, max_h=1024)   # early classification
    try:
        for c in (a, b):
            c.upload_frame(sq.frames[0], 1)
            c.set_state(sq.P0, sq.X0)
            c.set_candidates(np.zeros((0, 2)), np.zeros((0, 2)), np.zeros((0, 16)))
        pend = None
        for k in range(1, 13):
            for c in (a, b):
                c.upload_frame(sq.frames[k], k + 1)
            a.klt_track_landmarks(k, k + 1, add=pend)
            b.klt_track_landmarks(k, k + 1, add=pend, early_alpha=1.0)
            alpha = 1.0 if k % 3 else 0.4                         # every third frame: another threshold
            if k % 4 == 2:
                Ca, Cb = a.get_candidates(), b.get_candidates()   # looks at the candidates: taken back
                assert all(np.array_equal(u, v) for u, v in zip(Ca, Cb))
            oka, Ta, na = a.pnp_ransac(K, pnp)
            a.track_triangulate_promote(k, k + 1, K, Ta, alpha)
            if k % 2:
                okb, Tb, nb, _ = b.pnp_ransac_promote(K, pnp, alpha, early=True)
                b.promote_commit()
            else:
                okb, Tb, nb = b.pnp_ransac(K, pnp)
                b.track_triangulate_promote(k, k + 1, K, Tb, alpha)
            assert oka and okb and na == nb and np.array_equal(Ta, Tb), k
            pend = (sq.feeds[k].copy(), Ta.copy()) if len(sq.feeds[k]) else None
            ca, cb = a.counters(), b.counters()
            for key in ("n_landmarks", "n_candidates", "n_promoted", "n_candidates_lost", "cf_alias"):
                assert ca[key] == cb[key], (k, key)
            Pa, Xa = a.get_state(); Pb, Xb = b.get_state()
            assert np.array_equal(Pa, Pb) and np.array_equal(Xa, Xb), k
            for u, v in zip(a.get_candidates(), b.get_candidates()):
                assert np.array_equal(u, v)
        # no pose: three landmarks only
        for c in (a, b):
            P, X = c.get_state()
            c.set_state(P[:3], X[:3])
        k = 13
        Ta, ca = None, None
        a.upload_frame(sq.frames[k], k + 1)
        a.klt_track_landmarks(k, k + 1, add=pend)
        oka, _, _ = a.pnp_ransac(K, pnp)
        assert not oka
        Tb, cb = b.process_frame(sq.frames[k], k, k + 1, K, 1.0, None, pnp=pnp)   # (tracks with the early half)
        assert cb["pnp_ok"] == 0 and cb["n_promoted"] == 0
        ca = a.counters()
        assert ca["n_landmarks"] == cb["n_landmarks"]
        Pa, Xa = a.get_state(); Pb, Xb = b.get_state()
        assert np.array_equal(Pa, Pb) and np.array_equal(Xa, Xb)
    finally:
        a.close(); b.close()


def test_pipelined_frames_with_changing_pnp_arguments(pkg, scene_mod):
    """The next frame's PnP is enqueued ahead with the CURRENT call's camera matrix and parameters.  A next call
    that asks for other ones (iteration budget, threshold, a slightly different K) gets its own run: the run
    enqueued ahead is taken back (its pruning wrote into another buffer), poses and states stay those of the
    synchronous path with the same schedule; state reads in between see the rows the tracking step kept."""
    sq = get_sequence(scene_mod, "kitti_half")
    a = pkg.Context(device=0, max_w=2048, max_h=1024)
    b = pkg.Context(device=0, max_w=2048, max_h=1024)
    n = 13

    def args(k):
        p = pkg.pnp_defaults()
        K = sq.K.copy()
        if k % 4 == 2:
            p.max_iters = 300
        if k % 4 == 3:
            p.reproj_thresh = 1.5
        if k % 6 == 5:
            K[0, 2] += 0.25
        return K, p

    try:
        for c in (a, b):
            c.upload_frame(sq.frames[0], 1)
            c.set_state(sq.P0, sq.X0)
            c.set_candidates(np.zeros((0, 2)), np.zeros((0, 2)), np.zeros((0, 16)))
        for k in range(1, n):
            K, p = args(k)
            Ta, ca = a.process_frame(sq.frames[k], k, k + 1, K, 1.0, sq.feeds[k], pnp=p)
            if k + 1 < len(sq.frames):
                b.prefetch_frame(sq.frames[k + 1], k + 2)
            Tb, cb = b.process_frame(sq.frames[k], k, k + 1, K, 1.0, sq.feeds[k], pnp=p, pipelined=True)
            assert ca["pnp_ok"] == 1 and cb["pnp_ok"] == 1 and np.array_equal(Ta, Tb), k
            assert (ca["n_inliers"], ca["ransac_iters"]) == (cb["n_inliers"], cb["ransac_iters"]), k
            if k == 7:   # a state read while the step ahead (tracking + PnP) is in flight: it is taken back, then redone
                Pa, Xa = a.get_state(); Pb, Xb = b.get_state()
                assert np.array_equal(Pa, Pb) and np.array_equal(Xa, Xb)
        K, p = args(n)
        Ta, ca = a.process_frame(sq.frames[n], n, n + 1, K, 1.0, sq.feeds[n], pnp=p)
        Tb, cb = b.process_frame(sq.frames[n], n, n + 1, K, 1.0, sq.feeds[n], pnp=p)
        assert np.array_equal(Ta, Tb)
        Pa, Xa = a.get_state(); Pb, Xb = b.get_state()
        assert np.array_equal(Pa, Pb) and np.array_equal(Xa, Xb)
        for u, v in zip(a.get_candidates(), b.get_candidates()):
            assert np.array_equal(u, v)
    finally:
        a.close(); b.close()
