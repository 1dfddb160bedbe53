"""GPU tests of the round-2 additions to the per-frame path:

  - vo_append_landmarks / vo_queue_landmarks (reference vo_pipeline.py:318-320: np.vstack of externally
    triangulated landmarks onto P/X), incl. the device-side top-up rule, against numpy and the CPU port;
  - vo_process_frame_pipelined when no pose comes out of the PnP while the next frame's step is already
    enqueued: state identical to the synchronous call, bit for bit;
  - the forward-backward status check (BASELINE.json north_star) at unit level and inside the tracking
    step, against two oracle LK calls.
"""
import numpy as np
import pytest

from test_pipeline_oracle import get_sequence

pytestmark = pytest.mark.gpu


def _fresh(pkg, sq):
    c = pkg.Context(device=0, max_w=2048, max_h=1024)
    c.upload_frame(sq.frames[0], 1)
    c.set_state(sq.P0, sq.X0)
    c.set_candidates(np.zeros((0, 2)), np.zeros((0, 2)), np.zeros((0, 16)))
    return c


def test_append_landmarks_is_vstack(pkg, scene_mod):
    sq = get_sequence(scene_mod, "kitti_half")
    rng = np.random.default_rng(3)
    c = _fresh(pkg, sq)
    try:
        n0 = len(sq.P0)
        P1 = rng.uniform(0, 300, (37, 2)).astype(np.float32)
        X1 = rng.normal(size=(37, 3))
        c.append_landmarks(P1, X1)                      # all rows
        P, X = c.get_state()
        assert np.array_equal(P, np.vstack([sq.P0, P1])) and np.array_equal(X, np.vstack([sq.X0, X1]))
        assert c.counters()["n_appended"] == 37
        # top-up: only what is missing up to the target, decided on the device
        P2 = rng.uniform(0, 300, (50, 2)).astype(np.float32)
        X2 = rng.normal(size=(50, 3))
        c.append_landmarks(P2, X2, target_total=n0 + 37 + 12)
        P, X = c.get_state()
        assert len(P) == n0 + 49 and np.array_equal(P[-12:], P2[:12]) and np.array_equal(X[-12:], X2[:12])
        assert c.counters()["n_appended"] == 12
        c.append_landmarks(P2, X2, target_total=10)     # already above the target: nothing
        assert c.n_state() == n0 + 49 and c.counters()["n_appended"] == 0
        # many small pageable batches in a row (the staging ring is reused safely), then a large one (growth)
        ref_P, ref_X = [P], [X]
        for k in range(9):
            p = rng.uniform(0, 300, (5 + k, 2)).astype(np.float32); x = rng.normal(size=(5 + k, 3))
            c.append_landmarks(p, x)
            ref_P.append(p); ref_X.append(x)
        p = rng.uniform(0, 300, (9000, 2)).astype(np.float32); x = rng.normal(size=(9000, 3))
        c.append_landmarks(p, x)
        ref_P.append(p); ref_X.append(x)
        P, X = c.get_state()
        assert np.array_equal(P, np.vstack(ref_P)) and np.array_equal(X, np.vstack(ref_X))
        # the appended rows are tracked and used like any others: a tracking step right after
        c.set_state(sq.P0, sq.X0)
        c.append_landmarks(sq.P0[:40] + 0.0, sq.X0[:40])
        c.upload_frame(sq.frames[1], 2)
        n_kept = c.klt_track_landmarks(1, 2, sync=True)
        assert n_kept >= n0 + 30
        with pytest.raises(pkg.VoError):                # between a tracking step and its PnP: refused
            c.append_landmarks(P1, X1)
    finally:
        c.close()


def _run_with_batches(pkg, sq, pipelined, n, target, burst_at, rng_seed=5):
    """n frames with a candidate feed, a top-up batch every frame and an outlier batch on some frames."""
    rng = np.random.default_rng(rng_seed)
    c = _fresh(pkg, sq)
    K, pnp = sq.K, pkg.pnp_defaults()
    Ts, cnts = [], []
    try:
        for k in range(1, n):
            topP = rng.uniform(20, 300, (60, 2)).astype(np.float32)      # (not real landmarks: PnP throws them out again)
            topX = rng.normal(size=(60, 3)) + [0, 0, 20]
            c.append_landmarks(topP, topX, target_total=target, queue=True)
            if k in burst_at:
                c.append_landmarks(topP[:25] + 1.0, topX[:25] + 3.0, queue=True)
            if pipelined and k + 1 < len(sq.frames):
                c.prefetch_frame(sq.frames[k + 1], k + 2)
            T, cnt = c.process_frame(sq.frames[k], k, k + 1, K, 1.0, sq.feeds[k], pnp=pnp, pipelined=pipelined)
            assert cnt["pnp_ok"] == 1
            Ts.append(T); cnts.append(cnt)
        T, cnt = c.process_frame(sq.frames[n], n, n + 1, K, 1.0, sq.feeds[n], pnp=pnp)   # synchronous: final counters
        Ts.append(T); cnts.append(cnt)
        return Ts, cnts, c.get_state(), c.get_candidates()
    finally:
        c.close()


def test_queued_batches_sync_equals_pipelined_and_follow_the_rule(pkg, scene_mod):
    sq = get_sequence(scene_mod, "kitti_half")
    n0 = len(sq.P0)
    n, target, burst_at = 12, n0 + 5, (3, 7)
    Ta, ca, Sa, Ca = _run_with_batches(pkg, sq, False, n, target, burst_at)
    Tb, cb, Sb, Cb = _run_with_batches(pkg, sq, True, n, target, burst_at)
    assert all(np.array_equal(x, y) for x, y in zip(Ta, Tb))
    assert np.array_equal(Sa[0], Sb[0]) and np.array_equal(Sa[1], Sb[1])
    assert all(np.array_equal(u, v) for u, v in zip(Ca, Cb))
    # synchronous leg: after every frame |P| = inliers + promoted + appended, with the top-up rule
    for k, c in enumerate(ca[:-1], start=1):
        before = c["n_inliers"] + c["n_promoted"]
        top = int(np.clip(target - before, 0, 60))
        extra = 25 if k in burst_at else 0
        assert c["n_landmarks"] == before + top + extra, k
        assert c["n_appended"] == (extra if extra else top)       # the LAST batch's count


def test_pipelined_call_without_a_pose_equals_the_synchronous_call(pkg, scene_mod):
    """the steady pipelined state (the next frame's tracking and early classification already
    enqueued) when the PnP finds no model: tracking results, candidate buffers and scalars of the failing frame
    must survive, and the state must be the synchronous call's."""
    sq = get_sequence(scene_mod, "kitti_half")
    K, pnp = sq.K, pkg.pnp_defaults()
    a, b = _fresh(pkg, sq), _fresh(pkg, sq)
    try:
        for k in range(1, 7):       # steady state first: candidates present, frames overlapped on b
            if k + 1 < len(sq.frames):
                b.prefetch_frame(sq.frames[k + 1], k + 2)
            Ta, ca = a.process_frame(sq.frames[k], k, k + 1, K, 1.0, sq.feeds[k], pnp=pnp)
            Tb, cb = b.process_frame(sq.frames[k], k, k + 1, K, 1.0, sq.feeds[k], pnp=pnp, pipelined=True)
            assert np.array_equal(Ta, Tb)
        # three landmarks only: no model for frame 7.  set_state drops the step b enqueued ahead; re-establish it
        # with one more good frame on copies of the state, then fail
        k = 7
        for c in (a, b):
            P, X = c.get_state()
            c.set_state(P[:3], X[:3])
        b.prefetch_frame(sq.frames[k + 1], k + 2)       # the frame AFTER the failing one is already there
        Ta, ca = a.process_frame(sq.frames[k], k, k + 1, K, 1.0, sq.feeds[k], pnp=pnp)
        Tb, cb = b.process_frame(sq.frames[k], k, k + 1, K, 1.0, sq.feeds[k], pnp=pnp, pipelined=True)
        assert ca["pnp_ok"] == 0 and cb["pnp_ok"] == 0
        for c in (a, b):
            cc = c.counters()
            assert cc["n_promoted"] == 0
        ca, cb = a.counters(), b.counters()
        for key in ("n_landmarks", "n_candidates", "n_candidates_lost", "cf_alias"):
            assert ca[key] == cb[key], key
        Pa, Xa = a.get_state(); Pb, Xb = b.get_state()
        assert np.array_equal(Pa, Pb) and np.array_equal(Xa, Xb)
        for u, v in zip(a.get_candidates(), b.get_candidates()):
            assert np.array_equal(u, v)
        # and the stream goes on identically once landmarks are back (the reference re-bootstraps here)
        for c in (a, b):
            c.set_state(sq.P0, sq.X0)
            c.upload_frame(sq.frames[0], 500)
        Ta, ca = a.process_frame(sq.frames[1], 500, 501, K, 1.0, None, pnp=pnp)
        Tb, cb = b.process_frame(sq.frames[1], 500, 501, K, 1.0, None, pnp=pnp, pipelined=True)
        assert ca["pnp_ok"] and cb["pnp_ok"] and np.array_equal(Ta, Tb)
    finally:
        a.close(); b.close()


def test_pipelined_failure_in_the_steady_state(pkg, scene_mod):
    """The same with the failure arriving in the middle of the overlapped stream: the frame's landmarks all
    sit on a stale map (3-D points of other pixels), so PnP finds no model although KLT keeps the tracks."""
    sq = get_sequence(scene_mod, "kitti_half")
    K, pnp = sq.K, pkg.pnp_defaults()
    pnp.max_iters = 30
    a, b = _fresh(pkg, sq), _fresh(pkg, sq)
    try:
        for k in range(1, 5):
            b.prefetch_frame(sq.frames[k + 1], k + 2)
            Ta, _ = a.process_frame(sq.frames[k], k, k + 1, K, 1.0, sq.feeds[k], pnp=pnp)
            Tb, _ = b.process_frame(sq.frames[k], k, k + 1, K, 1.0, sq.feeds[k], pnp=pnp, pipelined=True)
            assert np.array_equal(Ta, Tb)
        Pa, Xa = a.get_state()
        bad = np.roll(Xa, 7, axis=0) * np.array([1.0, -1.0, 0.3])        # nothing consistent left
        for c in (a, b):
            c.set_state(Pa, bad)
        for k in range(5, 8):
            b.prefetch_frame(sq.frames[k + 1], k + 2)
            Ta, ca = a.process_frame(sq.frames[k], k, k + 1, K, 1.0, sq.feeds[k], pnp=pnp)
            Tb, cb = b.process_frame(sq.frames[k], k, k + 1, K, 1.0, sq.feeds[k], pnp=pnp, pipelined=True)
            assert ca["pnp_ok"] == cb["pnp_ok"] and np.array_equal(Ta, Tb), k
        Ta, ca = a.process_frame(sq.frames[8], 8, 9, K, 1.0, sq.feeds[8], pnp=pnp)
        Tb, cb = b.process_frame(sq.frames[8], 8, 9, K, 1.0, sq.feeds[8], pnp=pnp)
        for key in ("n_landmarks", "n_candidates", "n_promoted", "n_candidates_lost", "n_inliers", "cf_alias", "pnp_ok"):
            assert ca[key] == cb[key], key
        Pa, Xa = a.get_state(); Pb, Xb = b.get_state()
        assert np.array_equal(Pa, Pb) and np.array_equal(Xa, Xb)
        for u, v in zip(a.get_candidates(), b.get_candidates()):
            assert np.array_equal(u, v)
    finally:
        a.close(); b.close()


# ---- forward-backward check ------------------------------------------------------------------------------------
def _oracle_fb(oracle, prev, nxt, pts, thr):
    out, st, err = oracle.calc_optical_flow_pyr_lk(prev, nxt, pts)
    back, bst, _ = oracle.calc_optical_flow_pyr_lk(nxt, prev, out)
    d2 = ((back - pts) ** 2).sum(1)
    keep = (st == 1) & (bst == 1) & (d2 <= np.float32(thr * thr))
    return out, st, keep.astype(np.uint8), np.sqrt(d2)


@pytest.mark.parametrize("thr", [0.5, 1.0])
def test_fb_check_unit_level_vs_two_oracle_calls(pkg, ctx, oracle, thr):
    import synth
    prev, nxt, pts = synth.klt_pair(240, 320, 21, 600, shift=(-7.5, 5.2))   # ~100 tracks fail the way back
    ctx.upload_frame(prev, 7301)     # (keys identify frames: 7001 / 7002 belong to test_pipeline_gpu)
    ctx.upload_frame(nxt, 7302)
    prm = pkg.klt_defaults()
    out0, st0, _ = ctx.calc_optical_flow_pyr_lk(7301, 7302, pts, prm)
    prm.fb_check, prm.fb_thresh = 1, thr
    out1, st1, _ = ctx.calc_optical_flow_pyr_lk(7301, 7302, pts, prm)
    o_out, o_st, o_keep, o_d = _oracle_fb(oracle, prev, nxt, pts, thr)
    assert np.array_equal(out0, out1)                       # the check only clears status flags
    assert (st0 == o_st).mean() >= 0.995
    # tracks whose round-trip distance is not within 0.05 px of the threshold must agree exactly
    clear = np.abs(o_d - thr) > 0.05
    both = (st0 == 1) & (o_st == 1) & clear
    assert np.array_equal(st1[both], o_keep[both])
    assert (st1 == o_keep).mean() >= 0.995
    assert st1.sum() < st0.sum()                            # the painted patch really costs tracks
    assert not np.any((st1 == 1) & (st0 == 0))


def test_fb_check_inside_the_tracking_step(pkg, scene_mod, oracle):
    """klt_fb_check on the pipeline path (landmarks on the main stream, candidates on the side stream): kept rows
    are the ones two oracle LK calls keep."""
    sq = get_sequence(scene_mod, "kitti_half")
    c = _fresh(pkg, sq)
    try:
        prm = pkg.klt_defaults()
        prm.fb_check, prm.fb_thresh = 1, 0.5
        cand = sq.feeds[1][:80]
        c.set_candidates(cand.astype(np.float32), cand, np.tile(np.eye(4).ravel(), (len(cand), 1)))
        c.upload_frame(sq.frames[1], 2)
        n_kept = c.klt_track_landmarks(1, 2, prm, sync=True)
        _, _, keep, d = _oracle_fb(oracle, sq.frames[0], sq.frames[1], sq.P0, 0.5)
        unsure = int((np.abs(d - 0.5) <= 0.05).sum())
        assert abs(n_kept - int(keep.sum())) <= unsure + max(2, len(sq.P0) // 200)
        P, X = c.get_state()
        exp = sq.X0[keep == 1]
        if len(X) == len(exp):
            assert np.array_equal(X, exp)
        prev_c, next_c, st_c = c.candidate_tracks()
        _, _, keep_c, d_c = _oracle_fb(oracle, sq.frames[0], sq.frames[1], cand.astype(np.float32), 0.5)
        clear = np.abs(d_c - 0.5) > 0.05
        assert (st_c[clear] == keep_c[clear]).mean() >= 0.99
        # without the option nothing of this happens
        c2 = _fresh(pkg, sq)
        try:
            c2.upload_frame(sq.frames[1], 2)
            assert c2.klt_track_landmarks(1, 2, sync=True) >= n_kept
        finally:
            c2.close()
    finally:
        c.close()


# ---- the class surface one frame ahead (config frame_ahead) ---------------------------------------------------------
def _drive(vo, sq, n, frame_ahead, read_states, pinned=None):
    """main.py's loop with candidate feed, landmark batches, a frame handed over early, and the reads main.py does."""
    vo.config['frame_ahead'] = frame_ahead
    vo._reset_runtime()
    rng = np.random.default_rng(9)
    feeds = [f.copy() for f in sq.feeds]
    batches = [[(rng.uniform(20, 300, (40, 2)).astype(np.float32), rng.normal(size=(40, 3)) + [0, 0, 20], len(sq.P0) + 3)]
               + ([(rng.uniform(20, 300, (15, 2)).astype(np.float32), rng.normal(size=(15, 3)) + [0, 0, 9], 0)] if k % 4 == 2 else [])
               for k in range(n + 2)]
    vision = vo.BestVision(sq.K)
    vision.state = {"P": sq.P0.astype(np.float64), "X": sq.X0.copy()}
    vision.extended_state = {"C": np.array([]), "F": np.array([]), "T": np.array([])}
    associate = vo.KeypointsToLandmarksAssociator(sq.K, np.eye(4))
    pose_estimator = vo.PoseEstimator(sq.K)
    out = dict(T=[], prevX=[], midX=[], endX=[], nP=[])
    frames = sq.frames if pinned is None else pinned
    img1 = frames[0]
    for k in range(1, n):
        img2 = frames[k]
        vo.set_candidate_feed([feeds[k]])
        vo.set_landmark_feed([batches[k]])
        if k + 1 < len(frames):
            vo.prefetch_frame(frames[k + 1])
        new_state, lost = associate.associateKeypointsToLandmarks(img1, img2, vision.state)
        if read_states and k % 3 == 1:
            out["nP"].append(len(new_state["P"]))                     # tracked rows, before the pose (debug print of main.py:232)
        new_pose = pose_estimator.estimatePose(new_state, k)
        out["T"].append(new_pose.copy())
        if read_states:
            out["prevX"].append(np.array(vision.state["X"]))           # main.py:248: the PREVIOUS frame's final state
            if k % 2 == 0:
                out["midX"].append(np.array(new_state["X"]))           # the inlier rows of this frame
        tri = vo.LandmarkTriangulator(sq.K, None)
        vision.state, vision.extended_state, _ = tri.triangulate_landmark(img1, img2, new_state, vision.extended_state, new_pose)
        if read_states and k % 5 == 0:
            out["endX"].append(np.array(vision.state["X"]))            # this frame's final state, read before the next associate
        img1 = img2
    out["final"] = (np.array(vision.state["P"]), np.array(vision.state["X"]))
    out["ext"] = tuple(np.array(vision.extended_state[k]) for k in "CFT")
    return out


@pytest.mark.parametrize("read_states", [False, True])
def test_dropin_frame_ahead_equals_the_transparent_path(dropin, pkg, scene_mod, read_states):
    """config frame_ahead: estimatePose enqueues the rest of the frame and the next frame's tracking step before it
    waits for the pose.  Same poses, same states (also the ones read late, main.py:248, and in between), bit for bit."""
    sq = get_sequence(scene_mod, "kitti_half")
    n = 14
    h, w = sq.frames[0].shape
    pinned = []
    for f in sq.frames[:n + 1]:
        pa = pkg.PinnedArray((h, w), np.uint8)
        pa.array[...] = f
        pinned.append(pa)
    views = [p.array for p in pinned]
    a = _drive(dropin, sq, n, False, read_states, views)
    b = _drive(dropin, sq, n, True, read_states, views)
    assert len(a["T"]) == len(b["T"]) == n - 1
    for k, (x, y) in enumerate(zip(a["T"], b["T"])):
        assert np.array_equal(x, y), k
    for key in ("prevX", "midX", "endX"):
        assert len(a[key]) == len(b[key])
        for k, (x, y) in enumerate(zip(a[key], b[key])):
            assert np.array_equal(x, y), (key, k)
    assert a["nP"] == b["nP"]
    assert np.array_equal(a["final"][0], b["final"][0]) and np.array_equal(a["final"][1], b["final"][1])
    for u, v in zip(a["ext"], b["ext"]):
        assert np.array_equal(u, v)
    dropin.config['frame_ahead'] = False


def test_dropin_frame_ahead_refuses_other_arguments(dropin, scene_mod):
    sq = get_sequence(scene_mod, "kitti_half")
    dropin.config['frame_ahead'] = True
    dropin._reset_runtime()
    try:
        vision = dropin.BestVision(sq.K)
        state = {"P": sq.P0.astype(np.float64), "X": sq.X0.copy()}
        ext = {"C": np.array([]), "F": np.array([]), "T": np.array([])}
        associate = dropin.KeypointsToLandmarksAssociator(sq.K, np.eye(4))
        pe = dropin.PoseEstimator(sq.K)
        dropin.set_candidate_feed([sq.feeds[1], sq.feeds[2]])
        ns, _ = associate.associateKeypointsToLandmarks(sq.frames[0], sq.frames[1], state)
        T = pe.estimatePose(ns, 1)
        T2 = T.copy()
        T2[0, 3] += 0.5
        with pytest.raises(RuntimeError):
            dropin.LandmarkTriangulator(sq.K, None).triangulate_landmark(sq.frames[0], sq.frames[1], ns, ext, T2)
    finally:
        dropin.config['frame_ahead'] = False
