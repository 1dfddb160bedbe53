"""GPU tests (B200 box) of SURVEY 8f-4: vo_map_accumulate / vo_map_get / vo_map_render against
oracle/map_ref.py (main.py:239-255 restated with the reference's numpy calls; raster rules of
csrc/map.cu).  Gates: path positions 1e-12 (pos = -R^T t vs LAPACK inverse), landmark rows bit-exact and
in order (rows within 1e-9 of a threshold excepted), rendered image pixel-exact."""
import numpy as np
import pytest

from oracle import map_ref
from test_pipeline_oracle import get_sequence

pytestmark = pytest.mark.gpu


def test_map_accumulate_along_a_sequence(pkg, scene_mod):
    sq = get_sequence(scene_mod, "kitti_half")
    c = pkg.Context(device=0, max_w=2048, max_h=1024)
    try:
        c.upload_frame(sq.frames[0], 1)
        c.set_state(sq.P0, sq.X0)
        c.set_candidates(np.zeros((0, 2)), np.zeros((0, 2)), np.zeros((0, 16)))
        c.map_reset()
        # thresholds that cut through this scene's landmark cloud (the reference's 3 / 50 keep almost all)
        ref = map_ref.MapRef(r_min=6.0, r_max=14.0)
        near = 0
        for k in range(1, 12):
            T, cnt = c.process_frame(sq.frames[k], k, k + 1, sq.K, 1.0, sq.feeds[k])
            _, Xprev = c.get_prev_state()                       # vision.state['X'] at main.py:248
            pos, flt = ref.update(T, Xprev)
            d = np.linalg.norm(Xprev - pos, axis=1)
            near += int(((np.abs(d - 6.0) < 1e-9) | (np.abs(d - 14.0) < 1e-9)).sum())
            n = c.map_accumulate(None if k % 2 else T, 6.0, 14.0, use_prev_state=True, sync=(k % 3 == 0))
            if n is not None and near == 0:
                assert n == int(flt.sum())
        traj, pts, pos_d = c.map_get()
        assert traj.shape == ref.poses_plotting.shape and np.abs(traj - ref.poses_plotting).max() < 1e-12
        assert np.array_equal(pos_d, traj[-1])
        assert near == 0, "a landmark sits on a threshold: pick other radii"
        assert 0 < len(pts) < 11 * len(sq.X0)
        assert np.array_equal(pts, ref.X_plotting)
        # rendered plot: same data -> same pixels
        for (w, h, m, r) in ((1200, 600, 40, 3), (320, 200, 0, 1), (257, 131, 7, 0)):
            img, b = c.map_render(w, h, m, r)
            oimg, ob = map_ref.render(traj, pts, w, h, m, r)
            assert np.array_equal(b, ob), (b, ob)
            assert np.array_equal(img, oimg), (w, h, int((img != oimg).any(axis=2).sum()))
        # current (not previous) landmarks, host pose, reference radii
        Pc, Xc = c.get_state()
        ref2 = map_ref.MapRef()
        ref2.X_plotting, ref2.poses_plotting = ref.X_plotting.copy(), ref.poses_plotting.copy()
        ref2.update(T, Xc)
        c.map_accumulate(T, 3.0, 50.0, use_prev_state=False)
        traj2, pts2, _ = c.map_get()
        assert np.array_equal(pts2, ref2.X_plotting) and np.abs(traj2 - ref2.poses_plotting).max() < 1e-12
        c.map_reset()
        traj3, pts3, _ = c.map_get()
        assert len(traj3) == 0 and len(pts3) == 0
        img, b = c.map_render(64, 48, 4, 3)
        oimg, ob = map_ref.render(traj3, pts3, 64, 48, 4, 3)
        assert np.array_equal(img, oimg) and np.array_equal(b, ob)
    finally:
        c.close()


def test_map_growth_and_many_rows(pkg):
    """Buffers grow without losing rows; more landmark rows than one CTA pass (1024)."""
    rng = np.random.default_rng(5)
    c = pkg.Context(device=0, max_w=640, max_h=480)
    try:
        ref = map_ref.MapRef()
        for k in range(6):
            n = 3000 + 517 * k
            X = rng.uniform(-40, 40, size=(n, 3))
            c.set_state(rng.uniform(0, 100, size=(n, 2)).astype(np.float32), X)
            T = np.eye(4); T[:3, 3] = rng.normal(size=3)
            ref.update(T, X)
            c.map_accumulate(T, use_prev_state=False)
        traj, pts, _ = c.map_get()
        assert np.array_equal(pts, ref.X_plotting) and np.abs(traj - ref.poses_plotting).max() < 1e-12
        img, b = c.map_render(800, 400, 20, 2)
        oimg, ob = map_ref.render(traj, pts, 800, 400, 20, 2)
        assert np.array_equal(img, oimg) and np.array_equal(b, ob)
    finally:
        c.close()


def test_dropin_trajectory_map(tmp_path, monkeypatch, scene_mod):
    """TrajectoryMap in the drop-in == main.py:239-255 done on the host with the lazy state mirror."""
    import importlib, os, sys
    from conftest import CONFIG_YAML, PKG_NAME, ROOT
    (tmp_path / "config.yaml").write_text(CONFIG_YAML % "feed")
    monkeypatch.chdir(tmp_path)
    monkeypatch.syspath_prepend(os.path.join(ROOT, PKG_NAME))
    sys.modules.pop("vo_pipeline", None)
    vo = importlib.import_module("vo_pipeline")
    try:
        sq = get_sequence(scene_mod, "kitti_half")
        vo.set_candidate_feed([f.copy() for f in sq.feeds[1:]])
        vision = vo.BestVision(sq.K)
        vision.state = {"P": sq.P0.astype(np.float64), "X": sq.X0.copy()}
        vision.extended_state = {"C": np.array([]), "F": np.array([]), "T": np.array([])}
        associate = vo.KeypointsToLandmarksAssociator(sq.K, np.eye(4))
        pose_estimator = vo.PoseEstimator(sq.K)
        ref = map_ref.MapRef()
        img1 = sq.frames[0]
        tmap = None
        for k in range(1, 10):
            img2 = sq.frames[k]
            new_state, lost = associate.associateKeypointsToLandmarks(img1, img2, vision.state)
            new_pose = pose_estimator.estimatePose(new_state, k)
            if tmap is None:
                tmap = vo.TrajectoryMap()
            tmap.update(new_pose if k % 2 else None)
            ref.update(new_pose, vision.state["X"])              # main.py:248 on the host mirror
            tri = vo.LandmarkTriangulator(sq.K, None)
            vision.state, vision.extended_state, _ = tri.triangulate_landmark(img1, img2, new_state, vision.extended_state, new_pose)
            img1 = img2
        traj, pts = tmap.arrays()
        assert np.array_equal(pts, ref.X_plotting) and np.abs(traj - ref.poses_plotting).max() < 1e-12
        out = str(tmp_path / "plot.png")
        tmap.savefig(out, w=640, h=320)
        assert os.path.getsize(out) > 500
    finally:
        vo._reset_runtime()
        sys.modules.pop("vo_pipeline", None)
