"""CPU tests: the port of the reference's three per-frame methods (oracle/ref_pipeline.py) against
golden runs of the reference's OWN classes imported from /root/reference/src
(tests/golden/seq.npz, made by tests/golden/make_golden.py seq), plus host-side logic of the drop-in
that needs no GPU."""
import importlib
import os
import sys

import numpy as np
import pytest

from conftest import CONFIG_YAML, PKG_NAME, ROOT

GOLD = os.path.join(os.path.dirname(__file__), "golden")
SEQ_CASES = {  # must match tests/golden/make_golden.py
    "kitti_half": ("kitti", 0.5, 36, 600, 80, 3),
    "parking_half": ("parking", 0.5, 24, 300, 60, 4),
    "malaga_half": ("malaga", 0.5, 20, 900, 100, 5),
    # full resolution, the BASELINE configs' sizes
    "kitti_full": ("kitti", 1.0, 61, 2000, 150, 7),
    "malaga_full": ("malaga", 1.0, 31, 3000, 200, 8),
}
SHI_SEQ_CASES = {  # name: (shape, scale, n_frames, n_tracks, seed) - must match tests/golden/make_golden.py
    "kitti_half_shi": ("kitti", 0.5, 14, 300, 6),
}
_seq_cache = {}


def get_sequence(scene_mod, name):
    if name not in _seq_cache:
        shape, scale, nf, nt, nfeed, seed = SEQ_CASES[name]
        _seq_cache[name] = scene_mod.Sequence(shape, n_frames=nf, n_tracks=nt, n_feed=nfeed, seed=seed, scale=scale)
    return _seq_cache[name]


def get_shi_sequence(scene_mod, name):
    if name not in _seq_cache:
        shape, scale, nf, nt, seed = SHI_SEQ_CASES[name]
        _seq_cache[name] = scene_mod.Sequence(shape, n_frames=nf, n_tracks=nt, n_feed=1, seed=seed, scale=scale)
    return _seq_cache[name]


def run_port_shi(sq, variant):
    """config 'shi-mask': the port detects with cv2.goodFeaturesToTrack (shi_alpha 1.5, shi confidence)."""
    from oracle.ref_pipeline import RefPipeline
    rp = RefPipeline(sq.K, pnp=variant, confidence=0.999993, alpha=1.5, detector="shi")
    state = {"P": sq.P0.astype(np.float64), "X": sq.X0.copy()}
    ext = {"C": np.array([]), "F": np.array([]), "T": np.array([])}
    Ts, nP, nC = [], [], []
    for k in range(1, len(sq.frames)):
        state, ext, T = rp.frame(sq.frames[k - 1], sq.frames[k], state, ext, None)
        Ts.append(T); nP.append(len(state["P"])); nC.append(len(ext["C"]))
    return np.array(Ts), np.array(nP), np.array(nC), state, ext


def run_port(sq, variant, row_match="isin"):
    from oracle.ref_pipeline import RefPipeline
    rp = RefPipeline(sq.K, pnp=variant, row_match=row_match)
    state = {"P": sq.P0.astype(np.float64), "X": sq.X0.copy()}
    ext = {"C": np.array([]), "F": np.array([]), "T": np.array([])}
    Ts, nP, nC = [], [], []
    for k in range(1, len(sq.frames)):
        state, ext, T = rp.frame(sq.frames[k - 1], sq.frames[k], state, ext, sq.feeds[k].copy())
        Ts.append(T); nP.append(len(state["P"])); nC.append(len(ext["C"]))
    return np.array(Ts), np.array(nP), np.array(nC), state, ext


@pytest.mark.parametrize("name", list(SEQ_CASES))
@pytest.mark.parametrize("variant", ["literal", "p3p"])
def test_port_reproduces_reference_classes(scene_mod, name, variant):
    pytest.importorskip("cv2")
    g = np.load(os.path.join(GOLD, "seq.npz"))
    sq = get_sequence(scene_mod, name)
    if sq.checksum() != int(g[name + "_checksum"]):
        pytest.skip("rendered frames differ from the golden generator's on this machine (libm)")
    Ts, nP, nC, state, ext = run_port(sq, variant)
    assert np.array_equal(nP, g["%s_%s_nP" % (name, variant)])
    assert np.array_equal(nC, g["%s_%s_nC" % (name, variant)])
    assert np.allclose(Ts, g["%s_%s_T" % (name, variant)], rtol=0, atol=1e-9)
    if variant == "p3p":
        assert np.allclose(state["X"], g[name + "_p3p_X_last"], rtol=1e-9, atol=1e-9)


@pytest.mark.parametrize("name", list(SHI_SEQ_CASES))
@pytest.mark.parametrize("variant", ["literal", "p3p"])
def test_port_reproduces_reference_classes_shi_mask(scene_mod, name, variant):
    """The reference run with find_new_candidates_method 'shi-mask' (its own detector)."""
    pytest.importorskip("cv2")
    g = np.load(os.path.join(GOLD, "seq.npz"))
    sq = get_shi_sequence(scene_mod, name)
    if sq.checksum() != int(g[name + "_checksum"]):
        pytest.skip("rendered frames differ from the golden generator's on this machine (libm)")
    Ts, nP, nC, state, ext = run_port_shi(sq, variant)
    assert np.array_equal(nP, g["%s_%s_nP" % (name, variant)])
    assert np.array_equal(nC, g["%s_%s_nC" % (name, variant)])
    assert np.allclose(Ts, g["%s_%s_T" % (name, variant)], rtol=0, atol=1e-9)
    if variant == "p3p":
        assert np.array_equal(np.asarray(ext["C"], np.float64), g[name + "_p3p_C_last"])


def test_status_mask_equals_isin_row_matching(scene_mod):
    """App. B-4: the reference matches rows with an element-wise np.isin; the GPU path compacts by
    tracking status.  On the synthetic sequences both give the same state sizes frame by frame."""
    pytest.importorskip("cv2")
    sq = get_sequence(scene_mod, "kitti_half")
    a = run_port(sq, "p3p", "isin")
    b = run_port(sq, "p3p", "status")
    assert np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])


def test_scene_geometry(scene_mod):
    sc = scene_mod.CorridorScene("kitti", seed=5, scale=0.25, length=120.0)
    img = sc.render(3)
    assert img.shape == (sc.h, sc.w) and img.dtype == np.uint8 and img.std() > 10
    pts = scene_mod.corner_points(img, 200, 1, min_dist=4)
    X = sc.backproject(3, pts)
    back, z = sc.project(3, X)
    assert np.abs(back - pts).max() < 1e-6 and (z > 0).all()
    # ATE of the ground truth against itself, and against a scaled+rotated copy, is zero
    gt = [sc.pose(k) for k in range(12)]
    assert scene_mod.ate(gt, gt) < 1e-9
    S = np.eye(4); S[:3, :3] = np.array([[0, -1, 0], [1, 0, 0], [0, 0, 1.0]])
    moved = [T @ S for T in gt]
    assert scene_mod.ate(moved, gt) < 1e-6


def test_dropin_surface_without_gpu(tmp_path, monkeypatch):
    """The drop-in module imports without a GPU, loads config.yaml from the CWD at import, and
    exports the reference's names with the reference's signatures."""
    import inspect
    (tmp_path / "config.yaml").write_text(CONFIG_YAML % "sift-sift")
    monkeypatch.chdir(tmp_path)
    monkeypatch.syspath_prepend(os.path.join(ROOT, PKG_NAME))
    sys.modules.pop("vo_pipeline", None)
    vo = importlib.import_module("vo_pipeline")
    try:
        assert vo.config["find_new_candidates_method"] == "sift-sift" and vo.config_path == "config.yaml"
        for name in ("BestVision", "VOInitializer", "KeypointsToLandmarksAssociator", "KeypointsToLandmarksAssociation",
                     "PoseEstimator", "LandmarkTriangulator", "load_config", "np", "yaml", "os", "Dict"):
            assert hasattr(vo, name), name
        sig = lambda f: list(inspect.signature(f).parameters)
        assert sig(vo.KeypointsToLandmarksAssociator.__init__) == ["self", "K", "current_pose"]
        assert sig(vo.KeypointsToLandmarksAssociator.associateKeypointsToLandmarks) == ["self", "old_frame", "new_frame", "state"]
        assert sig(vo.PoseEstimator.estimatePose) == ["self", "associations", "img_idx"]
        assert sig(vo.LandmarkTriangulator.__init__) == ["self", "K", "old_des"]
        assert sig(vo.LandmarkTriangulator.triangulate_landmark) == ["self", "old_frame", "new_frame", "state", "extended_state", "new_pose"]
        assert sig(vo.VOInitializer.get_2D_3D_landmarks_association) == ["self", "kps_1", "kps_2", "T_img1_img2"]
        K = np.eye(3)
        bv = vo.BestVision(K)
        assert set(bv.state) == {"P", "X"} and set(bv.extended_state) == {"C", "F", "T"} and bv.K is K
        pe = vo.PoseEstimator(K)
        assert pe.REPOJ_THRESH == 2 and pe.CONFIDENCE == 0.99999
        tri = vo.LandmarkTriangulator(K, None)      # constructed every frame by main.py:280: must be trivial
        assert tri.old_des is None
        # no CPU fallback: the hot path fails loudly when there is no CUDA device
        import torch
        if not torch.cuda.is_available():
            with pytest.raises(Exception):
                vo.KeypointsToLandmarksAssociator(K, np.eye(4)).associateKeypointsToLandmarks(
                    np.zeros((64, 64), np.uint8), np.zeros((64, 64), np.uint8), {"P": np.zeros((4, 2)), "X": np.zeros((4, 3))})
    finally:
        sys.modules.pop("vo_pipeline", None)
