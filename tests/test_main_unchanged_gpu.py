"""The reference's own src/main.py, UNCHANGED, driven over the drop-in (BASELINE configs[0]):
a synthetic parking-shaped dataset is written in the reference's directory layout (parking/images/img_%05d.png,
parking/poses.txt), main.py is executed with runpy twice - with this package's vo_pipeline.py first on sys.path
(GPU) and with the reference's vo_pipeline.py (CPU / cv2) - and the two trajectories are compared.

The reference scripts are looked for in /root/reference/src (build container; no GPU there) or in baseline/_ref/src
(where __graft_entry__.build() stashes unmodified copies: git-ignored, but it travels to the GPU box).  Without
them the test is skipped and says so."""
import contextlib
import importlib
import io
import os
import runpy
import sys
import types

import numpy as np
import pytest

from conftest import PKG_NAME, ROOT, parity_report

pytestmark = pytest.mark.gpu

CONFIG = """dataset: 'parking'
find_new_candidates_method: 'shi-mask'
init_detector: 'sift'
init_descriptor: 'sift'
debug: False
visualization: False
animation: False
sift_alpha: 1
sift_ransac_confidence: 0.99999
shi_alpha: 1.5
shi_ransac_confidence: 0.999993
"""


def reference_src():
    for d in ("/root/reference/src", os.path.join(ROOT, "baseline", "_ref", "src")):
        if os.path.exists(os.path.join(d, "main.py")) and os.path.exists(os.path.join(d, "vo_pipeline.py")):
            return d
    return None


def stub_matplotlib():
    """matplotlib is not installed here; main.py and visual.py import it.  Returns the names to remove again."""
    def getter(k):
        if k.startswith("__"):
            raise AttributeError(k)
        return lambda *a, **kw: None
    added = []
    for name in ("matplotlib", "matplotlib.pyplot"):
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.__getattr__ = getter
            sys.modules[name] = m
            added.append(name)
    if added:
        sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    return added


def run_main(src, first_on_path, workdir):
    """runpy over the unmodified main.py with `first_on_path` ahead of the reference's src directory."""
    cwd = os.getcwd()
    saved = list(sys.path)
    for name in ("vo_pipeline", "visual"):
        sys.modules.pop(name, None)
    os.chdir(workdir)
    sys.path[:0] = [p for p in (first_on_path, src) if p]
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            g = runpy.run_path(os.path.join(src, "main.py"), run_name="__main__")
        info = {"state_class": g["vision"].state.__class__.__name__, "n_landmarks": int(len(g["vision"].state["P"])),
                "n_candidates": int(len(g["vision"].extended_state["C"])), "module": g["PoseEstimator"].__module__,
                "module_file": getattr(sys.modules.get("vo_pipeline"), "__file__", "")}
    finally:
        os.chdir(cwd)
        sys.path[:] = saved
        vo = sys.modules.pop("vo_pipeline", None)
        if vo is not None and hasattr(vo, "_reset_runtime"):
            vo._reset_runtime()
        sys.modules.pop("visual", None)
    return np.array(g["poses_plotting"]), info


def test_reference_main_py_runs_unchanged_over_the_dropin(tmp_path, scene_mod):
    src = reference_src()
    if src is None:
        pytest.skip("the reference's main.py is on neither /root/reference/src nor baseline/_ref/src on this machine")
    cv2 = pytest.importorskip("cv2")
    stubs = stub_matplotlib()
    try:
        _run_and_compare(src, tmp_path, scene_mod)
    finally:
        for name in stubs:
            sys.modules.pop(name, None)


def _run_and_compare(src, tmp_path, scene_mod):
    n = 44
    sq = scene_mod.Sequence("parking", n_frames=n, n_tracks=10, n_feed=1, seed=21, step=0.08, heading=float(np.pi / 2), z0=25.0)
    scene_mod.write_dataset(sq, str(tmp_path), "parking")
    (tmp_path / "config.yaml").write_text(CONFIG)
    ours, g_ours = run_main(src, os.path.join(ROOT, PKG_NAME), str(tmp_path))
    assert g_ours["state_class"] == "DeviceState" and PKG_NAME in g_ours["module_file"]      # it really was the drop-in
    ref, g_ref = run_main(src, None, str(tmp_path))
    assert g_ref["state_class"] == "dict" and g_ref["module_file"].startswith(src)           # ... and the reference's own module
    assert ours.shape == ref.shape == (n - 1 - 4, 3)
    # ground truth: camera centres of frames 4 .. n-2 relative to frame 0 (the bootstrap fixes frame 0 as the origin and the
    # scale by |t| = 1 between frames 0 and 4: compare after a similarity alignment, as scene.ate does)
    def as_T(c):
        T = np.eye(4); T[:3, 3] = -c
        return T
    gt = [sq.poses[k] for k in range(4, n - 1)]
    ate_ours = scene_mod.ate([as_T(c) for c in ours], gt)
    ate_ref = scene_mod.ate([as_T(c) for c in ref], gt)
    path = sq.scene.step * (n - 1 - 4)
    d = np.linalg.norm(ours - ref, axis=1)
    scale = np.linalg.norm(ref[-1] - ref[0])
    print("main.py unchanged, %d frames: ATE drop-in %.4f m, reference %.4f m (path %.2f m); max |pos diff| %.3g = %.2f %% of the distance travelled"
          % (len(ours), ate_ours, ate_ref, path, d.max(), 100 * d.max() / scale))
    parity_report("main_py_unchanged", "parking_640x480_shi_mask", dict(
        frames=int(len(ours)), ate_dropin_m=ate_ours, ate_reference_m=ate_ref, path_m=path, max_position_diff=float(d.max()),
        max_position_diff_rel=float(d.max() / scale), landmarks_dropin=g_ours["n_landmarks"],
        landmarks_reference=g_ref["n_landmarks"], candidates_dropin=g_ours["n_candidates"], candidates_reference=g_ref["n_candidates"],
        reference_src=src,
        note="reference: EPnP-5 RANSAC hypotheses (its literal call); drop-in: P3P hypotheses - two RANSAC realisations of one trajectory"))
    assert np.isfinite(ours).all()
    assert ate_ours <= 1.01 * ate_ref + 0.02 * path                                # north_star: ATE within 1 % (plus 2 % of the path: monocular scale from 4 frames)
    assert d.max() <= 0.05 * scale
