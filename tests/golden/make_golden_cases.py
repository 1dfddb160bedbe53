"""Case tables and input builders shared by make_golden.py and the tests (importable without cv2
and without /root/reference)."""
import importlib
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import synth  # noqa: E402

GFTT_CASES = {  # name: (h, w, seed, n_mask_points, kind) - SURVEY.md 8f-1
    "kitti": (376, 1241, 21, 400, "texture"),
    "parking": (480, 640, 22, 150, "texture"),
    "malaga": (768, 1024, 23, 600, "texture"),
    "odd": (75, 101, 24, 20, "texture"),
    "scene": (376, 1241, 25, 300, "scene"),
}


def gftt_image(name):
    h, w, seed, nmask, kind = GFTT_CASES[name]
    if kind == "scene":
        scene_mod = importlib.import_module("locally-consistent-visual-odometry-pipeline_b200.scene")
        img = scene_mod.Sequence("kitti", n_frames=2, n_tracks=50, n_feed=10, seed=seed).frames[1]
    else:
        img = synth.texture(h, w, seed)
    rng = np.random.default_rng(seed)
    pts = np.stack([rng.uniform(-3, w + 3, nmask), rng.uniform(-3, h + 3, nmask)], 1).astype(np.float32)
    return img, pts


def reference_mask(img, pts):
    """The mask find_new_candidates_shi builds (vo_pipeline.py:371-378), verbatim semantics."""
    mask = np.ndarray(shape=(img.shape[0], img.shape[1]), dtype=np.uint8)
    mask.fill(255)
    for i in range(len(pts)):
        y, x = int(pts[i][1]), int(pts[i][0])
        if 0 <= y < mask.shape[0] and 0 <= x < mask.shape[1]:
            mask[y, x] = 0
    return mask
