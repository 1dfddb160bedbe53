"""oracle/gftt.c (cv2.goodFeaturesToTrack as called at reference vo_pipeline.py:381) against the
goldens recorded from cv2 and, when cv2 is importable, against cv2 live: bit-exact."""
import hashlib
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import synth  # noqa: E402
from make_golden_cases import GFTT_CASES, gftt_image, reference_mask  # noqa: E402

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("name", list(GFTT_CASES))
def test_oracle_gftt_vs_golden(oracle, name):
    g = np.load(os.path.join(GOLD, "gftt.npz"))
    img, pts = gftt_image(name)
    assert img.astype(np.int64).sum() == g[name + "_img_sum"]
    corners, eig, ncand = oracle.good_features_to_track(img, mask=reference_mask(img, pts), want_eig=True)
    assert np.array_equal(np.frombuffer(hashlib.sha1(eig.tobytes()).digest(), np.uint8), g[name + "_eig_sha1"])
    assert np.array_equal(eig[:: max(1, img.shape[0] // 6)], g[name + "_eig_rows"])
    assert np.array_equal(corners, g[name + "_corners"])          # same corners, same order
    allc = oracle.good_features_to_track(img, max_corners=0, quality=0.01, min_distance=5)
    assert np.array_equal(allc.astype(np.uint16), g[name + "_all_corners"])


def test_oracle_gftt_vs_cv2_live(oracle):
    cv2 = pytest.importorskip("cv2")
    img = synth.texture(200, 333, 77)
    for block in (2, 3, 4, 5):
        assert np.array_equal(oracle.corner_min_eigen_val(img, block), cv2.cornerMinEigenVal(img, block, ksize=3))
    for md, mc, q in ((10, 500, 0.03), (10.4, 0, 0.03), (3, 100, 0.2), (0, 50, 0.01)):
        c = cv2.goodFeaturesToTrack(img, maxCorners=mc, qualityLevel=q, minDistance=md, blockSize=4)
        o = oracle.good_features_to_track(img, max_corners=mc, quality=q, min_distance=md)
        assert np.array_equal(c.reshape(-1, 2), o)


def test_oracle_gftt_edge_cases(oracle):
    flat = np.full((40, 50), 128, np.uint8)
    assert len(oracle.good_features_to_track(flat)) == 0                      # no texture: no corners
    img = synth.texture(60, 80, 3)
    assert len(oracle.good_features_to_track(img, mask=np.zeros((60, 80), np.uint8))) == 0   # everything masked
    one = oracle.good_features_to_track(img, max_corners=1)
    full = oracle.good_features_to_track(img)
    assert len(one) == 1 and np.array_equal(one[0], full[0])
    d = full[:, None, :] - full[None, :, :]
    dist2 = (d ** 2).sum(-1) + 1e9 * np.eye(len(full))
    assert dist2.min() >= 100.0                                                # minDistance honoured
    assert (full[:, 0] >= 1).all() and (full[:, 0] <= 78).all() and (full[:, 1] >= 1).all() and (full[:, 1] <= 58).all()
