"""CUDA Shi-Tomasi detector (csrc/det.cu) through the C-ABI against the oracle, the
cv2 goldens and cv2 live: the min-eigenvalue map bit for bit, the corner list identical in content
and order (reference call: vo_pipeline.py:381 with the mask of vo_pipeline.py:371-378)."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import synth  # noqa: E402
from make_golden_cases import GFTT_CASES, gftt_image, reference_mask  # noqa: E402

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
_key = [5000]


def upload(ctx, img):
    _key[0] += 1
    ctx.upload_frame(img, _key[0])
    return _key[0]


@pytest.mark.parametrize("use_cluster", [1, 0])
@pytest.mark.parametrize("name", list(GFTT_CASES))
def test_gftt_vs_golden_and_oracle(ctx, oracle, name, use_cluster):
    g = np.load(os.path.join(GOLD, "gftt.npz"))
    img, pts = gftt_image(name)
    k = upload(ctx, img)
    mask = reference_mask(img, pts)
    # the reference's mask given as the point list it is built from, and as an image
    dflt = ctx_params(ctx, use_cluster=use_cluster)
    c_pts, eig, ncand = ctx.good_features_to_track(k, exclude_pts=pts, params=dflt, want_eig=True, shape=img.shape)
    c_img = ctx.good_features_to_track(k, mask=mask, params=dflt)
    o_c, o_eig, o_ncand = oracle.good_features_to_track(img, mask=mask, want_eig=True)
    assert np.array_equal(eig, o_eig)                      # cv2.cornerMinEigenVal, bit for bit
    assert ncand == o_ncand
    assert np.array_equal(c_pts, g[name + "_corners"]) and np.array_equal(c_img, g[name + "_corners"])
    assert np.array_equal(c_pts, o_c)
    prm = ctx_params(ctx, max_corners=0, quality=0.01, min_distance=5, use_cluster=use_cluster)
    allc = ctx.good_features_to_track(k, params=prm, shape=img.shape)
    assert np.array_equal(allc.astype(np.uint16), g[name + "_all_corners"])


def ctx_params(ctx, max_corners=500, quality=0.03, min_distance=10.0, block=4, use_cluster=1):
    import importlib
    pkg = importlib.import_module("locally-consistent-visual-odometry-pipeline_b200")
    p = pkg.gftt_defaults()
    p.max_corners, p.quality_level, p.min_distance, p.block_size = max_corners, quality, min_distance, block
    p.use_cluster = use_cluster
    return p


def test_gftt_vs_cv2_live(ctx):
    cv2 = pytest.importorskip("cv2")
    img = synth.texture(200, 333, 77)
    k = upload(ctx, img)
    agree = []
    for md, mc, q, bs in ((10, 500, 0.03, 4), (10.4, 0, 0.03, 4), (3, 100, 0.2, 4), (0, 50, 0.01, 4), (7, 300, 0.05, 3), (10, 500, 0.03, 5)):
        c = cv2.goodFeaturesToTrack(img, maxCorners=mc, qualityLevel=q, minDistance=md, mask=None, blockSize=bs, gradientSize=3)
        c = c.reshape(-1, 2)
        o = ctx.good_features_to_track(k, params=ctx_params(ctx, mc, q, md, bs), shape=img.shape)
        agree.append(len(c) == len(o) and np.array_equal(c, o))
        if not agree[-1]:
            # another cv2 build than the goldens': its SIMD dispatch decides a few rounding details of the min-eigenvalue
            # map (oracle/gftt.c), which can swap near-ties in the greedy selection - but not move corners about
            d = np.abs(c[:, None, :] - o[None, :, :]).max(-1).min(1) if len(c) and len(o) else np.ones(1)
            assert abs(len(c) - len(o)) <= max(2, len(c) // 50) and (d <= 1.0).mean() >= 0.9, (md, mc, q, bs)
    # on the build the goldens were recorded with every case is identical
    assert np.mean(agree) >= 0.5
    if cv2.__version__ == str(np.load(os.path.join(GOLD, "gftt.npz"))["cv2_version"]):
        assert all(agree)


def test_gftt_large_candidate_sets_and_sort_paths(ctx, oracle):
    """> 16384 candidates: the selection kernel sorts in global memory instead of shared memory."""
    img = synth.texture(768, 1024, 31, radius=1, passes=1)
    k = upload(ctx, img)
    for md, q, mc in ((0, 0.0005, 0), (2, 0.0005, 0), (10, 0.0005, 0), (10, 0.0005, 3000), (3, 0.0005, 20000), (25, 0.0005, 500)):
        prm = ctx_params(ctx, mc, q, md)
        c, _, ncand = ctx.good_features_to_track(k, params=prm, want_eig=True, shape=img.shape)
        o, _, o_ncand = oracle.good_features_to_track(img, max_corners=mc, quality=q, min_distance=md, want_eig=True)
        assert ncand == o_ncand and ncand > 16384
        assert np.array_equal(c, o)


def test_gftt_edge_cases(ctx, oracle):
    flat = np.full((40, 50), 128, np.uint8)
    assert len(ctx.good_features_to_track(upload(ctx, flat))) == 0
    img = synth.texture(60, 80, 3)
    k = upload(ctx, img)
    assert len(ctx.good_features_to_track(k, mask=np.zeros((60, 80), np.uint8))) == 0
    one = ctx.good_features_to_track(k, params=ctx_params(ctx, 1))
    full = ctx.good_features_to_track(k)
    assert len(one) == 1 and np.array_equal(one[0], full[0])
    assert np.array_equal(full, oracle.good_features_to_track(img))
    tiny = synth.texture(5, 7, 9)
    assert np.array_equal(ctx.good_features_to_track(upload(ctx, tiny)), oracle.good_features_to_track(tiny))
    with pytest.raises(Exception):
        ctx.good_features_to_track(k, params=ctx_params(ctx, block=9))
    with pytest.raises(Exception):
        ctx.good_features_to_track(987654321)       # frame not resident


def test_detect_add_candidates_on_device(ctx, oracle):
    """find_new_candidates_shi + append without leaving the device (vo_pipeline.py:357-390, :564-591):
    corners masked at the PREVIOUS-frame positions of the candidates KLT kept become C = F rows with
    T = new_pose, after the promotion step of the same frame."""
    prev, nxt, pts = synth.klt_pair(376, 1241, 1, 700)
    ka, kb = upload(ctx, prev), upload(ctx, nxt)
    K = synth.K_KITTI
    C = pts.astype(np.float32)
    F = C.astype(np.float64)
    T0 = np.repeat(np.eye(4)[None], len(C), 0)
    Tn = np.eye(4)
    Tn[0, 3] = -0.3
    ctx.set_state(np.zeros((0, 2), np.float32), np.zeros((0, 3)))
    ctx.set_candidates(C, F, T0)
    ctx.track_candidates(ka, kb)
    prev_pts, next_pts, st = ctx.candidate_tracks()
    assert np.array_equal(prev_pts, C) and 0 < st.sum() < len(C)
    ctx.track_triangulate_promote(ka, kb, K, Tn, 90.0, sync=True)       # alpha 90 deg: nothing is promoted
    Cn0, Fn0, Tn0 = ctx.get_candidates()
    assert len(Cn0) == st.sum()
    ctx.detect_add_candidates_shi(kb, T=Tn)
    Cn, Fn, Tn1 = ctx.get_candidates()
    want = oracle.good_features_to_track(nxt, mask=reference_mask(nxt, prev_pts[st == 1]))
    assert len(want) == 500
    assert np.array_equal(Cn[:len(Cn0)], Cn0) and np.array_equal(Fn[:len(Cn0)], Fn0)
    assert np.array_equal(Cn[len(Cn0):], want) and np.array_equal(Fn[len(Cn0):], want.astype(np.float64))
    assert np.array_equal(Tn1[len(Cn0):], np.repeat(Tn[None], 500, 0))
    assert ctx.counters()["n_candidates"] == len(Cn0) + 500
