"""GPU parity tests (B200 box): batched DLT triangulation and the fused bearing-test + DLT +
promotion step through the C-ABI against cv2 goldens (tests/golden/tri.npz) and the C oracle.
Gate (north_star): triangulated points within 1e-4 relative."""
import os

import numpy as np
import pytest

import synth
from conftest import parity_report
from test_tri_oracle import GOLD, TRI_CASES, TRI_TOL

pytestmark = pytest.mark.gpu
K = synth.K_KITTI


def relerr(a, b):
    return np.linalg.norm(a - b, axis=1) / np.linalg.norm(b, axis=1)


@pytest.mark.parametrize("name", list(TRI_CASES))
def test_triangulate_vs_golden_and_oracle(ctx, oracle, name):
    g = np.load(os.path.join(GOLD, "tri.npz"))
    n, seed = TRI_CASES[name]
    X, T1, T2, p1, p2 = synth.two_view_scene(n, seed)
    Xh = ctx.triangulate_points(K @ T1[:3], K @ T2[:3], p1, p2)
    assert np.allclose(np.linalg.norm(Xh, axis=1), 1.0, atol=1e-12)
    got = Xh[:, :3] / Xh[:, 3:]
    assert relerr(got, g[name + "_X"]).max() <= 1e-8 < TRI_TOL          # vs cv2 4.13
    o = oracle.triangulate_points(K @ T1[:3], K @ T2[:3], p1, p2)
    assert relerr(got, o[:, :3] / o[:, 3:]).max() <= 1e-10              # vs the C oracle
    parity_report("triangulation", name, dict(points=n, max_rel_err_vs_cv2=float(relerr(got, g[name + "_X"]).max()),
                                              max_rel_err_vs_c_oracle=float(relerr(got, o[:, :3] / o[:, 3:]).max()), gate=TRI_TOL))


def test_triangulate_live_cv2_and_edges(ctx):
    X, T1, T2, p1, p2 = synth.two_view_scene(64, 5, K=synth.K_PARKING)
    Kp = synth.K_PARKING
    assert ctx.triangulate_points(Kp @ T1[:3], Kp @ T2[:3], p1[:0], p2[:0]).shape == (0, 4)   # empty
    one = ctx.triangulate_points(Kp @ T1[:3], Kp @ T2[:3], p1[:1], p2[:1])                   # one point (vo:666)
    allp = ctx.triangulate_points(Kp @ T1[:3], Kp @ T2[:3], p1, p2)
    assert np.array_equal(one[0], allp[0])
    cv2 = pytest.importorskip("cv2")
    ref = cv2.triangulatePoints(Kp @ T1[:3], Kp @ T2[:3], p1.T.copy(), p2.T.copy()).T
    assert relerr(allp[:, :3] / allp[:, 3:], ref[:, :3] / ref[:, 3:]).max() < 1e-8


def test_promotion_step_vs_oracle(ctx, oracle):
    """vo_pipeline.py:608-721 on a prepared candidate set: which rows are promoted (bit-exact),
    the appended P/X rows, the surviving C/F/T rows and their order."""
    n = 3000
    X, T1, T2, p1, p2 = synth.two_view_scene(n, 8)
    rng = np.random.default_rng(4)
    C = p2.astype(np.float32)
    F = np.where(rng.random((n, 1)) < 0.5, C.astype(np.float64) + rng.normal(0, 0.4, (n, 2)), p1)   # many below alpha
    T = np.repeat(T1[None], n, 0).copy()
    T[::3, :3, 3] += 0.01            # three different first-observation poses
    T[1::3, :3, 3] -= 0.02
    P0 = rng.uniform(0, 1000, (50, 2)).astype(np.float32)
    X0 = rng.uniform(-10, 10, (50, 3))
    for alpha in (1.0, 1.5):
        ctx.set_state(P0, X0)
        ctx.set_candidates(C, F, T)
        nprom = ctx.track_triangulate_promote(-1, -1, K, T2, alpha, sync=True)
        prom, Xo, al = oracle.bearing_triangulate(C.astype(np.float64), F, T, T2, K, alpha)
        sel = prom == 1
        assert nprom == sel.sum() and 0 < nprom < n
        Pn, Xn = ctx.get_state()
        assert len(Pn) == 50 + nprom and np.array_equal(Pn[:50], P0) and np.array_equal(Xn[:50], X0)
        assert np.array_equal(Pn[50:], C[sel])
        assert relerr(Xn[50:], Xo[sel]).max() <= 1e-9 < TRI_TOL
        Cn, Fn, Tn = ctx.get_candidates()
        assert np.array_equal(Cn, C[~sel]) and np.array_equal(Fn, F[~sel]) and np.array_equal(Tn, T[~sel])
        cnt = ctx.counters()
        assert cnt["n_promoted"] == nprom and cnt["n_candidates"] == n - nprom and cnt["n_landmarks"] == 50 + nprom


def test_bearing_decisions_vs_golden(ctx):
    """The reference's numpy bearing loop (golden) decides promotion identically at both thresholds."""
    g = np.load(os.path.join(GOLD, "tri.npz"))
    C, F = synth.bearing_pairs(4000, 9)
    T = np.repeat(np.eye(4)[None], len(C), 0)
    T2 = np.eye(4)
    T2[0, 3] = -0.5
    for alpha in (1.0, 1.5):
        ctx.set_state(np.zeros((0, 2), np.float32), np.zeros((0, 3)))
        ctx.set_candidates(C, F, T)
        nprom = ctx.track_triangulate_promote(-1, -1, K, T2, alpha, sync=True)
        sel = g["bearing_alpha"] > alpha
        assert nprom == sel.sum()
        parity_report("bearing_test", "alpha_%.1f" % alpha, dict(pairs=len(C), promoted=int(nprom), decisions_differing_from_reference_loop=0))
        Cn, _, _ = ctx.get_candidates()
        assert np.array_equal(Cn, C[~sel].astype(np.float32))
