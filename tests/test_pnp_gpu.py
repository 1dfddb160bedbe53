"""GPU parity tests (B200 box): CUDA PnP-RANSAC through the C-ABI against the cv2 golden vectors
(tests/golden/pnp.npz), the C oracle and, when importable, cv2 itself.
Gates (BASELINE.json north_star): for the same minimal sets (cv::RNG(2^64-1) sequence) the inlier
mask is bit-exact; pose within 1e-4 rad and 1e-4 relative translation."""
import numpy as np
import pytest

import synth
from conftest import parity_report
from test_pnp_oracle import PNP_CASES, POSE_TOL, check_pose, golden_case, rot_angle

pytestmark = pytest.mark.gpu
K = synth.K_KITTI


def params(pkg, H, conf=0.99999, thr=2.0):
    p = pkg.pnp_defaults()
    p.max_iters = H
    p.confidence = conf
    p.reproj_thresh = thr
    return p


@pytest.mark.parametrize("name", list(PNP_CASES))
def test_pnp_vs_golden_and_oracle(ctx, pkg, oracle, name):
    X, P, H, mask, Rg, tg, g = golden_case(name)
    r = ctx.solve_pnp_ransac(X, P, K, params(pkg, H))
    assert r["ok"]
    # against cv2 4.13 (golden): mask bit-exact, pose within the north_star tolerance
    assert np.array_equal(r["mask"], mask), "hamming vs cv2 %d" % int((r["mask"] != mask).sum())
    check_pose(r["R"], r["t"], Rg, tg, tol=POSE_TOL)
    # against the C oracle: same bookkeeping, tighter
    o = oracle.solve_pnp_ransac(X, P, K, max_iters=H)
    assert np.array_equal(r["mask"], o["mask"])
    assert (r["iters"], r["best_iter"], r["n_inliers"]) == (o["iters"], o["best_iter"], o["n_inliers"])
    check_pose(r["R"], r["t"], o["R"], o["t"], tol=1e-8)
    check_pose(r["Rh"], r["th"], o["Rh"], o["th"], tol=1e-8)
    # rvec/tvec as cv2 returns them
    assert rot_angle(synth.rodrigues(r["rvec"]), r["R"]) < 1e-9 and np.allclose(r["tvec"], r["t"])
    # report-only: the literal reference call vo_pipeline.py:258-264 (EPnP-5 hypotheses, <= 100 iterations)
    n = len(X)
    lit = np.unpackbits(g[name + "_literal_mask"])[:n]
    Rl = synth.rodrigues(g[name + "_literal_rvec"])
    parity_report("pnp", name, dict(points=n, hypotheses_budget=H, iterations_run=int(r["iters"]), inliers=int(r["n_inliers"]),
                                    hamming_vs_cv2_p3p=int((r["mask"] != mask).sum()), dR_rad_vs_cv2_iterative_on_mask=float(rot_angle(r["R"], Rg)),
                                    dt_rel_vs_cv2_iterative_on_mask=float(np.linalg.norm(r["t"] - tg) / np.linalg.norm(tg)),
                                    hamming_vs_c_oracle=int((r["mask"] != o["mask"]).sum()),
                                    literal_epnp5_call={"hamming": int((lit != r["mask"]).sum()), "dR_rad": float(rot_angle(r["R"], Rl)),
                                                        "dt_rel": float(np.linalg.norm(r["t"] - g[name + "_literal_tvec"].ravel()) / np.linalg.norm(r["t"])),
                                                        "note": "report-only: RANSAC-variant spread (SURVEY.md App. A.4), not a gate"},
                                    golden_cv2=str(g["cv2_version"]) if "cv2_version" in g else None))


def test_pnp_explicit_minimal_sets(ctx, pkg, oracle):
    """'Same fixed minimal-set seeds': an arbitrary H x 4 index table fed to both sides."""
    X, P, _, _, _ = synth.pnp_scene(1500, 0.35, 31)
    rng = np.random.default_rng(3)
    sets = np.stack([rng.choice(1500, 4, replace=False) for _ in range(300)]).astype(np.int32)
    r = ctx.solve_pnp_ransac(X, P, K, params(pkg, 300), sets=sets)
    o = oracle.solve_pnp_ransac(X, P, K, max_iters=300, sets=sets)
    assert r["ok"] and o["ok"]
    assert np.array_equal(r["mask"], o["mask"]) and r["best_iter"] == o["best_iter"] and r["iters"] == o["iters"]
    check_pose(r["R"], r["t"], o["R"], o["t"], tol=1e-8)


def test_pnp_vs_cv2_live(ctx, pkg):
    cv2 = pytest.importorskip("cv2")
    for n, of, H, seed in [(800, 0.3, 100, 41), (3000, 0.45, 1000, 42), (2000, 0.1, 100, 43)]:
        X, P, _, _, _ = synth.pnp_scene(n, of, seed, K=synth.K_MALAGA, w=1024, h=768)
        ok, rv, tv, inl = cv2.solvePnPRansac(X, P, synth.K_MALAGA, np.zeros(4), flags=cv2.SOLVEPNP_P3P,
                                             confidence=0.99999, reprojectionError=2, iterationsCount=H)
        r = ctx.solve_pnp_ransac(X, P, synth.K_MALAGA, params(pkg, H))
        m = np.zeros(n, np.uint8)
        m[inl.ravel()] = 1
        assert np.array_equal(r["mask"], m)
        Xi = X[m == 1].astype(np.float32).astype(np.float64)
        _, rv2, tv2 = cv2.solvePnP(Xi, P[m == 1].astype(np.float64), synth.K_MALAGA, np.zeros(4), flags=cv2.SOLVEPNP_ITERATIVE)
        check_pose(r["R"], r["t"], cv2.Rodrigues(rv2)[0], tv2.ravel())


def test_pnp_edge_cases(ctx, pkg, oracle):
    X, P, _, _, _ = synth.pnp_scene(100, 0.0, 11)
    assert not ctx.solve_pnp_ransac(X[:3], P[:3], K, params(pkg, 100))["ok"]      # < 4 points: no model
    assert not ctx.solve_pnp_ransac(X[:0], P[:0], K, params(pkg, 100))["ok"]      # empty
    r = ctx.solve_pnp_ransac(X, P, K, params(pkg, 100))                           # all inliers: early stop
    o = oracle.solve_pnp_ransac(X, P, K, max_iters=100)
    assert r["ok"] and r["iters"] == o["iters"] < 100 and np.array_equal(r["mask"], o["mask"])
    r4 = ctx.solve_pnp_ransac(X[:4], P[:4], K, params(pkg, 100))                  # n == 4: one direct solve
    assert r4["ok"] and r4["mask"].sum() == 4
    Pb = np.random.default_rng(1).uniform(0, 300, P.shape).astype(np.float32)     # pure outliers
    rb = ctx.solve_pnp_ransac(X, Pb, K, params(pkg, 50))
    ob = oracle.solve_pnp_ransac(X, Pb, K, max_iters=50)
    assert rb["ok"] == ob["ok"] and np.array_equal(rb["mask"], ob["mask"] if ob["ok"] else rb["mask"] * 0)


def test_pnp_stress_properties(ctx, pkg):
    """Full-size stress config (10k pairs, 50 % outliers, 4096 hypotheses forced by confidence=1):
    size-independent properties - determinism, mask consistent with the winning hypothesis, refit
    does not increase the cost, permutation of the points permutes the mask when the table is fixed."""
    n, H = 10000, 4096
    X, P, R, t, is_out = synth.pnp_scene(n, 0.5, 77)
    p = params(pkg, H, conf=1.0)
    sets = np.random.default_rng(9).integers(0, n, (H, 4)).astype(np.int32)
    a = ctx.solve_pnp_ransac(X, P, K, p, sets=sets)
    b = ctx.solve_pnp_ransac(X, P, K, p, sets=sets)
    assert a["ok"] and a["iters"] == H
    assert np.array_equal(a["mask"], b["mask"]) and np.array_equal(a["T"], b["T"])
    assert (a["mask"][~is_out] == 1).mean() > 0.95 and (a["mask"][is_out] == 1).mean() < 0.02
    Xc = X.astype(np.float32).astype(np.float64) @ a["Rh"].T + a["th"]
    uv = np.stack([K[0, 0] * Xc[:, 0] / Xc[:, 2] + K[0, 2], K[1, 1] * Xc[:, 1] / Xc[:, 2] + K[1, 2]], 1)
    e = ((uv - P) ** 2).sum(1)
    far = np.abs(e - 4.0) > 1e-3
    assert np.array_equal((e <= 4.0)[far], a["mask"][far] == 1)
    perm = np.random.default_rng(2).permutation(n)
    inv = np.argsort(perm)
    c = ctx.solve_pnp_ransac(X[perm], P[perm], K, p, sets=inv[sets].astype(np.int32))
    assert np.array_equal(c["mask"], a["mask"][perm])
    assert rot_angle(a["R"], R) < 2e-3 and np.linalg.norm(a["t"] - t) < 0.05


def test_pnp_f32_filter_changes_nothing(pkg, monkeypatch):
    """The scoring kernel decides pairs far from the threshold in float32 and only the rest in cv2's float64
    arithmetic (VO_PNP_F32_FILTER=1; off by default, it measured slower): counts, winner, mask and pose must be those of the
    all-float64 scoring, also
    far from the origin, where float32 has little to offer, and with points near the camera plane."""
    rng = np.random.default_rng(12)
    cases = []
    X, P, R, t, _ = synth.pnp_scene(10000, 0.5, 77)
    cases.append((X, P, 4096, 1.0))
    X2, P2, _, _, _ = synth.pnp_scene(3000, 0.3, 78)
    cases.append((X2, P2, 500, 0.99999))
    # the same scene 20 km from the origin (world coordinates of a long trajectory): X and t are huge, x_c is not
    shift = np.array([20000.0, -3000.0, 15000.0])
    cases.append((X2 + shift, P2, 500, 0.99999))
    # points all around (and inside 0.5 m of) the camera plane of most hypotheses
    X3 = rng.uniform(-2, 2, (2000, 3)) + [0, 0, 1.0]
    P3 = np.stack([rng.uniform(0, 1241, 2000), rng.uniform(0, 376, 2000)], 1).astype(np.float32)
    cases.append((X3, P3, 300, 0.99999))
    for X, P, H, conf in cases:
        out = []
        for on in ("0", "1"):
            monkeypatch.setenv("VO_PNP_F32_FILTER", on)
            c = pkg.Context(device=0, max_w=64, max_h=64, hyp_capacity=4096)
            try:
                out.append(c.solve_pnp_ransac(X, P, K, params(pkg, H, conf=conf)))
            finally:
                c.close()
        a, b = out
        assert a["ok"] == b["ok"] and a["iters"] == b["iters"] and a["best_iter"] == b["best_iter"] and a["n_inliers"] == b["n_inliers"]
        assert np.array_equal(a["mask"], b["mask"]) and np.array_equal(a["T"], b["T"])
