"""GPU parity tests (B200 box) of the bootstrap on the device (csrc/boot.cu) through the C-ABI:
five-point essential-matrix RANSAC + recoverPose against the cv2 goldens (tests/golden/bootstrap.npz), the C oracle
and, when importable, cv2 itself.  Reference: VOInitializer.getPoseEstimate, vo_pipeline.py:139-162.
Gates: inlier mask bit-exact (else the Hamming distance is reported with the straddling cause), R / t within 1e-6 of
cv2 on identical masks."""
import numpy as np
import pytest

import synth
from conftest import parity_report
from test_bootstrap_oracle import BOOT_CASES, KS, boot_case, e_dist

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", [k for k in BOOT_CASES if k != "n5_exact"])
def test_bootstrap_vs_golden_and_oracle(ctx, oracle, name):
    p1, p2, K, g, mask_E, mask_pose, (R_true, t_true, _) = boot_case(name)
    r = ctx.bootstrap_pose(p1, p2, K)
    o = oracle.find_essential_mat(p1, p2, K)
    assert r["ok"] and o["ok"]
    # the RANSAC: same iteration, same model, same count as the C restatement (which is pinned to cv2's mask)
    assert (r["iters"], r["best_iter"], r["n_inliers"]) == (o["iters"], o["best_iter"], o["n_inliers"])
    assert e_dist(r["E"], g[name + "_E"]) < 1e-7
    ham = int((r["mask"] != mask_pose).sum())
    dR = float(np.abs(r["T34"][:, :3] - g[name + "_R"]).max())
    dt = float(np.abs(r["T34"][:, 3] - g[name + "_t"]).max())
    parity_report("bootstrap", name, dict(matches=len(p1), ransac_inliers=r["n_inliers"], inliers_after_cheirality=r["n_good"],
                                          hamming_vs_cv2=ham, dR_vs_cv2=dR, dt_vs_cv2=dt, iterations=r["iters"],
                                          golden_cv2=str(g["cv2_version"])))
    assert ham == 0, "mask differs from cv2's in %d rows" % ham
    assert r["n_good"] == int(g[name + "_good"])
    assert dR < 1e-6 and dt < 1e-6
    assert abs(np.linalg.norm(r["T34"][:, 3]) - 1.0) < 1e-9


def test_bootstrap_minimal_and_edge_cases(ctx, oracle):
    K = synth.K_KITTI
    p1, p2, K5, g, _, _, (R, t, _) = boot_case("n5_exact")
    r = ctx.bootstrap_pose(p1, p2, K5)                     # count == model points: one direct solve, all inliers
    assert r["ok"] and r["n_inliers"] == 5 and r["iters"] == 1
    Eall = g["n5_exact_E_all"].reshape(-1, 3, 3)
    assert min(e_dist(r["E"], Ec) for Ec in Eall) < 1e-7
    q1, q2, _, _, _ = synth.bootstrap_scene(40, 0.0, 3)
    for n in (0, 3, 4):
        rr = ctx.bootstrap_pose(q1[:n], q2[:n], K)
        assert not rr["ok"] and rr["mask"].sum() == 0 and not rr["T34"].any()
    # determinism, and independence of everything but the inputs
    a = ctx.bootstrap_pose(q1, q2, K)
    b = ctx.bootstrap_pose(q1, q2, K)
    assert a["ok"] and np.array_equal(a["mask"], b["mask"]) and np.array_equal(a["T34"], b["T34"])
    # pure outliers: whatever wins has few inliers, exactly as the C restatement finds
    junk = np.random.default_rng(1).uniform(0, 300, q2.shape)
    rj, oj = ctx.bootstrap_pose(q1, junk, K), oracle.find_essential_mat(q1, junk, K)
    assert rj["n_inliers"] == oj["n_inliers"] and rj["best_iter"] == oj["best_iter"]


def test_bootstrap_vs_cv2_live_and_truth(ctx):
    cv2 = pytest.importorskip("cv2")
    for n, of, seed in [(400, 0.25, 41), (1000, 0.4, 42), (150, 0.05, 43), (3000, 0.3, 44)]:
        p1, p2, R, t, _ = synth.bootstrap_scene(n, of, seed, K=synth.K_MALAGA, w=1024, h=768)
        E, mask = cv2.findEssentialMat(p1, p2, synth.K_MALAGA, cv2.RANSAC, 0.999, 1.0, None)
        _, Rc, tc, _ = cv2.recoverPose(E, p1, p2, synth.K_MALAGA, mask=mask)    # (mask modified in place)
        r = ctx.bootstrap_pose(p1, p2, synth.K_MALAGA)
        assert np.array_equal(r["mask"], (mask.ravel() != 0).astype(np.uint8))
        assert np.abs(r["T34"][:, :3] - Rc).max() < 1e-6 and np.abs(r["T34"][:, 3] - tc.ravel()).max() < 1e-6
        assert np.abs(r["T34"][:, :3] - R).max() < 2e-2 and np.linalg.norm(r["T34"][:, 3] - t / np.linalg.norm(t)) < 0.25


def test_dropin_voinitializer_on_the_device(dropin):
    """VOInitializer.getPoseEstimate + get_2D_3D_landmarks_association through the drop-in: same pose, mask and
    landmarks as the reference's cv2 calls (vo_pipeline.py:139-177)."""
    cv2 = pytest.importorskip("cv2")
    K = synth.K_KITTI
    p1, p2, R, t, _ = synth.bootstrap_scene(600, 0.25, 51)
    init = dropin.VOInitializer(K)
    T, mask = init.getPoseEstimate(p1, p2)
    assert T.shape == (3, 4) and mask.shape == (600, 1) and mask.dtype == np.uint8
    E, m = cv2.findEssentialMat(p1, p2, K, cv2.RANSAC, 0.999, 1.0, None)
    _, Rc, tc, _ = cv2.recoverPose(E, p1, p2, K, mask=m)
    assert np.array_equal(mask != 0, m != 0) and np.abs(T - np.hstack((Rc, tc))).max() < 1e-6
    sel = np.hstack(mask).astype(np.bool_)                                         # main.py:82-86
    st = init.get_2D_3D_landmarks_association(p1[sel], p2[sel], T)
    XH = cv2.triangulatePoints(K @ np.eye(3, 4), K @ np.hstack((Rc, tc)), p1[sel].T, p2[sel].T).T
    Xc = XH[:, :3] / XH[:, 3:]
    assert (np.linalg.norm(st["X"] - Xc, axis=1) / np.linalg.norm(Xc, axis=1)).max() < 1e-4
    dropin.config["gpu_bootstrap"] = False
    try:
        T2, mask2 = init.getPoseEstimate(p1, p2)
        assert np.array_equal(mask2 != 0, mask != 0) and np.abs(T2 - T).max() < 1e-6
    finally:
        dropin.config.pop("gpu_bootstrap", None)
