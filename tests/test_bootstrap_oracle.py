"""CPU tests: the C restatement of the bootstrap's cv2 calls (oracle/fivept.c: findEssentialMat with the five-point
RANSAC, recoverPose) against golden cv2 outputs (tests/golden/bootstrap.npz, made by make_golden.py bootstrap) and,
when importable, cv2 itself.  Reference: VOInitializer.getPoseEstimate, /root/reference/src/vo_pipeline.py:139-162."""
import os

import numpy as np
import pytest

import synth

GOLD = os.path.join(os.path.dirname(__file__), "golden")
BOOT_CASES = {  # must match tests/golden/make_golden.py
    "n300_o20": (300, 0.2, 2, "kitti"), "n800_o30": (800, 0.3, 3, "kitti"), "n1500_o45": (1500, 0.45, 4, "kitti"),
    "n200_o0": (200, 0.0, 5, "kitti"), "n60_o10": (60, 0.1, 6, "parking"), "n2000_o20": (2000, 0.2, 7, "malaga"),
    "n500_o50": (500, 0.5, 8, "kitti"), "n5_exact": (5, 0.0, 9, "kitti"),
}
KS = {"kitti": synth.K_KITTI, "parking": synth.K_PARKING, "malaga": synth.K_MALAGA}


def boot_case(name):
    g = np.load(os.path.join(GOLD, "bootstrap.npz"))
    n, of, seed, kn = BOOT_CASES[name]
    K = KS[kn]
    p1, p2, R, t, is_out = synth.bootstrap_scene(n, of, seed, K=K, noise=0.0 if n == 5 else 0.3)
    assert abs(p1.sum() + p2.sum() - float(g[name + "_psum"])) < 1e-6, "synthetic input drifted from the golden generator"
    mask_E = np.unpackbits(g[name + "_mask_E"])[:n]
    mask_pose = np.unpackbits(g[name + "_mask_pose"])[:n]
    return p1, p2, K, g, mask_E, mask_pose, (R, t, is_out)


def e_dist(A, B):
    A = A / np.linalg.norm(A); B = B / np.linalg.norm(B)
    return min(np.linalg.norm(A - B), np.linalg.norm(A + B))


@pytest.mark.parametrize("name", list(BOOT_CASES))
def test_oracle_essential_ransac_vs_golden(oracle, name):
    p1, p2, K, g, mask_E, mask_pose, _ = boot_case(name)
    o = oracle.find_essential_mat(p1, p2, K)
    assert o["ok"]
    if len(p1) == 5:   # count == model points: one direct solve, cv2 returns every model stacked
        Eall = g[name + "_E_all"].reshape(-1, 3, 3)
        q1 = (p1 - K[:2, 2]) / [K[0, 0], K[1, 1]]
        q2 = (p2 - K[:2, 2]) / [K[0, 0], K[1, 1]]
        Es = oracle.five_point(q1, q2)
        assert len(Es) == len(Eall)
        for Ec in Eall:
            assert min(e_dist(E, Ec) for E in Es) < 1e-9
        return
    assert np.array_equal(o["mask"], mask_E), "hamming %d" % int((o["mask"] != mask_E).sum())
    assert e_dist(o["E"], g[name + "_E"]) < 1e-9
    good, R, t, m, which = oracle.recover_pose(g[name + "_E"], p1, p2, K, mask=mask_E)
    assert good == int(g[name + "_good"]) and np.array_equal(m, mask_pose)
    assert np.abs(R - g[name + "_R"]).max() < 1e-12 and np.abs(t - g[name + "_t"]).max() < 1e-12


def test_oracle_five_point_recovers_the_true_pose(oracle):
    p1, p2, R, t, _ = synth.bootstrap_scene(5, 0.0, 31, noise=0.0)
    K = synth.K_KITTI
    q1 = (p1 - K[:2, 2]) / [K[0, 0], K[1, 1]]
    q2 = (p2 - K[:2, 2]) / [K[0, 0], K[1, 1]]
    Es = oracle.five_point(q1, q2)
    tx = np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]])
    assert 1 <= len(Es) <= 10 and min(e_dist(E, tx @ R) for E in Es) < 1e-9
    for E in Es:   # every model satisfies the essential-matrix constraints and the five epipolar equations
        assert abs(np.linalg.det(E)) < 1e-10
        assert np.abs(2 * E @ E.T @ E - np.trace(E @ E.T) * E).max() < 1e-9
        for a, b in zip(q1, q2):
            assert abs(np.append(b, 1) @ E @ np.append(a, 1)) < 1e-10


def test_oracle_bootstrap_vs_cv2_live(oracle):
    cv2 = pytest.importorskip("cv2")
    for n, of, seed in [(400, 0.25, 41), (1000, 0.4, 42), (150, 0.05, 43)]:
        p1, p2, R, t, _ = synth.bootstrap_scene(n, of, seed, K=synth.K_MALAGA, w=1024, h=768)
        E, mask = cv2.findEssentialMat(p1, p2, synth.K_MALAGA, cv2.RANSAC, 0.999, 1.0, None)
        o = oracle.find_essential_mat(p1, p2, synth.K_MALAGA)
        assert np.array_equal(o["mask"], (mask.ravel() != 0).astype(np.uint8)) and e_dist(o["E"], E) < 1e-7   # (roots from two different polynomial solvers)
        m2 = mask.copy()
        good, Rc, tc, _ = cv2.recoverPose(E, p1, p2, synth.K_MALAGA, mask=m2)
        g2, Ro, to, mo, _ = oracle.recover_pose(E, p1, p2, synth.K_MALAGA, mask=mask)
        assert g2 == good and np.array_equal(mo, (m2.ravel() != 0).astype(np.uint8))
        assert np.abs(Ro - Rc).max() < 1e-12 and np.abs(to - tc.ravel()).max() < 1e-12
        # and the pose is the true one (translation up to scale)
        assert np.abs(Ro - R).max() < 2e-2 and np.linalg.norm(to - t / np.linalg.norm(t)) < 0.25


def test_oracle_bootstrap_edge_cases(oracle):
    K = synth.K_KITTI
    p1, p2, _, _, _ = synth.bootstrap_scene(40, 0.0, 3)
    assert not oracle.find_essential_mat(p1[:4], p2[:4], K)["ok"]          # fewer than 5 matches: no model (cv2 returns None)
    assert not oracle.find_essential_mat(p1[:0], p2[:0], K)["ok"]
    rng = np.random.default_rng(1)
    junk = rng.uniform(0, 300, p2.shape)
    o = oracle.find_essential_mat(p1, junk, K)                               # pure outliers: whatever wins has few inliers
    assert o["n_inliers"] < 20
