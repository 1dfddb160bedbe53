"""CPU tests: the LK oracle (oracle/lk.c) against cv2 golden vectors and, when cv2 is importable,
against cv2 itself.  Gates from BASELINE.json north_star: <=0.05 px on tracks both sides mark ok,
status equal on >=99.5 % (the restatement actually reaches 100 % / 2e-4 px)."""
import os

import numpy as np
import pytest

import synth

GOLD = os.path.join(os.path.dirname(__file__), "golden")

KLT_CASES = {  # name: (h, w, seed, n)
    "kitti": (376, 1241, 1, 2000),
    "parking": (480, 640, 2, 500),
    "malaga": (768, 1024, 3, 3000),
    "small": (120, 160, 4, 300),
    "tiny": (40, 60, 5, 100),
}


def check_klt(got, ref, pos_tol=0.05, status_frac=0.995):
    gp, gs = got[0], got[1].ravel()
    rp, rs = ref[0], ref[1].ravel()
    agree = (gs == rs).mean()
    both = (gs == 1) & (rs == 1)
    d = np.linalg.norm(gp - rp, axis=1)
    assert agree >= status_frac, "status agreement %.4f" % agree
    assert both.sum() > 0
    assert d[both].max() <= pos_tol, "max position error %.5f px" % d[both].max()
    return agree, d[both].max()


@pytest.mark.parametrize("name", list(KLT_CASES))
def test_oracle_lk_vs_golden(oracle, name):
    h, w, seed, n = KLT_CASES[name]
    g = np.load(os.path.join(GOLD, "klt_%s.npz" % name))
    prev, nxt, pts = synth.klt_pair(h, w, seed, n)
    assert np.array_equal(pts, g["pts"]), "synthetic input drifted from the golden generator"
    assert int(prev.astype(np.int64).sum()) == int(g["prev_sum"]) and int(nxt.astype(np.int64).sum()) == int(g["next_sum"])
    out, st, err = oracle.calc_optical_flow_pyr_lk(prev, nxt, pts)
    agree, dmax = check_klt((out, st), (g["next_pts"], g["status"]), pos_tol=1e-3, status_frac=1.0)
    ok = (st == 1) & (g["status"].ravel() == 1)
    assert np.abs(err - g["err"].ravel())[ok].max() < 5e-3


def test_oracle_pyrdown_vs_golden(oracle):
    g = np.load(os.path.join(GOLD, "pyrdown.npz"))
    img = synth.texture(75, 101, 11)
    lv = img
    for l in range(1, 4):
        lv = oracle.pyr_down(lv)
        assert np.array_equal(lv, g["l%d" % l])


def test_oracle_lk_vs_cv2_live(oracle):
    cv2 = pytest.importorskip("cv2")
    prev, nxt, pts = synth.klt_pair(240, 320, 21, 400, shift=(4.6, -2.2))
    ref = cv2.calcOpticalFlowPyrLK(prev, nxt, pts, None)
    out, st, err = oracle.calc_optical_flow_pyr_lk(prev, nxt, pts)
    check_klt((out, st), (ref[0], ref[1]), pos_tol=1e-3, status_frac=1.0)
