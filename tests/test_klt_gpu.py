"""GPU parity tests (run on the B200 box): CUDA KLT through the C-ABI against the golden vectors
(cv2 4.13 outputs), the C oracle and, when importable, cv2 itself.
Gates (BASELINE.json north_star): positions within 0.05 px on tracks both mark ok, status flags
equal on >= 99.5 % of tracks."""
import os

import numpy as np
import pytest

import synth
from conftest import parity_report
from test_klt_oracle import GOLD, KLT_CASES, check_klt

pytestmark = pytest.mark.gpu


def run_cuda(ctx, prev, nxt, pts, key0):
    ctx.upload_frame(prev, key0)
    ctx.upload_frame(nxt, key0 + 1)
    return ctx.calc_optical_flow_pyr_lk(key0, key0 + 1, pts, want_iters=True)


def test_pyramid_bit_exact(ctx, oracle):
    for i, (h, w) in enumerate([(376, 1241), (75, 101), (48, 64)]):
        img = synth.texture(h, w, 30 + i)
        ctx.upload_frame(img, 9000 + i)
        ref = img
        for level in range(4):
            got, der = ctx.pyramid_level(9000 + i, level)
            assert np.array_equal(got, ref), "level %d differs" % level
            assert np.array_equal(der, oracle.scharr(ref)), "Scharr plane of level %d differs" % level
            if min(ref.shape) <= 2:
                break
            ref = oracle.pyr_down(ref)


@pytest.mark.parametrize("name", list(KLT_CASES))
def test_klt_vs_golden_and_oracle(ctx, oracle, name):
    h, w, seed, n = KLT_CASES[name]
    g = np.load(os.path.join(GOLD, "klt_%s.npz" % name))
    prev, nxt, pts = synth.klt_pair(h, w, seed, n)
    out, st, err, iters = run_cuda(ctx, prev, nxt, pts, 100 + seed * 2)
    # against cv2 4.13 (golden)
    agree, dmax = check_klt((out, st), (g["next_pts"], g["status"]))
    # against the C oracle: same structure, so tighter
    o_out, o_st, o_err, o_it = oracle.calc_optical_flow_pyr_lk(prev, nxt, pts, want_iters=True)
    check_klt((out, st), (o_out, o_st), pos_tol=2e-3, status_frac=0.999)
    ok = (st == 1) & (o_st == 1)
    assert np.abs(err - o_err)[ok].max() < 5e-3
    assert abs(int(iters.sum()) - int(o_it.sum())) <= max(4, 0.002 * o_it.sum())
    print("%s: status agreement vs cv2 %.4f, max |dp| %.2e px" % (name, agree, dmax))
    o_agree, o_dmax = check_klt((out, st), (o_out, o_st), pos_tol=2e-3, status_frac=0.999)
    parity_report("klt", name, dict(size="%dx%d" % (w, h), tracks=n, status_agreement_vs_cv2=float(agree), max_dp_px_vs_cv2=float(dmax),
                                    status_agreement_vs_c_oracle=float(o_agree), max_dp_px_vs_c_oracle=float(o_dmax),
                                    golden_cv2=str(g["cv2_version"]), gates="0.05 px, 99.5 % status"))


def test_klt_vs_cv2_live(ctx):
    cv2 = pytest.importorskip("cv2")
    for i, (h, w, n, shift) in enumerate([(376, 1241, 2000, (5.4, -1.3)), (768, 1024, 3000, (-7.9, 3.2)),
                                          (480, 640, 500, (11.2, 0.4))]):
        prev, nxt, pts = synth.klt_pair(h, w, 50 + i, n, shift=shift)
        ref = cv2.calcOpticalFlowPyrLK(prev, nxt, pts, None)
        out, st, err, _ = run_cuda(ctx, prev, nxt, pts, 300 + 2 * i)
        check_klt((out, st), (ref[0], ref[1]))


def test_klt_edge_cases(ctx, oracle):
    prev, nxt, _ = synth.klt_pair(120, 160, 7, 1)
    ctx.upload_frame(prev, 500)
    ctx.upload_frame(nxt, 501)
    # empty input (cv2 returns (None, None, None); the C-ABI returns empty arrays)
    out, st, err = ctx.calc_optical_flow_pyr_lk(500, 501, np.zeros((0, 2), np.float32))
    assert out.shape == (0, 2) and st.shape == (0,)
    # points far outside the frame lose track exactly as in the reference
    pts = np.float32([[-40, -40], [500, 500], [80, 60], [-10.5, 30.25], [159.9, 119.9], [0, 0]])
    out, st, err = ctx.calc_optical_flow_pyr_lk(500, 501, pts)
    o_out, o_st, _ = oracle.calc_optical_flow_pyr_lk(prev, nxt, pts)
    assert np.array_equal(st, o_st)
    assert np.abs(out - o_out)[st == 1].max() < 2e-3 if (st == 1).any() else True
    # identical frames: zero motion
    out, st, err = ctx.calc_optical_flow_pyr_lk(500, 500, np.float32([[80, 60], [40, 90]]))
    assert np.abs(out - np.float32([[80, 60], [40, 90]])).max() < 1e-3
    # unknown frame key fails loudly
    with pytest.raises(Exception):
        ctx.calc_optical_flow_pyr_lk(500, 99999, pts)


def test_klt_saturation_linearity(ctx):
    """Full-size property test: 64k tracks = the 2k golden tracks tiled 32x must give 32 identical
    copies (the result of a track is independent of grid position / neighbours)."""
    h, w, seed, n = KLT_CASES["kitti"]
    prev, nxt, pts = synth.klt_pair(h, w, seed, n)
    big = np.tile(pts, (32, 1))
    out, st, err, _ = run_cuda(ctx, prev, nxt, big, 100 + seed * 2)
    out = out.reshape(32, n, 2)
    st = st.reshape(32, n)
    assert np.array_equal(out, np.broadcast_to(out[0], out.shape))
    assert np.array_equal(st, np.broadcast_to(st[0], st.shape))


def test_klt_tma_staging_is_bit_identical(pkg, monkeypatch):
    """VO_KLT_TMA=1: the window tile staged by one cp.async.bulk.tensor.2d per warp and level (64 B x 32 rows, origin on a
    16 B boundary, mbarrier wait) instead of the __ldg loop.  Same bytes in shared memory, so the same results - positions,
    status, error and iteration counts, bit for bit - on every golden shape, with points outside the frame and across slot
    reuse with another geometry."""
    c = pkg.Context(device=0, max_w=2048, max_h=1024)
    try:
        key = 40
        for name in ("kitti", "small", "malaga", "tiny", "parking", "kitti"):
            h, w, seed, n = KLT_CASES[name]
            prev, nxt, pts = synth.klt_pair(h, w, seed, n)
            c.upload_frame(prev, key)
            c.upload_frame(nxt, key + 1)
            res = []
            for on in ("0", "1"):
                monkeypatch.setenv("VO_KLT_TMA", on)
                res.append(c.calc_optical_flow_pyr_lk(key, key + 1, pts, want_iters=True))
            for a, b in zip(*res):
                assert np.array_equal(a, b), name
            key += 2
    finally:
        c.close()
