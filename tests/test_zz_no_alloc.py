"""vo_solve_pnp_ransac_noalloc (no counterpart in the reference): the
PnP-RANSAC entry point that allocates nothing - all device scratch in one static device array of the library -
against the existing vo_solve_pnp_ransac on the same inputs, through the C-ABI.  Same kernels, same arithmetic: masks
and bookkeeping must be identical, poses within the 1e-8 the PnP tests use between two implementations.

Collected last on purpose; passes alone in a fresh process and after the others."""
import numpy as np
import pytest

import synth
from test_pnp_oracle import check_pose

pytestmark = pytest.mark.gpu
K = synth.K_KITTI


def params(pkg, H):
    p = pkg.pnp_defaults()
    p.max_iters = H
    return p


def same(a, b):
    assert a["ok"] and b["ok"]
    assert np.array_equal(a["mask"], b["mask"])
    assert (a["iters"], a["best_iter"], a["n_inliers"], a["lm_iters"]) == (b["iters"], b["best_iter"], b["n_inliers"], b["lm_iters"])
    check_pose(a["R"], a["t"], b["R"], b["t"], tol=1e-8)
    check_pose(a["Rh"], a["th"], b["Rh"], b["th"], tol=1e-8)
    assert np.allclose(a["rvec"], b["rvec"], atol=1e-9) and np.allclose(a["tvec"], b["tvec"], atol=1e-9)


def test_noalloc_pnp_equals_the_allocating_entry_point_twice_in_a_row(ctx, pkg):
    # two calls in a row on DIFFERENT problems, the second one smaller and harder (more iterations): whatever the
    # first call left in the shared scratch (hypotheses, counts, mask beyond the second call's sizes) must not show
    cases = [(3000, 0.30, 51, 200), (900, 0.55, 52, 1000)]
    want = []
    for n, of, seed, H in cases:
        X, P, _, _, _ = synth.pnp_scene(n, of, seed)
        want.append((X, P, H, ctx.solve_pnp_ransac(X, P, K, params(pkg, H))))
    assert want[1][3]["iters"] > want[0][3]["iters"]
    for X, P, H, ref in want:
        same(ctx.solve_pnp_ransac(X, P, K, params(pkg, H), noalloc=True), ref)
    # ... and the first problem once more, right behind the second
    X, P, H, ref = want[0]
    same(ctx.solve_pnp_ransac(X, P, K, params(pkg, H), noalloc=True), ref)
    # the allocating entry point still works on its own buffers afterwards
    same(ctx.solve_pnp_ransac(X, P, K, params(pkg, H)), ref)


def test_noalloc_pnp_with_a_callers_minimal_sets(ctx, pkg):
    X, P, _, _, _ = synth.pnp_scene(1500, 0.35, 31)
    rng = np.random.default_rng(3)
    sets = np.stack([rng.choice(1500, 4, replace=False) for _ in range(300)]).astype(np.int32)
    same(ctx.solve_pnp_ransac(X, P, K, params(pkg, 300), sets=sets, noalloc=True),
         ctx.solve_pnp_ransac(X, P, K, params(pkg, 300), sets=sets))


def test_noalloc_pnp_at_its_limits_and_beyond(ctx, pkg):
    nmax, hmax = pkg.NOALLOC_MAX_POINTS, pkg.NOALLOC_MAX_ITERS
    X, P, _, _, _ = synth.pnp_scene(nmax, 0.5, 53)
    same(ctx.solve_pnp_ransac(X, P, K, params(pkg, hmax), noalloc=True), ctx.solve_pnp_ransac(X, P, K, params(pkg, hmax)))
    X2, P2, _, _, _ = synth.pnp_scene(nmax + 1, 0.2, 54)
    with pytest.raises(pkg.VoError) as e:
        ctx.solve_pnp_ransac(X2, P2, K, params(pkg, 100), noalloc=True)
    assert e.value.code == pkg.VO_ERR_CAPACITY and str(nmax) in str(e.value)
    with pytest.raises(pkg.VoError) as e:
        ctx.solve_pnp_ransac(X, P, K, params(pkg, hmax + 1), noalloc=True)
    assert e.value.code == pkg.VO_ERR_CAPACITY and str(hmax) in str(e.value)
    # a refused call leaves the context usable
    same(ctx.solve_pnp_ransac(X, P, K, params(pkg, 100), noalloc=True), ctx.solve_pnp_ransac(X, P, K, params(pkg, 100)))


def test_noalloc_pnp_does_not_touch_device_memory_accounting(ctx, pkg):
    """The call must not allocate: free device memory (cudaMemGetInfo through the runtime the library links) is the
    same before and after, on a problem larger than anything this fresh context has sized its own buffers for."""
    import ctypes
    try:
        rt = ctypes.CDLL("libcudart.so")
    except OSError:
        import glob, os
        cands = sorted(glob.glob("/usr/local/cuda/lib64/libcudart.so*"))
        if not cands:
            pytest.skip("libcudart not found for cudaMemGetInfo")
        rt = ctypes.CDLL(cands[-1])
    c2 = pkg.Context(device=0, max_w=640, max_h=480, track_capacity=256, hyp_capacity=128)
    try:
        X, P, _, _, _ = synth.pnp_scene(4000, 0.4, 55)
        c2.solve_pnp_ransac(X[:64], P[:64], K, params(pkg, 50), noalloc=True)   # set-up of the scratch (once per device)
        free0, free1, total = ctypes.c_size_t(), ctypes.c_size_t(), ctypes.c_size_t()
        assert rt.cudaMemGetInfo(ctypes.byref(free0), ctypes.byref(total)) == 0
        r = c2.solve_pnp_ransac(X, P, K, params(pkg, 1000), noalloc=True)
        assert rt.cudaMemGetInfo(ctypes.byref(free1), ctypes.byref(total)) == 0
        assert free1.value == free0.value
        same(r, ctx.solve_pnp_ransac(X, P, K, params(pkg, 1000)))
    finally:
        c2.close()
