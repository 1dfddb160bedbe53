"""Per-iteration latency of one KLT warp: one track (and one per SM) between unrelated noise frames, so that
it runs to the iteration cap on every level; compared with the same call between identical frames
(one step per level).  (t_many - t_one) / (iters_many - iters_one) = latency of one Gauss-Newton step."""
import importlib, sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("locally-consistent-visual-odometry-pipeline_b200")
cabi = importlib.import_module("locally-consistent-visual-odometry-pipeline_b200._cabi")
rng = np.random.default_rng(5)
import cv2
a = cv2.GaussianBlur(rng.integers(0, 256, (376, 1241), dtype=np.uint8), (0, 0), 2.0)
b = cv2.GaussianBlur(rng.integers(0, 256, (376, 1241), dtype=np.uint8), (0, 0), 2.0)
ctx = cabi.Context(0)
ctx.upload_frame(a, 1); ctx.upload_frame(b, 2)
def run(k0, k1, pts, reps=300):
    ts = []
    for _ in range(reps):
        t = time.perf_counter(); out = ctx.calc_optical_flow_pyr_lk(k0, k1, pts, want_iters=True); ts.append(time.perf_counter() - t)
    return np.median(ts) * 1e6, out[3]
for n in (1, 148, 592, 2000):
    pts = np.stack([rng.uniform(100, 1100, n), rng.uniform(60, 300, n)], 1).astype(np.float32)
    t1, i1 = run(1, 1, pts)
    t2, i2 = run(1, 2, pts)
    print("n=%d  same frames: %.1f us, iters max %d | noise frames: %.1f us, iters max %d mean %.1f | per step %.3f us"
          % (n, t1, i1.max(), t2, i2.max(), i2.mean(), (t2 - t1) / max(1, i2.max() - i1.max())))
