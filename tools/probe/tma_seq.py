import sys, importlib, os, numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import synth
pkg = importlib.import_module("locally-consistent-visual-odometry-pipeline_b200")
ctx = pkg.Context(device=0, max_w=2048, max_h=1024)
def cmp(prev, nxt, pts, k):
    ctx.upload_frame(prev, k); ctx.upload_frame(nxt, k + 1)
    os.environ["VO_KLT_TMA"] = "1"; a = ctx.calc_optical_flow_pyr_lk(k, k + 1, pts, want_iters=True)
    os.environ["VO_KLT_TMA"] = "0"; b = ctx.calc_optical_flow_pyr_lk(k, k + 1, pts, want_iters=True)
    print(prev.shape, "key", k, "equal:", np.array_equal(a[0], b[0]), np.array_equal(a[1], b[1]), "n diff", int((a[0] != b[0]).any(1).sum()), "of", len(pts))
k = 10
for shape in [(188, 620), (240, 320), (376, 1241), (240, 320), (75, 101), (240, 320), (768, 1024), (240, 320), (188, 620), (240, 320)]:
    prev, nxt, pts = synth.klt_pair(shape[0], shape[1], 21, 600, shift=(-7.5, 5.2))
    cmp(prev, nxt, pts, k); k += 2
