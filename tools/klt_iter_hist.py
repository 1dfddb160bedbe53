#!/usr/bin/env python
"""Distribution of Gauss-Newton iterations per KLT track on the bench sequence (unit-level API)."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
pkg = importlib.import_module(bench.PKG)
cfg = bench.CONFIGS['kitti']
sq = bench.build_stream(cfg, 11, 16)
ctx = pkg.Context(device=0, max_w=2048, max_h=1024)
ctx.upload_frame(sq.frames[10], 1); ctx.upload_frame(sq.frames[11], 2)
pts = np.concatenate([sq.top_P[10], sq.out_P[10][:1400], sq.feeds[10]]).astype(np.float32)
out, st, err, it = ctx.calc_optical_flow_pyr_lk(1, 2, pts, want_iters=True)
print("tracks", len(pts), "status ok", st.mean(), "iters mean %.1f" % it.mean())
print("percentiles 50/90/99/max:", np.percentile(it, [50, 90, 99]), it.max())
print("hist (bins of 10):", np.bincount(it // 10))
