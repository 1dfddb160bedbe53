#!/usr/bin/env python
"""Phases of the PnP stage inside the RANSAC kernels (ns, %globaltimer stamps of the finishing CTA), averaged
over the bench workload (normal frames and burst frames separately).  Usage: python tools/pnp_phases.py [config]"""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
pkg = importlib.import_module(bench.PKG)
cfg = bench.CONFIGS[sys.argv[1] if len(sys.argv) > 1 else "kitti"]
N = 120
sq = bench.build_stream(cfg, 11, N)
ctx = pkg.Context(device=0, max_w=2048, max_h=1024)
pnp = pkg.pnp_defaults(); pnp.max_iters = cfg["pnp_iters"]
ctx.upload_frame(sq.frames[0], 1); ctx.set_state(sq.P0, sq.X0)
ctx.set_candidates(np.zeros((0, 2)), np.zeros((0, 2)), np.zeros((0, 16)))
acc, heavy, info = [], [], []
for s in range(N):
    b = s + 1
    for bP, bX, target in bench.batches_for(cfg, sq, s, b):
        ctx.append_landmarks(bP, bX, target, queue=True)
    T, cnt = ctx.process_frame(sq.frames[b], s + 1, s + 2, sq.K, 1.0, sq.feeds[b], pnp=pnp)
    if s >= 24:
        st = np.array(ctx.pnp_stamps(), float)
        (heavy if cnt["ransac_iters"] >= 100 else acc).append(st)
        info.append((cnt["n_tracked"], cnt["ransac_iters"], cnt["lm_iters"]))
names = ("round 0: P3P (finishing CTA)", "round 0: scoring", "wait + select (+ later rounds)", "mask + prune + inlier pack (to refit start)", "LM refit (own kernel)")
for label, a in (("normal frames", acc), ("burst frames (>= 100 RANSAC iterations)", heavy)):
    if not a:
        continue
    a = np.array(a) / 1e3
    d = a[:, :5]     # (pnp_stamps already returns the differences between consecutive stamps)
    print("PnP phases of the finishing CTA, us (mean / median / max over %d %s)" % (len(a), label))
    for k, nme in enumerate(names):
        print("  %-46s %6.1f %6.1f %6.1f" % (nme, d[:, k].mean(), np.median(d[:, k]), d[:, k].max()))
    print("  %-46s %6.1f" % ("sum", d.sum(axis=1).mean()))
i = np.array(info)
print("points %.0f, iterations %.1f, LM iterations %.2f (means)" % (i[:, 0].mean(), i[:, 1].mean(), i[:, 2].mean()))
ctx.close()
