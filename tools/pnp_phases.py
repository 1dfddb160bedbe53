#!/usr/bin/env python
"""Phases of the PnP stage inside pnp_chunk_kernel (ns, %globaltimer stamps of the finishing CTA), averaged
over the bench sequence.  Usage: python tools/pnp_phases.py"""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
pkg = importlib.import_module(bench.PKG)
_, sq = bench.build_sequence(seed=11)
ctx = pkg.Context(device=0, max_w=2048, max_h=1024)
pnp = pkg.pnp_defaults(); pnp.max_iters = bench.PNP_ITERS
ctx.upload_frame(sq.frames[0], 1); ctx.set_state(sq.P0, sq.X0)
ctx.set_candidates(np.zeros((0, 2)), np.zeros((0, 2)), np.zeros((0, 16)))
order = bench.frame_order(len(sq.frames), 140)
acc = []; nl = nc = len(sq.P0), 0
n_land, n_cand = len(sq.P0), 0
for s in range(140):
    b = order[s + 1]
    feed = bench.pick_feed(sq, b, n_land, n_cand)
    T, cnt = ctx.process_frame(sq.frames[b], s + 1, s + 2, sq.K, 1.0, feed, pnp=pnp)
    n_land, n_cand = cnt["n_landmarks"], cnt["n_candidates"]
    if s >= 20:
        acc.append(ctx.pnp_stamps())
a = np.array(acc, float) / 1e3
names = ("P3P (this CTA)", "scoring", "wait for the other CTAs + select", "inlier mask", "LM refit", "pruning")
print("PnP phases of the finishing CTA, us (mean / median / max over %d frames)" % len(a))
for k, nme in enumerate(names):
    print("  %-34s %6.1f %6.1f %6.1f" % (nme, a[:, k].mean(), np.median(a[:, k]), a[:, k].max()))
print("  %-34s %6.1f" % ("sum", a.sum(axis=1).mean()))
ctx.close()
