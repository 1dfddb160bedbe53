#!/usr/bin/env python
"""Host-side cost (us) of the enqueue-only C-ABI calls, measured with the GPU idle before each call
(the call returns as soon as the work is on the stream).  Usage: python tools/host_cost.py"""
import importlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

pkg = importlib.import_module(bench.PKG)
_, sq = bench.build_sequence(seed=11)
ctx = pkg.Context(device=0, max_w=2048, max_h=1024)
h, w = sq.frames[0].shape
pinned = [pkg.PinnedArray((h, w), np.uint8) for _ in range(8)]
for p, f in zip(pinned, sq.frames):
    p.array[...] = f
ctx.upload_frame(sq.frames[0], 1)
ctx.set_state(sq.P0, sq.X0)
ctx.set_candidates(np.zeros((0, 2)), np.zeros((0, 2)), np.zeros((0, 16)))


def timeit(name, fn, n=200, setup=None):
    ts = []
    for i in range(n):
        if setup:
            setup(i)
        ctx.synchronize()
        t0 = time.perf_counter()
        fn(i)
        ts.append(time.perf_counter() - t0)
        ctx.synchronize()
    ts = np.array(ts[10:]) * 1e6
    print("%-44s median %6.1f us   min %6.1f   p90 %6.1f" % (name, np.median(ts), ts.min(), np.percentile(ts, 90)))


timeit("upload_frame (pinned src, main stream)", lambda i: ctx.upload_frame(pinned[i % 8].array, 1000 + i))
timeit("prefetch_frame (pinned src, ingest stream)", lambda i: ctx.prefetch_frame(pinned[i % 8].array, 5000 + i))
timeit("upload_frame (pageable src)", lambda i: ctx.upload_frame(sq.frames[i % 8], 9000 + i))
K = sq.K
pnp = pkg.pnp_defaults()
pnp.max_iters = 2000
keys = {}


def prep(i):
    ctx.upload_frame(pinned[(i + 1) % 8].array, 20000 + i)
    ctx.upload_frame(pinned[i % 8].array, 30000 + i)
    ctx.set_state(sq.P0, sq.X0)


timeit("klt_track_landmarks (enqueue only)", lambda i: ctx.klt_track_landmarks(30000 + i, 20000 + i), setup=prep)
feed = sq.feeds[1][:200].copy()
T = np.eye(4)
timeit("klt_track_landmarks_add (200 rows)", lambda i: ctx.klt_track_landmarks(30000 + i, 20000 + i, add=(feed, T)), setup=prep)
timeit("add_candidates (200 rows, pageable)", lambda i: ctx.add_candidates(feed, T))
timeit("has_frame", lambda i: ctx.has_frame(1))
timeit("counters(sync=False)", lambda i: ctx.counters(sync=False))
ctx.close()
