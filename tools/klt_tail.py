#!/usr/bin/env python
"""How much of the landmark KLT launch (latency regime) is the tail of slow-converging tracks?
Times the unit-level entry point (device buffers, CUDA events) on the bench frames for the landmark
set, and again without the tracks above an iteration percentile."""
import ctypes, importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import torch
pkg = importlib.import_module(bench.PKG)
_, sq = bench.build_sequence(seed=11)
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream)
ctx = pkg.Context(device=0, max_w=2048, max_h=1024, stream=stream.cuda_stream)
ctx.upload_frame(sq.frames[10], 1); ctx.upload_frame(sq.frames[11], 2)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
prm = pkg.klt_defaults()
ptr = lambda t: ctypes.c_void_p(t.data_ptr())

def run(pts, label, cold):
    n = len(pts)
    d_pts = torch.from_numpy(np.ascontiguousarray(pts, np.float32)).to(dev)
    d_out = torch.empty((n, 2), dtype=torch.float32, device=dev); d_st = torch.empty(n, dtype=torch.uint8, device=dev)
    d_it = torch.empty(n, dtype=torch.int32, device=dev)
    ms = []
    for r in range(12):
        if cold:
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        ctx.lib.vo_calc_optical_flow_pyr_lk(ctx.h, 1, 2, ptr(d_pts), n, ptr(d_out), ptr(d_st), None, ptr(d_it), ctypes.byref(prm))
        b.record(stream); b.synchronize()
        ms.append(a.elapsed_time(b))
    it = d_it.cpu().numpy()
    print("%-34s n=%5d %s  median %.1f us  min %.1f us   iters mean %.1f p99 %d max %d" % (
        label, n, "cold" if cold else "warm", 1e3 * np.median(ms[2:]), 1e3 * np.min(ms[2:]), it.mean(), np.percentile(it, 99), it.max()))
    return it

P = sq.P0[:400].astype(np.float32)
for cold in (True, False):
    it = run(P, "landmarks", cold)
    for cut in (60, 40, 25):
        run(P[it <= cut], "landmarks, iters <= %d" % cut, cold)
    run(P[:1], "one track", cold)
    run(P[np.argmax(it):np.argmax(it) + 1], "slowest track alone", cold)
print("hist (bins of 10):", np.bincount(it // 10))
ctx.close()
