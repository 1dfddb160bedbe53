#!/usr/bin/env python
"""Pipelined vs synchronous frames on the bench stream (device-resident inputs, as bench.run_gpu's value leg):
first step whose pose differs, with the counters around it.  Usage: python tools/pipe_vs_sync.py [steps]"""
import ctypes, importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import torch
pkg = importlib.import_module(bench.PKG)
cfg = bench.CONFIGS["kitti"]
N = int(sys.argv[1]) if len(sys.argv) > 1 else 144
sq = bench.build_stream(cfg, 11, N + 4)
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream)
ctx = pkg.Context(device=0, max_w=2048, max_h=1024, stream=stream.cuda_stream)
h, w = sq.frames[0].shape
d_frames = torch.from_numpy(np.stack(sq.frames)).to(dev)
up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev) if len(a) else None
d_feeds = [up(f) for f in sq.feeds]
d_top = [(up(p), up(x)) for p, x in zip(sq.top_P, sq.top_X)]
d_out = [(up(p), up(x)) for p, x in zip(sq.out_P, sq.out_X)]
lib, hnd = ctx.lib, ctx.h
pnp = pkg.pnp_defaults(); pnp.max_iters = cfg["pnp_iters"]; klt = pkg.klt_defaults()
Kc = np.ascontiguousarray(sq.K, np.float64); T44 = np.zeros((4, 4)); cnt = pkg.Counters()
vp = lambda t: ctypes.c_void_p(t.data_ptr())
order = bench.frame_order(len(sq.frames), N + 2)

def run(pipelined):
    ctx.drop_frames(); ctx.upload_frame(sq.frames[0], 1); ctx.set_state(sq.P0, sq.X0)
    ctx.set_candidates(np.zeros((0, 2)), np.zeros((0, 2)), np.zeros((0, 16)))
    out = []
    for s in range(N):
        b = order[s + 1]
        assert lib.vo_prefetch_frame(hnd, vp(d_frames[order[s + 2]]), w, h, w, s + 3) == 0
        tp, tx = d_top[b]
        if tp is not None:
            assert lib.vo_queue_landmarks(hnd, vp(tp), vp(tx), tp.shape[0], cfg["n_land"]) == 0
        if s % cfg["burst_every"] == cfg["burst_every"] - 1 and d_out[b][0] is not None:
            assert lib.vo_queue_landmarks(hnd, vp(d_out[b][0]), vp(d_out[b][1]), d_out[b][0].shape[0], 0) == 0
        fd = d_feeds[b]; nn = 0 if fd is None else fd.shape[0]
        fn = lib.vo_process_frame_pipelined if pipelined else lib.vo_process_frame
        r = fn(hnd, vp(d_frames[b]), w, h, w, s + 1, s + 2, Kc.ctypes.data, ctypes.byref(klt), ctypes.byref(pnp), bench.ALPHA,
               vp(fd) if nn else None, nn, T44.ctypes.data, ctypes.byref(cnt))
        assert r == 0, lib.vo_last_error(hnd).decode()
        out.append((T44.copy(), cnt.n_tracked, cnt.n_inliers, cnt.ransac_iters, cnt.n_landmarks, cnt.n_candidates, cnt.n_promoted, cnt.lm_iters))
    return out

a = run(True); b = run(False)
bad = [k for k in range(N) if not np.array_equal(a[k][0], b[k][0])]
print("steps", N, "differing poses", len(bad), "first", bad[:5])
for k in (bad[:1] and range(max(bad[0] - 2, 0), min(bad[0] + 3, N))):
    print(k, "pipe", a[k][1:], "sync", b[k][1:], "dT", float(np.abs(a[k][0] - b[k][0]).max()))
ctx.close()
