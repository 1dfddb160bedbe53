#!/usr/bin/env python
"""PnP-RANSAC stress shape: time of one vo_solve_pnp_ransac call against the hypothesis budget and the point count
(budget pinned with confidence 1).  Usage: python tools/pnp_stress_scan.py"""
import ctypes, importlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import synth
import torch
pkg = importlib.import_module("locally-consistent-visual-odometry-pipeline_b200")
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream)
ctx = pkg.Context(device=0, max_w=64, max_h=64, stream=stream.cuda_stream, hyp_capacity=4096)
K = np.ascontiguousarray(synth.K_KITTI)
T = np.zeros((4, 4)); info = np.zeros(4, np.int32); ok = ctypes.c_int(0)
for n in (2000, 10000):
    X, P, _, _, _ = synth.pnp_scene(n, 0.5, 5)
    dX, dP = torch.from_numpy(X).to(dev), torch.from_numpy(np.ascontiguousarray(P, np.float32)).to(dev)
    for H in (1, 128, 129, 256, 1152, 2176, 4096):
        prm = pkg.pnp_defaults(); prm.max_iters = H; prm.confidence = 1.0
        ms = []
        for rep in range(12):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            r = ctx.lib.vo_solve_pnp_ransac(ctx.h, ctypes.c_void_p(dX.data_ptr()), ctypes.c_void_p(dP.data_ptr()), n, K.ctypes.data, ctypes.byref(prm),
                                            None, T.ctypes.data, None, None, None, None, info.ctypes.data, ctypes.byref(ok))
            b.record(stream); b.synchronize()
            ms.append(a.elapsed_time(b))
        st = ctx.pnp_stamps() / 1e3
        print("n %5d H %4d: %.3f ms (min %.3f); stamps us: p3p %.1f score %.1f rounds+select %.1f mask/prune %.1f LM %.1f; iters %d inliers %d"
              % (n, H, np.median(ms[2:]), min(ms), st[0], st[1], st[2], st[3], st[4], info[0], info[2]))
ctx.close()
