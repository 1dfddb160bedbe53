#!/usr/bin/env python
"""bench_micro.py - saturation micro-benchmarks of the hot path's kernels (regime 2).

bench.py measures the single-stream latency regime (~2 000 tracks per frame: less than one wave on
148 SMs).  This file measures the same kernels, through the same C-ABI entry points, at sizes that
fill the machine, so that a roofline fraction means something:

  klt        cv2.calcOpticalFlowPyrLK replacement, 262 144 tracks on one 1241x376 pair (KLT tracks/s -
             the second half of BASELINE.json's metric), algorithmic bytes per track as bench.py
  pyramid    upload + 4-level pyramid + Scharr planes per frame (HBM-bound: bytes written per frame)
  ransac     cv2.solvePnPRansac replacement, stress config: 10 000 correspondences, 50 % outliers,
             4096 P3P hypotheses, budget pinned (confidence 1.0) => hypothesis-point pairs/s
  dlt        cv2.triangulatePoints replacement, 1 Mi points

Inputs and outputs are device buffers (the C-ABI copies with cudaMemcpyDefault, so device pointers
are legal); each launch is bracketed by CUDA events on the ctx stream; one JSON line per kernel.
"""
import argparse
import ctypes
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
PKG = "locally-consistent-visual-odometry-pipeline_b200"


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


def timed(torch, stream, fn, reps, warmup=2):
    for _ in range(warmup):
        fn()
    ms = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        fn()
        b.record(stream)
        b.synchronize()
        ms.append(a.elapsed_time(b))
    return float(np.median(ms)), float(np.min(ms))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tracks", type=int, default=262144)
    ap.add_argument("--points", type=int, default=1 << 20)
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench_micro.py needs a CUDA device (no CPU path)")
    import synth
    pkg = importlib.import_module(PKG)
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    ctx = pkg.Context(device=0, max_w=2048, max_h=1024, stream=stream.cuda_stream, hyp_capacity=4096)
    lib, h = ctx.lib, ctx.h
    pk = peaks()
    hbm = pk.get("hbm_gbs", 6650.0)
    sm_mhz = pk.get("sm_max_mhz", 1965.0)
    fp32_peak = 148 * 128 * 2 * sm_mhz * 1e6 / 1e12      # TFLOP/s, CUDA-core FMA
    ptr = lambda t: ctypes.c_void_p(t.data_ptr())
    want = set(args.only.split(",")) if args.only else {"klt", "pyramid", "ransac", "dlt", "gftt", "map"}

    def check(r):
        if r != 0:
            raise RuntimeError(lib.vo_last_error(h).decode())

    if "klt" in want:
        n = args.tracks
        prev, nxt, _ = synth.klt_pair(376, 1241, 1, 16)
        rng = np.random.default_rng(7)
        pts = np.stack([rng.uniform(-5, 1241 + 5, n), rng.uniform(-5, 376 + 5, n)], 1).astype(np.float32)
        ctx.upload_frame(prev, 1)
        ctx.upload_frame(nxt, 2)
        d_pts = torch.from_numpy(pts).to(dev)
        d_out = torch.empty((n, 2), dtype=torch.float32, device=dev)
        d_st = torch.empty(n, dtype=torch.uint8, device=dev)
        d_err = torch.empty(n, dtype=torch.float32, device=dev)
        d_it = torch.empty(n, dtype=torch.int32, device=dev)
        prm = pkg.klt_defaults()

        def run():
            check(lib.vo_calc_optical_flow_pyr_lk(h, 1, 2, ptr(d_pts), n, ptr(d_out), ptr(d_st), ptr(d_err), ptr(d_it), ctypes.byref(prm)))
        med, best = timed(torch, stream, run, args.reps)
        iters = int(d_it.sum().item())
        ok = float(d_st.float().mean().item())
        alg = (4 * 576 + 505) * n + 484 * iters
        print(json.dumps({"kernel": "klt_track_kernel<21>", "regime": "saturation", "tracks": n, "ms": med, "ms_min": best,
                          "klt_tracks_per_s": n / (med * 1e-3), "lk_iterations_per_track": iters / n, "status_ok_frac": ok,
                          "algorithmic_bytes": alg, "achieved_gbs": alg / (med * 1e-3) / 1e9, "hbm_peak_gbs": hbm,
                          "frac_of_hbm": alg / (med * 1e-3) / 1e9 / hbm,
                          "note": "window gathers are L1/L2-resident (both pyramids are 1.6 MB); the GB/s figure is gather "
                                  "throughput expressed against the HBM copy peak, the binding unit is L1/LSU + integer issue"}))

    if "pyramid" in want:
        frames = [synth.texture(376, 1241, s) for s in range(4)]
        d_frames = [torch.from_numpy(f).to(dev) for f in frames]
        key = [100]

        def run():
            for f in d_frames:
                key[0] += 1
                check(lib.vo_upload_frame(h, ptr(f), 1241, 376, 1241, key[0]))
        med, best = timed(torch, stream, run, args.reps)
        # bytes per frame: raw read + D2D copy write, per level padded u8 write + short2 derivative write + source re-read
        tot = 0
        w, hh = 1241, 376
        for l in range(4):
            stride = ((w + 64 + 63) // 64) * 64
            tot += stride * (hh + 64) * (1 + 4) + w * hh
            w, hh = (w + 1) // 2, (hh + 1) // 2
        tot += 2 * 1241 * 376
        per = med / len(d_frames)
        print(json.dumps({"kernel": "build_level_kernel (4 levels) + frame copy", "regime": "back-to-back", "ms_per_frame": per,
                          "frames_per_s": 1e3 / per, "algorithmic_bytes_per_frame": tot, "achieved_gbs": tot / (per * 1e-3) / 1e9,
                          "hbm_peak_gbs": hbm, "frac_of_hbm": tot / (per * 1e-3) / 1e9 / hbm,
                          "note": "4 dependent launches of <=1148 CTAs on a 5.6 MB working set: launch-latency bound, not HBM bound"}))

    if "ransac" in want:
        n, H = 10000, 4096
        X, P, _, _, _ = synth.pnp_scene(n, 0.5, 5)
        d_X = torch.from_numpy(np.ascontiguousarray(X, np.float64)).to(dev)
        d_P = torch.from_numpy(np.ascontiguousarray(P, np.float32)).to(dev)
        d_mask = torch.empty(n, dtype=torch.uint8, device=dev)
        Kc = np.ascontiguousarray(synth.K_KITTI, np.float64)
        prm = pkg.pnp_defaults()
        prm.max_iters = H
        prm.confidence = 1.0     # the adaptive budget never shrinks below H: every hypothesis is scored
        info = np.zeros(4, np.int32)
        ok = ctypes.c_int()
        T44 = np.zeros(16)

        def run():
            check(lib.vo_solve_pnp_ransac(h, ptr(d_X), ptr(d_P), n, Kc.ctypes.data, ctypes.byref(prm), None, T44.ctypes.data,
                                          None, None, None, ptr(d_mask), info.ctypes.data, ctypes.byref(ok)))
        med, best = timed(torch, stream, run, args.reps)
        pairs = float(info[0]) * n
        flops = pairs * 35 + info[0] * 2000.0
        print(json.dumps({"kernel": "pnp_ransac_kernel (P3P + scoring + select + mask + prune) + pnp_refit_kernel", "regime": "stress config 4", "n": n,
                          "hypotheses_scored": int(info[0]), "inliers": int(info[2]), "lm_iters": int(info[3]), "ms": med, "ms_min": best,
                          "pairs_per_s": pairs / (med * 1e-3), "achieved_tflops": flops / (med * 1e-3) / 1e12,
                          "fp32_peak_tflops": fp32_peak, "frac_of_fp32_peak": flops / (med * 1e-3) / 1e12 / fp32_peak,
                          "note": "projection is FP64 in cv::projectPoints order (bit-exact masks), so the FP64 pipe, not FP32, bounds scoring"}))

    if "dlt" in want:
        n = args.points
        Xw, T1, T2, p1, p2 = synth.two_view_scene(4096, 3)
        reps_ = (n + 4095) // 4096
        p1 = np.tile(p1, (reps_, 1))[:n]
        p2 = np.tile(p2, (reps_, 1))[:n]
        M1 = np.ascontiguousarray(synth.K_KITTI @ T1[:3])
        M2 = np.ascontiguousarray(synth.K_KITTI @ T2[:3])
        d1 = torch.from_numpy(np.ascontiguousarray(p1, np.float64)).to(dev)
        d2 = torch.from_numpy(np.ascontiguousarray(p2, np.float64)).to(dev)
        dX = torch.empty((n, 4), dtype=torch.float64, device=dev)

        def run():
            check(lib.vo_triangulate_points(h, M1.ctypes.data, M2.ctypes.data, 0, ptr(d1), ptr(d2), n, ptr(dX)))
        med, best = timed(torch, stream, run, args.reps)
        byt = n * (32 + 32) * 2      # p1,p2 in + X4 out, each staged once more by the entry point's D2D copies
        print(json.dumps({"kernel": "triangulate_kernel (4x4 Jacobi SVD, FP64)", "regime": "saturation", "points": n, "ms": med,
                          "ms_min": best, "points_per_s": n / (med * 1e-3), "achieved_gflops_fp64": n * 3000.0 / (med * 1e-3) / 1e9,
                          "algorithmic_bytes": byt, "achieved_gbs": byt / (med * 1e-3) / 1e9, "hbm_peak_gbs": hbm}))
    if "gftt" in want:
        # cv2.goodFeaturesToTrack replacement on a KITTI-shaped frame, mask of 1500 points,
        # corners left on the device (the pipeline appends them there); cv2 on the host cores beside it
        import time
        scene = importlib.import_module(PKG + ".scene")
        img = scene.Sequence("kitti", n_frames=2, n_tracks=50, n_feed=10, seed=25).frames[1]
        rng = np.random.default_rng(3)
        pts = np.stack([rng.uniform(0, 1241, 1500), rng.uniform(0, 376, 1500)], 1).astype(np.float32)
        ctx.upload_frame(img, 900)
        d_pts = torch.from_numpy(pts).to(dev)
        d_out = torch.empty((500, 2), dtype=torch.float32, device=dev)
        prm = pkg.gftt_defaults()
        n = ctypes.c_int()
        ginfo = np.zeros(4, np.int32)

        def run():
            check(lib.vo_good_features_to_track(h, 900, None, ptr(d_pts), len(pts), ctypes.byref(prm), ptr(d_out), 500,
                                                ctypes.byref(n), None, ginfo.ctypes.data))
        med, best = timed(torch, stream, run, args.reps)
        line = {"kernel": "shi_eig + shi_nms + shi_select (Shi-Tomasi, 500 corners, minDistance 10)", "regime": "one 1241x376 frame",
                "ms": med, "ms_min": best, "corners": n.value, "candidates": int(ginfo[0]), "selection_rounds": int(ginfo[2]),
                "selection_kcycles": {"sort": int(ginfo[3] % 1000), "bucketing": int(ginfo[3] // 1000 % 1000), "rounds": int(ginfo[3] // 1000000)}, "frames_per_s": 1e3 / med,
                "algorithmic_bytes": 1241 * 376 * (1 + 1 + 4 + 4 + 1), "note": "u8 frame + mask in, float32 map out and back in for the NMS"}
        try:
            import cv2
            mask = np.full(img.shape, 255, np.uint8)
            mask[pts[:, 1].astype(int), pts[:, 0].astype(int)] = 0
            ts = []
            for _ in range(12):
                t0 = time.perf_counter()
                c = cv2.goodFeaturesToTrack(img, maxCorners=500, qualityLevel=0.03, minDistance=10, mask=mask, blockSize=4,
                                            gradientSize=3, useHarrisDetector=False, k=0.04)
                ts.append((time.perf_counter() - t0) * 1e3)
            line["cv2_ms"] = float(np.median(ts[2:]))
            line["cv2_threads"] = cv2.getNumThreads()
            line["same_corners_as_cv2"] = bool(np.array_equal(c.reshape(-1, 2), d_out[:n.value].cpu().numpy()))
        except ImportError:
            pass
        print(json.dumps(line))
    if "map" in want:
        # trajectory / map bookkeeping: main.py:239-255 per frame on the device, and the
        # headless plot; numpy (the reference's own calls, oracle/map_ref.py) on the host beside it
        import time
        from oracle import map_ref
        rng = np.random.default_rng(9)
        nl = 2000
        X = rng.uniform(-40, 40, size=(nl, 3))
        P = rng.uniform(0, 1000, size=(nl, 2)).astype(np.float32)
        ctx.set_state(P, X)
        T = np.eye(4); T[:3, 3] = (0.3, -0.1, 2.0)
        d_T = torch.from_numpy(T).to(dev)
        check(lib.vo_map_reset(h))

        def run():
            check(lib.vo_map_accumulate(h, ptr(d_T), 3.0, 50.0, 0, None))
        med, best = timed(torch, stream, run, 200)
        ref = map_ref.MapRef()
        ts = []
        for _ in range(200):
            t0 = time.perf_counter()
            ref.update(T, X)
            ts.append((time.perf_counter() - t0) * 1e3)
        traj, pts, _ = ctx.map_get()
        line = {"kernel": "map_accumulate_kernel", "regime": "2000 landmarks per frame, %d rows accumulated" % len(pts),
                "ms": med, "ms_min": best, "algorithmic_bytes": nl * 24 + int(len(pts) / 202) * 24,
                "numpy_ms": float(np.median(ts)), "note": "one launch per frame, no D2H; the reference reads all landmarks back and filters with numpy"}
        print(json.dumps(line))
        img = torch.empty((600, 1200, 3), dtype=torch.uint8, device=dev)
        bounds = np.zeros(4)

        def run2():
            check(lib.vo_map_render(h, ptr(img), 1200, 600, 40, 3, bounds.ctypes.data))
        med, best = timed(torch, stream, run2, 20)
        t0 = time.perf_counter()
        oimg, ob = map_ref.render(traj, pts, 1200, 600, 40, 3)
        t_np = (time.perf_counter() - t0) * 1e3
        print(json.dumps({"kernel": "map_bounds + map_clear + map_points + map_path", "regime": "1200x600 plot of %d landmarks, %d poses" % (len(pts), len(traj)),
                          "ms": med, "ms_min": best, "algorithmic_bytes": 1200 * 600 * 3 + len(pts) * 24,
                          "numpy_ms": t_np, "same_pixels_as_oracle": bool(np.array_equal(oimg, img.cpu().numpy()))}))
    ctx.close()


if __name__ == "__main__":
    main()
