#!/usr/bin/env python
"""bench.py - VO frames/s of the per-frame hot path on the workloads BASELINE.json names.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--config kitti|parking|malaga|pnp_stress] [--impl reference]

A step is one frame through the whole hot path: KLT of landmarks and candidates (4-level pyramid, 21x21
window), PnP-RANSAC (P3P hypotheses, LM refit, inlier pruning), candidate promotion (bearing test + DLT
triangulation), the append of that frame's candidate feed and of its landmark batches.

Workload (SURVEY.md 8a sizing, default --config kitti = BASELINE configs[1]): a synthetic forward drive
(ray-cast corridor, known trajectory), |P| held at n_land landmarks in a STATIONARY regime: after every
frame the landmarks are topped up to n_land from a prior map (true 3-D points of fresh corners of that frame:
map-based relocalisation, the reference's own np.vstack of re-bootstrap landmarks, vo_pipeline.py:318-320);
every `burst_every`-th frame a stale part of that map comes with it (burst_n rows whose 3-D points belong to
other pixels), so that PnP-RANSAC has ~47 % outliers to reject on that share of the frames and runs >100 of
its iterations there.  Candidates come from a pre-generated feed (detectors are outside the timed region,
SURVEY.md 8d) and are promoted by the bearing test + DLT as in the reference.  All arms (this GPU path, the
drop-in classes, the CPU port) get identical frames, feeds and batches; how many top-up rows are taken is
decided by the same rule in each (on the device for the GPU path).  `preroll` untimed frames bring the state
into the steady regime before the warm-up, in every arm.

  value  frames/s, frames / feed / batches already resident in HBM, vo_process_frame_pipelined per step (the
         host waits for each frame's pose; the tracking of the next, already ingested frame is enqueued behind
         the current frame's promotion), CUDA events per step on the ctx stream, L2 flushed between steps
  e2e    frames/s through the drop-in vo_pipeline classes in main.py's order, frames in pinned HOST memory
         (H2D every step), feed and batches in pageable host memory, pose read back every step; wall clock
         between syncs; MEDIAN of `e2e_windows` consecutive windows of K steps
  parity the same frames through the CPU port (oracle/ref_pipeline.py over cv2) in the same process: per-frame
         pose / inlier count / |P| / |C| against the GPU path (synchronous leg)
  roofline   the KLT kernel (dominant launch): algorithmic bytes per launch / its CUDA-event time
  cpu_baseline / --impl reference   the port of the reference's three methods over cv2 on the host cores

N > 1 (torchrun): one independent stream per GPU, no collective on the data path (weak scaling); the
reference arm runs N independent streams on the host cores (cores // N threads each).
"""
import argparse
import gc
import importlib
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "locally-consistent-visual-odometry-pipeline_b200"

ALPHA = 1.0
PREROLL = 24           # untimed frames that bring the state into the steady regime (every arm)
F_MAX = 384            # rendered frames per stream; longer runs traverse them ping-pong
E2E_WINDOWS = 9

CONFIGS = {
    # BASELINE.json configs[1] - the configuration the metric is quoted on
    "kitti": dict(workload="kitti_1241x376_P2000_pyr4_win21_pnp2000", shape="kitti", n_land=2000, n_feed=150,
                  burst_every=8, burst_n=1700, topup_max=600, pnp_iters=2000, scene={}),
    # configs[0]: parking-shaped, sideways motion in front of a wall
    "parking": dict(workload="parking_640x480_P500_pyr4_win21_pnp100", shape="parking", n_land=500, n_feed=60,
                    burst_every=8, burst_n=420, topup_max=200, pnp_iters=100,
                    scene=dict(step=0.08, heading=float(np.pi / 2), z0=25.0, rate_lo=0.02)),
    # configs[2]: Malaga-shaped, 3k tracks with continuous triangulation
    "malaga": dict(workload="malaga_1024x768_P3000_pyr4_win21_pnp2000", shape="malaga", n_land=3000, n_feed=250,
                   burst_every=8, burst_n=2500, topup_max=800, pnp_iters=2000, scene={}),
}
STRESS = dict(workload="pnp_stress_10000pts_50pct_outliers_4096hyp", n=10000, outliers=0.5, hyp=4096)


def frame_order(n_frames, steps):
    """0,1,...,F-1,F-2,...,1,0,1,...: consecutive frames are always 1 step of camera motion apart."""
    idx, k, d = [], 0, 1
    for _ in range(steps + 1):
        idx.append(k)
        if k + d < 0 or k + d >= n_frames:
            d = -d
        k += d
    return idx


def stream_seed(rank):
    """One independent sequence per rank: the stream a rank processes is a function
    of its rank only, so no rank needs anything from another one on the data path."""
    return 11 + rank


def build_stream(cfg, seed, n_steps, workers=None):
    scene = importlib.import_module(PKG + ".scene")
    nf = min(n_steps + 3, F_MAX)
    sq = scene.SteadyStream(cfg["shape"], n_frames=nf, n_tracks=cfg["n_land"], n_feed=cfg["n_feed"],
                            n_burst=max(cfg["burst_n"], cfg["topup_max"]), seed=seed, workers=workers, **cfg["scene"])
    # top-up rows: fresh corners of each frame with their TRUE 3-D points (a prior map); burst rows: the same
    # pixels with the 3-D points of other pixels (a stale part of the map)
    sq.top_P = [np.ascontiguousarray(p[:cfg["topup_max"]], np.float32) for p in sq.burst_P]
    sq.top_X = [np.ascontiguousarray(sq.scene.backproject(k, p), np.float64) for k, p in enumerate(sq.top_P)]
    sq.out_P = [np.ascontiguousarray(p[:cfg["burst_n"]], np.float32) for p in sq.burst_P]
    sq.out_X = [np.ascontiguousarray(x[:cfg["burst_n"]], np.float64) for x in sq.burst_X]
    return sq


def batches_for(cfg, sq, s, b):
    """Landmark batches appended after step s (frame index b): top-up to n_land, plus the stale rows on burst frames."""
    out = [(sq.top_P[b], sq.top_X[b], cfg["n_land"])]
    if cfg["burst_every"] > 0 and s % cfg["burst_every"] == cfg["burst_every"] - 1:
        out.append((sq.out_P[b], sq.out_X[b], 0))
    return out


class ClockSampler(threading.Thread):
    """SM clock / throttle reasons under load (NVML), sampled every ~1 ms while `active` (the device-timed legs)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None
        self.active = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
            nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
        }
        while not self.stop_flag:
            if self.active:
                try:
                    self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                    for bit, name in names.items():
                        if r & bit:
                            self.reasons.add(name)
                except Exception:
                    break
            time.sleep(0.001 if self.active else 0.005)

    def summary(self):
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def pose_delta(Ta, Tb):
    """(rotation angle between the two poses in rad, |dt| / |t|)."""
    R = Ta[:3, :3] @ Tb[:3, :3].T
    ang = float(np.arccos(np.clip((np.trace(R) - 1.0) * 0.5, -1.0, 1.0)))
    return ang, float(np.linalg.norm(Ta[:3, 3] - Tb[:3, 3]) / max(np.linalg.norm(Tb[:3, 3]), 1e-12))


# ---- CPU arm: the reference's three methods (port over cv2) ------------------------------------------
def run_port(cfg, sq, steps, warmup, threads, trace=None, raw_sample=None):
    """The port over `preroll + warmup + steps` frames; times the last `steps`.  trace: list that receives
    (T, n_inliers, |P|, |C|, ransac n/a) per frame from the first one; raw_sample: dict that receives the arrays of
    one timed frame for the raw-cv2 baseline."""
    import cv2
    from oracle import ref_pipeline as rp_mod
    cv2.setNumThreads(threads)
    rp = rp_mod.RefPipeline(sq.K, pnp="p3p", max_iters=cfg["pnp_iters"], alpha=ALPHA, row_match="isin")
    state = {"P": sq.P0.astype(np.float64), "X": sq.X0.copy()}
    ext = {"C": np.array([]), "F": np.array([]), "T": np.array([])}
    total = PREROLL + warmup + steps
    order = frame_order(len(sq.frames), total)
    t0 = None
    tracks = 0
    for s in range(total):
        if s == PREROLL + warmup:
            t0 = time.perf_counter()
            tracks = 0
        a, b = order[s], order[s + 1]
        tracks += len(state["P"]) + len(ext["C"])
        feed = sq.feeds[b].copy()
        if raw_sample is not None and s == PREROLL + warmup + 7 and not raw_sample:
            raw_sample.update(prev=sq.frames[a], next=sq.frames[b], P=state["P"].astype(np.float32).copy(), X=state["X"].copy())
        if trace is not None:    # staged, to keep the PnP input of every frame for the parity block
            assoc, _ = rp.associate(sq.frames[a], sq.frames[b], state)
            pnp_in = (assoc["P"].copy(), assoc["X"].copy())
            T = rp.estimate(assoc)
            state, ext = rp.triangulate(sq.frames[a], sq.frames[b], assoc, ext, T, feed)
        else:
            state, ext, T = rp.frame(sq.frames[a], sq.frames[b], state, ext, feed)   # (< 4 inliers would raise: not on this scene)
        n_inl = rp.info["n_inliers"]
        for bP, bX, target in batches_for(cfg, sq, s, b):
            state = rp_mod.append_landmarks(state, bP, bX, target)
        if trace is not None:
            trace.append((T.copy(), n_inl, len(state["P"]), len(ext["C"]), rp.info["n_promoted"], len(pnp_in[0]), pnp_in))
    dt = time.perf_counter() - t0
    return steps / dt, dt, tracks / steps, cv2.getNumThreads()


def _port_worker(args, q=None, idx=0):
    cfg, seed, steps, warmup, threads = args
    sq = build_stream(cfg, seed, PREROLL + warmup + steps, workers=max(threads, 1))
    fps, dt, tracks, thr = run_port(cfg, sq, steps, warmup, threads)
    if q is not None:
        q.put((idx, dt, tracks, thr))
    return dt, tracks, thr


def raw_cv2_baseline(cfg, sq, sample, cores):
    """Baseline B: the raw cv2 calls of the path on one timed frame's arrays,
    1 thread and all threads, median and min of 30 repetitions after 5 warm-up calls."""
    import cv2
    if not sample:
        return None
    P, X = sample["P"], sample["X"]
    K = sq.K
    out = {"n_points": int(len(P)), "reps": 30}
    for thr in (1, cores):
        cv2.setNumThreads(thr)
        res = {}

        def bench(fn):
            for _ in range(5):
                fn()
            ts = []
            for _ in range(30):
                t = time.perf_counter()
                fn()
                ts.append((time.perf_counter() - t) * 1e3)
            return {"median_ms": float(np.median(ts)), "min_ms": float(np.min(ts))}
        res["calcOpticalFlowPyrLK"] = bench(lambda: cv2.calcOpticalFlowPyrLK(sample["prev"], sample["next"], P, None))
        p1, st, _ = cv2.calcOpticalFlowPyrLK(sample["prev"], sample["next"], P, None)
        ok = st.ravel() == 1
        Pn, Xn = p1[ok], X[ok]
        res["solvePnPRansac_P3P"] = bench(lambda: cv2.solvePnPRansac(Xn, Pn, K, np.zeros(4), flags=cv2.SOLVEPNP_P3P, confidence=0.99999,
                                                                      reprojectionError=2.0, iterationsCount=cfg["pnp_iters"]))
        res["solvePnPRansac_literal"] = bench(lambda: cv2.solvePnPRansac(Xn, Pn, K, np.zeros(4), flags=cv2.SOLVEPNP_ITERATIVE,
                                                                          confidence=0.99999, reprojectionError=2.0))
        M1 = K @ np.eye(3, 4)
        M2 = K @ np.hstack([np.eye(3), np.array([[0.8], [0.0], [0.0]])])
        pts = np.ascontiguousarray(Pn[:256].T.astype(np.float64))
        res["triangulatePoints_256"] = bench(lambda: cv2.triangulatePoints(M1, M2, pts, pts + 1.0))
        out["threads_%d" % thr] = res
    cv2.setNumThreads(cores)
    return out


# ---- GPU arm ---------------------------------------------------------------------------------------
def barrier(world):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()


def run_gpu(args, cfg, rank, world, local_rank, sq):
    import ctypes
    import torch
    pkg = importlib.import_module(PKG)
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    ctx = pkg.Context(device=local_rank, max_w=2048, max_h=1024, stream=stream.cuda_stream)
    K = sq.K
    F = len(sq.frames)
    h, w = sq.frames[0].shape
    pnp = pkg.pnp_defaults()
    pnp.max_iters = cfg["pnp_iters"]
    klt = pkg.klt_defaults()
    steps, warmup = args.steps, args.warmup
    total = PREROLL + warmup + steps

    # inputs resident in HBM: all frames, the candidate feed, the landmark batches
    d_frames = torch.from_numpy(np.stack(sq.frames)).to(dev)
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev) if len(a) else None
    d_feeds = [up(f) for f in sq.feeds]
    d_top = [(up(p), up(x)) for p, x in zip(sq.top_P, sq.top_X)]
    d_out = [(up(p), up(x)) for p, x in zip(sq.out_P, sq.out_X)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)      # > 126 MB L2
    lib, hnd = ctx.lib, ctx.h
    T44 = np.zeros((4, 4))
    cnt = pkg.Counters()
    Kc = np.ascontiguousarray(K, np.float64)
    vp = lambda t: ctypes.c_void_p(t.data_ptr())

    def check(r):
        if r != 0:
            raise RuntimeError(lib.vo_last_error(hnd).decode())

    def reset_state():
        ctx.drop_frames()
        ctx.upload_frame(sq.frames[0], 1)
        ctx.set_state(sq.P0, sq.X0)
        ctx.set_candidates(np.zeros((0, 2)), np.zeros((0, 2)), np.zeros((0, 16)))

    def dev_step(s, order, pipelined):
        b = order[s + 1]
        if args.prefetch:
            # frame ingest: the NEXT step's frame goes to the ingest stream now, after this step's start event has
            # been recorded; on the GPU it runs under the chain that is executing at that moment
            check(lib.vo_prefetch_frame(hnd, vp(d_frames[order[s + 2]]), w, h, w, s + 3))
        tp, tx = d_top[b]
        if tp is not None:
            check(lib.vo_queue_landmarks(hnd, vp(tp), vp(tx), tp.shape[0], cfg["n_land"]))
        if cfg["burst_every"] > 0 and s % cfg["burst_every"] == cfg["burst_every"] - 1 and d_out[b][0] is not None:
            check(lib.vo_queue_landmarks(hnd, vp(d_out[b][0]), vp(d_out[b][1]), d_out[b][0].shape[0], 0))
        fd = d_feeds[b]
        nn = 0 if fd is None else fd.shape[0]
        fn = lib.vo_process_frame_pipelined if pipelined else lib.vo_process_frame
        check(fn(hnd, vp(d_frames[b]), w, h, w, s + 1, s + 2, Kc.ctypes.data, ctypes.byref(klt), ctypes.byref(pnp), ALPHA,
                 vp(fd) if nn else None, nn, T44.ctypes.data, ctypes.byref(cnt)))
        if not cnt.pnp_ok:
            raise RuntimeError("PnP failed at step %d" % s)

    order = frame_order(F, total + 2)
    sampler = ClockSampler(local_rank)
    sampler.start()

    # ---- value: device-resident inputs, per-step CUDA events, L2 flush between steps -------------------
    reset_state()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    poses_pipe = []
    launches0 = t_wall0 = None
    barrier(world)
    torch.cuda.synchronize()
    for s in range(total):
        t = s - (PREROLL + warmup)
        if t == 0:
            barrier(world)
            torch.cuda.synchronize()
            sampler.active = True
            launches0 = ctx.launch_count()
            t_wall0 = time.perf_counter()
        flush.zero_()
        if t >= 0:
            ev[t][0].record(stream)
        dev_step(s, order, args.prefetch)
        if t >= 0:
            ev[t][1].record(stream)
        poses_pipe.append(T44.copy())
    torch.cuda.synchronize()
    barrier(world)
    t_wall = time.perf_counter() - t_wall0
    launches = ctx.launch_count() - launches0
    step_ms = np.array([a.elapsed_time(b) for a, b in ev])
    total_ms = float(step_ms.sum())

    # ---- synchronous leg over the same steps: stage timers (CUDA events inside the library), final per-frame
    # counters and poses (what the parity block compares with the CPU port) -----------------------------------
    reset_state()
    ctx.set_profiling(False)
    trace = []
    acc = dict(land=0, cand=0, it_land=0, it_cand=0, lm=0, ransac=0, inl=0, prom=0, app=0, tracked=0, heavy=0)
    for s in range(total):
        if s == PREROLL + warmup:
            ctx.set_profiling(True)
        flush.zero_()
        n_land0, n_cand0 = (len(sq.P0), 0) if s == 0 else (cnt.n_landmarks, cnt.n_candidates)
        dev_step(s, order, False)
        trace.append((T44.copy(), cnt.n_inliers, cnt.n_landmarks, cnt.n_candidates, cnt.n_promoted, cnt.n_tracked))
        if s >= PREROLL + warmup:
            acc["land"] += n_land0; acc["cand"] += n_cand0
            acc["it_land"] += cnt.klt_iters; acc["it_cand"] += cnt.klt_iters_candidates
            acc["lm"] += cnt.lm_iters; acc["ransac"] += cnt.ransac_iters; acc["inl"] += cnt.n_inliers
            acc["prom"] += cnt.n_promoted; acc["app"] += cnt.n_appended; acc["tracked"] += cnt.n_tracked
            acc["heavy"] += cnt.ransac_iters >= 100
    stage_ms, nfr = ctx.stage_times()
    ctx.set_profiling(False)
    sampler.active = False
    sampler.stop_flag = True
    final_counts = (cnt.n_landmarks, cnt.n_candidates)
    pipe_equals_sync = all(np.array_equal(a, b[0]) for a, b in zip(poses_pipe, trace))
    ctx.close()
    del d_frames, flush

    # ---- e2e: the drop-in classes in main.py's order, frames in pinned host memory ---------------------
    e2e = run_e2e(pkg, cfg, sq, steps, warmup, local_rank, world, args.prefetch)
    nfr = max(nfr, 1)
    return dict(total_ms=total_ms, step_ms=step_ms, t_wall=t_wall, launches=launches, stage_ms=stage_ms,
                klt=dict(land_tracks=acc["land"] / nfr, land_iters=acc["it_land"] / nfr,
                         cand_tracks=acc["cand"] / nfr, cand_iters=acc["it_cand"] / nfr),
                clocks=sampler.summary(), e2e=e2e, final_counts=final_counts, trace=trace, pipe_equals_sync=pipe_equals_sync,
                per_frame=dict(landmarks_tracked_into_frame=acc["land"] / nfr, candidates_tracked_into_frame=acc["cand"] / nfr,
                               pnp_points=acc["tracked"] / nfr, pnp_inliers=acc["inl"] / nfr, ransac_iters=acc["ransac"] / nfr,
                               frames_with_100plus_ransac_iters=acc["heavy"] / nfr, lm_iters=acc["lm"] / nfr,
                               promoted=acc["prom"] / nfr, appended=acc["app"] / nfr))


def job_throughput(total_ms, e2e_fps, steps, world, device=None):
    """Whole-job frames/s from per-rank timings: the slowest rank defines the job's time (MAX over
    ranks of the device time and of the e2e time per frame), the job's work is world x steps frames.
    The only collective of the multi-GPU path; it is outside the timed region."""
    if world > 1:
        import torch
        import torch.distributed as dist
        t = torch.tensor([total_ms, 1.0 / e2e_fps], dtype=torch.float64, device=device or "cpu")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, e2e_fps = float(t[0]), 1.0 / float(t[1])
    return world * steps / (total_ms * 1e-3), world * e2e_fps, total_ms


def run_e2e(pkg, cfg, sq, steps, warmup, local_rank, world, prefetch=True):
    """Public API end to end: KeypointsToLandmarksAssociator / PoseEstimator / LandmarkTriangulator exactly as
    main.py:229-294 calls them; every step uploads its frame from pinned host memory, hands over the step's
    candidate feed and landmark batches from pageable host memory and reads the pose back (np.linalg.inv of it,
    as main.py:239 does).  E2E_WINDOWS consecutive windows of `steps` frames; the median window is reported."""
    import tempfile
    tmp = tempfile.mkdtemp(prefix="vo_bench_")
    with open(os.path.join(tmp, "config.yaml"), "w") as f:
        f.write("dataset: 'kitti'\nfind_new_candidates_method: 'feed'\ninit_detector: 'sift'\ninit_descriptor: 'sift'\n"
                "debug: False\nvisualization: False\nanimation: False\nsift_alpha: 1\nsift_ransac_confidence: 0.99999\n"
                "shi_alpha: 1.5\nshi_ransac_confidence: 0.999993\npnp_max_iters: %d\ndevice: %d\nframe_ahead: %s\n"
                % (cfg["pnp_iters"], local_rank, "True" if prefetch else "False"))
    cwd = os.getcwd()
    os.chdir(tmp)
    sys.path.insert(0, os.path.join(ROOT, PKG))
    try:
        sys.modules.pop("vo_pipeline", None)
        vo = importlib.import_module("vo_pipeline")
    finally:
        os.chdir(cwd)
    F = len(sq.frames)
    h, w = sq.frames[0].shape
    pinned = [pkg.PinnedArray((h, w), np.uint8) for _ in range(F)]
    for p, f in zip(pinned, sq.frames):
        p.array[...] = f
    total = PREROLL + warmup + steps * E2E_WINDOWS
    order = frame_order(F, total + 2)
    vision = vo.BestVision(sq.K)
    vision.state = {"P": sq.P0.astype(np.float64), "X": sq.X0.copy()}
    vision.extended_state = {"C": np.array([]), "F": np.array([]), "T": np.array([])}
    associate = vo.KeypointsToLandmarksAssociator(sq.K, np.eye(4))
    pose_estimator = vo.PoseEstimator(sq.K)
    h2d = d2h = 0
    img2 = None
    marks = []
    trace = [] if os.environ.get("VO_BENCH_TRACE") else None   # per-step host times to stderr (diagnostic)
    for s in range(total):
        t = s - (PREROLL + warmup)
        if t >= 0 and t % steps == 0:
            vo._runtime().ctx.synchronize()
            barrier(world)
            if t == 0:
                # as timeit does: no cyclic-GC pass inside the timed region (one full collection of an interpreter
                # with torch and cv2 loaded costs several ms and made this leg bimodal)
                gc.collect()
                gc.disable()
                h2d = d2h = 0
            marks.append(time.perf_counter())
        a, b = order[s], order[s + 1]
        # a fresh array object per step, as cv2.imread gives main.py a new frame every iteration
        img1 = pinned[a].array if s == 0 else img2
        img2 = pinned[b].array[:]
        feed = sq.feeds[b]
        bat = batches_for(cfg, sq, s, b)
        vo.set_candidate_feed([feed])
        vo.set_landmark_feed([bat])
        if prefetch:   # frame ingest: hand the next iteration's frame over before this one is processed
            vo.prefetch_frame(pinned[order[s + 2]].array)
        new_state, lost = associate.associateKeypointsToLandmarks(img1, img2, vision.state)
        new_pose = pose_estimator.estimatePose(new_state, s)
        pos = np.linalg.inv(new_pose)[0:3, 3]
        tri = vo.LandmarkTriangulator(sq.K, None)
        vision.state, vision.extended_state, _ = tri.triangulate_landmark(img1, img2, new_state, vision.extended_state, new_pose)
        h2d += w * h + feed.nbytes + 128 + sum(p.nbytes + x.nbytes for p, x, _ in bat)
        d2h += 256
        if trace is not None:
            trace.append(time.perf_counter())
    vo._runtime().ctx.synchronize()
    barrier(world)
    marks.append(time.perf_counter())
    gc.enable()
    if trace:
        d = np.diff(np.array(trace[PREROLL + warmup:])) * 1e6
        print("e2e step us: median %.1f p90 %.1f p99 %.1f max %.1f, steps > 2x median: %d"
              % (np.median(d), np.percentile(d, 90), np.percentile(d, 99), d.max(), int((d > 2 * np.median(d)).sum())), file=sys.stderr)
    vo._reset_runtime()
    sys.modules.pop("vo_pipeline", None)
    sys.path.remove(os.path.join(ROOT, PKG))
    win = steps / np.diff(np.array(marks))
    n = steps * E2E_WINDOWS
    return dict(fps=float(np.median(win)), windows=[float(x) for x in win], h2d=h2d / n, d2h=d2h / n)


def run_files_leg(pkg, cfg, sq, local_rank, n_frames=160, workers=None):
    """PNG files on disk -> poses, through the drop-in classes with FrameIngest(workers=k) decoding ahead
    (cv2.imread on k threads, in-order hand-off into a pinned ring).  Reported beside the headline, not as it: the
    decode, not the VO core, sets this number."""
    import shutil
    import tempfile
    import cv2
    scene = importlib.import_module(PKG + ".scene")
    workers = workers or max(min((os.cpu_count() or 2) - 2, 16), 1)
    tmp = tempfile.mkdtemp(prefix="vo_files_")
    n = min(n_frames, len(sq.frames))
    paths = scene.write_dataset(sq, tmp, "kitti", n)
    with open(os.path.join(tmp, "config.yaml"), "w") as f:
        f.write("dataset: 'kitti'\nfind_new_candidates_method: 'feed'\ninit_detector: 'sift'\ninit_descriptor: 'sift'\n"
                "debug: False\nvisualization: False\nanimation: False\nsift_alpha: 1\nsift_ransac_confidence: 0.99999\n"
                "shi_alpha: 1.5\nshi_ransac_confidence: 0.999993\npnp_max_iters: %d\ndevice: %d\nframe_ahead: True\n" % (cfg["pnp_iters"], local_rank))
    cwd = os.getcwd()
    os.chdir(tmp)
    sys.path.insert(0, os.path.join(ROOT, PKG))
    try:
        sys.modules.pop("vo_pipeline", None)
        vo = importlib.import_module("vo_pipeline")
    finally:
        os.chdir(cwd)
    cv2.setNumThreads(1)
    try:
        t0 = time.perf_counter()
        for _ in vo.FrameIngest(paths, prefetch=False, workers=workers):
            pass
        decode_fps = n / (time.perf_counter() - t0)
        t0 = time.perf_counter()
        for p in paths[:24]:
            cv2.imread(p, cv2.IMREAD_GRAYSCALE)
        one_thread_fps = 24 / (time.perf_counter() - t0)
        vision = vo.BestVision(sq.K)
        vision.state = {"P": sq.P0.astype(np.float64), "X": sq.X0.copy()}
        vision.extended_state = {"C": np.array([]), "F": np.array([]), "T": np.array([])}
        associate = vo.KeypointsToLandmarksAssociator(sq.K, np.eye(4))
        pose_estimator = vo.PoseEstimator(sq.K)
        img1 = None
        t0 = None
        for b, img2 in enumerate(vo.FrameIngest(paths, workers=workers)):
            if b == 0:
                img1 = img2
                continue
            if b == 9:     # the first frames fill the decode pipeline and warm the kernels up
                vo._runtime().ctx.synchronize()
                t0 = time.perf_counter()
            vo.set_candidate_feed([sq.feeds[b]])
            vo.set_landmark_feed([batches_for(cfg, sq, b - 1, b)])
            new_state, lost = associate.associateKeypointsToLandmarks(img1, img2, vision.state)
            new_pose = pose_estimator.estimatePose(new_state, b)
            pos = np.linalg.inv(new_pose)[0:3, 3]
            tri = vo.LandmarkTriangulator(sq.K, None)
            vision.state, vision.extended_state, _ = tri.triangulate_landmark(img1, img2, new_state, vision.extended_state, new_pose)
            img1 = img2
        vo._runtime().ctx.synchronize()
        fps = (n - 9) / (time.perf_counter() - t0)
    finally:
        vo._reset_runtime()
        sys.modules.pop("vo_pipeline", None)
        sys.path.remove(os.path.join(ROOT, PKG))
        shutil.rmtree(tmp, ignore_errors=True)
    return {"frames_per_s": fps, "decode_only_frames_per_s": decode_fps, "cv2_imread_one_thread_frames_per_s": one_thread_fps,
            "decode_workers": workers, "frames": n, "format": "PNG %dx%d gray" % (sq.frames[0].shape[1], sq.frames[0].shape[0]),
            "note": "FrameIngest(paths, workers=%d) -> associate / estimatePose / triangulate_landmark; main.py:216-223 decodes one frame per "
                    "iteration on one thread" % workers}


def parity_block(cfg, sq, gpu_trace, port_trace):
    """GPU path (synchronous leg) against the CPU port on the same frames, same process."""
    scene = importlib.import_module(PKG + ".scene")
    n = min(len(gpu_trace), len(port_trace))
    first_div = None
    max_dr = max_dt = 0.0
    for k in range(n):
        g, p = gpu_trace[k], port_trace[k]
        if (g[1], g[2], g[3]) != (p[1], p[2], p[3]):
            first_div = k
            break
        dr, dt = pose_delta(g[0], p[0])
        max_dr, max_dt = max(max_dr, dr), max(max_dt, dt)
    cause = None
    if first_div is not None:
        g, p = gpu_trace[first_div], port_trace[first_div]
        if g[5] != p[5]:
            cause = {"kind": "KLT status flag", "n_tracked_gpu": int(g[5]), "n_tracked_cpu": int(p[5]),
                     "note": "north_star allows 0.5 % of the status flags to differ from cv2"}
        elif g[1] != p[1]:
            # same PnP input size: an inlier decision fell on the other side of the 2 px threshold.  How close to it
            # do the port's errors come?  (winning hypothesis from the C restatement on the port's PnP input)
            cause = {"kind": "inlier decision at the reprojection threshold", "n_inliers_gpu": int(g[1]), "n_inliers_cpu": int(p[1])}
            try:
                import oracle
                Pi, Xi = p[6]
                o = oracle.solve_pnp_ransac(Xi, Pi, sq.K, max_iters=cfg["pnp_iters"])
                Xc = Xi.astype(np.float32).astype(np.float64) @ o["Rh"].T + o["th"]
                uv = np.stack([sq.K[0, 0] * Xc[:, 0] / Xc[:, 2] + sq.K[0, 2], sq.K[1, 1] * Xc[:, 1] / Xc[:, 2] + sq.K[1, 2]], 1)
                e = np.sqrt(((uv - Pi) ** 2).sum(1))
                cause["closest_error_to_threshold_px"] = float(np.abs(e - 2.0).min())
                cause["oracle_inliers"] = int(o["n_inliers"])
            except Exception as ex:   # pragma: no cover
                cause["note"] = "margin not computed: %s" % ex
        else:
            cause = {"kind": "state size", "landmarks": [int(g[2]), int(p[2])], "candidates": [int(g[3]), int(p[3])]}
    all_dr, all_dt = zip(*[pose_delta(gpu_trace[k][0], port_trace[k][0]) for k in range(n)])
    order = frame_order(len(sq.frames), n + 1)
    gt = [sq.poses[order[k + 1]] for k in range(n)]
    return {"frames_compared": n, "first_diverging_frame": first_div, "first_divergence_cause": cause,
            "identical_prefix": {"frames": n if first_div is None else first_div, "max_dR_rad": max_dr, "max_dt_rel": max_dt},
            "all_frames": {"max_dR_rad": float(max(all_dr)), "max_dt_rel": float(max(all_dt)),
                           "inlier_count_max_abs_diff": int(max(abs(gpu_trace[k][1] - port_trace[k][1]) for k in range(n))),
                           "landmark_count_max_abs_diff": int(max(abs(gpu_trace[k][2] - port_trace[k][2]) for k in range(n))),
                           "candidate_count_max_abs_diff": int(max(abs(gpu_trace[k][3] - port_trace[k][3]) for k in range(n)))},
            "ate_m": {"gpu": scene.ate([t[0] for t in gpu_trace[:n]], gt), "cpu_port": scene.ate([t[0] for t in port_trace[:n]], gt)},
            "tolerances": {"dR_rad": 1e-4, "dt_rel": 1e-4, "note": "north_star gates on frames with identical inlier masks; "
                           "counts diverge where a reprojection error straddles the 2 px threshold in float32"}}


# ---- PnP stress (BASELINE configs[3]) --------------------------------------------------------------------
def stress_inputs(seed):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import synth
    X, P, _, _, _ = synth.pnp_scene(STRESS["n"], STRESS["outliers"], seed)
    return X, np.ascontiguousarray(P, np.float32), synth.K_KITTI


def run_stress_gpu(args, local_rank, world):
    import ctypes
    import torch
    pkg = importlib.import_module(PKG)
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    ctx = pkg.Context(device=local_rank, max_w=64, max_h=64, stream=stream.cuda_stream, hyp_capacity=STRESS["hyp"])
    X, P, K = stress_inputs(5 + local_rank)
    dX, dP = torch.from_numpy(X).to(dev), torch.from_numpy(P).to(dev)
    hX, hP = pkg.PinnedArray(X.shape, np.float64), pkg.PinnedArray(P.shape, np.float32)
    hX.array[...] = X
    hP.array[...] = P
    prm = pkg.pnp_defaults()
    prm.max_iters = STRESS["hyp"]
    prm.confidence = 1.0          # budget pinned: all 4096 hypotheses are solved and scored
    Kc = np.ascontiguousarray(K, np.float64)
    T = np.zeros((4, 4))
    info = np.zeros(4, np.int32)
    ok = ctypes.c_int(0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def call(xp, pp):
        r = ctx.lib.vo_solve_pnp_ransac(ctx.h, xp, pp, len(X), Kc.ctypes.data, ctypes.byref(prm), None, T.ctypes.data, None, None, None,
                                        None, info.ctypes.data, ctypes.byref(ok))
        if r != 0 or not ok.value:
            raise RuntimeError("stress PnP failed")
    steps, warmup = args.steps, args.warmup
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = 0
    for s in range(warmup + steps):
        if s == warmup:
            barrier(world)
            torch.cuda.synchronize()
            sampler.active = True
            launches0 = ctx.launch_count()
        flush.zero_()
        if s >= warmup:
            ev[s - warmup][0].record(stream)
        call(ctypes.c_void_p(dX.data_ptr()), ctypes.c_void_p(dP.data_ptr()))
        if s >= warmup:
            ev[s - warmup][1].record(stream)
    torch.cuda.synchronize()
    launches = ctx.launch_count() - launches0
    total_ms = float(sum(a.elapsed_time(b) for a, b in ev))
    wins = []
    for _ in range(E2E_WINDOWS):   # host buffers: H2D of the correspondences and D2H of pose + counters inside
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for s in range(steps):
            call(ctypes.c_void_p(hX.array.ctypes.data), ctypes.c_void_p(hP.array.ctypes.data))
        wins.append(steps / (time.perf_counter() - t0))
    sampler.active = False
    sampler.stop_flag = True
    ctx.close()
    pairs = STRESS["n"] * STRESS["hyp"]
    return dict(total_ms=total_ms, launches=launches, clocks=sampler.summary(), iters=int(info[0]), inliers=int(info[2]),
                e2e=dict(fps=float(np.median(wins)), windows=wins, h2d=X.nbytes + P.nbytes + 72, d2h=256), pairs=pairs)


def run_stress_port(steps, threads):
    """cv2.solvePnPRansac refuses confidence = 1 (it asserts confidence < 1), so its budget cannot be pinned at the
    4096 hypotheses the config names; the CPU arm is therefore the C restatement of that routine (oracle/pnp.c,
    scalar, one thread - as OpenCV's loop is) with the budget pinned exactly like the GPU arm's.  cv2's own adaptive
    run on the same data is timed beside it for orientation."""
    import cv2
    import oracle
    cv2.setNumThreads(threads)
    X, P, K = stress_inputs(5)
    oracle.solve_pnp_ransac(X, P, K, max_iters=64, confidence=1.0)
    t0 = time.perf_counter()
    for _ in range(steps):
        r = oracle.solve_pnp_ransac(X, P, K, max_iters=STRESS["hyp"], confidence=1.0)
    fps = steps / (time.perf_counter() - t0)
    conf = float(np.nextafter(1.0, 0.0))
    t0 = time.perf_counter()
    ok, rv, tv, inl = cv2.solvePnPRansac(X, P, K, np.zeros(4), flags=cv2.SOLVEPNP_P3P, confidence=conf, reprojectionError=2.0,
                                         iterationsCount=STRESS["hyp"])
    cv_ms = (time.perf_counter() - t0) * 1e3
    note = ("oracle/pnp.c: %d hypotheses solved and scored, %d inliers; cv2.solvePnPRansac(P3P) at the largest confidence it accepts "
            "(1 - 2^-53) stops adaptively and takes %.1f ms on the same data" % (r["iters"], r["n_inliers"], cv_ms))
    return fps, note


def pin_rank_to_cores(local_rank, world):
    """Each rank gets its own slice of the host cores; library thread pools are kept out of the GPU arm."""
    cores = sorted(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else list(range(os.cpu_count() or 1))
    if world > 1 and len(cores) >= world:
        per = len(cores) // world
        mine = cores[local_rank * per:(local_rank + 1) * per]
        try:
            os.sched_setaffinity(0, mine)
        except Exception:
            pass
        return mine
    return cores


_OUT_FD = None


def claim_stdout():
    """stdout carries exactly ONE JSON line: whatever libraries write to fd 1 meanwhile (NCCL's version banner) goes to stderr."""
    global _OUT_FD
    if _OUT_FD is None:
        sys.stdout.flush()
        _OUT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _OUT_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_OUT_FD, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="kitti", choices=list(CONFIGS) + ["pnp_stress"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-prefetch", dest="prefetch", action="store_false",
                    help="upload + pyramid of a frame inside its own step (no frame ingest one step ahead)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    claim_stdout()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    cores = os.cpu_count() or 1
    stress = args.config == "pnp_stress"
    cfg = None if stress else CONFIGS[args.config]

    if stress:
        config = {"workload": STRESS["workload"], "correspondences": STRESS["n"], "outlier_fraction": STRESS["outliers"],
                  "hypotheses_per_frame": STRESS["hyp"], "pnp_minimal_solver": "P3P", "confidence": 1.0,
                  "refit": "LM on the winner's inlier set"}
        metric = "PnP-RANSAC frames/s (10k 2D-3D correspondences, 50% outliers, 4096 P3P hypotheses per frame)"
    else:
        hh, ww = {"kitti": (376, 1241), "parking": (480, 640), "malaga": (768, 1024)}[cfg["shape"]]
        config = {"workload": cfg["workload"], "frame": "%dx%d u8" % (ww, hh), "landmarks_held": cfg["n_land"],
                  "candidate_feed_per_frame": cfg["n_feed"], "burst": "every %d-th frame +%d stale-map landmarks (outliers)" % (cfg["burst_every"], cfg["burst_n"]),
                  "pyramid_levels": 4, "window": 21, "pnp_max_iters": cfg["pnp_iters"], "pnp_minimal_solver": "P3P",
                  "alpha_deg": ALPHA, "preroll_frames": PREROLL,
                  "sequence": "ray-cast corridor, forward drive (ping-pong beyond %d frames), one stream per GPU" % F_MAX}
        metric = "VO frames/s at %dx%d (KLT+PnP-RANSAC+triangulate); KLT tracks/s" % (ww, hh)
    base = {"metric": metric, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "data": "synthetic", "config": config}

    if args.impl == "reference":
        if rank != 0:
            return
        import cv2
        n_streams = max(args.gpus, 1)
        thr = max(cores // n_streams, 1)
        if stress:
            steps = min(args.steps, 8)
            fps, note = run_stress_port(steps, cores)    # (the RANSAC loop is single-threaded, in OpenCV too)
            sample = "%d runs over the stress correspondences; %s" % (steps, note)
            tracks = 0.0
        else:
            steps = min(args.steps, 120)   # bounded sample: the CPU path runs at tens of ms per frame
            warm = min(args.warmup, 5)
            jobs = [(cfg, stream_seed(r), steps, warm, thr) for r in range(n_streams)]
            if n_streams == 1:
                res = [_port_worker(jobs[0])]
            else:   # the same N streams the GPU arm runs, one process each, cores // N OpenCV threads each
                import multiprocessing as mp
                mpc = mp.get_context("spawn")
                q = mpc.Queue()
                procs = [mpc.Process(target=_port_worker, args=(j, q, i)) for i, j in enumerate(jobs)]
                for p in procs:
                    p.start()
                got = [q.get(timeout=3000) for _ in procs]
                for p in procs:
                    p.join()
                res = [g[1:] for g in sorted(got)]
            slowest = max(r[0] for r in res)
            fps = n_streams * steps / slowest
            tracks = float(np.mean([r[1] for r in res]))
            sample = ("%d stream(s) x %d consecutive frames (after %d pre-roll + %d warm-up), oracle/ref_pipeline.py over cv2 %s, "
                      "%d OpenCV thread(s) per stream on %d cores (only LK is multi-threaded inside OpenCV)"
                      % (n_streams, steps, PREROLL, warm, cv2.__version__, thr, cores))
        line = dict(base)
        line.update({"impl": "reference", "value": fps, "ms_per_step": 1e3 * n_streams / fps, "steps": steps,
                     "dtype": "u8/f32/f64 (OpenCV)", "klt_tracks_per_s": tracks * fps,
                     "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": 1 if stress else (cores if n_streams > 1 else thr), "kind": "port", "sample": sample},
                     "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
        emit(line)
        return

    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: this package has no CPU path (use --impl reference for the CPU arm)")
    my_cores = pin_rank_to_cores(local_rank, world)
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    torch.set_num_threads(1)
    try:
        import cv2
        cv2.setNumThreads(1)
    except Exception:
        pass

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass

    if stress:
        r = run_stress_gpu(args, local_rank, world)
        fps, e2e_job_fps, total_ms = job_throughput(r["total_ms"], r["e2e"]["fps"], args.steps, world, "cuda")
        if rank == 0:
            sm_mhz = peaks.get("sm_max_mhz", 1965.0)
            fp64_peak = 148 * 64 * 2 * sm_mhz * 1e6 / 1e12   # TFLOP/s: 64 FP64 lanes per SM (DFMA)
            flops = r["pairs"] * 35.0 + STRESS["hyp"] * 2000.0
            ach = flops / (total_ms / args.steps * 1e-3) / 1e12
            line = dict(base)
            line.update({"value": fps, "ms_per_step": total_ms / args.steps, "dtype": "f64 (P3P, projection), f32 (inlier test)",
                         "e2e": {"value": e2e_job_fps, "unit": "frames/s", "h2d_bytes_per_step": r["e2e"]["h2d"], "d2h_bytes_per_step": r["e2e"]["d2h"],
                                 "windows": r["e2e"]["windows"]},
                         "gpu_launches": int(r["launches"]), "clocks": r["clocks"],
                         "roofline": {"bound": "tensor", "kernel": "pnp_ransac_kernel (P3P + scoring)", "achieved": ach, "peak": fp64_peak,
                                      "unit": "TFLOP/s", "frac": ach / fp64_peak, "traffic": None,
                                      "peak_source": "FP64 CUDA-core pipe: 148 SMs x 64 DFMA lanes x 2 x %.0f MHz (no tensor cores on this path; "
                                                     "'tensor' is the contract's name for the compute roof)" % sm_mhz,
                                      "algorithmic_flops_per_launch": flops},
                         "state": {"ransac_iters": r["iters"], "inliers": r["inliers"]}, "notes": {"l2": "flushed between steps"}})
            if world == 1 and not args.no_cpu_baseline:
                n = 4
                cf, note = run_stress_port(n, cores)
                line["cpu_baseline"] = {"value": cf, "unit": "frames/s", "cores": 1, "kind": "port",
                                        "sample": "%d runs over the same correspondences; %s" % (n, note)}
            emit(line)
        if world > 1:
            torch.distributed.destroy_process_group()
        return

    need = PREROLL + args.warmup + args.steps * E2E_WINDOWS
    sq = build_stream(cfg, stream_seed(rank), need, workers=max(len(my_cores), 1))
    r = run_gpu(args, cfg, rank, world, local_rank, sq)
    fps, e2e_job_fps, total_ms = job_throughput(r["total_ms"], r["e2e"]["fps"], args.steps, world, "cuda")
    if rank != 0:
        if world > 1:
            torch.distributed.destroy_process_group()
        return

    peak, peak_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (6650.0, "fallback")
    # KLT algorithmic bytes per track: 4 levels x (w+3)^2 template+derivative window fetch, (w+1)^2 per
    # Gauss-Newton step, one (w+1)^2 error pass, 21 B of point I/O.  Reported for the landmark launch - the head of the
    # chain the pose waits for; tracks and iterations are the kernel's own counters.
    k = r["klt"]
    land_ms = r["stage_ms"]["klt_landmarks"]
    land_bytes = (4 * 576 + 505) * k["land_tracks"] + 484 * k["land_iters"]
    cand_ms = r["stage_ms"]["klt_candidates"]
    cand_bytes = (4 * 576 + 505) * k["cand_tracks"] + 484 * k["cand_iters"]
    achieved = land_bytes / (land_ms * 1e-3) / 1e9 if land_ms > 0 else 0.0
    prof = {}
    try:
        prof = json.load(open(os.path.join(ROOT, "profiles", "klt_traffic.json")))
    except Exception:
        pass
    line = dict(base)
    line.update({
        "value": fps, "ms_per_step": total_ms / args.steps, "dtype": "u8/i32 (KLT), f64 (P3P, projection, DLT), f32 (inlier test)",
        "e2e": {"value": e2e_job_fps, "unit": "frames/s", "h2d_bytes_per_step": r["e2e"]["h2d"], "d2h_bytes_per_step": r["e2e"]["d2h"],
                "windows": r["e2e"]["windows"], "statistic": "median of %d consecutive windows of %d steps (max over ranks)" % (E2E_WINDOWS, args.steps)},
        "gpu_launches": int(r["launches"]),
        "klt_tracks_per_s": (k["land_tracks"] + k["cand_tracks"]) * fps / world,
        "clocks": r["clocks"],
        "roofline": {"bound": "hbm", "kernel": "klt_track_kernel<21> (landmark launch, main stream)", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": prof.get("dram_bytes_per_launch"), "traffic_capture": prof.get("capture"),
                     "issue_active_pct": prof.get("issue_active_pct"), "l1_pct": prof.get("l1_pct"), "dram_pct": prof.get("dram_pct"),
                     "binding_unit": "dependent-iteration latency (issue slots / L1), not HBM: the window gathers are served by shared memory and L1/L2",
                     "peak_source": peak_src, "algorithmic_bytes_per_launch": land_bytes, "kernel_ms": land_ms,
                     "tracks_per_launch": k["land_tracks"], "lk_iterations_per_launch": k["land_iters"],
                     "candidate_launch": {"kernel_ms": cand_ms, "tracks": k["cand_tracks"], "lk_iterations": k["cand_iters"],
                                          "achieved": cand_bytes / (cand_ms * 1e-3) / 1e9 if cand_ms > 0 else 0.0}},
        "stage_ms": r["stage_ms"], "wall_ms_per_step_incl_flush": 1e3 * r["t_wall"] / args.steps,
        "state": {"landmarks": r["final_counts"][0], "candidates": r["final_counts"][1], **r["per_frame"]},
        "pipelined_poses_equal_synchronous": bool(r["pipe_equals_sync"]),
        "notes": {"l2": "flushed between steps (256 MiB memset outside the per-step events)",
                  "ingest": ("frame k+1 uploaded and its pyramid built on the ingest stream during step k (vo_prefetch_frame), every step still "
                             "copies and builds exactly one frame; frames overlapped (vo_process_frame_pipelined)") if args.prefetch
                  else "in-step upload + pyramid, synchronous vo_process_frame",
                  "host": "rank pinned to cores %s, torch/cv2 thread pools at 1 in the GPU arm" % (
                      "%d-%d" % (my_cores[0], my_cores[-1]) if my_cores else "?")},
    })
    if world == 1 and args.config == "kitti":
        try:
            line["files_to_poses"] = run_files_leg(importlib.import_module(PKG), cfg, sq, local_rank)
        except Exception as ex:   # (an extra leg: never let it take the headline line down)
            line["files_to_poses"] = {"error": repr(ex)}
    if world == 1 and not args.no_cpu_baseline:
        n = 60
        n_par = PREROLL + min(args.warmup, 5) + n
        port_trace, raw = [], {}
        cfps, dt, tracks, thr = run_port(cfg, sq, n, min(args.warmup, 5), cores, trace=port_trace, raw_sample=raw)
        line["cpu_baseline"] = {"value": cfps, "unit": "frames/s", "cores": thr, "kind": "port",
                                "sample": "%d consecutive frames of the same stream (after %d pre-roll + %d warm-up) through oracle/ref_pipeline.py "
                                          "(the reference's three methods over cv2, all host threads; %d cores on this box)"
                                          % (n, PREROLL, min(args.warmup, 5), cores),
                                "raw_cv2_calls": raw_cv2_baseline(cfg, sq, raw, cores)}
        line["parity"] = parity_block(cfg, sq, r["trace"][:n_par], port_trace)
    emit(line)
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
