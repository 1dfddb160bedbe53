#!/usr/bin/env python
"""Wall time per phase of the e2e loop (bench.run_e2e re-stated with timestamps).  VO_TRACE=1 adds the
library's own report of the time the host spends blocked on the pose.
Usage: VO_TRACE=1 python tools/e2e_phases.py [key=value ...]   e.g. speculative_tracking=0"""
import importlib, os, sys, tempfile, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
pkg = importlib.import_module(bench.PKG)
_, sq = bench.build_sequence(seed=11)
over = dict(a.split("=") for a in sys.argv[1:] if "=" in a)
tmp = tempfile.mkdtemp()
open(os.path.join(tmp, "config.yaml"), "w").write(
    "dataset: 'kitti'\nfind_new_candidates_method: 'feed'\ninit_detector: 'sift'\ninit_descriptor: 'sift'\ndebug: False\n"
    "visualization: False\nanimation: False\nsift_alpha: 1\nsift_ransac_confidence: 0.99999\nshi_alpha: 1.5\n"
    "shi_ransac_confidence: 0.999993\npnp_max_iters: %d\n" % bench.PNP_ITERS + "".join("%s: %s\n" % (k, "True" if v == "1" else "False") for k, v in over.items() if k != "prefetch"))
os.chdir(tmp); sys.path.insert(0, os.path.join(ROOT, bench.PKG))
vo = importlib.import_module("vo_pipeline")
use_pf = over.get("prefetch", "1") == "1"
F = len(sq.frames); h, w = sq.frames[0].shape
pinned = [pkg.PinnedArray((h, w), np.uint8) for _ in range(F)]
for p, f in zip(pinned, sq.frames): p.array[...] = f
steps, warm = 400, 30
order = bench.frame_order(F, steps + warm + 1)
vision = vo.BestVision(sq.K)
vision.state = {"P": sq.P0.astype(np.float64), "X": sq.X0.copy()}
vision.extended_state = {"C": np.array([]), "F": np.array([]), "T": np.array([])}
associate = vo.KeypointsToLandmarksAssociator(sq.K, np.eye(4)); pe = vo.PoseEstimator(sq.K)
n_land, n_cand = len(sq.P0), 0
acc = np.zeros(6); img2 = None; pc = time.perf_counter
for s in range(steps + warm):
    if s == warm:
        vo._runtime().ctx.synchronize(); acc[:] = 0; T0 = pc()
    t0 = pc()
    a, b = order[s], order[s + 1]
    img1 = pinned[a].array if s == 0 else img2
    img2 = pinned[b].array[:]
    feed = bench.pick_feed(sq, b, n_land, n_cand); vo.set_candidate_feed([feed])
    if use_pf: vo.prefetch_frame(pinned[order[s + 2]].array)
    t1 = pc()
    new_state, lost = associate.associateKeypointsToLandmarks(img1, img2, vision.state)
    t2 = pc()
    new_pose = pe.estimatePose(new_state, s)
    t3 = pc()
    pos = np.linalg.inv(new_pose)[0:3, 3]
    tri = vo.LandmarkTriangulator(sq.K, None)
    t4 = pc()
    vision.state, vision.extended_state, _ = tri.triangulate_landmark(img1, img2, new_state, vision.extended_state, new_pose)
    t5 = pc()
    c = vo._runtime().ctx.counters(sync=False); n_land, n_cand = c["n_landmarks"], c["n_candidates"] + len(feed)
    t6 = pc()
    acc += [t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4, t6 - t5]
vo._runtime().ctx.synchronize()
dt = pc() - T0
print("overrides", over, "fps %.0f  us/step %.1f" % (steps / dt, 1e6 * dt / steps))
for n, v in zip(("loop head", "associate", "estimatePose", "inv+ctor", "triangulate", "loop tail"), acc / steps * 1e6):
    print("  %-14s %6.1f us" % (n, v))
