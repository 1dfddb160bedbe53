"""N > 1 host logic on CPU (gloo, world_size 2): one independent stream per rank, no collective on
the data path, whole-job frames/s = world x steps / (slowest rank's time).  The per-rank "stream"
here is the CPU port of the reference's three methods on a small sequence (test infrastructure);
on the GPU box bench.py runs the CUDA path per rank with the same partitioning and aggregation."""
import importlib
import os
import socket
import sys
import time

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG_NAME = "locally-consistent-visual-odometry-pipeline_b200"


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch
    import torch.distributed as dist
    import bench
    from oracle.ref_pipeline import RefPipeline
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        scene = importlib.import_module(PKG_NAME + ".scene")
        sq = scene.Sequence("kitti", n_frames=6, n_tracks=200, n_feed=30, seed=bench.stream_seed(rank), scale=0.25)
        rp = RefPipeline(sq.K, pnp="p3p", max_iters=100, alpha=1.0, row_match="isin")
        state = {"P": sq.P0.astype(np.float64), "X": sq.X0.copy()}
        ext = {"C": np.array([]), "F": np.array([]), "T": np.array([])}
        steps = 4
        bench.barrier(world)
        t0 = time.perf_counter()
        poses = []
        for s in range(steps):
            state, ext, T = rp.frame(sq.frames[s], sq.frames[s + 1], state, ext, sq.feeds[s + 1][:20].copy())
            poses.append(T)
        if rank == 1:
            time.sleep(0.25)        # a slow rank must define the job's time
        ms = (time.perf_counter() - t0) * 1e3
        fps, e2e, total_ms = bench.job_throughput(ms, steps / (ms * 1e-3), steps, world)
        # what every rank saw, for the parent's checks (files, not collectives: the data path has none)
        np.savez(os.path.join(out_dir, "rank%d.npz" % rank), ms=ms, fps=fps, e2e=e2e, total_ms=total_ms,
                 poses=np.array(poses), P0=sq.P0)
    finally:
        dist.destroy_process_group()


def test_two_streams_gloo(tmp_path):
    import torch.multiprocessing as mp
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    r = [np.load(tmp_path / ("rank%d.npz" % k)) for k in range(world)]
    slow = max(float(x["ms"]) for x in r)
    for x in r:
        assert abs(float(x["total_ms"]) - slow) < 1e-9                 # MAX over ranks, same on all ranks
        assert abs(float(x["fps"]) - world * 4 / (slow * 1e-3)) < 1e-6  # whole-job aggregate
        assert abs(float(x["e2e"]) - world * 4 / (slow * 1e-3)) < 1e-6
    assert float(r[1]["ms"]) >= 250.0 and slow == float(r[1]["ms"])
    # the two ranks processed different streams (different seeds => different scenes and tracks)
    assert r[0]["P0"].shape != r[1]["P0"].shape or not np.allclose(r[0]["P0"], r[1]["P0"])
    for x in r:                                                         # and each stream really ran
        assert x["poses"].shape == (4, 4, 4) and np.isfinite(x["poses"]).all()


def test_single_rank_is_identity():
    sys.path.insert(0, ROOT)
    import bench
    fps, e2e, total_ms = bench.job_throughput(200.0, 123.0, 100, 1)
    assert fps == pytest.approx(500.0) and e2e == pytest.approx(123.0) and total_ms == 200.0
    assert bench.stream_seed(0) != bench.stream_seed(1)
    order = bench.frame_order(5, 12)
    assert len(order) == 13 and all(abs(a - b) == 1 for a, b in zip(order, order[1:])) and max(order) == 4
