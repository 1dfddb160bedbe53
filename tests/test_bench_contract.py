"""CPU tests of the bench contract: the reference arm (`bench.py --impl reference`) runs here and prints
the line the driver expects, ranks other than 0 stay silent, and the committed B200 bench line
(profiles/) carries every key of the contract."""
import glob
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
             "vs_baseline", "dtype", "data", "config", "e2e", "cpu_baseline"}


def run_bench(args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, capture_output=True, text=True, env=e,
                       timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    return r.stdout.strip()


def test_reference_arm_line():
    pytest.importorskip("cv2")
    out = run_bench(["--impl", "reference", "--steps", "2", "--warmup", "1"])
    line = json.loads(out.splitlines()[-1])
    assert BASE_KEYS <= set(line), BASE_KEYS - set(line)
    assert line["impl"] == "reference" and line["unit"] == "frames/s" and line["higher_is_better"] is True
    assert line["config"]["workload"] == "kitti_1241x376_P2000_pyr4_win21_pnp2000" and line["config"]["landmarks_held"] == 2000
    assert line["value"] > 0 and line["e2e"] == {"value": line["value"], "unit": "frames/s", "h2d_bytes_per_step": 0,
                                                 "d2h_bytes_per_step": 0}
    cb = line["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == line["value"] and "sample" in cb
    # ranks other than 0 exit without work and without output
    assert run_bench(["--impl", "reference", "--gpus", "2"], {"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}) == ""


def test_reference_arm_runs_n_streams():
    """--gpus N: the reference arm runs the same N independent streams the GPU arm does (one process each) and
    reports the job's frames/s = N x steps / slowest stream."""
    pytest.importorskip("cv2")
    out = run_bench(["--impl", "reference", "--gpus", "2", "--steps", "2", "--warmup", "1", "--config", "parking"])
    line = json.loads(out.splitlines()[-1])
    assert line["n_gpus"] == 2 and line["value"] > 0 and "2 stream(s)" in line["cpu_baseline"]["sample"]
    assert line["config"]["workload"].startswith("parking_640x480")


def test_reference_arm_stress_config():
    pytest.importorskip("cv2")
    out = run_bench(["--impl", "reference", "--steps", "1", "--warmup", "1", "--config", "pnp_stress"])
    line = json.loads(out.splitlines()[-1])
    assert line["config"]["workload"] == "pnp_stress_10000pts_50pct_outliers_4096hyp" and line["value"] > 0


def test_committed_b200_line_has_the_contract_keys():
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r2_bench_?.json")))
    assert files, "no committed bench line under profiles/"
    line = json.loads(open(files[-1]).read().strip().splitlines()[-1])
    assert BASE_KEYS | {"roofline", "clocks", "gpu_launches"} <= set(line)
    rf = line["roofline"]
    assert {"bound", "achieved", "peak", "unit", "frac", "traffic"} <= set(rf) and rf["bound"] in ("hbm", "tensor")
    assert abs(rf["frac"] - rf["achieved"] / rf["peak"]) < 1e-9
    assert line["gpu_launches"] > 0 and line["e2e"]["h2d_bytes_per_step"] >= 1241 * 376
    assert line["config"]["workload"] == "kitti_1241x376_P2000_pyr4_win21_pnp2000" and "model" not in line["config"]
    assert line["state"]["landmarks_tracked_into_frame"] >= 1800 and "parity" in line and "raw_cv2_calls" in line["cpu_baseline"]
    assert not set(line["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
