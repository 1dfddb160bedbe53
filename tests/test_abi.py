"""The C-ABI boundary without a GPU: libvo_b200.so loads, exports every symbol include/vo_b200.h
declares (and nothing undeclared is bound), the ctypes table mirrors the header, and a context
cannot be created without a CUDA device (no CPU fallback).  No compute calls."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "vo_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"//[^\n]*", "", src)
    body = src[src.index('extern "C"'):]
    names = re.findall(r"^\s*(?:const\s+)?(?:int64_t|int|void|char\s*\*|const char\s*\*)\s*\*?\s*(vo_[a-z0-9_]+)\s*\(", body, flags=re.M)
    return sorted(set(names))


def test_header_declares_the_path():
    names = declared_symbols()
    # one entry point per row of SURVEY.md 8a, plus the unit-level mirrors of the three cv2 calls
    for must in ("vo_create", "vo_destroy", "vo_last_error", "vo_upload_frame", "vo_klt_track_landmarks",
                 "vo_pnp_ransac", "vo_track_triangulate_promote", "vo_add_candidates", "vo_process_frame",
                 "vo_calc_optical_flow_pyr_lk", "vo_solve_pnp_ransac", "vo_triangulate_points",
                 "vo_set_state", "vo_get_state", "vo_set_candidates", "vo_get_candidates"):
        assert must in names, must
    text = open(HEADER).read()
    assert "vo_pipeline.py:215" in text and "vo_pipeline.py:258" in text and "vo_pipeline.py:666" in text
    code = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    assert "torch" not in code.lower() and "Tensor" not in code and "#include <cuda" not in code


def test_ctypes_table_mirrors_header(pkg):
    cabi = __import__(pkg.__name__ + "._cabi", fromlist=["SIGNATURES"])
    assert sorted(cabi.SIGNATURES) == declared_symbols()


def test_library_exports_every_declared_symbol(pkg):
    path = pkg.build(force=False)
    lib = ctypes.CDLL(path)
    for name in declared_symbols():
        assert hasattr(lib, name), "libvo_b200.so does not export %s" % name
    # -fvisibility=hidden: nothing but the declared C entry points is exported from our sources
    out = subprocess.run(["nm", "-D", "--defined-only", path], capture_output=True, text=True).stdout
    exported = sorted(l.split()[-1] for l in out.splitlines() if " T " in l)
    ours = [s for s in exported if s.startswith("vo_")]
    assert ours == declared_symbols()
    lib.vo_abi_version.restype = ctypes.c_int
    m = re.search(r"#define\s+VO_B200_ABI_VERSION\s+(\d+)", open(HEADER).read())
    assert lib.vo_abi_version() == int(m.group(1))


def test_library_is_sm100a_native(pkg):
    path = pkg.build(force=False)
    r = subprocess.run(["cuobjdump", "-lelf", path], capture_output=True, text=True)
    if r.returncode != 0:
        pytest.skip("cuobjdump not available")
    archs = set(re.findall(r"sm_\d+a?", r.stdout))
    assert archs == {"sm_100a"}, archs


def test_defaults_match_reference_calls(pkg):
    """cv2.calcOpticalFlowPyrLK defaults (vo_pipeline.py:215 passes none) and the PoseEstimator
    constants (vo_pipeline.py:239-243, 258-264)."""
    k = pkg.klt_defaults()
    assert (k.win, k.max_level, k.max_iters, k.epsilon, k.min_eig_thr, k.fb_check) == (21, 3, 30, 0.01, 1e-4, 0)
    p = pkg.pnp_defaults()
    assert (p.max_iters, p.confidence, p.reproj_thresh) == (100, 0.99999, 2.0)


def test_no_cpu_fallback(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(Exception) as e:
        pkg.Context(device=0)
    assert "CUDA" in str(e.value) or "CPU" in str(e.value)
