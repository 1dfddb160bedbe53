"""B200-native per-frame core of the locally-consistent visual odometry pipeline.

Layout
  csrc/            hand-written sm_100a CUDA kernels + the C-ABI (include/vo_b200.h)
  _cabi.py         ctypes binding of libvo_b200.so
  vo_pipeline.py   drop-in for the reference's src/vo_pipeline.py class surface

The directory name carries a hyphen (it mirrors the upstream repository name), so import it with
`importlib.import_module("locally-consistent-visual-odometry-pipeline_b200")`, or put this
directory on sys.path ahead of the reference's src/ to get the drop-in `vo_pipeline` module.
"""
from . import _build  # noqa: F401
from ._cabi import (Context, Counters, GfttParams, KltParams, PnpParams, PinnedArray, VoError, gftt_defaults,  # noqa: F401
                    klt_defaults, pnp_defaults, load, lib_path, NOALLOC_MAX_POINTS, NOALLOC_MAX_ITERS)

VO_ERR_CAPACITY = -5   # include/vo_b200.h

__all__ = ["Context", "KltParams", "PnpParams", "PinnedArray", "VoError", "load", "lib_path", "build"]


def build(force=False, verbose=False):
    """Compile libvo_b200.so for sm_100a (nvcc cross-compiles without a GPU)."""
    return _build.build(force=force, verbose=verbose)
