import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

PKG_NAME = "locally-consistent-visual-odometry-pipeline_b200"


# (resolved at import: fixtures change the working directory later)
_PARITY_PATH = os.path.abspath(os.environ["VO_PARITY_REPORT"]) if os.environ.get("VO_PARITY_REPORT") else None


def parity_report(section, name, values):
    """Every magnitude the parity tests print, kept as JSON too when VO_PARITY_REPORT names a file
    (tools/gpu_round.sh sets it; the result is committed as profiles/parity_r2.json)."""
    import json
    path = _PARITY_PATH
    if not path:
        return
    try:
        rep = json.load(open(path))
    except Exception:
        rep = {}
    rep.setdefault(section, {})[name] = values
    with open(path, "w") as f:
        json.dump(rep, f, indent=1, sort_keys=True)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    return importlib.import_module(PKG_NAME)


@pytest.fixture(scope="session")
def oracle():
    import oracle as o
    o.build()
    return o


@pytest.fixture(scope="session")
def ctx(pkg):
    """One shared vo_ctx on cuda:0.  Fails loudly (no CPU fallback) if there is no GPU."""
    c = pkg.Context(device=0, max_w=2048, max_h=1024)
    yield c
    c.close()


CONFIG_YAML = """dataset: 'kitti'
find_new_candidates_method: '%s'
init_detector: 'sift'
init_descriptor: 'sift'
debug: False
visualization: False
animation: False
sift_alpha: 1
sift_ransac_confidence: 0.99999
shi_alpha: 1.5
shi_ransac_confidence: 0.999993
"""


@pytest.fixture(scope="session")
def scene_mod():
    return importlib.import_module(PKG_NAME + ".scene")


@pytest.fixture()
def dropin_shi(tmp_path, monkeypatch):
    """The drop-in configured as the reference's 'shi-mask' mode (Shi-Tomasi candidates every frame)."""
    (tmp_path / "config.yaml").write_text(CONFIG_YAML % "shi-mask")
    monkeypatch.chdir(tmp_path)
    monkeypatch.syspath_prepend(os.path.join(ROOT, PKG_NAME))
    sys.modules.pop("vo_pipeline", None)
    mod = importlib.import_module("vo_pipeline")
    yield mod
    mod._reset_runtime()
    sys.modules.pop("vo_pipeline", None)


@pytest.fixture()
def dropin(tmp_path, monkeypatch):
    """The drop-in module imported the way main.py does it: as top-level `vo_pipeline`, with the
    package directory ahead on sys.path and config.yaml in the CWD (method 'feed' = candidate feed)."""
    (tmp_path / "config.yaml").write_text(CONFIG_YAML % "feed")
    monkeypatch.chdir(tmp_path)
    monkeypatch.syspath_prepend(os.path.join(ROOT, PKG_NAME))
    sys.modules.pop("vo_pipeline", None)
    mod = importlib.import_module("vo_pipeline")
    yield mod
    mod._reset_runtime()
    sys.modules.pop("vo_pipeline", None)
