"""The documents name things the code must actually have: every environment switch read by the library is listed in
INTEGRATION.md (and nothing else is), every entry point declared in include/vo_b200.h appears in INTEGRATION.md's
entry-point table or in its prose, and the limits of the no-allocation entry point are the same in the header, the
binding and the documents."""
import glob
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "locally-consistent-visual-odometry-pipeline_b200")


def _read(*parts):
    with open(os.path.join(*parts)) as f:
        return f.read()


def test_environment_switches_are_documented():
    code = set()
    for path in glob.glob(os.path.join(PKG, "csrc", "*.cu")) + glob.glob(os.path.join(PKG, "csrc", "*.cuh")):
        code |= set(re.findall(r'getenv\("(VO_[A-Z0-9_]+)"\)', _read(path)))
    doc = set(re.findall(r"`(VO_[A-Z0-9_]+)(?:=[^`]*)?`", _read(ROOT, "INTEGRATION.md").split("## 4. Environment switches")[1]))
    assert code - doc == set(), "switches read by the library but missing from INTEGRATION.md: %s" % sorted(code - doc)
    assert doc - code == set(), "switches documented but not read anywhere: %s" % sorted(doc - code)


def test_every_entry_point_is_mentioned_in_the_integration_notes():
    header = _read(ROOT, "include", "vo_b200.h")
    names = set(re.findall(r"^(?:int|void|const char\*)\s+(vo_[a-z0-9_]+)\s*\(", header, re.M))
    assert len(names) > 40
    text = _read(ROOT, "INTEGRATION.md") + _read(ROOT, "DESIGN.md")
    # housekeeping calls that need no reference line: creation, errors, defaults, counters, debug
    housekeeping = {n for n in names if re.search(r"_default$|^vo_(create|destroy|last_error|device_count|synchronize|launch_count|"
                                                  r"get_counters|peek_counters|debug_|set_profiling|get_stage_times|has_frame|drop_frames|"
                                                  r"get_pyramid_level|promote_pending|candidates_tracked|get_lost|map_reset|version)", n)}
    missing = sorted(n for n in names - housekeeping if n not in text)
    assert not missing, "declared in include/vo_b200.h but mentioned in neither INTEGRATION.md nor DESIGN.md: %s" % missing


def test_noalloc_limits_agree():
    header = _read(ROOT, "include", "vo_b200.h")
    pts = int(re.search(r"#define VO_NOALLOC_MAX_POINTS (\d+)", header).group(1))
    its = int(re.search(r"#define VO_NOALLOC_MAX_ITERS (\d+)", header).group(1))
    cabi = _read(PKG, "_cabi.py")
    assert int(re.search(r"NOALLOC_MAX_POINTS = (\d+)", cabi).group(1)) == pts
    assert int(re.search(r"NOALLOC_MAX_ITERS = (\d+)", cabi).group(1)) == its
    for doc in ("DESIGN.md", "INTEGRATION.md"):
        t = _read(ROOT, doc).replace(" ", " ").replace(" ", " ")
        assert re.search(r"4[ ,]?096", t) and re.search(r"1[ ,]?024", t), doc
