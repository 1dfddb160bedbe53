#!/usr/bin/env python
"""Host-side profile of the e2e leg of bench.py (drop-in classes in main.py's order): cProfile of
the Python layer plus wall time per stage method.  Usage: python tools/profile_e2e.py [steps]"""
import cProfile
import importlib
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 300
pkg = importlib.import_module(bench.PKG)
cfg = bench.CONFIGS['kitti']
sq = bench.build_stream(cfg, 11, bench.PREROLL + 20 + steps * bench.E2E_WINDOWS)
pr = cProfile.Profile()
PF = "--no-prefetch" not in sys.argv
r0 = bench.run_e2e(pkg, cfg, sq, steps, 20, 0, 1, PF)       # warm (also un-profiled reference number)
pr.enable()
r = bench.run_e2e(pkg, cfg, sq, steps, 20, 0, 1, PF)
pr.disable()
print("e2e fps un-profiled %.1f, profiled %.1f" % (r0["fps"], r["fps"]))
st = pstats.Stats(pr)
st.strip_dirs().sort_stats("tottime").print_stats(28)
