"""A/B helper: run a script of this repo against another build of the library.
   python tools/with_lib.py build_ab/libvo_old.so bench.py --no-cpu-baseline"""
import importlib, os, runpy, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root)
b = importlib.import_module("locally-consistent-visual-odometry-pipeline_b200._build")
b.LIB = os.path.abspath(sys.argv[1]); b.needs_build = lambda: False
sys.argv = sys.argv[2:]
runpy.run_path(os.path.join(root, sys.argv[0]), run_name="__main__")
