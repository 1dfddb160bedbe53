"""Build helper: compiles csrc/*.cu into libvo_b200.so with nvcc for sm_100a (in-tree).

The library has no torch dependency: it is a plain CUDA runtime shared object loaded with ctypes.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libvo_b200.so")
OBJ = os.path.join(HERE, "build")
HEADER = os.path.join(HERE, "..", "include", "vo_b200.h")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "--expt-relaxed-constexpr",
]


def _nvcc():
    cand = os.environ.get("NVCC")
    if cand:
        return cand
    if os.path.exists("/usr/local/cuda/bin/nvcc"):
        return "/usr/local/cuda/bin/nvcc"
    return "nvcc"


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _headers_mtime():
    files = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if not f.endswith(".cu")]
    files += [HEADER, os.path.abspath(__file__)]
    return max(os.path.getmtime(f) for f in files)


def needs_build():
    if not os.path.exists(LIB):
        return True
    newest = max([_headers_mtime()] + [os.path.getmtime(s) for s in sources()])
    return os.path.getmtime(LIB) < newest


def build(force=False, verbose=False, extra_flags=()):
    """Compile every .cu for sm_100a and link libvo_b200.so next to this file."""
    if not force and not needs_build():
        return LIB
    os.makedirs(OBJ, exist_ok=True)
    nvcc = _nvcc()
    hdr_mtime = _headers_mtime()

    def compile_one(src):
        obj = os.path.join(OBJ, os.path.basename(src)[:-3] + ".o")
        if (not force and os.path.exists(obj)
                and os.path.getmtime(obj) >= max(os.path.getmtime(src), hdr_mtime)):
            return obj
        cmd = [nvcc, *NVCC_FLAGS, *extra_flags, "-c", src, "-o", obj]
        if verbose:
            print(" ".join(cmd), file=sys.stderr)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
        if verbose and r.stderr:
            print(r.stderr, file=sys.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(compile_one, sources()))
    tmp = LIB + ".tmp.%d" % os.getpid()     # link under another name, then rename: nobody dlopens a half-written library
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", tmp, *objs, "-cudart", "shared"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    os.replace(tmp, LIB)
    return LIB


if __name__ == "__main__":
    flags = ["-Xptxas", "-v"] if "--ptxas" in sys.argv else []
    print(build(force="--force" in sys.argv or bool(flags), verbose=True, extra_flags=flags))
