"""Build libtafl_b200.so in-tree with nvcc for sm_100a (no JIT cache, no torch extension).

    python -m alphazeroforhnefatafl_b200.build [--force] [--verbose]
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
ROOT = os.path.dirname(HERE)
LIB = os.path.join(HERE, "libtafl_b200.so")
SOURCES = ["tafl_engine.cu", "tafl_mcts.cu", "tafl_comm.cu", "tafl_wide.cu", "tafl_host.cpp"]
HEADERS = ["tafl_bb.cuh", "tafl_core.cuh", "tafl_select.cuh", "tafl_ctx.hpp", "tafl_engine.hpp",
           "tafl_warp.cuh", "tafl_mcts.cuh", "tafl_wide.cuh"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
         "-fmad=false",                      # MCTS PUCT must round like the f64 oracle (no FMA contraction)
         "-Xcompiler", "-fPIC",
         "--expt-relaxed-constexpr", "-Xcudafe", "--diag_suppress=177"]


def _stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [os.path.join(ROOT, "include", "tafl.h"),
                                                                 os.path.abspath(__file__)]
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, defines=(), suffix=""):
    """defines / suffix: an A/B variant of the library (e.g. defines=("TAFL_COUNT_NESTED",), suffix="_v1" ->
    libtafl_b200_v1.so, loaded with TAFL_LIB=...); the product library is the one without either."""
    lib = LIB if not suffix else os.path.join(HERE, "libtafl_b200%s.so" % suffix)
    if not suffix and not force and not _stale():
        return LIB
    objdir = os.path.join(HERE, "build" + suffix)
    os.makedirs(objdir, exist_ok=True)

    def compile_one(src):
        obj = os.path.join(objdir, os.path.splitext(src)[0] + ".o")
        cmd = [NVCC] + FLAGS + ["-D" + d for d in defines] + (["-Xptxas", "-v"] if verbose else []) + ["-x", "cu", "-c", os.path.join(CSRC, src), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    cmd = [NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", lib] + objs + ["-ldl"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    return lib


if __name__ == "__main__":
    defs = [a[2:] for a in sys.argv[1:] if a.startswith("-D")]
    suf = [a.split("=", 1)[1] for a in sys.argv[1:] if a.startswith("--suffix=")]
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv, defines=defs, suffix=suf[0] if suf else ""))
