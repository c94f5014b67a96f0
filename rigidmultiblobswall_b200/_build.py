"""Builds librmb_mobility.so in-tree with nvcc for sm_100a (no JIT cache: the
built file travels to the GPU box with the repo snapshot)."""
import os
import shutil
import subprocess

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
SO = os.path.join(PKG, "librmb_mobility.so")
SOURCES = ["rmb_mobility.cu", "rmb_source_target.cu", "rmb_rigid.cu"]
HEADERS = ["pair_math.cuh", "mf_core.cuh", "tile_pipe.cuh", "st_math.cuh", "sym_schedule.cuh", os.path.join("..", "..", "include", "rmb_mobility.h"),
           os.path.join("..", "..", "include", "rmb_rigid.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-shared", "-Xcompiler", "-fPIC"]


def _nvcc():
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: librmb_mobility.so cannot be built")


def have_nvcc():
    try:
        _nvcc()
        return True
    except RuntimeError:
        return False


def stale():
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS]
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not stale():
        return SO
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", SO] + srcs
    subprocess.check_call(cmd, cwd=CSRC)
    return SO
