"""TEST INFRASTRUCTURE ONLY -- CPU oracle for the blob mobility product.

ctypes binding of oracle/mobility_oracle.c (a plain-C restatement of
/root/reference/mobility/mobility_numba.py + the wrappers of mobility.py).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this package.  The product
(rigidmultiblobswall_b200/) never does.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libmobility_oracle.so")
BLOCKS = {"UF": 0, "UT": 1, "WF": 2, "WT": 3}
_lib = None


def _cpu_signature():
    """-march=native code must not travel between machines: rebuild when the host CPU changes."""
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    return str(hash(line.strip()))
    except OSError:
        pass
    return "unknown"


def build(force=False):
    src = os.path.join(_HERE, "mobility_oracle.c")
    sig_file = _SO + ".cpu"
    sig = _cpu_signature()
    old = open(sig_file).read() if os.path.exists(sig_file) else ""
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src) or old != sig:
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B"])
        with open(sig_file, "w") as f:
            f.write(sig)
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_SO)
        dp = ctypes.POINTER(ctypes.c_double)
        _lib.rmb_oracle_block.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_long, dp, dp,
                                          ctypes.c_double, ctypes.c_double, dp,
                                          ctypes.c_long, ctypes.c_long, dp]
        _lib.rmb_oracle_wall_wrap.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_long, dp, dp,
                                              ctypes.c_double, ctypes.c_double, dp,
                                              ctypes.c_long, ctypes.c_long, dp]
        _lib.rmb_oracle_uf_wall_fast.argtypes = [ctypes.c_long, dp, dp, ctypes.c_double, ctypes.c_double,
                                                 ctypes.c_long, ctypes.c_long, dp]
        _lib.rmb_oracle_threads.restype = ctypes.c_int
        _lib.rmb_oracle_set_threads.argtypes = [ctypes.c_int]
        _lib.rmb_oracle_source_target.argtypes = [ctypes.c_int, ctypes.c_long, ctypes.c_long, dp, dp, dp, dp, dp,
                                                  ctypes.c_double, dp, dp]
        _lib.rmb_oracle_source_target_wall_wrap.argtypes = [ctypes.c_long, ctypes.c_long, dp, dp, dp, dp, dp,
                                                            ctypes.c_double, dp, dp]
    return _lib


def _p(x):
    return x.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


def threads():
    return int(lib().rmb_oracle_threads())


def host_cores():
    """CPUs this process may run on (torchrun does not restrict the affinity, it only exports OMP_NUM_THREADS=1)."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def set_threads(n=None):
    """Use n (default: all host cores) OpenMP threads whatever OMP_NUM_THREADS says."""
    lib().rmb_oracle_set_threads(int(n or host_cores()))
    return threads()


def mobility_product(block, wall, r_vectors, vec, eta, a, periodic_length=None,
                     in_plane=False, targets=None):
    """block in UF/UT/WF/WT; wall in {'single_wall','no_wall'}.

    Mirrors mobility.py's `<wall>_mobility_<block>_numba(r, vec, eta, a,
    periodic_length=L)` including the height clamp and damping B for the wall
    forms.  targets=(begin, end) restricts the computed rows (bounded sample).
    """
    r = np.ascontiguousarray(np.asarray(r_vectors, dtype=np.float64).reshape(-1, 3))
    v = np.ascontiguousarray(np.asarray(vec, dtype=np.float64).reshape(-1, 3))
    n = r.shape[0]
    L = np.zeros(3) if periodic_length is None else np.ascontiguousarray(periodic_length, dtype=np.float64)
    b, e = (0, n) if targets is None else targets
    out = np.empty(3 * (e - b))
    if wall == "single_wall":
        lib().rmb_oracle_wall_wrap(BLOCKS[block], int(in_plane), n, _p(r), _p(v), eta, a, _p(L), b, e, _p(out))
    elif wall == "no_wall":
        assert not in_plane
        lib().rmb_oracle_block(BLOCKS[block], 0, n, _p(r), _p(v), eta, a, _p(L), b, e, _p(out))
    elif wall == "free_surface":
        assert not in_plane and block == "UF"
        lib().rmb_oracle_block(BLOCKS[block], 2, n, _p(r), _p(v), eta, a, _p(L), b, e, _p(out))
    else:
        raise ValueError(wall)
    return out


def uf_wall_fast(r_vectors, force, eta, a, targets=None):
    """Vectorised timing variant of single-wall UF without images (the CPU baseline of bench.py)."""
    r = np.ascontiguousarray(np.asarray(r_vectors, dtype=np.float64).reshape(-1, 3))
    v = np.ascontiguousarray(np.asarray(force, dtype=np.float64).reshape(-1, 3))
    n = r.shape[0]
    b, e = (0, n) if targets is None else targets
    out = np.empty(3 * (e - b))
    lib().rmb_oracle_uf_wall_fast(n, _p(r), _p(v), eta, a, b, e, _p(out))
    return out


def source_target(wall, source, target, force, radius_source, radius_target, eta, periodic_length=None):
    """Mirrors mobility.py's `<wall>_mobility_trans_times_force_source_target_numba(source, target, force,
    radius_source, radius_target, eta, periodic_length=L)` (:551, :588, :1407); wall in
    {'single_wall','no_wall','free_surface'}.  The single-wall form includes the per-radius clamp and damping."""
    src = np.ascontiguousarray(np.asarray(source, dtype=np.float64).reshape(-1, 3))
    trg = np.ascontiguousarray(np.asarray(target, dtype=np.float64).reshape(-1, 3))
    f = np.ascontiguousarray(np.asarray(force, dtype=np.float64).reshape(-1, 3))
    rs = np.ascontiguousarray(radius_source, dtype=np.float64)
    rt = np.ascontiguousarray(radius_target, dtype=np.float64)
    L = np.zeros(3) if periodic_length is None else np.ascontiguousarray(periodic_length, dtype=np.float64)
    ns, nt = src.shape[0], trg.shape[0]
    assert f.shape[0] == ns and rs.size == ns and rt.size == nt
    out = np.zeros(3 * nt)
    if wall == "single_wall":
        lib().rmb_oracle_source_target_wall_wrap(nt, ns, _p(trg), _p(src), _p(f), _p(rt), _p(rs), eta, _p(L), _p(out))
    elif wall in ("no_wall", "free_surface"):
        lib().rmb_oracle_source_target(0 if wall == "no_wall" else 2, nt, ns, _p(trg), _p(src), _p(f), _p(rt), _p(rs),
                                       eta, _p(L), _p(out))
    else:
        raise ValueError(wall)
    return out
