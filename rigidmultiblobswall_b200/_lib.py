"""ctypes binding of librmb_mobility.so (C ABI: include/rmb_mobility.h).

There is no CPU fallback: if the library is missing and cannot be built, or no
CUDA device is present, calls raise."""
import ctypes
import os

from . import _build

UF, UT, WF, WT, UF_UT = 0, 1, 2, 3, 4
NO_WALL, SINGLE_WALL, FREE_SURFACE = 0, 1, 2
FP64, FP32 = 0, 1
IN_PLANE, NO_TMA, SYMMETRIC, DETERMINISTIC = 1, 2, 4, 8
SYMMETRIC_MIN_N = 8000       # = RMB_SYMMETRIC_MIN_N (include/rmb_mobility.h)

_lib = None


class RmbError(RuntimeError):
    pass


def load():
    global _lib
    if _lib is not None:
        return _lib
    so = os.environ.get("RMB_LIB")                   # RMB_LIB: tuning builds (tools/build_tune.py)
    if not so:
        so = _build.SO
        # rebuild when a source is newer than the library (no-op when fresh); a machine without nvcc (the .so is
        # shipped prebuilt to the GPU box) uses what is there, and a missing library with no compiler raises
        if _build.stale() and (_build.have_nvcc() or not os.path.exists(so)):
            _build.build()
    lib = ctypes.CDLL(so)
    c_dp = ctypes.c_void_p
    lib.rmb_version.restype = ctypes.c_int
    lib.rmb_last_error.restype = ctypes.c_char_p
    lib.rmb_device_sm_count.restype = ctypes.c_int
    lib.rmb_workspace_bytes.restype = ctypes.c_size_t
    lib.rmb_workspace_bytes.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_long, ctypes.c_long,
                                        ctypes.POINTER(ctypes.c_double)]
    lib.rmb_launches_per_product.restype = ctypes.c_int
    lib.rmb_launches_per_product.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_long,
                                             ctypes.c_long, ctypes.POINTER(ctypes.c_double)]
    lib.rmb_mobility_product.restype = ctypes.c_int
    lib.rmb_mobility_product.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                         ctypes.c_long, ctypes.c_long, ctypes.c_long,
                                         c_dp, c_dp, c_dp, ctypes.c_double, ctypes.c_double,
                                         ctypes.POINTER(ctypes.c_double), c_dp, c_dp, ctypes.c_size_t, c_dp]
    lib.rmb_static_workspace_max.restype = ctypes.c_size_t
    lib.rmb_mobility_product_static.restype = ctypes.c_int
    lib.rmb_mobility_product_static.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                                ctypes.c_long, ctypes.c_long, ctypes.c_long,
                                                c_dp, c_dp, c_dp, ctypes.c_double, ctypes.c_double,
                                                ctypes.POINTER(ctypes.c_double), c_dp, c_dp]
    lib.rmb_source_target_workspace_bytes.restype = ctypes.c_size_t
    lib.rmb_source_target_workspace_bytes.argtypes = [ctypes.c_int, ctypes.c_long, ctypes.c_long,
                                                      ctypes.POINTER(ctypes.c_double)]
    lib.rmb_mobility_source_target.restype = ctypes.c_int
    lib.rmb_mobility_source_target.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_long, ctypes.c_long,
                                               c_dp, c_dp, c_dp, c_dp, c_dp, ctypes.c_double,
                                               ctypes.POINTER(ctypes.c_double), c_dp, c_dp, ctypes.c_size_t, c_dp]
    lib.rmb_symmetric_acc_len.restype = ctypes.c_size_t
    lib.rmb_symmetric_acc_len.argtypes = [ctypes.c_long]
    lib.rmb_symmetric_partial.restype = ctypes.c_int
    lib.rmb_symmetric_partial.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_long, ctypes.c_int, ctypes.c_int, c_dp,
                                          c_dp, c_dp, ctypes.c_double, ctypes.c_double, ctypes.POINTER(ctypes.c_double),
                                          c_dp, c_dp, ctypes.c_size_t, c_dp]
    lib.rmb_symmetric_finish.restype = ctypes.c_int
    lib.rmb_symmetric_finish.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_long, ctypes.c_long, ctypes.c_long,
                                         c_dp, c_dp, c_dp, c_dp, ctypes.c_double, ctypes.c_double, c_dp, c_dp]
    lib.rmb_fma_peak.restype = ctypes.c_int
    lib.rmb_fma_peak.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_double),
                                 ctypes.POINTER(ctypes.c_double)]
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise RmbError("librmb_mobility: error %d: %s" % (rc, load().rmb_last_error().decode()))


def fma_peak(precision=FP64, iters=200000):
    """Measured FMA-pipe peak in FLOP/s (roofline denominator), elapsed ms."""
    fl, ms = ctypes.c_double(), ctypes.c_double()
    check(load().rmb_fma_peak(precision, iters, ctypes.byref(fl), ctypes.byref(ms)))
    return fl.value, ms.value
