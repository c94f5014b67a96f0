"""The triangular schedule of the pair-symmetric kernel, enumerated on the CPU from the product's own
csrc/sym_schedule.cuh (compiled with g++): over all ranks, row groups, row parts and warps every unordered pair
of 32-blob blocks is met exactly once and every diagonal block exactly once."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def lib(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("s") / "host_sym_schedule.so")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-DRMB_HOST_ONLY", "-shared", "-fPIC",
                           os.path.join(HERE, "host_sym_schedule.cpp"), "-o", so])
    lib = ctypes.CDLL(so)
    ip = ctypes.POINTER(ctypes.c_int)
    lib.host_sym_schedule.argtypes = [ctypes.c_long, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                      ctypes.c_long, ctypes.c_int, ip, ip, ip]
    return lib


@pytest.mark.parametrize("nt", [1, 2])
@pytest.mark.parametrize("world", [1, 2, 3, 8])
@pytest.mark.parametrize("n", [1, 255, 256, 257, 5000, 20000])
def test_every_block_pair_once(lib, n, world, nt):
    tile, tpb = 256, 128
    nb = ((n + tile - 1) // tile) * tile // 32
    sym = np.zeros((nb, nb), dtype=np.int32)
    diag = np.zeros(nb, dtype=np.int32)
    split = np.zeros(world, dtype=np.int32)
    ip = ctypes.POINTER(ctypes.c_int)
    rc = lib.host_sym_schedule(n, world, nt, tpb, tile, 148 * 4, 10, sym.ctypes.data_as(ip), diag.ctypes.data_as(ip),
                               split.ctypes.data_as(ip))
    assert rc == nb
    assert (diag == 1).all()
    assert (sym[np.triu_indices(nb, 1)] == 1).all() and (sym[np.tril_indices(nb)] == 0).all()
    assert (split >= 1).all()
    if n == 20000:
        assert split.max() > 1          # small launches are split over blockIdx.y
