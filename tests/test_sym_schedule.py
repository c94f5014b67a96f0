"""The triangular schedule of the pair-symmetric kernel, enumerated on the CPU from the product's own
csrc/sym_schedule.cuh (compiled with g++): over all ranks, work items (row group x chunk of tiles) and warps every
unordered pair of 32-blob blocks is met exactly once and every diagonal block exactly once; the items of a row tile
it without gaps; the rectangular grid wastes few cells."""
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
                                      ctypes.c_long, ctypes.c_int, ctypes.c_int, ctypes.c_int, ip, ip, ip]
    return lib


@pytest.mark.parametrize("nt", [1, 2])
@pytest.mark.parametrize("world", [1, 2, 3, 8])
@pytest.mark.parametrize("n", [1, 255, 256, 257, 5000, 20000])
def test_every_block_pair_once(lib, n, world, nt):
    tile, tpb = 256, 128
    nb = ((n + tile - 1) // tile) * tile // 32
    sym = np.zeros((nb, nb), dtype=np.int32)
    diag = np.zeros(nb, dtype=np.int32)
    info = np.zeros((world, 5), dtype=np.int32)
    ip = ctypes.POINTER(ctypes.c_int)
    rc = lib.host_sym_schedule(n, world, nt, tpb, tile, 148 * 5, 128, 1, 16, sym.ctypes.data_as(ip),
                               diag.ctypes.data_as(ip), info.ctypes.data_as(ip))
    assert rc == nb
    assert (diag == 1).all()
    assert (sym[np.triu_indices(nb, 1)] == 1).all() and (sym[np.tril_indices(nb)] == 0).all()
    active = info[:, 0] > 0
    assert ((info[active, 2] >= 1) & (info[active, 2] <= 16)).all()
    if n == 20000:
        assert info[:, 3].min() > 1          # small launches are cut into several items


@pytest.mark.parametrize("world", [1, 8])
def test_items_are_equal_sized_and_the_grid_is_tight(lib, world):
    """BASELINE config 4 (N = 1M; the grid walk only, no count matrix): every rank's launch has > 100 items per
    resident CTA slot, (almost) all of them full chunks, < 3 % of the grid cells invalid, and the ranks' shares of
    the pairs equal to 1 %."""
    n, tile, tpb = 1000000, 256, 128
    info = np.zeros((world, 5), dtype=np.int32)
    ip = ctypes.POINTER(ctypes.c_int)
    rc = lib.host_sym_schedule(n, world, 1, tpb, tile, 148 * 5, 128, 1, 16, None, None, info.ctypes.data_as(ip))
    assert rc > 0
    for gx, gy, chunk, items, tile_rows in info:
        assert items >= 0.97 * gx * gy
        assert tile_rows >= 0.95 * items * chunk
        assert items >= 100 * 148 * 5
    per_rank = info[:, 4].astype(float)
    assert per_rank.max() / per_rank.min() < 1.01
