"""rmb_mobility_product_static: the allocation-free entry point (scratch in a file-scope `static __device__` array,
no cudaMalloc; include/rmb_mobility.h) against rmb_mobility_product on the same inputs, through the C ABI, two calls
in a row.  No counterpart in the reference.  The file is named so that
pytest collects it last (a failure here must not hide another test)."""
import ctypes

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
TOL64 = 1e-12
UF, WT, UF_UT = 0, 3, 4
SINGLE_WALL, FP64, FP32 = 1, 0, 1
RMB_E_WORKSPACE = -2


@pytest.fixture(scope="module")
def lib():
    from rigidmultiblobswall_b200 import _lib
    return _lib.load()


def cloud(n, seed):
    rng = np.random.default_rng(seed)
    a, eta = 0.13, 7.0
    side = (n * 4 * np.pi / 3 / 0.1) ** (1.0 / 3.0) * a
    r = side * rng.random((n, 3))                  # heights from 0: some blobs below z = a (clamp + damping)
    return (torch.from_numpy(r).cuda().reshape(-1), torch.from_numpy(rng.standard_normal(3 * n)).cuda(),
            torch.from_numpy(rng.standard_normal(3 * n)).cuda(), a, eta)


def both(lib, block, precision, n, r, va, vb, a, eta, L, flags=0):
    """(existing entry point with a caller-owned workspace, new entry point) on the same inputs"""
    Lc = None if L is None else (ctypes.c_double * 3)(*L)
    stream = torch.cuda.current_stream().cuda_stream
    vbp = vb.data_ptr() if block == UF_UT else None
    nbytes = lib.rmb_workspace_bytes(block, precision, n, n, Lc)
    ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    ref = torch.full((3 * n,), float("nan"), dtype=torch.float64, device="cuda")
    out = torch.full((3 * n,), float("nan"), dtype=torch.float64, device="cuda")
    rc = lib.rmb_mobility_product(block, SINGLE_WALL, precision, flags, n, n, 0, r.data_ptr(), va.data_ptr(), vbp,
                                  eta, a, Lc, ref.data_ptr(), ws.data_ptr(), ws.numel(), stream)
    assert rc == 0, lib.rmb_last_error()
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    rc = lib.rmb_mobility_product_static(block, SINGLE_WALL, precision, flags, n, n, 0, r.data_ptr(), va.data_ptr(),
                                         vbp, eta, a, Lc, out.data_ptr(), stream)
    assert rc == 0, lib.rmb_last_error()
    torch.cuda.synchronize()
    return ref, out, free0 - torch.cuda.mem_get_info()[0]


def relerr(x, ref):
    return float(torch.linalg.norm(x - ref) / torch.linalg.norm(ref))


@pytest.mark.parametrize("n,block,L", [(20000, UF, None),                 # pair-symmetric kernel: accumulator must restart at 0
                                       (3000, WT, None),                  # row kernel with split sources: partial sums
                                       (14000, UF_UT, (40.0, 40.0, 0.0)),  # fused block, 9 image boxes
                                       (257, UF, None)])
def test_static_workspace_matches_the_workspace_entry_point_twice_in_a_row(lib, n, block, L):
    assert lib.rmb_static_workspace_max() == 64 << 20
    r, f1, t1, a, eta = cloud(n, seed=n)
    _, f2, t2, _, _ = cloud(n, seed=n + 1)
    grown = []
    for va, vb in ((f1, t1), (f2, t2), (f1, t1)):             # the shared array keeps the previous call's contents
        ref, out, used = both(lib, block, FP64, n, r, va, vb, a, eta, L)
        assert relerr(out, ref) < TOL64
        grown.append(used)
    assert grown[1] <= 0 and grown[2] <= 0                    # after set-up a call takes no device memory
    # a different product in between (other size, other kernel) leaves nothing behind that matters
    r2, g, _, a2, eta2 = cloud(5000, seed=3)
    both(lib, UF, FP64, 5000, r2, g, g, a2, eta2, None)
    ref, out, _ = both(lib, block, FP64, n, r, f2, t2, a, eta, L)
    assert relerr(out, ref) < TOL64


def test_static_workspace_fp32_and_deterministic_flag(lib):
    n = 20000
    r, f, t, a, eta = cloud(n, seed=9)
    ref, out, _ = both(lib, UF, FP32, n, r, f, t, a, eta, None)            # default: fp32 pair-symmetric kernel (atomics)
    assert relerr(out, ref) < 1e-6
    ref, out, _ = both(lib, UF, FP32, n, r, f, t, a, eta, None, flags=8)   # RMB_DETERMINISTIC: fp32 row kernel
    assert torch.equal(out, ref)                              # the row kernel is bit-reproducible
    ref, out, _ = both(lib, UF, FP64, n, r, f, t, a, eta, None, flags=8)      # RMB_DETERMINISTIC
    assert torch.equal(out, ref)


def test_sizes_beyond_the_static_array_are_rejected(lib):
    n = 1000000
    assert lib.rmb_workspace_bytes(UF, FP64, n, n, None) > lib.rmb_static_workspace_max()
    p = ctypes.c_void_p(256)                                  # never dereferenced: the size check comes first
    rc = lib.rmb_mobility_product_static(UF, SINGLE_WALL, FP64, 0, n, n, 0, p, p, None, 1.0, 1.0, None, p, None)
    assert rc == RMB_E_WORKSPACE
    msg = lib.rmb_last_error().decode()
    assert "static" in msg and "rmb_mobility_product" in msg
    assert lib.rmb_mobility_product_static(9, SINGLE_WALL, FP64, 0, 10, 10, 0, p, p, None, 1.0, 1.0, None, p, None) == -1
    assert lib.rmb_mobility_product_static(UF, SINGLE_WALL, FP64, 0, 0, 0, 0, None, None, None, 1.0, 1.0, None, None, None) == 0
