"""GPU parity tests of the pair-symmetric kernels for the free-space (no_wall) rotation blocks UT / WF / WT and the
fused no_wall product u = UF f + UT t: golden vectors of the reference's numba functions, ragged sizes against the
oracle, and at n >= RMB_SYMMETRIC_MIN_N the default path against the row kernel and the multi-rank form."""
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
TOL64 = 1e-12
BLOCK_ID = {"UF": 0, "UT": 1, "WF": 2, "WT": 3}


def relerr(x, ref, floor=0.0):
    """norm-wise relative error; `floor` is the magnitude below which the reference itself is cancellation noise
    (a lone blob in free space: UT / WF give exactly 0, and its periodic images cancel pairwise to ~1e-18)"""
    return np.linalg.norm(np.asarray(x) - ref) / max(np.linalg.norm(ref), floor, 1e-300)


@pytest.fixture(scope="module")
def mb():
    from rigidmultiblobswall_b200 import mobility
    return mobility


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(HERE, "golden", "mf_golden.npz"))


@pytest.mark.parametrize("blk", ["UT", "WF", "WT"])
@pytest.mark.parametrize("lname", ["open", "xy", "x", "xyz"])
@pytest.mark.parametrize("cloud", ["dense", "dilute"])
def test_golden(mb, golden, cloud, lname, blk):
    r, v = golden[f"{cloud}/r"], golden[f"{cloud}/v"]
    a, eta = golden[f"{cloud}/a_eta"]
    L = golden[f"{cloud}/L_{lname}"]                 # xyz: periodic in z -> the library takes the row kernel
    out = mb.mobility_product(BLOCK_ID[blk], mb.NO_WALL, r, v, eta, a, periodic_length=L, flags=mb.SYMMETRIC)
    assert relerr(out, golden[f"{cloud}/{lname}/no_wall/{blk}"]) < TOL64


@pytest.mark.parametrize("n", [1, 2, 33, 255, 257, 1000, 3001])
def test_ragged_vs_oracle(mb, n):
    import oracle
    rng = np.random.default_rng(n + 29)
    a, eta = 0.21, 1.3
    side = (2.0 + 0.3 * n ** (1 / 3)) * a
    r = side * rng.random((n, 3)) - 0.5 * side                     # no wall: heights of either sign
    v, t = rng.standard_normal((n, 3)), rng.standard_normal((n, 3))
    for L in (None, np.array([side, 1.2 * side, 0.0])):
        if L is not None and n > 1000:
            continue
        ref = {blk: oracle.mobility_product(blk, "no_wall", r, v, eta, a, periodic_length=L) for blk in ("UT", "WF", "WT")}
        for blk in ("UT", "WF", "WT"):
            out = mb.mobility_product(BLOCK_ID[blk], mb.NO_WALL, r, v, eta, a, periodic_length=L, flags=mb.SYMMETRIC)
            assert relerr(out, ref[blk], floor=np.linalg.norm(v) / (8 * np.pi * eta * a ** 2)) < TOL64, (blk, L)
        fused = oracle.mobility_product("UF", "no_wall", r, v, eta, a, periodic_length=L) + \
            oracle.mobility_product("UT", "no_wall", r, t, eta, a, periodic_length=L)
        out = mb.mobility_product(mb.UF_UT, mb.NO_WALL, r, v, eta, a, periodic_length=L, vec_b=t, flags=mb.SYMMETRIC)
        assert relerr(out, fused) < TOL64, ("UF_UT", L)


@pytest.mark.parametrize("blk", ["UT", "WF", "WT", "UF_UT"])
def test_large_default_and_multi_rank(mb, blk):
    from rigidmultiblobswall_b200 import _lib
    lib = _lib.load()
    rng = np.random.default_rng(47)
    n, a, eta = 20000, 0.13, 7.0
    b = mb.UF_UT if blk == "UF_UT" else BLOCK_ID[blk]
    assert lib.rmb_launches_per_product(b, mb.NO_WALL, mb.FP64, 0, 10 ** 7, 10 ** 7, None) == 3
    side = (n * 4 * np.pi / 3 / 0.1) ** (1 / 3) * a
    r = torch.from_numpy(side * rng.random((n, 3))).cuda()
    v = torch.from_numpy(rng.standard_normal((n, 3))).cuda()
    t = torch.from_numpy(rng.standard_normal((n, 3))).cuda() if blk == "UF_UT" else None
    for L in (None, np.array([side, side, 0.0])):
        det = mb.mobility_product(b, mb.NO_WALL, r, v, eta, a, periodic_length=L, vec_b=t, flags=mb.DETERMINISTIC)
        sym = mb.mobility_product(b, mb.NO_WALL, r, v, eta, a, periodic_length=L, vec_b=t)
        assert float(torch.linalg.norm(sym - det) / torch.linalg.norm(det)) < 1e-13
        # self-comparison is not parity: 192 random rows of both against the oracle
        from test_mobility_gpu import oracle_rows, rows3
        idx, orc = oracle_rows("UF" if blk == "UF_UT" else blk, "no_wall", r, v, eta, a, L, seed=9)
        if blk == "UF_UT":
            orc = orc + oracle_rows("UT", "no_wall", r, t, eta, a, L, seed=9)[1]
        assert relerr(rows3(det, idx), orc) < TOL64 and relerr(rows3(sym, idx), orc) < TOL64
        world = 3
        acc = sum(mb.symmetric_partial(r, v, eta, a, rank=k, world=world, block=b, wall=mb.NO_WALL, vec_b=t,
                                       periodic_length=L).clone() for k in range(world))
        out = torch.cat([mb.symmetric_finish(r, v, acc, eta, a, targets=tg, block=b, wall=mb.NO_WALL, vec_b=t)
                         for tg in ((0, 9000), (9000, 11000))])
        assert float(torch.linalg.norm(out - det) / torch.linalg.norm(det)) < 1e-13
