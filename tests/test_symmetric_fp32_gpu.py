"""The fp32 variant of the pair-symmetric kernel (fp32 pair arithmetic, fp64 sums: per block of 32 partners the
contributions are added in fp32, then to fp64 registers / global atomics): golden vectors of the reference's numba
functions, ragged sizes against the oracle, and the default path at n >= RMB_SYMMETRIC_MIN_N, all at north_star's
fp32 bar of 1e-5 (measured: ~5e-7)."""
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
TOL32 = 1e-5
BLOCK_ID = {"UF": 0, "UT": 1, "WF": 2, "WT": 3}


def relerr(x, ref):
    x = x.detach().cpu().numpy() if hasattr(x, "detach") else np.asarray(x)
    return float(np.linalg.norm(x.reshape(-1) - np.asarray(ref).reshape(-1)) / np.linalg.norm(ref))


@pytest.fixture(scope="module")
def mb():
    from rigidmultiblobswall_b200 import mobility
    return mobility


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(HERE, "golden", "mf_golden.npz"))


@pytest.mark.parametrize("blk", ["UF", "UT", "WF", "WT"])
@pytest.mark.parametrize("wname", ["single_wall", "no_wall"])
@pytest.mark.parametrize("lname", ["open", "x", "xy"])
@pytest.mark.parametrize("cloud", ["dense", "dilute"])
def test_golden(mb, golden, cloud, lname, wname, blk):
    r, v = golden[f"{cloud}/r"], golden[f"{cloud}/v"]
    a, eta = golden[f"{cloud}/a_eta"]
    if wname == "single_wall":
        r = r + np.array([0.0, 0.0, a])      # fp32 bar: keep clear of the z < a damping edge cases of the dense cloud
        import oracle
        ref = oracle.mobility_product(blk, wname, r, v, eta, a, golden[f"{cloud}/L_{lname}"])
    else:
        ref = golden[f"{cloud}/{lname}/{wname}/{blk}"]
    wall = mb.SINGLE_WALL if wname == "single_wall" else mb.NO_WALL
    out = mb.mobility_product(BLOCK_ID[blk], wall, r, v, eta, a, periodic_length=golden[f"{cloud}/L_{lname}"],
                              precision=mb.FP32, flags=mb.SYMMETRIC)
    assert relerr(out, ref) < TOL32


@pytest.mark.parametrize("n", [1, 2, 31, 32, 33, 127, 128, 129, 255, 256, 257, 513, 3001])
def test_ragged_sizes_vs_oracle(mb, n):
    import oracle
    rng = np.random.default_rng(n)
    a, eta = 0.21, 1.3
    r = (2.0 + 0.3 * n ** (1 / 3)) * a * rng.random((n, 3))      # dense: overlaps and blobs below z = a
    f = rng.standard_normal((n, 3))
    ref = oracle.mobility_product("UF", "single_wall", r, f, eta, a)
    out = mb.mobility_product(mb.UF, mb.SINGLE_WALL, r, f, eta, a, precision=mb.FP32, flags=mb.SYMMETRIC)
    assert relerr(out, ref) < TOL32
    t = rng.standard_normal((n, 3))
    ref = ref + oracle.mobility_product("UT", "single_wall", r, t, eta, a)
    out = mb.mobility_product(mb.UF_UT, mb.SINGLE_WALL, r, f, eta, a, vec_b=t, precision=mb.FP32, flags=mb.SYMMETRIC)
    assert relerr(out, ref) < TOL32


def test_default_selection_large(mb):
    """n >= RMB_SYMMETRIC_MIN_N: the fp32 wrappers take the pair-symmetric kernel (3 launches), deterministic=True the
    row kernel (bit-reproducible); both within 1e-5 of 192 random oracle rows, open and with 9 image boxes."""
    from rigidmultiblobswall_b200 import _lib
    from test_mobility_gpu import oracle_rows, rows3
    lib = _lib.load()
    rng = np.random.default_rng(5)
    n, a, eta = 30000, 0.13, 7.0
    assert lib.rmb_launches_per_product(mb.UF, mb.SINGLE_WALL, mb.FP32, 0, n, n, None) == 3
    assert lib.rmb_launches_per_product(mb.UF, mb.SINGLE_WALL, mb.FP32, mb.DETERMINISTIC, 10 ** 7, 10 ** 7, None) == 2
    side = (n * 4 * np.pi / 3 / 0.1) ** (1 / 3) * a
    r = torch.from_numpy(side * rng.random((n, 3)) + np.array([0, 0, a])).cuda()
    f = torch.from_numpy(rng.standard_normal((n, 3))).cuda()
    for L in (None, np.array([side, side, 0.0])):
        sym = mb.single_wall_mobility_trans_times_force_b200_fp32(r, f, eta, a, periodic_length=L)
        row = mb.single_wall_mobility_trans_times_force_b200_fp32(r, f, eta, a, periodic_length=L, deterministic=True)
        row2 = mb.single_wall_mobility_trans_times_force_b200_fp32(r, f, eta, a, periodic_length=L, deterministic=True)
        assert torch.equal(row, row2)
        idx, orc = oracle_rows("UF", "single_wall", r, f, eta, a, L, seed=11)
        assert relerr(rows3(sym, idx), orc) < TOL32 and relerr(rows3(row, idx), orc) < TOL32
        assert not torch.equal(sym, row)                      # two different kernels did run
        for blk in ("UT", "WF", "WT"):
            out = mb.mobility_product(BLOCK_ID[blk], mb.SINGLE_WALL, r, f, eta, a, periodic_length=L, precision=mb.FP32)
            idx, orc = oracle_rows(blk, "single_wall", r, f, eta, a, L, seed=12)
            assert relerr(rows3(out, idx), orc) < TOL32, blk


def test_deterministic_keyword_on_the_wrappers_and_the_sharded_product(mb):
    """deterministic=True (an extra keyword of the *_b200 wrappers, and of parallel.ShardedMobility) keeps the
    one-thread-per-target kernel: bit-reproducible run to run, fp64 and fp32, above the 8,000-blob threshold."""
    from rigidmultiblobswall_b200 import parallel
    rng = np.random.default_rng(8)
    n, a, eta = 12000, 0.13, 7.0
    side = (n * 4 * np.pi / 3 / 0.1) ** (1 / 3) * a
    r = torch.from_numpy(side * rng.random((n, 3))).cuda()
    f = torch.from_numpy(rng.standard_normal((n, 3))).cuda()
    t = torch.from_numpy(rng.standard_normal((n, 3))).cuda()
    for fn in (mb.single_wall_mobility_trans_times_force_b200, mb.single_wall_mobility_trans_times_force_b200_fp32,
               mb.no_wall_mobility_rot_times_torque_b200):
        d1, d2 = fn(r, f, eta, a, deterministic=True), fn(r, f, eta, a, deterministic=True)
        assert torch.equal(d1, d2)
        s = fn(r, f, eta, a)
        assert not torch.equal(s, d1) and relerr(s, d1.cpu().numpy()) < (1e-5 if fn.__name__.endswith("fp32") else 1e-13)
    d1 = mb.single_wall_mobility_trans_times_force_torque_b200(r, f, t, eta, a, deterministic=True)
    assert torch.equal(d1, mb.single_wall_mobility_trans_times_force_torque_b200(r, f, t, eta, a, deterministic=True))
    sm_d = parallel.ShardedMobility(r, eta, a, group=False, deterministic=True)
    sm_s = parallel.ShardedMobility(r, eta, a, group=False)
    assert not sm_d.symmetric and sm_s.symmetric
    assert torch.equal(sm_d(f.reshape(-1)), sm_d(f.reshape(-1)))
    assert relerr(sm_s(f.reshape(-1)), sm_d(f.reshape(-1)).cpu().numpy()) < 1e-13
