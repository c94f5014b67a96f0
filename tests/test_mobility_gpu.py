"""GPU parity tests: the CUDA path (through the C ABI) against
 (1) the committed golden outputs of the reference's numba implementation,
 (2) the CPU oracle (oracle/mobility_oracle.c) on seeded inputs,
 (3) size-independent properties at large N.
Tolerances are BASELINE.json's: 1e-12 relative (fp64), 1e-5 (fp32), norm-wise."""
import ctypes
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
TOL64, TOL32 = 1e-12, 1e-5
BLOCK_ID = {"UF": 0, "UT": 1, "WF": 2, "WT": 3}


def _cases():
    g = np.load(os.path.join(HERE, "golden", "mf_golden.npz"))
    return [str(c) for c in g["cases"]]


def relerr(x, ref):
    return np.linalg.norm(np.asarray(x) - ref) / np.linalg.norm(ref)


def oracle_rows(block, wall, r, v, eta, a, L=None, m=192, seed=0, in_plane=False):
    """(idx, oracle rows of M v for m RANDOM targets): self-comparison between two of our kernels is not parity, so
    every such test also checks both against the oracle.  The product is equivariant under a permutation of the
    blobs: the oracle's contiguous `targets` range is applied to a copy with the sampled blobs moved to the front."""
    import oracle
    r = (r.detach().cpu().numpy() if hasattr(r, "detach") else np.asarray(r)).reshape(-1, 3)
    v = (v.detach().cpu().numpy() if hasattr(v, "detach") else np.asarray(v)).reshape(-1, 3)
    n = r.shape[0]
    rng = np.random.default_rng(seed)
    idx = np.sort(rng.choice(n, min(m, n), replace=False))
    perm = np.concatenate([idx, np.setdiff1d(np.arange(n), idx, assume_unique=True)])
    ref = oracle.mobility_product(block, wall, r[perm], v[perm], eta, a, L, in_plane=in_plane, targets=(0, len(idx)))
    return idx, ref


def rows3(u, idx):
    u = u.detach().cpu().numpy() if hasattr(u, "detach") else np.asarray(u)
    return u.reshape(-1, 3)[idx].reshape(-1)


@pytest.fixture(scope="module")
def mb():
    from rigidmultiblobswall_b200 import mobility
    assert torch.cuda.is_available()
    return mobility


def _run(mb, case, golden, precision, flags=0):
    cloud, lname, wall, blk = case.split("/")
    r, v = golden[f"{cloud}/r"], golden[f"{cloud}/v"]
    a, eta = golden[f"{cloud}/a_eta"]
    L = golden[f"{cloud}/L_{lname}"]
    fl = flags | (mb.IN_PLANE if wall == "in_plane" else 0)
    w = {"no_wall": mb.NO_WALL, "free_surface": mb.FREE_SURFACE}.get(wall, mb.SINGLE_WALL)
    return mb.mobility_product(BLOCK_ID[blk], w, r, v, eta, a, periodic_length=L, precision=precision, flags=fl)


@pytest.mark.parametrize("case", _cases())
def test_golden_fp64(mb, golden, case):
    out = _run(mb, case, golden, mb.FP64)
    assert out.shape == golden[case].shape
    assert relerr(out, golden[case]) < TOL64


@pytest.mark.parametrize("case", [c for c in _cases() if "/open/" in c or "/xy/" in c])
def test_golden_fp64_plain_loads(mb, golden, case):
    """same kernels with the non-TMA tile loader (RMB_NO_TMA)"""
    out = _run(mb, case, golden, mb.FP64, flags=mb.NO_TMA)
    assert relerr(out, golden[case]) < TOL64


@pytest.mark.parametrize("case", [c for c in _cases() if c.startswith(("dilute/", "c1/"))])
def test_golden_fp32(mb, golden, case):
    out = _run(mb, case, golden, mb.FP32)
    assert relerr(out, golden[case]) < TOL32


def test_named_wrappers_match_reference_signature(mb, golden):
    r, v = golden["dense/r"], golden["dense/v"]
    a, eta = golden["dense/a_eta"]
    L = golden["dense/L_xy"]
    u = mb.single_wall_mobility_trans_times_force_b200(r, v, eta, a, periodic_length=L)
    assert isinstance(u, np.ndarray) and u.shape == (r.size,)
    assert relerr(u, golden["dense/xy/single_wall/UF"]) < TOL64
    # flat r_vectors, non-contiguous force slice, extra positional args are ignored like the reference
    big = np.zeros((r.shape[0], 6)); big[:, :3] = v
    u2 = mb.single_wall_mobility_trans_times_force_b200(r.flatten(), big[:, :3], eta, a, "ignored", periodic_length=L)
    assert np.array_equal(u, u2)
    w = mb.no_wall_mobility_rot_times_torque_b200(r, v, eta, a)
    assert relerr(w, golden["dense/open/no_wall/WT"]) < TOL64
    u32 = mb.single_wall_mobility_trans_times_force_b200_fp32(golden["dilute/r"], golden["dilute/v"],
                                                              golden["dilute/a_eta"][1], golden["dilute/a_eta"][0])
    assert relerr(u32, golden["dilute/open/single_wall/UF"]) < TOL32


def test_inputs_not_mutated_and_torch_stays_on_device(mb, golden):
    r, v = golden["dense/r"].copy(), golden["dense/v"].copy()
    a, eta = golden["dense/a_eta"]
    r0, v0 = r.copy(), v.copy()
    mb.single_wall_mobility_trans_times_force_b200(r, v, eta, a)
    assert np.array_equal(r, r0) and np.array_equal(v, v0)
    rt, vt = torch.from_numpy(r).cuda(), torch.from_numpy(v).cuda()
    ut = mb.single_wall_mobility_trans_times_force_b200(rt, vt, eta, a)
    assert isinstance(ut, torch.Tensor) and ut.is_cuda and ut.dtype == torch.float64
    assert relerr(ut.cpu().numpy(), golden["dense/open/single_wall/UF"]) < TOL64
    assert torch.equal(rt.cpu(), torch.from_numpy(r0))


@pytest.mark.parametrize("wall", ["single_wall", "no_wall"])
@pytest.mark.parametrize("lname", ["open", "xy"])
def test_fused_force_torque(mb, golden, wall, lname):
    """fused UF f + UT t (mobility_pycuda.py:1266): oracle = sum of the two golden-checked products"""
    r, f = golden["dense/r"], golden["dense/v"]
    a, eta = golden["dense/a_eta"]
    L = golden[f"dense/L_{lname}"]
    t = f[::-1].copy()
    import oracle
    ref = golden[f"dense/{lname}/{wall}/UF"] + oracle.mobility_product("UT", wall, r, t, eta, a, L)
    fn = getattr(mb, f"{wall}_mobility_trans_times_force_torque_b200")
    assert relerr(fn(r, f, t, eta, a, periodic_length=L), ref) < TOL64


@pytest.mark.parametrize("n", [1, 2, 31, 127, 128, 129, 255, 256, 257, 1000, 3001])
def test_ragged_sizes_vs_oracle(mb, n):
    """sizes around the CTA (128) and tile (256) boundaries, incl. N=1 (self term only)"""
    import oracle
    rng = np.random.default_rng(n)
    a, eta = 0.21, 1.3
    side = max(2.0, (n * 4 * np.pi / 3 / 0.15) ** (1 / 3)) * a
    r = side * rng.random((n, 3)); r[:, 2] += 0.3 * a          # some blobs below z = a
    v = rng.standard_normal((n, 3))
    for blk in ("UF", "UT", "WF", "WT"):
        ref = oracle.mobility_product(blk, "single_wall", r, v, eta, a)
        out = mb.mobility_product(BLOCK_ID[blk], mb.SINGLE_WALL, r, v, eta, a)
        assert relerr(out, ref) < TOL64, (blk, n)


def test_empty_input(mb):
    out = mb.single_wall_mobility_trans_times_force_b200(np.zeros((0, 3)), np.zeros((0, 3)), 1.0, 1.0)
    assert out.shape == (0,)


def test_source_split_path_vs_oracle(mb):
    """N = 5000: too few target CTAs to fill the chip -> sources are split over blockIdx.y + finish kernel"""
    import oracle
    rng = np.random.default_rng(7)
    n, a, eta = 5000, 0.5, 1.0
    r = 40 * a * rng.random((n, 3)); r[:, 2] += 0.5 * a
    v = rng.standard_normal((n, 3))
    L = np.array([45 * a, 50 * a, 0.0])
    for blk in ("UF", "WT"):
        ref = oracle.mobility_product(blk, "single_wall", r, v, eta, a, L)
        out = mb.mobility_product(BLOCK_ID[blk], mb.SINGLE_WALL, r, v, eta, a, periodic_length=L)
        assert relerr(out, ref) < TOL64


def test_target_shards_reassemble(mb, golden):
    """multi-GPU decomposition on one device: target ranges computed separately give the full product"""
    r, v = golden["c1/r"], golden["c1/v"]
    a, eta = golden["c1/a_eta"]
    full = mb.single_wall_mobility_trans_times_force_b200(r, v, eta, a)
    parts = [mb.mobility_product(mb.UF, mb.SINGLE_WALL, r, v, eta, a, targets=(o, c))
             for o, c in ((0, 300), (300, 450), (750, 250))]
    assert np.array_equal(np.concatenate(parts), full)
    assert relerr(full, golden["c1/open/single_wall/UF"]) < TOL64


def test_large_n_properties_and_sampled_oracle(mb):
    """N = 60k (3.6e9 pairs): sampled rows against the oracle, linearity, UF symmetry f.(M g) = g.(M f)"""
    import oracle
    rng = np.random.default_rng(11)
    n, a, eta = 60000, 0.13, 7.0
    side = (n * 4 * np.pi / 3 / 0.1) ** (1 / 3) * a
    r = side * rng.random((n, 3)); r[:, 2] += a
    f, g = rng.standard_normal((n, 3)), rng.standard_normal((n, 3))
    rt, ft, gt = (torch.from_numpy(x).cuda() for x in (r, f, g))
    fn = mb.single_wall_mobility_trans_times_force_b200
    uf, ug = fn(rt, ft, eta, a), fn(rt, gt, eta, a)
    ufg = fn(rt, 2.0 * ft - 0.5 * gt, eta, a)
    assert float(torch.linalg.norm(ufg - (2.0 * uf - 0.5 * ug)) / torch.linalg.norm(ufg)) < 1e-13
    sym = abs(float(torch.dot(gt.reshape(-1), uf) - torch.dot(ft.reshape(-1), ug)))
    assert sym / abs(float(torch.dot(gt.reshape(-1), uf))) < 1e-11
    b, e = 31337, 31337 + 64
    ref = oracle.mobility_product("UF", "single_wall", r, f, eta, a, targets=(b, e))
    assert relerr(uf[3 * b:3 * e].cpu().numpy(), ref) < TOL64


def test_transposition_identity_ut_wf(mb, golden):
    """UT = WF^T: f.(UT t) = t.(WF f), with wall and images"""
    r, f = golden["dense/r"], golden["dense/v"]
    a, eta = golden["dense/a_eta"]
    L = golden["dense/L_xy"]
    z_ok = r.copy(); z_ok[:, 2] += a          # identity holds for the undamped operator
    t = f[::-1].copy()
    ut = mb.single_wall_mobility_trans_times_torque_b200(z_ok, t, eta, a, periodic_length=L)
    wf = mb.single_wall_mobility_rot_times_force_b200(z_ok, f, eta, a, periodic_length=L)
    lhs, rhs = np.dot(f.flatten(), ut), np.dot(t.flatten(), wf)
    assert abs(lhs - rhs) / abs(lhs) < 1e-11


def test_rsqrt_seed_accuracy_via_far_pairs(mb):
    """two blobs far from each other and from the wall: u ~ Oseen; checks the MUFU.RSQ64H + cubic step end to end"""
    import oracle
    a, eta = 1.0, 1.0
    rng = np.random.default_rng(3)
    for scale in (1e0, 1e3, 1e6, 1e-2):
        r = np.array([[0.0, 0.0, 5.0], [3.0 * scale + 3.0, 1.0, 7.0 * scale + 5.0]])
        v = rng.standard_normal((2, 3))
        ref = oracle.mobility_product("UF", "single_wall", r, v, eta, a)
        out = mb.single_wall_mobility_trans_times_force_b200(r, v, eta, a)
        assert np.max(np.abs(out - ref) / np.abs(ref)) < 1e-13


# ---- pair-symmetric path (RMB_SYMMETRIC: every unordered pair once, both blobs updated) ----
@pytest.mark.parametrize("flags", [4, 4 | 2])
@pytest.mark.parametrize("cloud", ["dense", "dilute", "c1"])
def test_symmetric_golden(mb, golden, cloud, flags):
    r, v = golden[f"{cloud}/r"], golden[f"{cloud}/v"]
    a, eta = golden[f"{cloud}/a_eta"]
    out = mb.mobility_product(mb.UF, mb.SINGLE_WALL, r, v, eta, a, flags=flags)
    assert relerr(out, golden[f"{cloud}/open/single_wall/UF"]) < TOL64


@pytest.mark.parametrize("n", [1, 2, 31, 32, 33, 127, 128, 129, 255, 256, 257, 513, 1000, 3001])
def test_symmetric_ragged_sizes_vs_oracle(mb, n):
    import oracle
    rng = np.random.default_rng(n)
    a, eta = 0.21, 1.3
    r = (2.0 + 0.3 * n ** (1 / 3)) * a * rng.random((n, 3))      # dense: overlaps and blobs below z = a
    f = rng.standard_normal((n, 3))
    ref = oracle.mobility_product("UF", "single_wall", r, f, eta, a)
    out = mb.mobility_product(mb.UF, mb.SINGLE_WALL, r, f, eta, a, flags=mb.SYMMETRIC)
    assert relerr(out, ref) < TOL64


def test_symmetric_large_matches_default_and_in_plane(mb):
    rng = np.random.default_rng(77)
    n, a, eta = 50000, 0.13, 7.0
    side = (n * 4 * np.pi / 3 / 0.1) ** (1.0 / 3.0) * a
    r = side * rng.random((n, 3))
    f = rng.standard_normal((n, 3))
    ref = mb.mobility_product(mb.UF, mb.SINGLE_WALL, r, f, eta, a)
    out = mb.mobility_product(mb.UF, mb.SINGLE_WALL, r, f, eta, a, flags=mb.SYMMETRIC)
    assert relerr(out, ref) < 1e-13
    idx, orc = oracle_rows("UF", "single_wall", r, f, eta, a, seed=1)
    assert relerr(rows3(out, idx), orc) < TOL64 and relerr(rows3(ref, idx), orc) < TOL64
    ref = mb.mobility_product(mb.UF, mb.SINGLE_WALL, r, f, eta, a, flags=mb.IN_PLANE)
    out = mb.mobility_product(mb.UF, mb.SINGLE_WALL, r, f, eta, a, flags=mb.IN_PLANE | mb.SYMMETRIC)
    assert relerr(out, ref) < 1e-13
    idx, orc = oracle_rows("UF", "single_wall", r, f, eta, a, seed=2, in_plane=True)
    assert relerr(rows3(out, idx), orc) < TOL64 and relerr(rows3(ref, idx), orc) < TOL64
    # a request the symmetric kernel does not cover (target subset) silently takes the default path
    part = mb.mobility_product(mb.UF, mb.SINGLE_WALL, r, f, eta, a, flags=mb.SYMMETRIC, targets=(100, 300))
    assert relerr(part, ref_full := mb.mobility_product(mb.UF, mb.SINGLE_WALL, r, f, eta, a)[300:1200]) < 1e-13


def test_default_selection_and_deterministic_flag(mb):
    """n >= RMB_SYMMETRIC_MIN_N: the pair-symmetric kernel is the default; RMB_DETERMINISTIC keeps the row kernel,
    which is bit-reproducible call by call and whose target shards agree with the full product to rounding; the multi-rank form (partial / finish with world = 3 on one
    device, sums added by hand) gives the same product."""
    from rigidmultiblobswall_b200 import _lib
    lib = _lib.load()
    rng = np.random.default_rng(21)
    n, a, eta = 20000, 0.13, 7.0
    assert lib.rmb_launches_per_product(mb.UF, mb.SINGLE_WALL, mb.FP64, 0, n, n, None) == 3
    side = (n * 4 * np.pi / 3 / 0.1) ** (1 / 3) * a
    r = torch.from_numpy(side * rng.random((n, 3))).cuda()
    f = torch.from_numpy(rng.standard_normal((n, 3))).cuda()
    det = mb.mobility_product(mb.UF, mb.SINGLE_WALL, r, f, eta, a, flags=mb.DETERMINISTIC)
    det2 = mb.mobility_product(mb.UF, mb.SINGLE_WALL, r, f, eta, a, flags=mb.DETERMINISTIC)
    assert torch.equal(det, det2)
    # target shards: the same rows to rounding (the source split, hence the summation order of a row, follows the
    # number of target CTAs of the call, so shards are reproducible call by call but not bit-equal to the full product)
    parts = [mb.mobility_product(mb.UF, mb.SINGLE_WALL, r, f, eta, a, targets=(o, c))
             for o, c in ((0, 7000), (7000, 6000), (13000, 7000))]
    parts2 = [mb.mobility_product(mb.UF, mb.SINGLE_WALL, r, f, eta, a, targets=(o, c))
              for o, c in ((0, 7000), (7000, 6000), (13000, 7000))]
    assert torch.equal(torch.cat(parts), torch.cat(parts2))
    assert float(torch.linalg.norm(torch.cat(parts) - det) / torch.linalg.norm(det)) < 1e-14
    sym = mb.mobility_product(mb.UF, mb.SINGLE_WALL, r, f, eta, a)
    assert float(torch.linalg.norm(sym - det) / torch.linalg.norm(det)) < 1e-13
    idx, orc = oracle_rows("UF", "single_wall", r, f, eta, a, seed=3)
    assert relerr(rows3(det, idx), orc) < TOL64 and relerr(rows3(sym, idx), orc) < TOL64
    world = 3
    acc = sum(mb.symmetric_partial(r, f, eta, a, rank=k, world=world).clone() for k in range(world))
    # the finish step reads its arguments only: unrelated products on the same stream between the two steps
    # (a bigger cloud, another block, a source-target call) must not change the result
    r_big = torch.from_numpy(rng.random((n + 5000, 3)) + 1.0).cuda()
    mb.mobility_product(mb.WT, mb.SINGLE_WALL, r_big, r_big, eta, a)
    mb.mobility_source_target(mb.NO_WALL, r_big[:100], r, r_big[:100], torch.ones(100).cuda().double(),
                              torch.ones(n).cuda().double(), eta)
    out = torch.cat([mb.symmetric_finish(r, f, acc, eta, a, targets=t) for t in ((0, 9000), (9000, 11000))])
    assert float(torch.linalg.norm(out - det) / torch.linalg.norm(det)) < 1e-13


def test_sharded_mobility_single_rank_uses_symmetric_kernel(mb):
    from rigidmultiblobswall_b200 import parallel
    rng = np.random.default_rng(22)
    n, a, eta = 17000, 0.2, 1.0
    r = torch.from_numpy(9.0 * rng.random((n, 3))).cuda()
    f = torch.from_numpy(rng.standard_normal((n, 3))).cuda().reshape(-1)
    sm = parallel.ShardedMobility(r, eta, a, group=False)
    assert sm.symmetric
    ref = mb.mobility_product(mb.UF, mb.SINGLE_WALL, r, f, eta, a, flags=mb.DETERMINISTIC)
    assert float(torch.linalg.norm(sm(f) - ref) / torch.linalg.norm(ref)) < 1e-13
    Lx = np.array([9.0, 0, 0])
    smx = parallel.ShardedMobility(r, eta, a, group=False, periodic_length=Lx)
    assert smx.symmetric
    refx = mb.mobility_product(mb.UF, mb.SINGLE_WALL, r, f, eta, a, periodic_length=Lx, flags=mb.DETERMINISTIC)
    assert float(torch.linalg.norm(smx(f) - refx) / torch.linalg.norm(refx)) < 1e-13
    assert not parallel.ShardedMobility(r, eta, a, group=False, periodic_length=np.array([9.0, 0, 9.0])).symmetric


@pytest.mark.parametrize("n", [33, 1000, 3001, 40000])
def test_symmetric_no_wall(mb, golden, n):
    """free-space twin of the pair-symmetric kernel (no_wall UF): forced at small n, default at 40000"""
    import oracle
    rng = np.random.default_rng(n + 5)
    a, eta = 0.21, 1.3
    r = (2.0 + 0.5 * n ** (1 / 3)) * a * rng.random((n, 3)) - 1.0      # no wall: negative heights are legal
    f = rng.standard_normal((n, 3))
    out = mb.mobility_product(mb.UF, mb.NO_WALL, r, f, eta, a, flags=mb.SYMMETRIC if n < 20000 else 0)
    if n <= 3001:
        assert relerr(out, oracle.mobility_product("UF", "no_wall", r, f, eta, a)) < TOL64
    else:
        det = mb.mobility_product(mb.UF, mb.NO_WALL, r, f, eta, a, flags=mb.DETERMINISTIC)
        assert relerr(out, det) < 1e-13
        ref = oracle.mobility_product("UF", "no_wall", r, f, eta, a, targets=(777, 777 + 64))
        assert relerr(out[3 * 777:3 * (777 + 64)], ref) < TOL64
    for cloud in ("dense", "dilute"):
        rr, vv = golden[f"{cloud}/r"], golden[f"{cloud}/v"]
        aa, ee = golden[f"{cloud}/a_eta"]
        o = mb.mobility_product(mb.UF, mb.NO_WALL, rr, vv, ee, aa, flags=mb.SYMMETRIC)
        assert relerr(o, golden[f"{cloud}/open/no_wall/UF"]) < TOL64


# ---- pair-symmetric kernel with pseudo-periodic images in x / y (box b seen from i is box -b seen from j) ----
@pytest.mark.parametrize("wall", ["single_wall", "no_wall"])
@pytest.mark.parametrize("lname", ["xy", "x", "xyz"])
@pytest.mark.parametrize("cloud", ["dense", "dilute"])
def test_symmetric_periodic_golden(mb, golden, cloud, lname, wall):
    """xyz (periodic z) is outside the symmetric kernel: the flag must then fall back to the row kernel"""
    r, v = golden[f"{cloud}/r"], golden[f"{cloud}/v"]
    a, eta = golden[f"{cloud}/a_eta"]
    L = golden[f"{cloud}/L_{lname}"]
    w = mb.SINGLE_WALL if wall == "single_wall" else mb.NO_WALL
    out = mb.mobility_product(mb.UF, w, r, v, eta, a, periodic_length=L, flags=mb.SYMMETRIC)
    assert relerr(out, golden[f"{cloud}/{lname}/{wall}/UF"]) < TOL64
    out = mb.mobility_product(mb.UF, w, r, v, eta, a, periodic_length=L, flags=mb.SYMMETRIC | mb.IN_PLANE)
    if wall == "single_wall":
        assert relerr(out, golden[f"{cloud}/{lname}/in_plane/UF"]) < TOL64


@pytest.mark.parametrize("n", [1, 31, 33, 255, 257, 1000, 2049])
@pytest.mark.parametrize("per", [(1, 0), (0, 1), (1, 1)])
def test_symmetric_periodic_ragged_vs_oracle(mb, n, per):
    """blobs outside the primary cell (the wrap truncates), overlaps across the cell boundary, blobs below z = a"""
    import oracle
    rng = np.random.default_rng(1000 * n + 2 * per[0] + per[1])
    a, eta = 0.21, 1.3
    side = (2.0 + 0.3 * n ** (1 / 3)) * a
    r = side * (1.6 * rng.random((n, 3)) - 0.3)
    r[:, 2] = side * rng.random(n)
    f = rng.standard_normal((n, 3))
    L = np.array([side * per[0], 1.1 * side * per[1], 0.0])
    for wall, w in (("single_wall", mb.SINGLE_WALL), ("no_wall", mb.NO_WALL)):
        ref = oracle.mobility_product("UF", wall, r, f, eta, a, periodic_length=L)
        out = mb.mobility_product(mb.UF, w, r, f, eta, a, periodic_length=L, flags=mb.SYMMETRIC)
        assert relerr(out, ref) < TOL64


def test_symmetric_periodic_large_default_and_multi_rank(mb):
    """n >= RMB_SYMMETRIC_MIN_N with xy images: the symmetric kernel is the default and agrees with the row kernel;
    the multi-rank form (world = 3 on one device, sums added by hand) gives the same product"""
    from rigidmultiblobswall_b200 import _lib
    lib = _lib.load()
    rng = np.random.default_rng(31)
    n, a, eta = 20000, 0.13, 7.0
    side = (n * 4 * np.pi / 3 / 0.1) ** (1 / 3) * a
    L = np.array([side, side, 0.0])
    Lc = (ctypes.c_double * 3)(*L)
    assert lib.rmb_launches_per_product(mb.UF, mb.SINGLE_WALL, mb.FP64, 0, n, n, Lc) == 3
    Lz = (ctypes.c_double * 3)(side, 0.0, side)
    # periodic z: row kernel whatever the flags (its launch count depends on the source split only)
    assert lib.rmb_launches_per_product(mb.UF, mb.SINGLE_WALL, mb.FP64, mb.SYMMETRIC, 10 ** 7, 10 ** 7, Lz) == 2
    assert lib.rmb_launches_per_product(mb.UF, mb.SINGLE_WALL, mb.FP64, mb.SYMMETRIC, 10 ** 7, 10 ** 7, Lc) == 3
    r = torch.from_numpy(side * rng.random((n, 3))).cuda()
    f = torch.from_numpy(rng.standard_normal((n, 3))).cuda()
    for w in (mb.SINGLE_WALL, mb.NO_WALL):
        det = mb.mobility_product(mb.UF, w, r, f, eta, a, periodic_length=L, flags=mb.DETERMINISTIC)
        sym = mb.mobility_product(mb.UF, w, r, f, eta, a, periodic_length=L)
        assert float(torch.linalg.norm(sym - det) / torch.linalg.norm(det)) < 1e-13
        idx, orc = oracle_rows("UF", "single_wall" if w == mb.SINGLE_WALL else "no_wall", r, f, eta, a, L, seed=4)
        assert relerr(rows3(det, idx), orc) < TOL64 and relerr(rows3(sym, idx), orc) < TOL64
        world = 3
        acc = sum(mb.symmetric_partial(r, f, eta, a, rank=k, world=world, wall=w, periodic_length=L).clone()
                  for k in range(world))
        out = torch.cat([mb.symmetric_finish(r, f, acc, eta, a, targets=t, wall=w)
                         for t in ((0, 9000), (9000, 11000))])
        assert float(torch.linalg.norm(out - det) / torch.linalg.norm(det)) < 1e-13
    import oracle
    ref = oracle.mobility_product("UF", "single_wall", r.cpu().numpy(), f.cpu().numpy(), eta, a, periodic_length=L,
                                  targets=(5000, 5032))
    sym = mb.mobility_product(mb.UF, mb.SINGLE_WALL, r, f, eta, a, periodic_length=L)
    assert relerr(sym[3 * 5000:3 * 5032].cpu().numpy(), ref) < TOL64


# ---- pair-symmetric kernel for the rotation blocks (UT / WF / WT, single wall) ----
@pytest.mark.parametrize("blk", ["UT", "WF", "WT"])
@pytest.mark.parametrize("lname", ["open", "xy", "x", "xyz"])
@pytest.mark.parametrize("cloud", ["dense", "dilute"])
def test_symmetric_rotation_blocks_golden(mb, golden, cloud, lname, blk):
    r, v = golden[f"{cloud}/r"], golden[f"{cloud}/v"]
    a, eta = golden[f"{cloud}/a_eta"]
    L = golden[f"{cloud}/L_{lname}"]
    out = mb.mobility_product(BLOCK_ID[blk], mb.SINGLE_WALL, r, v, eta, a, periodic_length=L, flags=mb.SYMMETRIC)
    assert relerr(out, golden[f"{cloud}/{lname}/single_wall/{blk}"]) < TOL64
    if blk == "UT" and f"{cloud}/{lname}/in_plane/UT" in golden:
        out = mb.mobility_product(BLOCK_ID[blk], mb.SINGLE_WALL, r, v, eta, a, periodic_length=L,
                                  flags=mb.SYMMETRIC | mb.IN_PLANE)
        assert relerr(out, golden[f"{cloud}/{lname}/in_plane/UT"]) < TOL64


@pytest.mark.parametrize("n", [1, 2, 33, 255, 257, 1000, 3001])
def test_symmetric_rotation_blocks_ragged_vs_oracle(mb, n):
    import oracle
    rng = np.random.default_rng(n + 17)
    a, eta = 0.21, 1.3
    side = (2.0 + 0.3 * n ** (1 / 3)) * a
    r = side * rng.random((n, 3))                                  # dense: overlaps and blobs below z = a
    v = rng.standard_normal((n, 3))
    for L in (None, np.array([side, 1.2 * side, 0.0])):
        if L is not None and n > 1000:
            continue
        for blk in ("UT", "WF", "WT"):
            ref = oracle.mobility_product(blk, "single_wall", r, v, eta, a, periodic_length=L)
            out = mb.mobility_product(BLOCK_ID[blk], mb.SINGLE_WALL, r, v, eta, a, periodic_length=L, flags=mb.SYMMETRIC)
            assert relerr(out, ref) < TOL64, (blk, L)


@pytest.mark.parametrize("blk", ["UT", "WF", "WT"])
def test_symmetric_rotation_blocks_large_default_and_multi_rank(mb, blk):
    """n >= RMB_SYMMETRIC_MIN_N: default = symmetric kernel, equal to the row kernel (RMB_DETERMINISTIC) and to the
    multi-rank form (world = 3 on one device, sums added by hand); the no-wall twins: tests/test_symmetric_free_gpu.py"""
    from rigidmultiblobswall_b200 import _lib
    lib = _lib.load()
    rng = np.random.default_rng(41)
    n, a, eta = 20000, 0.13, 7.0
    b = BLOCK_ID[blk]
    assert lib.rmb_launches_per_product(b, mb.SINGLE_WALL, mb.FP64, 0, 10 ** 7, 10 ** 7, None) == 3
    assert lib.rmb_launches_per_product(b, mb.NO_WALL, mb.FP64, mb.DETERMINISTIC, 10 ** 7, 10 ** 7, None) == 2
    side = (n * 4 * np.pi / 3 / 0.1) ** (1 / 3) * a
    r = torch.from_numpy(side * rng.random((n, 3))).cuda()
    v = torch.from_numpy(rng.standard_normal((n, 3))).cuda()
    for L in (None, np.array([side, side, 0.0])):
        det = mb.mobility_product(b, mb.SINGLE_WALL, r, v, eta, a, periodic_length=L, flags=mb.DETERMINISTIC)
        sym = mb.mobility_product(b, mb.SINGLE_WALL, r, v, eta, a, periodic_length=L)
        assert float(torch.linalg.norm(sym - det) / torch.linalg.norm(det)) < 1e-13
        idx, orc = oracle_rows(blk, "single_wall", r, v, eta, a, L, seed=5)
        assert relerr(rows3(det, idx), orc) < TOL64 and relerr(rows3(sym, idx), orc) < TOL64
        world = 3
        acc = sum(mb.symmetric_partial(r, v, eta, a, rank=k, world=world, block=b, periodic_length=L).clone()
                  for k in range(world))
        out = torch.cat([mb.symmetric_finish(r, v, acc, eta, a, targets=t, block=b)
                         for t in ((0, 9000), (9000, 11000))])
        assert float(torch.linalg.norm(out - det) / torch.linalg.norm(det)) < 1e-13
    nw = mb.mobility_product(b, mb.NO_WALL, r, v, eta, a, flags=mb.SYMMETRIC)
    nw_det = mb.mobility_product(b, mb.NO_WALL, r, v, eta, a, flags=mb.DETERMINISTIC)
    assert float(torch.linalg.norm(nw - nw_det) / torch.linalg.norm(nw_det)) < 1e-13
    idx, orc = oracle_rows(blk, "no_wall", r, v, eta, a, seed=6)
    assert relerr(rows3(nw, idx), orc) < TOL64 and relerr(rows3(nw_det, idx), orc) < TOL64


# ---- pair-symmetric kernel for the fused product u = UF f + UT t (single wall) ----
@pytest.mark.parametrize("lname", ["open", "xy", "x", "xyz"])
@pytest.mark.parametrize("cloud", ["dense", "dilute"])
def test_symmetric_fused_force_torque_golden(mb, golden, cloud, lname):
    r, f = golden[f"{cloud}/r"], golden[f"{cloud}/v"]
    a, eta = golden[f"{cloud}/a_eta"]
    L = golden[f"{cloud}/L_{lname}"]
    t = f[::-1].copy()
    import oracle
    ref = golden[f"{cloud}/{lname}/single_wall/UF"] + oracle.mobility_product("UT", "single_wall", r, t, eta, a, L)
    out = mb.mobility_product(mb.UF_UT, mb.SINGLE_WALL, r, f, eta, a, periodic_length=L, vec_b=t, flags=mb.SYMMETRIC)
    assert relerr(out, ref) < TOL64


@pytest.mark.parametrize("n", [1, 33, 257, 1000, 3001])
def test_symmetric_fused_ragged_vs_oracle(mb, n):
    import oracle
    rng = np.random.default_rng(n + 23)
    a, eta = 0.21, 1.3
    side = (2.0 + 0.3 * n ** (1 / 3)) * a
    r = side * rng.random((n, 3))
    f, t = rng.standard_normal((n, 3)), rng.standard_normal((n, 3))
    for L in (None, np.array([side, 0.0, 0.0])):
        ref = (oracle.mobility_product("UF", "single_wall", r, f, eta, a, periodic_length=L) +
               oracle.mobility_product("UT", "single_wall", r, t, eta, a, periodic_length=L))
        out = mb.mobility_product(mb.UF_UT, mb.SINGLE_WALL, r, f, eta, a, periodic_length=L, vec_b=t, flags=mb.SYMMETRIC)
        assert relerr(out, ref) < TOL64


def test_symmetric_fused_large_default_and_multi_rank(mb):
    rng = np.random.default_rng(43)
    n, a, eta = 20000, 0.13, 7.0
    side = (n * 4 * np.pi / 3 / 0.1) ** (1 / 3) * a
    r = torch.from_numpy(side * rng.random((n, 3))).cuda()
    f = torch.from_numpy(rng.standard_normal((n, 3))).cuda()
    t = torch.from_numpy(rng.standard_normal((n, 3))).cuda()
    for L in (None, np.array([side, side, 0.0])):
        det = mb.mobility_product(mb.UF_UT, mb.SINGLE_WALL, r, f, eta, a, periodic_length=L, vec_b=t, flags=mb.DETERMINISTIC)
        two = (mb.mobility_product(mb.UF, mb.SINGLE_WALL, r, f, eta, a, periodic_length=L, flags=mb.DETERMINISTIC) +
               mb.mobility_product(mb.UT, mb.SINGLE_WALL, r, t, eta, a, periodic_length=L, flags=mb.DETERMINISTIC))
        assert float(torch.linalg.norm(two - det) / torch.linalg.norm(det)) < 1e-13
        sym = mb.single_wall_mobility_trans_times_force_torque_b200(r, f, t, eta, a, periodic_length=L)   # default
        assert float(torch.linalg.norm(sym - det) / torch.linalg.norm(det)) < 1e-13
        idx, o1 = oracle_rows("UF", "single_wall", r, f, eta, a, L, seed=7)
        _, o2 = oracle_rows("UT", "single_wall", r, t, eta, a, L, seed=7)            # same seed: same rows
        assert relerr(rows3(det, idx), o1 + o2) < TOL64 and relerr(rows3(sym, idx), o1 + o2) < TOL64
        world = 3
        acc = sum(mb.symmetric_partial(r, f, eta, a, rank=k, world=world, block=mb.UF_UT, vec_b=t,
                                       periodic_length=L).clone() for k in range(world))
        out = torch.cat([mb.symmetric_finish(r, f, acc, eta, a, targets=tg, block=mb.UF_UT, vec_b=t)
                         for tg in ((0, 9000), (9000, 11000))])
        assert float(torch.linalg.norm(out - det) / torch.linalg.norm(det)) < 1e-13


@pytest.mark.parametrize("flags", [4, 8])
@pytest.mark.parametrize("blk", [0, 1, 2, 3, 4])
@pytest.mark.parametrize("periodic", [False, True])
def test_workspace_is_exactly_what_rmb_workspace_bytes_says(mb, blk, periodic, flags):
    """C ABI called with a workspace of exactly rmb_workspace_bytes() followed by a guard region: the product
    (pair-symmetric, flags = 4, and row kernel, flags = 8) must leave the guard untouched"""
    from rigidmultiblobswall_b200 import _lib
    lib = _lib.load()
    rng = np.random.default_rng(5)
    n, a, eta = 3000, 0.2, 1.0
    r = torch.from_numpy(8.0 * rng.random((n, 3))).cuda()
    va = torch.from_numpy(rng.standard_normal((n, 3))).cuda()
    vb = torch.from_numpy(rng.standard_normal((n, 3))).cuda()
    Lc = (ctypes.c_double * 3)(8.0, 8.0, 0.0) if periodic else None
    nbytes = lib.rmb_workspace_bytes(blk, mb.FP64, n, n, Lc)
    guard = 1 << 16
    ws = torch.full((nbytes + guard,), 0xAB, dtype=torch.uint8, device="cuda")
    out = torch.empty(3 * n, dtype=torch.float64, device="cuda")
    rc = lib.rmb_mobility_product(blk, mb.SINGLE_WALL, mb.FP64, flags, n, n, 0, r.data_ptr(), va.data_ptr(),
                                  vb.data_ptr(), eta, a, Lc, out.data_ptr(), ws.data_ptr(), nbytes,
                                  torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.rmb_last_error()
    torch.cuda.synchronize()
    assert bool((ws[nbytes:] == 0xAB).all())
    assert bool(torch.isfinite(out).all())
    # one byte less is refused
    rc = lib.rmb_mobility_product(blk, mb.SINGLE_WALL, mb.FP64, flags, n, n, 0, r.data_ptr(), va.data_ptr(),
                                  vb.data_ptr(), eta, a, Lc, out.data_ptr(), ws.data_ptr(), nbytes - 1,
                                  torch.cuda.current_stream().cuda_stream)
    assert rc != 0
