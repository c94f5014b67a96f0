"""GPU parity tests of the source-target UF product with per-blob radii
(rmb_mobility_source_target, through the C ABI) against
 (1) golden outputs of the reference's numba kernels (tests/golden/st_golden.npz),
 (2) the CPU oracle on seeded ragged sizes (tile / CTA boundaries, the source-split path),
 (3) the equal-radius kernel at a size no CPU check is needed for.
Tolerances: 1e-12 (fp64) / 1e-5 (fp32) norm-wise, BASELINE.json."""
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
TOL64, TOL32 = 1e-12, 1e-5


def relerr(x, ref):
    return np.linalg.norm(np.asarray(x) - ref) / np.linalg.norm(ref)


@pytest.fixture(scope="module")
def mb():
    from rigidmultiblobswall_b200 import mobility
    assert torch.cuda.is_available()
    return mobility


@pytest.fixture(scope="module")
def g():
    return np.load(os.path.join(HERE, "golden", "st_golden.npz"))


def _cases():
    return [str(c) for c in np.load(os.path.join(HERE, "golden", "st_golden.npz"))["cases"]]


def _inputs(g, case):
    name, lname, wall = case.split("/")
    src, trg, f = g[f"{name}/source"], g[f"{name}/target"], g[f"{name}/force"]
    rs, rt = g[f"{name}/radius_source"], g[f"{name}/radius_target"]
    if wall == "radii_single_wall":
        trg, rt, wall = src, rs, "single_wall"
    return wall, src, trg, f, rs, rt, float(g[f"{name}/eta"]), g[f"{name}/L_{lname}"]


@pytest.mark.parametrize("flags", [0, 2])   # 2 = RMB_NO_TMA (plain tile loader)
@pytest.mark.parametrize("case", _cases())
def test_golden_fp64(mb, g, case, flags):
    wall, src, trg, f, rs, rt, eta, L = _inputs(g, case)
    fn = getattr(mb, wall + "_mobility_trans_times_force_source_target_b200")
    if flags:
        w = {"no_wall": mb.NO_WALL, "single_wall": mb.SINGLE_WALL, "free_surface": mb.FREE_SURFACE}[wall]
        out = mb.mobility_source_target(w, src, trg, f, rs, rt, eta, periodic_length=L, flags=flags)
    else:
        out = fn(src, trg, f, rs, rt, eta, periodic_length=L)
    assert out.shape == g[case].shape
    assert relerr(out, g[case]) < TOL64


@pytest.mark.parametrize("case", [c for c in _cases() if c.startswith("dilute/")])
def test_golden_fp32(mb, g, case):
    wall, src, trg, f, rs, rt, eta, L = _inputs(g, case)
    out = getattr(mb, wall + "_mobility_trans_times_force_source_target_b200_fp32")(src, trg, f, rs, rt, eta,
                                                                                    periodic_length=L)
    assert relerr(out, g[case]) < TOL32


def test_mobility_radii_wrapper(mb, g):
    """mobility.py:1369 call shape: function passed in, sources == targets, self pairs inside."""
    case = "mixed/xy/radii_single_wall"
    _, src, _, f, rs, _, eta, L = _inputs(g, case)
    out = mb.mobility_radii_trans_times_force(src, f, eta, 0.0, rs,
                                              mb.single_wall_mobility_trans_times_force_source_target_b200,
                                              periodic_length=L)
    assert relerr(out, g[case]) < TOL64


def _cloud(rng, n, side, rlo, rhi):
    r = side * rng.random((n, 3))
    return r, rlo + (rhi - rlo) * rng.random(n)


@pytest.mark.parametrize("nt,ns", [(1, 1), (1, 700), (127, 255), (128, 256), (129, 257), (300, 1025), (2500, 40),
                                   (50, 6000)])
@pytest.mark.parametrize("wall", ["single_wall", "no_wall", "free_surface"])
def test_ragged_sizes_vs_oracle(mb, wall, nt, ns):
    """sizes around the 128-target CTA and 256-source tile; (50, 6000) and (1, 700) take the source-split path"""
    import oracle
    rng = np.random.default_rng(nt * 7919 + ns)
    side = 0.25 * max(nt, ns) ** (1.0 / 3.0) + 0.5
    src, rs = _cloud(rng, ns, side, 0.05, 0.25)
    trg, rt = _cloud(rng, nt, side, 0.05, 0.25)
    f = rng.standard_normal((ns, 3))
    L = np.array([1.4 * side, 0.0, 0.0]) if (nt + ns) % 2 else np.zeros(3)
    ref = oracle.source_target(wall, src, trg, f, rs, rt, 1.7, L)
    out = getattr(mb, wall + "_mobility_trans_times_force_source_target_b200")(src, trg, f, rs, rt, 1.7,
                                                                               periodic_length=L)
    assert relerr(out, ref) < TOL64


def test_empty_sets(mb):
    z = np.zeros((0, 3))
    trg = np.random.default_rng(1).random((5, 3))
    out = mb.no_wall_mobility_trans_times_force_source_target_b200(z, trg, z, np.zeros(0), np.full(5, 0.1), 1.0)
    assert out.shape == (15,) and np.all(out == 0)
    out = mb.single_wall_mobility_trans_times_force_source_target_b200(trg, z, trg, np.full(5, 0.1), np.zeros(0), 1.0)
    assert out.shape == (0,)


def test_device_tensors_stay_on_device_and_inputs_unchanged(mb, g):
    wall, src, trg, f, rs, rt, eta, L = _inputs(g, "mixed/open/single_wall")
    dev = torch.device("cuda")
    args = [torch.from_numpy(np.ascontiguousarray(x)).to(dev) for x in (src, trg, f, rs, rt)]
    keep = [a.clone() for a in args]
    out = mb.single_wall_mobility_trans_times_force_source_target_b200(*args, eta)
    assert isinstance(out, torch.Tensor) and out.is_cuda
    assert relerr(out.cpu().numpy(), g["mixed/open/single_wall"]) < TOL64
    for a, k in zip(args, keep):
        assert torch.equal(a, k)


@pytest.mark.parametrize("wall", ["single_wall", "no_wall", "free_surface"])
def test_equal_radii_matches_same_radius_kernel_large(mb, wall):
    """N = 20,000, all radii equal, z >= a: the source-target product must agree with the same-radius kernel
    (the survey probe found 5e-16 between the two numba kernels)."""
    rng = np.random.default_rng(5)
    n, a, eta = 20000, 0.13, 7.0
    side = (n * 4 * np.pi / 3 / 0.1) ** (1.0 / 3.0) * a
    r = side * rng.random((n, 3))
    r[:, 2] += a
    f = rng.standard_normal((n, 3))
    rad = np.full(n, a)
    w = {"no_wall": mb.NO_WALL, "single_wall": mb.SINGLE_WALL, "free_surface": mb.FREE_SURFACE}[wall]
    ref = mb.mobility_product(mb.UF, w, r, f, eta, a)
    out = mb.mobility_source_target(w, r, r, f, rad, rad, eta)
    assert relerr(out, ref) < TOL64


def test_linearity_and_target_subset_large(mb):
    rng = np.random.default_rng(9)
    ns, nt = 30000, 4000
    src, rs = _cloud(rng, ns, 12.0, 0.05, 0.3)
    trg, rt = _cloud(rng, nt, 12.0, 0.05, 0.3)
    f1, f2 = rng.standard_normal((ns, 3)), rng.standard_normal((ns, 3))
    fn = mb.single_wall_mobility_trans_times_force_source_target_b200
    u1, u2 = fn(src, trg, f1, rs, rt, 1.0), fn(src, trg, f2, rs, rt, 1.0)
    u12 = fn(src, trg, 2.0 * f1 - 3.0 * f2, rs, rt, 1.0)
    assert relerr(u12, 2.0 * u1 - 3.0 * u2) < 1e-11
    part = fn(src, trg[1000:1777], f1, rs, rt[1000:1777], 1.0)
    assert relerr(part, u1[3000:3 * 1777]) < 1e-13
