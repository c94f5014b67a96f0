"""Parity at the sizes of BASELINE.json's configs (C2 .. C5), through the public API / C ABI, against
  * golden answers the REFERENCE produced at that size (tests/golden/configs_golden.npz, generator
    tests/golden/make_golden_configs.py: multi_bodies.py + SciPy GMRES, stochastic_forcing_lanczos), and
  * oracle rows for RANDOM targets (not a contiguous range): the product is equivariant under a permutation of the
    blobs, so the oracle is run on a copy of the cloud with the sampled blobs moved to the front.
Tolerances: 1e-12 norm-wise for a product (north_star); for a Krylov solve both sides stop at a relative residual of
`tol`, so solutions agree to ~ tol * cond (1e-6 asserted for tol = 1e-8) and iteration counts to +-1."""
import os

import numpy as np
import pytest

import bench

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def relerr(x, ref):
    x = x.detach().cpu().numpy() if hasattr(x, "detach") else np.asarray(x)
    return float(np.linalg.norm(x.reshape(-1) - np.asarray(ref).reshape(-1)) / np.linalg.norm(ref))


@pytest.fixture(scope="module")
def cg():
    return np.load(os.path.join(HERE, "golden", "configs_golden.npz"))


def sampled_rows(n, m, seed):
    """m random target indices and the permutation that moves them to the front"""
    rng = np.random.default_rng(seed)
    idx = np.sort(rng.choice(n, m, replace=False))
    rest = np.setdiff1d(np.arange(n), idx, assume_unique=True)
    return idx, np.concatenate([idx, rest])


def rows3(u, idx):
    u = u.detach().cpu().numpy() if hasattr(u, "detach") else np.asarray(u)
    return u.reshape(-1, 3)[idx].reshape(-1)


# ---------------------------------------------------------------------------------------------------- C2 ----
def test_c2_mobility_solve_1000_boomerangs_against_the_reference(cg):
    """multi_bodies_utilities.py:414-540 (GMRES at :516) at BASELINE size: 1,000 boomerang_N_15 bodies."""
    from rigidmultiblobswall_b200 import rigid
    ref_cfg, loc, quat, FT, a, eta, tol = bench.make_c2(1000)
    bodies = rigid.RigidBodies(ref_cfg, loc, quat)
    assert bodies.n_blobs == 15000
    vel, lam, its, info = rigid.solve_mobility_problem(bodies, FT, eta, a, tolerance=tol)
    assert info == 0
    assert abs(its - int(cg["c2/iterations"][0])) <= 1
    assert relerr(vel, cg["c2/velocity"]) < 1e-6
    lam_s = lam.reshape(-1)[torch.from_numpy(cg["c2/lambda_index"]).cuda()]
    assert relerr(lam_s, cg["c2/lambda_sample"]) < 1e-5
    assert abs(float(torch.linalg.norm(lam)) / float(cg["c2/lambda_norm"][0]) - 1) < 1e-6


# ---------------------------------------------------------------------------------------------------- C3 ----
def test_c3_rollers_100k_all_blocks_random_rows_against_the_oracle():
    """The block table of quaternion_integrator_rollers.py:59-91 on the inputs of
    examples/Mobility_Prod_Timing/Multiblob.inputfile:17-30 scaled to 100k rollers: UF, UT, WF, WT and the fused
    UF f + UT t, pseudo-periodic in xy (9 image boxes), 256 random target rows each."""
    import oracle
    from rigidmultiblobswall_b200 import mobility as mb
    n = 100000
    r, f, t, a, eta, L = bench.make_c3(n)
    idx, perm = sampled_rows(n, 256, seed=31)
    rp, fp, tp = r[perm], f[perm], t[perm]
    rd, fd, td = (torch.from_numpy(x).cuda().reshape(-1) for x in (r, f, t))
    refs = {}
    for name, blk, vec, vp in (("UF", mb.UF, fd, fp), ("UT", mb.UT, td, tp), ("WF", mb.WF, fd, fp), ("WT", mb.WT, td, tp)):
        refs[name] = oracle.mobility_product(name, "single_wall", rp, vp, eta, a, L, targets=(0, 256))
        for flags in (0, mb.DETERMINISTIC):                    # default (pair-symmetric) kernel, row kernel
            u = mb.mobility_product(blk, mb.SINGLE_WALL, rd, vec, eta, a, periodic_length=L, flags=flags)
            assert relerr(rows3(u, idx), refs[name]) < 1e-12, (name, flags)
    u = mb.single_wall_mobility_trans_times_force_torque_b200(rd, fd, td, eta, a, periodic_length=L)
    assert relerr(rows3(u, idx), refs["UF"] + refs["UT"]) < 1e-12
    # the NumPy drop-in boundary (what QuaternionIntegratorRollers calls) returns the same rows
    u_np = mb.single_wall_mobility_rot_times_torque_b200(r, t, eta, a, periodic_length=L)
    assert isinstance(u_np, np.ndarray) and relerr(rows3(u_np, idx), refs["WT"]) < 1e-12


# ---------------------------------------------------------------------------------------------------- C4 ----
def test_c4_one_million_blobs_1000_random_targets_against_the_oracle():
    """N = 1M single-wall UF, 1,000 random targets (10^9 oracle pairs)."""
    import oracle
    from rigidmultiblobswall_b200 import mobility as mb
    n = 1000000
    r, f, a, eta = bench.make_cloud(n)
    idx, perm = sampled_rows(n, 1000, seed=41)
    ref = oracle.uf_wall_fast(r[perm], f[perm], eta, a, targets=(0, 1000))
    rd, fd = torch.from_numpy(r).cuda().reshape(-1), torch.from_numpy(f).cuda().reshape(-1)
    u = mb.single_wall_mobility_trans_times_force_b200(rd, fd, eta, a)
    assert relerr(rows3(u, idx), ref) < 1e-12
    # the fp32 variant at this size: fp32 pair arithmetic, fp64 running sum over the source tiles (north_star: 1e-5;
    # a plain fp32 sum over 10^6 sources measures 2e-5, profiles/r02j_fp32_error_vs_n.json)
    u32 = mb.single_wall_mobility_trans_times_force_b200_fp32(rd, fd, eta, a)
    assert relerr(rows3(u32, idx), ref) < 1e-5
    # size-independent properties at full size: the symmetry of M (f2 . M f1 = f1 . M f2) and linearity
    f2 = torch.from_numpy(np.random.default_rng(5).standard_normal(3 * n)).cuda()
    u2 = mb.single_wall_mobility_trans_times_force_b200(rd, f2, eta, a)
    s12, s21 = float(torch.dot(f2, u)), float(torch.dot(fd, u2))
    assert abs(s12 - s21) < 1e-11 * (float(torch.linalg.norm(f2)) * float(torch.linalg.norm(u)))
    u3 = mb.single_wall_mobility_trans_times_force_b200(rd, 0.5 * fd - 2.0 * f2, eta, a)
    assert relerr(u3, (0.5 * u - 2.0 * u2).cpu().numpy()) < 1e-12


# ---------------------------------------------------------------------------------------------------- C5 ----
def _brownian_step(n_shells, cg, noise_tol, vel_tol):
    """quaternion_integrator_multi_bodies.py:626-735 (stochastic slip + mobility solve) with
    stochastic_forcing.py:112-264 (preconditioned Lanczos), single rank."""
    from rigidmultiblobswall_b200 import parallel
    key = "c5_%d" % n_shells
    shell, a, eta, kT, dt, loc, quat, FT, z = bench.make_c5(n_shells)
    ltol, gtol = cg[key + "/tolerances_lanczos_gmres"]
    system = parallel.ShardedRigidSystem(shell, loc, quat, eta, a, group=False)
    assert system.n_blobs_total == 162 * n_shells
    if system.n_blobs_total >= 8000:
        assert system.mobility.symmetric                      # the flagship (pair-symmetric) kernel is what runs
    z_d = torch.from_numpy(z).cuda()
    noise, it_l = system.stochastic_noise(z_d, factor=np.sqrt(2 * kT / dt), tolerance=ltol)
    assert abs(it_l - int(cg[key + "/lanczos_iterations"][0])) <= 1
    nidx = torch.from_numpy(cg[key + "/noise_index"]).cuda()
    assert relerr(noise[nidx], cg[key + "/noise_sample"]) < noise_tol
    assert abs(float(torch.linalg.norm(noise)) / float(cg[key + "/noise_norm"][0]) - 1) < noise_tol
    U, lam, its, info = system.solve_mobility(system.local_slice(FT), tolerance=gtol, slip_loc=-noise)
    assert info == 0
    assert abs(its - int(cg[key + "/gmres_iterations"][0])) <= 1
    assert relerr(U, cg[key + "/velocity"]) < vel_tol
    assert relerr(lam.reshape(-1)[nidx], cg[key + "/lambda_sample"]) < 10 * vel_tol


def test_c5_brownian_step_128_shells_against_the_reference(cg):
    _brownian_step(128, cg, noise_tol=1e-6, vel_tol=1e-6)


def test_c5_brownian_step_1000_shells_against_the_reference(cg):
    """162,000 blobs: the largest C5 reduction the reference finishes in the build container (about an hour)."""
    if "c5_1000/velocity" not in cg:
        pytest.skip("tests/golden/configs_golden.npz holds no 1,000-shell answer")
    _brownian_step(1000, cg, noise_tol=1e-6, vel_tol=1e-6)
