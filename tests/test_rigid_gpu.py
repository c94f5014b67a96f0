"""GPU parity tests of the rigid-body rows against golden outputs of the reference's
multi_bodies.py / body.py / stochastic_forcing.py (tests/golden/rigid_golden.npz)."""
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def relerr(x, ref):
    x = x.detach().cpu().numpy() if hasattr(x, "detach") else np.asarray(x)
    return np.linalg.norm(x.reshape(-1) - np.asarray(ref).reshape(-1)) / np.linalg.norm(ref)


@pytest.fixture(scope="module")
def rg():
    return np.load(os.path.join(HERE, "golden", "rigid_golden.npz"))


@pytest.fixture(scope="module")
def setup(rg):
    from rigidmultiblobswall_b200 import rigid
    bodies = rigid.RigidBodies(rg["struct/boomerang_N_15"], rg["rigid/location"], rg["rigid/quaternion"])
    return rigid, bodies


def dev(x):
    return torch.from_numpy(np.ascontiguousarray(x)).cuda()


def test_blob_positions(rg, setup):
    rigid, bodies = setup
    assert bodies.n_blobs == 180 and bodies.n_bodies == 12 and bodies.uniform
    assert relerr(bodies.get_r_vectors(), rg["rigid/r_vectors"]) < 1e-15


def test_K_and_KT(rg, setup):
    rigid, bodies = setup
    assert relerr(rigid.K_matrix_vector_prod(bodies, dev(rg["rigid/U"])), rg["rigid/K_U"]) < 1e-14
    kt = rigid.K_matrix_T_vector_prod(bodies, dev(rg["rigid/lambda"]))
    assert tuple(kt.shape) == (24, 3)
    assert relerr(kt, rg["rigid/KT_lambda"]) < 1e-14
    # fused update out = beta*out + alpha*K U
    out = dev(rg["rigid/lambda"]).reshape(-1).clone()
    rigid.K_matrix_vector_prod(bodies, dev(rg["rigid/U"]), out=out, alpha=-1.0, beta=1.0)
    assert relerr(out, rg["rigid/lambda"] - rg["rigid/K_U"]) < 1e-14


def test_linear_operator_rigid(rg, setup):
    rigid, bodies = setup
    a, eta, _ = rg["rigid/a_eta_tol"]
    r = bodies.get_r_vectors()
    Ax = rigid.linear_operator_rigid(dev(rg["rigid/x"]), bodies, r, eta, a, periodic_length=np.zeros(3))
    assert relerr(Ax, rg["rigid/A_x"]) < 1e-12


def test_body_mobility_blocks_and_preconditioner(rg, setup):
    rigid, bodies = setup
    a, eta, _ = rg["rigid/a_eta_tol"]
    PC = rigid.build_block_diagonal_preconditioner(bodies, bodies.get_r_vectors(), bodies.n_blobs, eta, a)
    assert relerr(PC.M_blobs[0], rg["rigid/M_body0"]) < 1e-12
    assert relerr(PC.M_blobs[5], rg["rigid/M_body5"]) < 1e-12      # body with a blob below z = a (clamp + damping)
    assert relerr(PC(dev(rg["rigid/x"])), rg["rigid/PC_x"]) < 1e-9   # two inversions of cond ~1e3 blocks


def test_mobility_problem_gmres(rg, setup):
    rigid, bodies = setup
    a, eta, tol = rg["rigid/a_eta_tol"]
    res = []
    vel, lam, its, info = rigid.solve_mobility_problem(bodies, rg["rigid/FT"], eta, a, tolerance=tol, callback=res.append)
    assert info == 0
    assert abs(its - int(rg["rigid/gmres_iterations"][0])) <= 1
    sol = torch.cat([lam.reshape(-1), vel.reshape(-1)])
    # both solves stop at relative residual 1e-8 of a unit-norm RHS: solutions agree to ~tol * cond
    assert relerr(sol, rg["rigid/solution"]) < 1e-6
    # true residual of OUR solution with the REFERENCE-checked operator
    r = bodies.get_r_vectors()
    rhs = torch.cat([torch.zeros(540, dtype=torch.float64, device="cuda"), -dev(rg["rigid/FT"])])
    Ax = rigid.linear_operator_rigid(sol, bodies, r, eta, a)
    assert float(torch.linalg.norm(Ax - rhs) / torch.linalg.norm(rhs)) < 5 * tol


def test_ragged_bodies_preconditioner_matches_uniform_ones(rg):
    from rigidmultiblobswall_b200 import rigid
    a, eta = 0.25, 1.0
    shapes = [rg["struct/boomerang_N_15"], rg["struct/shell_N_12"], rg["struct/boomerang_N_15"]]
    loc = np.array([[0.0, 0, 3.0], [6.0, 1.0, 2.5], [-5.0, 4.0, 4.0]])
    q = np.array([[1.0, 0, 0, 0], [0.5, 0.5, 0.5, 0.5], [0.0, 1.0, 0, 0]])
    bodies = rigid.RigidBodies(shapes, loc, q)
    assert not bodies.uniform and bodies.nb_max == 15 and bodies.n_blobs == 42
    PC = rigid.BlockDiagonalPreconditioner(bodies, bodies.get_r_vectors(), eta, a)
    rng = np.random.default_rng(5)
    x = rng.standard_normal(3 * 42 + 18)
    y = PC(dev(x)).cpu().numpy()
    off = bodies.body_offset_host
    for k in range(3):
        one = rigid.RigidBodies([shapes[k]], loc[k:k + 1], q[k:k + 1])
        PC1 = rigid.BlockDiagonalPreconditioner(one, one.get_r_vectors(), eta, a)
        xk = np.concatenate([x[3 * off[k]:3 * off[k + 1]], x[126 + 6 * k:126 + 6 * k + 6]])
        yk = PC1(dev(xk)).cpu().numpy()
        got = np.concatenate([y[3 * off[k]:3 * off[k + 1]], y[126 + 6 * k:126 + 6 * k + 6]])
        assert np.linalg.norm(got - yk) / np.linalg.norm(yk) < 1e-10
        # for a single body the saddle-point operator IS the block: A(PC(x)) = x
        Ay = rigid.linear_operator_rigid(dev(yk), one, one.get_r_vectors(), eta, a).cpu().numpy()
        assert np.linalg.norm(Ay - xk) / np.linalg.norm(xk) < 1e-9


def test_lanczos_on_device(rg):
    from rigidmultiblobswall_b200 import mobility as mb, solvers
    a, eta, _ = rg["rigid/a_eta_tol"]
    rsub = dev(rg["rigid/r_vectors"][:60]).reshape(-1)
    tol, factor = rg["lanczos/tolerance_factor"]
    mult = lambda v: mb.single_wall_mobility_trans_times_force_b200(rsub, v, eta, a)
    noise, it = solvers.lanczos(factor=factor, tolerance=tol, dim=180, mobility_mult=mult, z=dev(rg["lanczos/z"]))
    assert abs(it - int(rg["lanczos/iterations"][0])) <= 1
    assert relerr(noise, rg["lanczos/noise"]) < 1e-8


def test_stochastic_preconditioner_and_preconditioned_lanczos(rg, setup):
    """multi_bodies.py:480-618 + quaternion_integrator_multi_bodies.py:703: P^T M P, P^-1 and the noise"""
    from rigidmultiblobswall_b200 import solvers
    rigid, bodies = setup
    a, eta, _ = rg["rigid/a_eta_tol"]
    r = bodies.get_r_vectors()
    PC, mobility_pc, P_inv_mult = rigid.build_block_diagonal_preconditioners_det_stoch(
        bodies, r, bodies.n_blobs, eta, a, periodic_length=np.zeros(3))
    w = dev(rg["stoch/w"])
    assert relerr(mobility_pc(w), rg["stoch/mobility_pc_w"]) < 1e-10
    assert relerr(P_inv_mult(w), rg["stoch/P_inv_w"]) < 1e-12
    assert relerr(PC(dev(rg["rigid/x"])), rg["stoch/PC_x"]) < 1e-9
    tol, factor = rg["stoch/tolerance_factor"]
    noise, it = solvers.lanczos(factor=factor, tolerance=tol, dim=540, mobility_mult=mobility_pc, L_mult=P_inv_mult,
                                z=dev(rg["stoch/z"]))
    assert abs(it - int(rg["stoch/iterations"][0])) <= 1
    assert relerr(noise, rg["stoch/noise"]) < 1e-8


def test_sharded_rigid_system_single_rank_matches_direct_solve(rg):
    """parallel.ShardedRigidSystem on one rank == rigid.solve_mobility_problem / the preconditioned Lanczos
    (the world-2 plumbing is covered on CPU by tests/test_parallel_gloo.py)."""
    from rigidmultiblobswall_b200 import parallel, rigid
    a, eta, tol = rg["rigid/a_eta_tol"]
    system = parallel.ShardedRigidSystem(rg["struct/boomerang_N_15"], rg["rigid/location"], rg["rigid/quaternion"],
                                         eta, a, group=False)
    assert system.ranges == [(0, 180)] and (system.b0, system.b1) == (0, 12)
    U, lam, its, info = system.solve_mobility(system.local_slice(rg["rigid/FT"]), tolerance=tol)
    bodies = rigid.RigidBodies(rg["struct/boomerang_N_15"], rg["rigid/location"], rg["rigid/quaternion"])
    U0, lam0, its0, info0 = rigid.solve_mobility_problem(bodies, rg["rigid/FT"], eta, a, tolerance=tol)
    assert info == 0 and its == its0
    assert relerr(U, U0.cpu().numpy()) < 1e-12 and relerr(lam, lam0.cpu().numpy()) < 1e-12
    z = torch.from_numpy(np.random.default_rng(4).standard_normal(540)).cuda()
    noise, it_l = system.stochastic_noise(z, factor=0.7, tolerance=1e-8)
    assert noise.shape == (540,) and 1 <= it_l < 40
    # same numbers as the golden-checked single-process path (rigid.build_block_diagonal_preconditioners_det_stoch)
    from rigidmultiblobswall_b200 import solvers
    r = bodies.get_r_vectors()
    _, mobility_pc, P_inv_mult = rigid.build_block_diagonal_preconditioners_det_stoch(bodies, r, 180, eta, a)
    ref, it_ref = solvers.lanczos(factor=0.7, tolerance=1e-8, mobility_mult=mobility_pc, L_mult=P_inv_mult, z=z)
    assert it_l == it_ref and relerr(noise, ref.cpu().numpy()) < 1e-12
