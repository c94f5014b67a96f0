"""GPU parity tests of the `resistance` and `body_mobility` schemes (multi_bodies_utilities.py:549-600) against
golden outputs of the reference's own dense computation (tests/golden/resistance_golden.npz, generator
tests/golden/make_golden_resistance.py).  The product path is matrix-free (pair kernel inside GMRES); the reference
inverts the dense blob mobility, so the two answers are independent."""
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
def gold():
    return (np.load(os.path.join(HERE, "golden", "rigid_golden.npz")),
            np.load(os.path.join(HERE, "golden", "resistance_golden.npz")))


@pytest.fixture(scope="module")
def setup(gold):
    from rigidmultiblobswall_b200 import rigid
    rg, _ = gold
    bodies = rigid.RigidBodies(rg["struct/boomerang_N_15"], rg["rigid/location"], rg["rigid/quaternion"])
    a, eta, _ = rg["rigid/a_eta_tol"]
    PC = rigid.BlockDiagonalPreconditioner(bodies, bodies.get_r_vectors(), eta, a)
    return rigid, bodies, PC, float(eta), float(a)


@pytest.mark.parametrize("key", ["noslip", "slip"])
def test_resistance_problem(gold, setup, key):
    _, g = gold
    rigid, bodies, PC, eta, a = setup
    slip = g["slip"] if key == "slip" else None
    force, lam, its, info = rigid.solve_resistance_problem(bodies, g["velocity"], eta, a, tolerance=1e-10, slip=slip,
                                                           PC=PC)
    assert info == 0 and 0 < its <= 60
    assert tuple(force.shape) == (12, 6) and tuple(lam.shape) == (180, 3)
    # GMRES stops at relative residual 1e-10, cond(M) = 73 (stored in the golden file)
    assert relerr(lam, g[f"resistance/{key}/force_blobs"]) < 1e-8
    assert relerr(force, g[f"resistance/{key}/force"]) < 1e-8


def test_resistance_problem_builds_its_own_preconditioner(gold, setup):
    _, g = gold
    rigid, bodies, _, eta, a = setup
    force, _, _, info = rigid.solve_resistance_problem(bodies, torch.from_numpy(g["velocity"]).cuda(), eta, a,
                                                       tolerance=1e-9)
    assert info == 0
    assert relerr(force, g["resistance/noslip/force"]) < 1e-7


def test_blob_preconditioner_is_the_block_inverse(setup):
    rigid, bodies, PC, eta, a = setup
    w = torch.randn(3 * bodies.n_blobs, dtype=torch.float64, device="cuda",
                    generator=torch.Generator(device="cuda").manual_seed(5))
    y = PC.M_inv_mult(w).reshape(12, 45, 1)
    assert relerr((PC.M_blobs @ y).reshape(-1), w.cpu().numpy()) < 1e-11


def test_body_mobility(gold, setup):
    _, g = gold
    rigid, bodies, PC, eta, a = setup
    N = rigid.body_mobility(bodies, eta, a, tolerance=1e-10, PC=PC)
    assert tuple(N.shape) == (72, 72)
    assert relerr(N, g["body_mobility/N"]) < 1e-8
    assert relerr(N, N.T.cpu().numpy()) < 1e-8


def test_mobility_problem_first_guess(gold, setup):
    """x0 = a previous solution (what the reference's integrators pass): fewer iterations, same answer"""
    rg, _ = gold
    rigid, bodies, PC, eta, a = setup
    vel, lam, its, info = rigid.solve_mobility_problem(bodies, rg["rigid/FT"], eta, a, tolerance=1e-8, PC=PC)
    guess = torch.cat([lam.reshape(-1), vel.reshape(-1)]) * (1 + 1e-3)
    vel2, lam2, its2, info2 = rigid.solve_mobility_problem(bodies, rg["rigid/FT"], eta, a, tolerance=1e-8, PC=PC, x0=guess)
    assert info == 0 and info2 == 0 and its2 < its
    assert relerr(torch.cat([lam2.reshape(-1), vel2.reshape(-1)]), rg["rigid/solution"]) < 1e-6


def test_prescribed_kinematics(gold):
    """bodies 2 and 7 move as prescribed (mixed mobility / resistance problem): operator, preconditioner and the
    `mobility` scheme against the reference (tests/golden/prescribed_golden.npz, make_golden_prescribed.py)"""
    from rigidmultiblobswall_b200 import rigid
    rg, _ = gold
    g = np.load(os.path.join(HERE, "golden", "prescribed_golden.npz"))
    a, eta, _ = rg["rigid/a_eta_tol"]
    presc = [g["prescribed_velocity"][k] if g["prescribed"][k] else None for k in range(12)]
    bodies = rigid.RigidBodies(rg["struct/boomerang_N_15"], rg["rigid/location"], rg["rigid/quaternion"],
                               prescribed_velocity=presc)
    assert bodies.any_prescribed
    r = bodies.get_r_vectors()
    x = torch.from_numpy(g["x"]).cuda()
    assert relerr(rigid.linear_operator_rigid(x, bodies, r, eta, a), g["A_x"]) < 1e-12
    PC = rigid.BlockDiagonalPreconditioner(bodies, r, eta, a)
    assert relerr(PC(x), g["PC_x"]) < 1e-9
    vel, lam, its, info = rigid.solve_mobility_problem(bodies, g["FT"], eta, a, tolerance=1e-10, slip=g["slip"], PC=PC)
    assert info == 0 and abs(its - int(g["gmres_iterations"][0])) <= 1
    assert relerr(vel, g["velocity"]) < 1e-8 and relerr(lam, g["lambda"]) < 1e-8
    assert relerr(rigid.K_matrix_T_vector_prod(bodies, lam, r_vectors=r).reshape(-1, 6), g["force"]) < 1e-8
