"""Host-logic tests of the Krylov solvers (pure torch: they run on CPU tensors
too).  GMRES against scipy.sparse.linalg.gmres driven the way the reference's
general_application_utils.gmres does; Lanczos against the reference's golden
output (tests/golden/rigid_golden.npz) with the CPU oracle as the operator."""
import os

import numpy as np
import pytest
import scipy.sparse.linalg as spla
import torch

from rigidmultiblobswall_b200 import solvers

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def rg():
    return np.load(os.path.join(HERE, "golden", "rigid_golden.npz"))


def test_gmres_matches_scipy_right_preconditioned():
    rng = np.random.default_rng(0)
    n = 300
    A = rng.standard_normal((n, n)) / np.sqrt(n) + 3.0 * np.eye(n)
    P = np.diag(1.0 / (np.diag(A) * (1 + 0.1 * rng.random(n))))
    b = rng.standard_normal(n)
    its = [0]
    ylo = spla.LinearOperator((n, n), matvec=lambda y: A @ (P @ y), dtype="float64")
    y, info = spla.gmres(ylo, b, rtol=1e-10, atol=0, restart=20, maxiter=100,
                         callback=lambda r: its.__setitem__(0, its[0] + 1), callback_type="pr_norm")
    x_ref = P @ y
    At, Pt, bt = torch.from_numpy(A), torch.from_numpy(P), torch.from_numpy(b)
    res = []
    x, info2, it2 = solvers.gmres(lambda v: At @ v, bt, tol=1e-10, restart=20, maxiter=100, M=lambda v: Pt @ v,
                                  callback=res.append)
    assert info == 0 and info2 == 0
    assert abs(it2 - its[0]) <= 1 and len(res) == it2
    assert np.linalg.norm(x.numpy() - x_ref) / np.linalg.norm(x_ref) < 1e-8
    assert np.linalg.norm(A @ x.numpy() - b) / np.linalg.norm(b) < 1e-9


def test_gmres_zero_rhs_and_no_preconditioner():
    A = torch.eye(5, dtype=torch.float64) * 2
    x, info, it = solvers.gmres(lambda v: A @ v, torch.zeros(5, dtype=torch.float64))
    assert info == 0 and it == 0 and float(x.abs().sum()) == 0
    b = torch.arange(5, dtype=torch.float64)
    x, info, it = solvers.gmres(lambda v: A @ v, b, tol=1e-12)
    assert info == 0 and torch.allclose(x, b / 2)


def test_gmres_initial_guess():
    rng = np.random.default_rng(3)
    n = 200
    A = torch.from_numpy(rng.standard_normal((n, n)) / np.sqrt(n) + 3.0 * np.eye(n))
    P = torch.diag(1.0 / torch.diag(A))
    b = torch.from_numpy(rng.standard_normal(n))
    x, info, its = solvers.gmres(lambda v: A @ v, b, tol=1e-10, restart=30, M=lambda v: P @ v)
    assert info == 0
    # exact solution as the guess: nothing to do; a perturbed one: fewer iterations, same answer, same target
    xe, info_e, its_e = solvers.gmres(lambda v: A @ v, b, x0=x, tol=1e-9, restart=30, M=lambda v: P @ v)
    assert info_e == 0 and its_e == 0 and torch.equal(xe, x)
    res = []
    x0 = x + 1e-4 * torch.from_numpy(rng.standard_normal(n))
    x2, info2, its2 = solvers.gmres(lambda v: A @ v, b, x0=x0, tol=1e-10, restart=30, M=lambda v: P @ v, callback=res.append)
    assert info2 == 0 and 0 < its2 < its and len(res) == its2
    assert float(torch.linalg.norm(A @ x2 - b) / torch.linalg.norm(b)) < 2e-10
    assert res[-1] <= 1e-10 and res[0] < 1e-3        # callbacks report residuals relative to ||b||
    assert float(torch.linalg.norm(x2 - x) / torch.linalg.norm(x)) < 1e-9


def test_gmres_reports_non_convergence():
    rng = np.random.default_rng(1)
    A = torch.from_numpy(rng.standard_normal((50, 50)))
    b = torch.from_numpy(rng.standard_normal(50))
    x, info, it = solvers.gmres(lambda v: A @ v, b, tol=1e-14, restart=3, maxiter=2)
    assert info == it == 6


def test_lanczos_matches_reference_golden(rg):
    import oracle
    a, eta, _ = rg["rigid/a_eta_tol"]
    rsub = rg["rigid/r_vectors"][:60]
    z = rg["lanczos/z"]
    tol, factor = rg["lanczos/tolerance_factor"]
    mult = lambda v: torch.from_numpy(oracle.mobility_product("UF", "single_wall", rsub, v.numpy(), eta, a))
    noise, it = solvers.lanczos(factor=factor, tolerance=tol, dim=180, mobility_mult=mult, z=torch.from_numpy(z))
    ref = rg["lanczos/noise"]
    assert abs(it - int(rg["lanczos/iterations"][0])) <= 1
    assert np.linalg.norm(noise.numpy() - ref) / np.linalg.norm(ref) < 1e-8


def test_lanczos_exact_on_diagonal_and_with_L_mult():
    d = torch.linspace(1.0, 10.0, 10, dtype=torch.float64)
    z = torch.ones(10, dtype=torch.float64)
    noise, it = solvers.lanczos(factor=2.0, tolerance=1e-12, mobility_mult=lambda v: d * v, z=z)
    assert torch.allclose(noise, 2.0 * torch.sqrt(d), atol=1e-10)
    noise2, _ = solvers.lanczos(factor=2.0, tolerance=1e-12, mobility_mult=lambda v: d * v, z=z, L_mult=lambda v: 3 * v)
    assert torch.allclose(noise2, 3 * noise, atol=1e-10)
    zero, it0 = solvers.lanczos(factor=0.0, dim=10, z=z)
    assert it0 == 0 and float(zero.abs().sum()) == 0


def test_dense_noise_functions_follow_the_reference():
    """stochastic_forcing_{eig, eig_symm, cholesky} (stochastic_forcing.py:7-110) on the reference's own dense
    single-wall mobility (tests/golden/dense_golden.npz); compared with the reference functions themselves when the
    checkout is present, and always through their defining property  noise = factor * S z  with  S S^T = M."""
    g = np.load(os.path.join(HERE, "golden", "dense_golden.npz"))
    M = g["single_wall"]
    rng = np.random.default_rng(8)
    z = rng.standard_normal(M.shape[0])
    Mt, zt = torch.from_numpy(M), torch.from_numpy(z)
    eye = torch.eye(M.shape[0], dtype=torch.float64)
    for fn in (solvers.stochastic_forcing_eig, solvers.stochastic_forcing_eig_symm, solvers.stochastic_forcing_cholesky):
        S = torch.stack([fn(Mt, 1.0, eye[:, c]) for c in range(M.shape[0])], dim=1)       # the square-root factor
        assert np.linalg.norm((S @ S.T).numpy() - M) < 1e-12 * np.linalg.norm(M)
        assert torch.allclose(fn(Mt, 1.7, zt), 1.7 * (S @ zt), rtol=1e-12, atol=0)
        w = fn(Mt, 0.5, None, generator=torch.Generator().manual_seed(1))
        assert w.shape == (M.shape[0],) and torch.equal(w, fn(Mt, 0.5, None, generator=torch.Generator().manual_seed(1)))
    Ssym = torch.stack([solvers.stochastic_forcing_eig_symm(Mt, 1.0, eye[:, c]) for c in range(M.shape[0])], dim=1)
    assert torch.allclose(Ssym, Ssym.T, rtol=0, atol=1e-14)
    ref_dir = "/root/reference/stochastic_forcing"
    if os.path.exists(os.path.join(ref_dir, "stochastic_forcing.py")):
        import importlib.util
        spec = importlib.util.spec_from_file_location("ref_stochastic_forcing", os.path.join(ref_dir, "stochastic_forcing.py"))
        ref = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(ref)
        # symmetric root and Cholesky factor are unique; the plain eig form depends on eigenvector signs, so it is
        # compared through its covariance above only
        assert np.allclose(solvers.stochastic_forcing_eig_symm(Mt, 1.3, zt).numpy(), ref.stochastic_forcing_eig_symm(M, 1.3, z),
                           rtol=1e-10, atol=1e-14)
        assert np.allclose(solvers.stochastic_forcing_cholesky(Mt, 1.3, zt).numpy(), ref.stochastic_forcing_cholesky(M, 1.3, z),
                           rtol=1e-10, atol=1e-14)
    # a singular mobility (eigenvalue 0) is legal for the eig forms, as in the reference
    sing = torch.diag(torch.tensor([2.0, 0.0, 0.5], dtype=torch.float64))
    assert torch.allclose(solvers.stochastic_forcing_eig_symm(sing, 1.0, torch.ones(3, dtype=torch.float64)),
                          torch.tensor([2.0 ** 0.5, 0.0, 0.5 ** 0.5], dtype=torch.float64))
