"""Host logic of rigidmultiblobswall_b200/rigid.py on CPU tensors: the saddle-point operator, the block-diagonal
preconditioner's apply, the mobility / resistance / body_mobility schemes and prescribed kinematics, with the three
CUDA entry points they call (K, K^T, blob mobility) replaced by dense NumPy stand-ins and the CPU oracle.  Answers
are the reference's own (tests/golden/{rigid,resistance,prescribed}_golden.npz).  The same functions run against the
real kernels in tests/test_rigid_gpu.py and tests/test_schemes_gpu.py."""
import os

import numpy as np
import pytest
import torch

import oracle
from rigidmultiblobswall_b200 import rigid

HERE = os.path.dirname(os.path.abspath(__file__))
NB_BLOBS = 15


def rel(x, ref):
    x = x.detach().numpy() if hasattr(x, "detach") else np.asarray(x)
    return np.linalg.norm(x.reshape(-1) - np.asarray(ref).reshape(-1)) / np.linalg.norm(ref)


@pytest.fixture(scope="module")
def gold():
    return {k: np.load(os.path.join(HERE, "golden", k + "_golden.npz")) for k in ("rigid", "resistance", "prescribed")}


@pytest.fixture()
def host(gold, monkeypatch):
    """(bodies stand-in, preconditioner, blob mobility product) on the CPU"""
    rg = gold["rigid"]
    a, eta, _ = rg["rigid/a_eta_tol"]
    r, loc = rg["rigid/r_vectors"], rg["rigid/location"]
    nb, n = len(loc), len(r)
    K = np.zeros((3 * n, 6 * nb))
    for i in range(n):
        b = i // NB_BLOBS
        d = r[i] - loc[b]
        K[3 * i:3 * i + 3, 6 * b:6 * b + 3] = np.eye(3)
        K[3 * i:3 * i + 3, 6 * b + 3:6 * b + 6] = -np.array([[0, -d[2], d[1]], [d[2], 0, -d[0]], [-d[1], d[0], 0]])
    assert np.allclose(K @ rg["rigid/U"].reshape(-1), rg["rigid/K_U"].reshape(-1))
    Kt = torch.from_numpy(K)

    class Bodies:
        device = torch.device("cpu")
        n_blobs, n_bodies, nb_max, uniform = n, nb, NB_BLOBS, True
        set_prescribed = rigid.RigidBodies.set_prescribed

        def get_r_vectors(self):
            return torch.from_numpy(r.reshape(-1).copy())

    bodies = Bodies()
    bodies.set_prescribed(None)

    def K_prod(bodies, vector, Nblobs=None, K_bodies=None, r_vectors=None, out=None, alpha=1.0, beta=0.0):
        v = alpha * (Kt @ vector.reshape(-1))
        if out is None:
            return v.view(-1, 3)
        out.mul_(beta).add_(v)
        return out.view(-1, 3)

    def KT_prod(bodies, vector, Nblobs=None, K_bodies=None, r_vectors=None, alpha=1.0):
        return (alpha * (Kt.T @ vector.reshape(-1))).view(-1, 3)

    monkeypatch.setattr(rigid, "K_matrix_vector_prod", K_prod)
    monkeypatch.setattr(rigid, "K_matrix_T_vector_prod", KT_prod)

    def mvp(rv, v, eta_, a_, periodic_length=None):
        return torch.from_numpy(oracle.mobility_product("UF", "single_wall", rv.numpy().reshape(-1, 3), v.numpy(), eta_, a_))

    m = 3 * NB_BLOBS
    eye = np.eye(m)
    PC = rigid.BlockDiagonalPreconditioner.__new__(rigid.BlockDiagonalPreconditioner)     # skip the CUDA build
    PC.bodies, PC.m, PC.pad_index = bodies, m, None
    M = np.array([np.array([oracle.mobility_product("UF", "single_wall", r[NB_BLOBS * b:NB_BLOBS * (b + 1)], eye[c], eta, a)
                            for c in range(m)]).T for b in range(nb)])
    PC.M_blobs = torch.from_numpy(M)
    PC.chol = torch.linalg.cholesky(PC.M_blobs)
    PC.M_inv = torch.cholesky_inverse(PC.chol)
    PC.K = torch.from_numpy(np.array([K[m * b:m * (b + 1), 6 * b:6 * b + 6] for b in range(nb)]))
    PC.N = torch.linalg.pinv(PC.K.transpose(1, 2) @ PC.M_inv @ PC.K, hermitian=True)
    return bodies, PC, mvp, float(eta), float(a)


def test_operator_preconditioner_and_mobility_scheme(gold, host):
    rg = gold["rigid"]
    bodies, PC, mvp, eta, a = host
    x = torch.from_numpy(rg["rigid/x"])
    assert rel(rigid.linear_operator_rigid(x, bodies, bodies.get_r_vectors(), eta, a, mvp), rg["rigid/A_x"]) < 1e-13
    assert rel(PC(x), rg["rigid/PC_x"]) < 1e-10
    vel, lam, its, info = rigid.solve_mobility_problem(bodies, rg["rigid/FT"], eta, a, tolerance=1e-8, PC=PC,
                                                       mobility_vector_prod=mvp)
    assert info == 0 and abs(its - int(rg["rigid/gmres_iterations"][0])) <= 1
    sol = torch.cat([lam.reshape(-1), vel.reshape(-1)])
    assert rel(sol, rg["rigid/solution"]) < 1e-6
    # a first guess saves iterations and lands on the same solution
    vel2, lam2, its2, info2 = rigid.solve_mobility_problem(bodies, rg["rigid/FT"], eta, a, tolerance=1e-8, PC=PC,
                                                           mobility_vector_prod=mvp, x0=sol * (1 + 1e-3))
    assert info2 == 0 and its2 < its
    assert rel(torch.cat([lam2.reshape(-1), vel2.reshape(-1)]), rg["rigid/solution"]) < 1e-6


@pytest.mark.parametrize("key", ["noslip", "slip"])
def test_resistance_scheme(gold, host, key):
    g = gold["resistance"]
    bodies, PC, mvp, eta, a = host
    force, lam, its, info = rigid.solve_resistance_problem(bodies, g["velocity"], eta, a, tolerance=1e-10, PC=PC,
                                                           slip=g["slip"] if key == "slip" else None,
                                                           mobility_vector_prod=mvp)
    assert info == 0
    assert rel(lam, g[f"resistance/{key}/force_blobs"]) < 1e-8
    assert rel(force, g[f"resistance/{key}/force"]) < 1e-8


def test_body_mobility_scheme(gold, host):
    g = gold["resistance"]
    bodies, PC, mvp, eta, a = host
    N = rigid.body_mobility(bodies, eta, a, tolerance=1e-10, PC=PC, mobility_vector_prod=mvp)
    assert rel(N, g["body_mobility/N"]) < 1e-8


def test_prescribed_kinematics(gold, host):
    g = gold["prescribed"]
    bodies, PC, mvp, eta, a = host
    presc = [g["prescribed_velocity"][k] if g["prescribed"][k] else None for k in range(bodies.n_bodies)]
    bodies.set_prescribed(presc)
    assert bodies.any_prescribed and int(bodies.prescribed.sum()) == 2
    x = torch.from_numpy(g["x"])
    assert rel(rigid.linear_operator_rigid(x, bodies, bodies.get_r_vectors(), eta, a, mvp), g["A_x"]) < 1e-13
    assert rel(PC(x), g["PC_x"]) < 1e-10
    vel, lam, its, info = rigid.solve_mobility_problem(bodies, g["FT"], eta, a, tolerance=1e-10, slip=g["slip"], PC=PC,
                                                       mobility_vector_prod=mvp)
    assert info == 0 and abs(its - int(g["gmres_iterations"][0])) <= 1
    assert rel(vel, g["velocity"]) < 1e-8 and rel(lam, g["lambda"]) < 1e-8
    assert torch.equal(vel[bodies.prescribed], bodies.prescribed_velocity[bodies.prescribed])
    assert rel(rigid.K_matrix_T_vector_prod(bodies, lam).reshape(-1, 6), g["force"]) < 1e-8
    with pytest.raises(ValueError):
        bodies.set_prescribed([None])
