"""Golden vectors for the rigid-body rows from the REFERENCE's own
multi_bodies.py / body.py / stochastic_forcing.py, imported in place.

    python tests/golden/make_golden_rigid.py     (build container only)

Harness-side stubs: `imp` (removed from Python 3.12) and the
reference's quaternion_integrator/gmres.py (imports a private SciPy module
that no longer exists); the GMRES oracle is scipy.sparse.linalg.gmres driven
exactly as general_application_utils.gmres:600-635 does (right preconditioning).
Also stores the reference geometry files needed at run time on the GPU box
(boomerang_N_15, shell_N_162, blob) as arrays.
"""
import os
import sys
import types

import numpy as np
import scipy.sparse.linalg as spla

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

REF = "/root/reference"


def import_multi_bodies():
    import importlib.util
    imp = types.ModuleType("imp")

    def find_module(name, path=None):
        if importlib.util.find_spec(name) is None:
            raise ImportError(name)
        return (None, None, None)

    imp.find_module = find_module
    sys.modules.setdefault("imp", imp)
    for name in ("quaternion_integrator.gmres", "gmres"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.path.insert(0, REF)
    sys.path.insert(0, os.path.join(REF, "multi_bodies"))
    import multi_bodies
    return multi_bodies


def read_vertex(name):
    path = os.path.join(REF, "multi_bodies", "Structures", name)
    with open(path) as f:
        lines = [l for l in f if not l.startswith("#") and l.strip()]
    head = lines[0].split()
    n = int(head[0])
    radius = float(head[1]) if len(head) > 1 else np.nan
    xyz = np.array([[float(x) for x in l.split()[:3]] for l in lines[1:1 + n]])
    return xyz, radius


def main():
    mbod = import_multi_bodies()
    from body import body
    from quaternion_integrator.quaternion import Quaternion
    from mobility import mobility as mb
    import stochastic_forcing.stochastic_forcing as stochastic

    rng = np.random.default_rng(42)
    store = {}
    for key, fn in (("boomerang_N_15", "boomerang_N_15.vertex"), ("shell_N_162", "shell_N_162_Rg_0_9497_Rh_1.vertex"),
                    ("shell_N_12", "shell_N_12_Rg_0_7921_Rh_1.vertex")):
        xyz, rad = read_vertex(fn)
        store[f"struct/{key}"] = xyz
        store[f"struct/{key}_radius"] = np.array([rad])

    a, eta, tol = 0.25, 1.0, 1e-8                       # multi_bodies/inputfile.dat
    ref_cfg = store["struct/boomerang_N_15"]
    nbod = 12
    loc = np.array([[4.0 * (k % 4), 4.0 * (k // 4), 3.0 + rng.random()] for k in range(nbod)])
    loc[5, 2] = 1.2                                     # one body close to the wall: blobs with z < a
    quat = rng.standard_normal((nbod, 4))
    quat /= np.linalg.norm(quat, axis=1)[:, None]
    bodies = []
    for k in range(nbod):
        b = body.Body(loc[k], Quaternion(quat[k]), ref_cfg, a)
        b.mobility_blobs = mbod.set_mobility_blobs("python")
        b.ID = "boomerang"
        bodies.append(b)
    Nblobs = sum(b.Nblobs for b in bodies)
    r_vectors = mbod.get_blobs_r_vectors(bodies, Nblobs)
    print("blobs below z=a:", int((r_vectors[:, 2] < a).sum()), "min z", r_vectors[:, 2].min())
    mbod.mobility_vector_prod = mbod.set_mobility_vector_prod("numba")

    U = rng.standard_normal((2 * nbod, 3))
    lam = rng.standard_normal((Nblobs, 3))
    store.update({"rigid/a_eta_tol": np.array([a, eta, tol]), "rigid/location": loc, "rigid/quaternion": quat,
                  "rigid/r_vectors": r_vectors, "rigid/U": U, "rigid/lambda": lam})
    store["rigid/K_U"] = mbod.K_matrix_vector_prod(bodies, U, Nblobs)
    store["rigid/KT_lambda"] = mbod.K_matrix_T_vector_prod(bodies, lam, Nblobs)

    n = 3 * Nblobs + 6 * nbod
    x = rng.standard_normal(n)
    L = np.zeros(3)
    lin = lambda v: mbod.linear_operator_rigid(v, bodies, [], r_vectors, eta, a, periodic_length=L)
    store["rigid/x"] = x
    store["rigid/A_x"] = lin(x)
    PC = mbod.build_block_diagonal_preconditioner(bodies, [], r_vectors, Nblobs, eta, a, step=0, update_PC=1)
    store["rigid/PC_x"] = PC(x)
    store["rigid/M_body0"] = bodies[0].calc_mobility_blobs(eta, a)
    store["rigid/M_body5"] = bodies[5].calc_mobility_blobs(eta, a)

    # mobility problem: RHS = [slip = 0 ; -force_torque], normalised (multi_bodies_utilities.py:477-512)
    FT = rng.standard_normal(6 * nbod)
    RHS = np.concatenate([np.zeros(3 * Nblobs), -FT])
    nrm = np.linalg.norm(RHS)
    RHSn = RHS / nrm
    its = [0]
    A_LO = spla.LinearOperator((n, n), matvec=lambda y: lin(PC(y)), dtype="float64")
    y, info = spla.gmres(A_LO, RHSn, rtol=tol, atol=0, restart=60, maxiter=1000,
                         callback=lambda r: its.__setitem__(0, its[0] + 1), callback_type="pr_norm")
    sol = PC(y) * nrm
    res = np.linalg.norm(lin(sol) - RHS) / nrm
    print("gmres info", info, "iterations", its[0], "true residual", res)
    store.update({"rigid/FT": FT, "rigid/solution": sol, "rigid/gmres_iterations": np.array([its[0]])})

    # Lanczos M^{1/2} z on the blobs of the first 4 bodies (stochastic_forcing/test.py recipe)
    rsub = r_vectors[:60]
    z = rng.standard_normal(180)
    mult = lambda v: mb.single_wall_mobility_trans_times_force_numba(rsub, v, eta, a)
    noise, it_l = stochastic.stochastic_forcing_lanczos(factor=1.3, tolerance=1e-10, dim=180, mobility_mult=mult, z=z)
    M = mb.single_wall_fluid_mobility(rsub, eta, a)
    w, Vv = np.linalg.eigh(M)
    exact = 1.3 * (Vv * np.sqrt(w)) @ (Vv.T @ z)
    print("lanczos iterations", it_l, "err vs eig", np.linalg.norm(noise - exact) / np.linalg.norm(exact))
    store.update({"lanczos/z": z, "lanczos/noise": noise, "lanczos/iterations": np.array([it_l]),
                  "lanczos/tolerance_factor": np.array([1e-10, 1.3])})

    # stochastic block-diagonal preconditioner (multi_bodies.py:480-618): P^T M P and P^-1 products
    PCd, mobility_pc, P_inv_mult = mbod.build_block_diagonal_preconditioners_det_stoch(
        bodies, r_vectors, Nblobs, eta, a, step=0, update_PC=1, periodic_length=L)
    w = rng.standard_normal(3 * Nblobs)
    store["stoch/w"] = w
    store["stoch/mobility_pc_w"] = mobility_pc(w.copy())
    store["stoch/P_inv_w"] = P_inv_mult(w.copy())
    store["stoch/PC_x"] = PCd(x)
    # preconditioned Lanczos noise for the whole 12-body suspension (quaternion_integrator_multi_bodies.py:703)
    z2 = rng.standard_normal(3 * Nblobs)
    noise2, it2 = stochastic.stochastic_forcing_lanczos(factor=0.7, tolerance=1e-10, dim=3 * Nblobs,
                                                        mobility_mult=mobility_pc, L_mult=P_inv_mult, z=z2)
    print("preconditioned lanczos iterations", it2)
    store.update({"stoch/z": z2, "stoch/noise": noise2, "stoch/iterations": np.array([it2]),
                  "stoch/tolerance_factor": np.array([1e-10, 0.7])})

    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "rigid_golden.npz")
    np.savez_compressed(path, **store)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
