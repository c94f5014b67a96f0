"""Golden vectors for bodies with PRESCRIBED kinematics (mixed mobility / resistance problem) from the REFERENCE's
own multi_bodies.py: linear_operator_rigid (:458-464), the block-diagonal preconditioner (:850-862) and the `mobility`
scheme's right-hand side and solve (multi_bodies_utilities.py:477-527), on the 12-boomerang suspension of
rigid_golden.npz with bodies 2 and 7 prescribed.

    python tests/golden/make_golden_prescribed.py     (build container only)
"""
import os
import sys

import numpy as np
import scipy.sparse.linalg as spla

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from make_golden_rigid import import_multi_bodies  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    mbod = import_multi_bodies()
    from body import body
    from quaternion_integrator.quaternion import Quaternion

    rg = np.load(os.path.join(HERE, "rigid_golden.npz"))
    a, eta, tol = rg["rigid/a_eta_tol"]
    ref_cfg, loc, quat = rg["struct/boomerang_N_15"], rg["rigid/location"], rg["rigid/quaternion"]
    nbod = len(loc)
    rng = np.random.default_rng(777)
    prescribed = np.zeros(nbod, dtype=bool)
    prescribed[[2, 7]] = True
    U_presc = np.zeros((nbod, 6))
    U_presc[prescribed] = rng.standard_normal((2, 6))
    bodies = []
    for k in range(nbod):
        b = body.Body(loc[k], Quaternion(quat[k]), ref_cfg, a)
        b.mobility_blobs = mbod.set_mobility_blobs("python")
        b.ID = "boomerang"
        if prescribed[k]:
            b.prescribed_kinematics = True
            b.prescribed_velocity = U_presc[k].copy()
        bodies.append(b)
    Nblobs = sum(b.Nblobs for b in bodies)
    r_vectors = mbod.get_blobs_r_vectors(bodies, Nblobs)
    mbod.mobility_vector_prod = mbod.set_mobility_vector_prod("numba")
    n = 3 * Nblobs + 6 * nbod
    L = np.zeros(3)
    lin = lambda v: mbod.linear_operator_rigid(v, bodies, [], r_vectors, eta, a, periodic_length=L)
    x = rng.standard_normal(n)
    PC = mbod.build_block_diagonal_preconditioner(bodies, [], r_vectors, Nblobs, eta, a, step=0, update_PC=1)
    store = {"prescribed": prescribed, "prescribed_velocity": U_presc, "x": x, "A_x": lin(x), "PC_x": PC(x)}

    # the `mobility` scheme with prescribed bodies (multi_bodies_utilities.py:477-527)
    FT = rng.standard_normal((nbod, 6))
    slip = 0.2 * rng.standard_normal((Nblobs, 3))
    RHS = np.concatenate([slip.flatten(), -FT.flatten()])
    offset = 0
    for k, b in enumerate(bodies):
        if b.prescribed_kinematics is True:
            KU = np.dot(b.calc_K_matrix(), b.calc_prescribed_velocity())
            RHS[3 * offset: 3 * (offset + b.Nblobs)] += KU.flatten()
            RHS[3 * Nblobs + k * 6: 3 * Nblobs + (k + 1) * 6] = 0.0
        offset += b.Nblobs
    nrm = np.linalg.norm(RHS)
    its = [0]
    A_LO = spla.LinearOperator((n, n), matvec=lambda y: lin(PC(y)), dtype="float64")
    y, info = spla.gmres(A_LO, RHS / nrm, rtol=1e-10, atol=0, restart=60, maxiter=1000,
                         callback=lambda r: its.__setitem__(0, its[0] + 1), callback_type="pr_norm")
    sol = PC(y) * nrm
    print("gmres info", info, "iterations", its[0], "true residual", np.linalg.norm(lin(sol) - RHS) / nrm)
    lam = sol[:3 * Nblobs].reshape(Nblobs, 3)
    body_rows = sol[3 * Nblobs:].reshape(nbod, 6).copy()
    velocity = body_rows.copy()
    velocity[prescribed] = U_presc[prescribed]                    # :524-527
    force = np.reshape(mbod.K_matrix_T_vector_prod(bodies, lam, Nblobs), (nbod, 6))     # :536
    print("constraint force rows equal K^T lambda on the prescribed bodies:",
          np.abs(body_rows[prescribed] - force[prescribed]).max())
    store.update({"FT": FT, "slip": slip, "RHS": RHS, "lambda": lam, "velocity": velocity, "force": force,
                  "gmres_iterations": np.array([its[0]])})
    path = os.path.join(HERE, "prescribed_golden.npz")
    np.savez_compressed(path, **store)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
