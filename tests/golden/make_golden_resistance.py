"""Golden vectors for the `resistance` and `body_mobility` schemes of the
REFERENCE's multi_bodies_utilities.py (:549-604), computed with the reference's
own multi_bodies.py functions (imported in place) on the 12-boomerang
suspension stored in rigid_golden.npz.

    python tests/golden/make_golden_resistance.py     (build container only)

The reference forms the dense blob mobility (multi_bodies.mobility_blobs ->
mobility.single_wall_fluid_mobility) and solves with LAPACK; the product path
is matrix-free, so these are independent answers.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from make_golden_rigid import import_multi_bodies  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    mbod = import_multi_bodies()
    from body import body
    from quaternion_integrator.quaternion import Quaternion

    rg = np.load(os.path.join(HERE, "rigid_golden.npz"))
    a, eta, _ = rg["rigid/a_eta_tol"]
    ref_cfg, loc, quat = rg["struct/boomerang_N_15"], rg["rigid/location"], rg["rigid/quaternion"]
    nbod = len(loc)
    bodies = []
    for k in range(nbod):
        b = body.Body(loc[k], Quaternion(quat[k]), ref_cfg, a)
        b.mobility_blobs = mbod.set_mobility_blobs("python")
        bodies.append(b)
    Nblobs = sum(b.Nblobs for b in bodies)
    r_vectors = mbod.get_blobs_r_vectors(bodies, Nblobs)
    assert np.array_equal(r_vectors, rg["rigid/r_vectors"])
    mbod.mobility_blobs = mbod.set_mobility_blobs("python")

    rng = np.random.default_rng(4242)
    velocity = rng.standard_normal((2 * nbod, 3))
    slip = 0.3 * rng.standard_normal((Nblobs, 3))
    store = {"velocity": velocity, "slip": slip}

    # scheme == 'resistance' (multi_bodies_utilities.py:550-574), once without and once with slip
    M = mbod.mobility_blobs(r_vectors, eta, a)
    for key, s in (("noslip", np.zeros((Nblobs, 3))), ("slip", slip)):
        rhs = (s + mbod.K_matrix_vector_prod(bodies, velocity, Nblobs)).reshape(-1)
        force_blobs = np.linalg.solve(M, rhs)
        store[f"resistance/{key}/force_blobs"] = force_blobs
        store[f"resistance/{key}/force"] = np.reshape(mbod.K_matrix_T_vector_prod(bodies, force_blobs, Nblobs), (nbod, 6))

    # scheme == 'body_mobility' (multi_bodies_utilities.py:596-600)
    K = mbod.calc_K_matrix(bodies, Nblobs)
    R = np.linalg.inv(M)
    N = np.linalg.pinv(K.T @ R @ K)
    store["body_mobility/N"] = N
    store["body_mobility/cond_M"] = np.array([np.linalg.cond(M)])
    print("cond(M) =", store["body_mobility/cond_M"][0], " |N - N^T| =", np.abs(N - N.T).max())

    path = os.path.join(HERE, "resistance_golden.npz")
    np.savez_compressed(path, **store)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
