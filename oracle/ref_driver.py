"""TEST / BASELINE INFRASTRUCTURE ONLY -- drives the UNMODIFIED reference (staged by oracle/make_ref.py) the way its own
scripts do, for golden generation (tests/golden/make_golden_configs.py), the executed drop-in test
(tests/test_reference_dropin_gpu.py) and the CPU legs of bench.py.  Never imported by the product."""
import os
import time

import numpy as np
import scipy.sparse.linalg as spla

from . import make_ref


def use_all_cores():
    """torchrun exports OMP_NUM_THREADS=1; numba reads NUMBA_NUM_THREADS at import.  Call before importing numba."""
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    os.environ["NUMBA_NUM_THREADS"] = str(n)
    os.environ["OMP_NUM_THREADS"] = str(n)
    return n


def numba_threads():
    import numba
    return int(numba.get_num_threads())


def build_bodies(mbod, ref_cfg, loc, quat, a, blobs_implementation="python"):
    """The reference's own list of body.Body (multi_bodies_utilities.py:283-319 without the file parsing)."""
    from body import body
    from quaternion_integrator.quaternion import Quaternion
    bodies = []
    for k in range(len(loc)):
        cfg = ref_cfg[k] if isinstance(ref_cfg, (list, tuple)) else ref_cfg      # one geometry per body, or one for all
        b = body.Body(loc[k], Quaternion(quat[k]), cfg, a)
        b.mobility_blobs = mbod.set_mobility_blobs(blobs_implementation)
        b.ID = "body"
        bodies.append(b)
    return bodies


def gmres_right(lin, PC, RHS, tol, restart=60, maxiter=1000):
    """general_application_utils.gmres:600-635 restated for the installed SciPy (it passes tol=, SciPy >= 1.14 only
    knows rtol=): right preconditioning, A' = A P^-1, x = P^-1 y, RHS scaled to unit norm as
    multi_bodies_utilities.py:510-512 does.  Returns (solution, info, iterations, true relative residual)."""
    n = RHS.size
    nrm = np.linalg.norm(RHS)
    its = [0]
    A_LO = spla.LinearOperator((n, n), matvec=lambda y: lin(PC(y)), dtype="float64")
    y, info = spla.gmres(A_LO, RHS / nrm, rtol=tol, atol=0, restart=restart, maxiter=maxiter,
                         callback=lambda r: its.__setitem__(0, its[0] + 1), callback_type="pr_norm")
    sol = PC(y) * nrm
    return sol, info, its[0], float(np.linalg.norm(lin(sol) - RHS) / nrm)


def mobility_problem(ref_cfg, loc, quat, FT, a, eta, tol, implementation="numba", mbod=None, slip=None):
    """The `mobility` scheme of multi_bodies_utilities.py:414-540 with the reference's own operator and
    preconditioner; `implementation` is the string set_mobility_vector_prod resolves (multi_bodies.py:233-287).
    Returns dict(solution, info, iterations, residual, seconds_pc, seconds_gmres, n_blobs)."""
    mbod = mbod or make_ref.import_multi_bodies()
    mbod.mobility_vector_prod = mbod.set_mobility_vector_prod(implementation)
    bodies = build_bodies(mbod, ref_cfg, loc, quat, a)
    Nblobs = sum(b.Nblobs for b in bodies)
    r_vectors = mbod.get_blobs_r_vectors(bodies, Nblobs)
    L = np.zeros(3)
    lin = lambda v: mbod.linear_operator_rigid(v, bodies, [], r_vectors, eta, a, periodic_length=L)
    lin(np.zeros(3 * Nblobs + 6 * len(bodies)))                       # JIT / first-call costs outside the timing
    t0 = time.perf_counter()
    PC = mbod.build_block_diagonal_preconditioner(bodies, [], r_vectors, Nblobs, eta, a, step=0, update_PC=1)
    t_pc = time.perf_counter() - t0
    RHS = np.concatenate([np.zeros(3 * Nblobs) if slip is None else np.asarray(slip).reshape(-1), -np.asarray(FT).reshape(-1)])
    t0 = time.perf_counter()
    sol, info, its, res = gmres_right(lin, PC, RHS, tol)
    t_g = time.perf_counter() - t0
    return {"solution": sol, "info": int(info), "iterations": int(its), "residual": res, "seconds_pc": t_pc,
            "seconds_gmres": t_g, "n_blobs": Nblobs, "bodies": bodies, "r_vectors": r_vectors}


def numba_uf_rate(n, make_cloud, reps=1, mb_ref=None):
    """pairs/s of the reference's public single-wall UF (mobility.py:1132 -> mobility_numba.py:124) on make_cloud(n),
    best of `reps` after the JIT call."""
    mb_ref = mb_ref or make_ref.import_mobility()
    r, f, a, eta = make_cloud(n)
    L = np.zeros(3)
    mb_ref.single_wall_mobility_trans_times_force_numba(r[:64].copy(), f[:64].copy(), eta, a, periodic_length=L)
    best = 1e300
    for _ in range(reps):
        t0 = time.perf_counter()
        mb_ref.single_wall_mobility_trans_times_force_numba(r, f, eta, a, periodic_length=L)
        best = min(best, time.perf_counter() - t0)
    return float(n) * n / best, best


def patch_utils_gmres(utils_module=None):
    """Third harness stub: general_application_utils.gmres passes `tol=` to SciPy (:598, :627), which
    SciPy >= 1.14 no longer accepts.  The staged FILE is not edited; at run time the module's function is replaced by
    the same call with `rtol=` (no preconditioner / left preconditioner branch, :596-598, the one the rollers use)."""
    if utils_module is None:
        make_ref.import_multi_bodies()
        import general_application_utils as utils_module

    def gmres(A, b, x0=None, tol=1e-05, restart=None, maxiter=None, xtype=None, M=None, callback=None, restrt=None,
              PC_side="right"):
        if M is not None and PC_side == "right":
            raise NotImplementedError("harness stub: right-preconditioned calls go through ref_driver.gmres_right")
        return spla.gmres(A, b, M=M, x0=x0, rtol=tol, atol=0, maxiter=maxiter, restart=restart, callback=callback,
                          callback_type="pr_norm")

    utils_module.gmres = gmres
    return utils_module


def rollers_integrator(mbod, r_vectors, force, omega, eta, a, L, tolerance, implementation="numba", domain="single_wall"):
    """The reference's own QuaternionIntegratorRollers (quaternion_integrator_rollers.py:23-125) set up for
    compute_deterministic_velocity_and_torque (:905): one single-blob body per roller, prescribed angular velocity."""
    from quaternion_integrator.quaternion_integrator_rollers import QuaternionIntegratorRollers
    from quaternion_integrator.quaternion import Quaternion
    from body import body
    patch_utils_gmres()
    blob = np.zeros((1, 3))
    bodies = [body.Body(np.array(x), Quaternion(np.array([1.0, 0.0, 0.0, 0.0])), blob, a) for x in np.asarray(r_vectors)]
    integ = QuaternionIntegratorRollers(bodies, len(bodies), "deterministic_forward_euler_rollers", tolerance=tolerance,
                                        domain=domain, mobility_vector_prod_implementation=implementation)
    integ.eta, integ.a, integ.periodic_length = eta, a, L
    integ.free_kinematics = "False"
    integ.print_residual = False
    integ.get_omega_one_roller = lambda: np.asarray(omega, dtype=np.float64)
    f = np.asarray(force, dtype=np.float64).reshape(-1)
    integ.calc_one_blob_forces = lambda r, blob_radius=None, blob_mass=None: f.copy()
    integ.calc_blob_blob_forces = lambda r, blob_radius=None: np.zeros_like(f)
    return integ
