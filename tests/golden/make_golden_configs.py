"""Golden answers of the REFERENCE at BASELINE-config size (build container only; minutes of CPU time).

    python tests/golden/make_golden_configs.py [--shells 128] [--skip-c2] [--skip-c5]

Writes tests/golden/configs_golden.npz:
  c2/*   BASELINE config 2, exactly bench.make_c2(1000): 1,000 boomerang_N_15 bodies (15,000 blobs), a = 0.25,
         eta = 1, tolerance 1e-8.  The reference's own linear_operator_rigid + build_block_diagonal_preconditioner
         (multi_bodies/multi_bodies.py:424-471, 752-903) with the numba product, under scipy.sparse.linalg.gmres
         driven as general_application_utils.gmres:600-635 does (right preconditioning) -- the mobility scheme of
         multi_bodies_utilities.py:414-540.  Stored: the 6,000 body velocities, 3,000 sampled blob forces, the
         iteration count, the reference's seconds on this machine.
  c5_<shells>/*  BASELINE config 5 reduced to `--shells` shell_N_162 spheres (bench.make_c5): the Brownian slip +
         solve of quaternion_integrator_multi_bodies.py:626-735 -- build_block_diagonal_preconditioners_det_stoch
         (multi_bodies.py:480), stochastic_forcing_lanczos on P^T M P mapped back by P^-1
         (stochastic_forcing.py:112-264), then the mobility solve with slip = -noise.  Stored: sampled noise entries,
         all body velocities, both iteration counts.
The reference is imported from oracle/_ref (byte-identical staged copy of /root/reference, oracle/make_ref.py).
"""
import argparse
import os
import sys
import time

import numpy as np
import scipy.sparse.linalg as spla

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import make_ref  # noqa: E402
from oracle.ref_driver import build_bodies, gmres_right  # noqa: E402
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shells", type=int, default=128)
    ap.add_argument("--skip-c2", action="store_true")
    ap.add_argument("--skip-c5", action="store_true")
    args = ap.parse_args()
    path = os.path.join(HERE, "configs_golden.npz")
    mbod = make_ref.import_multi_bodies()
    import stochastic_forcing.stochastic_forcing as stochastic
    mbod.mobility_vector_prod = mbod.set_mobility_vector_prod("numba")
    L = np.zeros(3)
    rng = np.random.default_rng(2025)
    new = {}

    if not args.skip_c2:
        ref_cfg, loc, quat, FT, a, eta, tol = bench.make_c2(1000)
        bodies = build_bodies(mbod, ref_cfg, loc, quat, a)
        Nblobs = sum(b.Nblobs for b in bodies)
        r_vectors = mbod.get_blobs_r_vectors(bodies, Nblobs)
        lin = lambda v: mbod.linear_operator_rigid(v, bodies, [], r_vectors, eta, a, periodic_length=L)
        lin(np.zeros(3 * Nblobs + 6 * len(bodies)))                       # numba JIT outside the timing
        t0 = time.perf_counter()
        PC = mbod.build_block_diagonal_preconditioner(bodies, [], r_vectors, Nblobs, eta, a, step=0, update_PC=1)
        t_pc = time.perf_counter() - t0
        RHS = np.concatenate([np.zeros(3 * Nblobs), -FT])
        t0 = time.perf_counter()
        sol, info, its, res = gmres_right(lin, PC, RHS, tol)
        t_g = time.perf_counter() - t0
        print("c2: info", info, "iterations", its, "true residual", res, "seconds pc %.2f gmres %.2f" % (t_pc, t_g))
        idx = np.sort(rng.choice(3 * Nblobs, 3000, replace=False))
        new.update({"c2/velocity": sol[3 * Nblobs:], "c2/lambda_index": idx, "c2/lambda_sample": sol[idx],
                    "c2/lambda_norm": np.array([np.linalg.norm(sol[:3 * Nblobs])]),
                    "c2/iterations": np.array([its]), "c2/true_residual": np.array([res]),
                    "c2/reference_seconds_pc_gmres_threads": np.array([t_pc, t_g, os.cpu_count()])})

    if not args.skip_c5:
        nb = args.shells
        shell, a, eta, kT, dt, loc, quat, FT, z = bench.make_c5(nb)
        bodies = build_bodies(mbod, shell, loc, quat, a)
        Nblobs = sum(b.Nblobs for b in bodies)
        r_vectors = mbod.get_blobs_r_vectors(bodies, Nblobs)
        lin = lambda v: mbod.linear_operator_rigid(v, bodies, [], r_vectors, eta, a, periodic_length=L)
        t0 = time.perf_counter()
        mbod.build_block_diagonal_preconditioners_det_stoch.initialized = []
        mbod.build_block_diagonal_preconditioners_det_stoch.mobility_bodies = []
        PC, mobility_pc, P_inv_mult = mbod.build_block_diagonal_preconditioners_det_stoch(
            bodies, r_vectors, Nblobs, eta, a, step=0, update_PC=1, periodic_length=L)
        t_pc = time.perf_counter() - t0
        factor, ltol = np.sqrt(2 * kT / dt), 1e-8
        t0 = time.perf_counter()
        noise, it_l = stochastic.stochastic_forcing_lanczos(factor=factor, tolerance=ltol, dim=3 * Nblobs,
                                                            mobility_mult=mobility_pc, L_mult=P_inv_mult, z=z.copy())
        t_l = time.perf_counter() - t0
        tol = 1e-8
        RHS = np.concatenate([-noise, -FT.reshape(-1)])
        t0 = time.perf_counter()
        sol, info, its, res = gmres_right(lin, PC, RHS, tol)
        t_g = time.perf_counter() - t0
        print("c5 (%d shells, %d blobs): lanczos its %d, gmres info %d its %d residual %.2e; seconds pc %.1f lanczos %.1f "
              "gmres %.1f" % (nb, Nblobs, it_l, info, its, res, t_pc, t_l, t_g))
        idx = np.sort(rng.choice(3 * Nblobs, 3000, replace=False))
        key = "c5_%d" % nb
        new.update({key + "/noise_index": idx, key + "/noise_sample": noise[idx],
                    key + "/noise_norm": np.array([np.linalg.norm(noise)]),
                    key + "/velocity": sol[3 * Nblobs:], key + "/lambda_sample": sol[idx],
                    key + "/lanczos_iterations": np.array([it_l]), key + "/gmres_iterations": np.array([its]),
                    key + "/tolerances_lanczos_gmres": np.array([ltol, tol]),
                    key + "/true_residual": np.array([res]),
                    key + "/reference_seconds_pc_lanczos_gmres_threads": np.array([t_pc, t_l, t_g, os.cpu_count()])})

    store = dict(np.load(path)) if os.path.exists(path) else {}     # re-read: another run may have written meanwhile
    store.update(new)
    np.savez_compressed(path, **store)
    print("wrote", path, os.path.getsize(path), "bytes; keys:", sorted(store))


if __name__ == "__main__":
    main()
