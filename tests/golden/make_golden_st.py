"""Golden vectors for the source-target kernels with per-blob radii, from the REFERENCE's own numba code.

Run in the build container only (needs /root/reference and numba):

    python tests/golden/make_golden_st.py

Calls the public wrappers of /root/reference/mobility/mobility.py
  single_wall_mobility_trans_times_force_source_target_numba   (:551)
  no_wall_mobility_trans_times_force_source_target_numba       (:588)
  free_surface_mobility_trans_times_force_source_target_numba  (:1407)
  mobility_radii_trans_times_force                             (:1369)
on seeded inputs and stores inputs + outputs in tests/golden/st_golden.npz.
"""
import os

import numpy as np

from make_golden import import_reference_mobility


def main():
    mb = import_reference_mobility()
    rng = np.random.default_rng(20261105)
    store, cases = {}, []
    eta = 2.3

    def cloud(n, side, rmin, rmax, zlo):
        r = side * rng.random((n, 3))
        r[:, 2] = zlo + (side - zlo) * rng.random(n)
        return r, rmin + (rmax - rmin) * rng.random(n)

    # "mixed": radii spread over a factor 6 in a dense cloud, so that all three RPY branches (far, partial
    # overlap, complete overlap r <= |a-b|) occur, with blobs below their own radius (clamp + damping).
    # "dilute": almost only far pairs.  Targets are a different set than the sources.
    sets = {
        "mixed": (cloud(150, 1.0, 0.05, 0.30, 0.01), cloud(90, 1.0, 0.05, 0.30, 0.01)),
        "dilute": (cloud(170, 6.0, 0.10, 0.20, 0.25), cloud(110, 6.0, 0.10, 0.20, 0.25)),
    }
    for name, ((src, rs), (trg, rt)) in sets.items():
        f = rng.standard_normal((src.shape[0], 3))
        span = np.maximum(src.max(axis=0), trg.max(axis=0)) - np.minimum(src.min(axis=0), trg.min(axis=0))
        store[f"{name}/source"], store[f"{name}/target"], store[f"{name}/force"] = src, trg, f
        store[f"{name}/radius_source"], store[f"{name}/radius_target"] = rs, rt
        store[f"{name}/eta"] = np.array(eta)
        Ls = {"open": np.zeros(3), "xy": np.array([1.3 * span[0], 1.6 * span[1], 0.0]),
              "xyz": np.array([1.3 * span[0], 1.6 * span[1], 2.5 * span[2]])}
        for lname, L in Ls.items():
            store[f"{name}/L_{lname}"] = L
            for wall in ("single_wall", "no_wall", "free_surface"):
                fn = getattr(mb, f"{wall}_mobility_trans_times_force_source_target_numba")
                key = f"{name}/{lname}/{wall}"
                store[key] = np.asarray(fn(src, trg, f, rs, rt, eta, periodic_length=L))
                cases.append(key)
            # sources == targets: mobility_radii_trans_times_force (mobility.py:1369) with the self pairs inside
            key = f"{name}/{lname}/radii_single_wall"
            store[key] = np.asarray(mb.mobility_radii_trans_times_force(
                src, f, eta, 0.0, rs, mb.single_wall_mobility_trans_times_force_source_target_numba, periodic_length=L))
            cases.append(key)

    # equal radii, z >= a: must agree with the same-radius kernel (SURVEY 8d parity recipe)
    src, rs = sets["dilute"][0]
    a = 0.15
    f = store["dilute/force"]
    u_st = mb.mobility_radii_trans_times_force(src, f, eta, a, np.full(src.shape[0], a),
                                               mb.single_wall_mobility_trans_times_force_source_target_numba)
    u_eq = mb.single_wall_mobility_trans_times_force_numba(src, f, eta, a)
    err = np.linalg.norm(u_st - u_eq) / np.linalg.norm(u_eq)
    print("source-target (equal radii) vs same-radius numba kernel:", err)
    assert err < 1e-13

    store["cases"] = np.array(cases)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "st_golden.npz")
    np.savez_compressed(path, **store)
    print("wrote", path, len(cases), "cases", os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
