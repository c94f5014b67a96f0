"""Generate golden vectors from the REFERENCE's own numba implementation.

Run in the build container only (needs /root/reference and numba):

    python tests/golden/make_golden.py

It imports /root/reference/mobility in place (nothing is copied), calls the
public wrappers of mobility/mobility.py (the drop-in boundary,
mobility.py:1119-1341) on seeded synthetic inputs and stores inputs + outputs
in tests/golden/mf_golden.npz.  The GPU box has no /root/reference, so the
tests read only the committed .npz.

Python 3.12 removed the stdlib module `imp` that mobility.py:7 imports; a
harness-side stub is registered first.
"""
import importlib.util
import os
import sys
import types

import numpy as np

REF = "/root/reference"


def import_reference_mobility():
    imp = types.ModuleType("imp")

    def find_module(name, path=None):
        if importlib.util.find_spec(name) is None:
            raise ImportError(name)
        return (None, None, None)

    imp.find_module = find_module
    sys.modules.setdefault("imp", imp)
    sys.path.insert(0, os.path.join(REF, "mobility"))
    import mobility as mb  # /root/reference/mobility/mobility.py; picks up mobility_numba (:33-37)
    return mb


BLOCKS = {
    "UF": "trans_times_force",
    "UT": "trans_times_torque",
    "WF": "rot_times_force",
    "WT": "rot_times_torque",
}


def clouds(rng):
    """name -> (r, a, eta).  'dense' is mobility/test_blobs.py:30-34 (z in
    [0,5a): blobs below z=a and overlapping pairs); 'dilute' is the 10% volume
    fraction cloud of SURVEY 8d (all z >= a, almost all far pairs)."""
    out = {}
    a, eta = 0.13, 7.0
    n = 160
    out["dense"] = (5 * a * rng.random((n, 3)), a, eta)
    n = 200
    side = (n * 4 * np.pi / 3 / 0.1) ** (1.0 / 3.0) * a
    r = side * rng.random((n, 3))
    r[:, 2] += a
    out["dilute"] = (r, a, eta)
    return out


def main():
    mb = import_reference_mobility()
    rng = np.random.default_rng(20261104)
    store = {}
    cases = []
    cl = clouds(rng)
    for cname, (r, a, eta) in cl.items():
        n = r.shape[0]
        span = r.max(axis=0) - r.min(axis=0)
        v = rng.standard_normal((n, 3))
        store[f"{cname}/r"] = r
        store[f"{cname}/v"] = v
        store[f"{cname}/a_eta"] = np.array([a, eta])
        # L chosen so that the minimum-image wrap is active for many pairs
        Ls = {
            "open": np.zeros(3),
            "xy": np.array([1.3 * span[0], 1.6 * span[1], 0.0]),
            "xyz": np.array([1.3 * span[0], 1.6 * span[1], 2.5 * span[2]]),
            "x": np.array([1.2 * span[0], 0.0, 0.0]),
        }
        for lname, L in Ls.items():
            store[f"{cname}/L_{lname}"] = L
            for blk, stem in BLOCKS.items():
                for wall in ("single_wall", "no_wall"):
                    fn = getattr(mb, f"{wall}_mobility_{stem}_numba")
                    out = fn(r, v, eta, a, periodic_length=L)
                    key = f"{cname}/{lname}/{wall}/{blk}"
                    store[key] = np.asarray(out)
                    cases.append(key)
            for blk, stem in (("UF", "trans_times_force"), ("UT", "trans_times_torque")):
                fn = getattr(mb, f"in_plane_mobility_{stem}_numba")
                out = fn(r, v, eta, a, periodic_length=L)
                key = f"{cname}/{lname}/in_plane/{blk}"
                store[key] = np.asarray(out)
                cases.append(key)

    # BASELINE config 1 word for word: N=1000, eta=7, a=0.13, r=5a*rand (test_blobs.py:30-44)
    n, a, eta = 1000, 0.13, 7.0
    r = 5 * a * rng.random((n, 3))
    f = rng.standard_normal((n, 3))
    store["c1/r"], store["c1/v"], store["c1/a_eta"] = r, f, np.array([a, eta])
    store["c1/L_open"] = np.zeros(3)
    store["c1/open/single_wall/UF"] = mb.single_wall_mobility_trans_times_force_numba(r, f, eta, a)
    cases.append("c1/open/single_wall/UF")
    # cross-check the numba path against the reference's dense matrix (mobility.py:1018)
    rr = mb.shift_heights(r[:200], a)
    dense = mb.single_wall_fluid_mobility(rr, eta, a)
    u_d = dense.dot(f[:200].flatten())
    u_n = mb.mobility_numba.single_wall_mobility_trans_times_force_numba(rr, f[:200], eta, a, np.zeros(3))
    err = np.linalg.norm(u_d - u_n) / np.linalg.norm(u_d)
    print("numba vs dense single_wall_fluid_mobility:", err)
    assert err < 1e-13

    # free surface (mobility.py:1390 -> mobility_numba.py:1770): no clamp, all z > 0
    for cname in ("dense", "dilute"):
        r, a, eta = cl[cname]
        v = store[f"{cname}/v"]
        for lname in ("open", "xy"):
            L = store[f"{cname}/L_{lname}"]
            key = f"{cname}/{lname}/free_surface/UF"
            store[key] = np.asarray(mb.free_surface_mobility_trans_times_force_numba(r, v, eta, a, periodic_length=L))
            cases.append(key)

    store["cases"] = np.array(cases)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "mf_golden.npz")
    np.savez_compressed(path, **store)
    print("wrote", path, len(cases), "cases", os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
