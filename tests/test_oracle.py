"""Pins oracle/mobility_oracle.c against the reference's numba outputs
(tests/golden/mf_golden.npz, made by tests/golden/make_golden.py)."""
import numpy as np
import pytest

import oracle


def _cases():
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "mf_golden.npz"))
    return [str(c) for c in g["cases"]]


def relerr(x, ref):
    return np.linalg.norm(x - ref) / np.linalg.norm(ref)


@pytest.mark.parametrize("case", _cases())
def test_oracle_matches_reference_numba(golden, case):
    cloud, lname, wall, blk = case.split("/")
    r, v = golden[f"{cloud}/r"], golden[f"{cloud}/v"]
    a, eta = golden[f"{cloud}/a_eta"]
    L = golden[f"{cloud}/L_{lname}"]
    in_plane = wall == "in_plane"
    out = oracle.mobility_product(blk, "single_wall" if in_plane else wall, r, v, eta, a, L, in_plane=in_plane)
    ref = golden[case]
    assert out.shape == ref.shape
    # numba is fastmath (not bit-reproducible); 1e-13 norm-wise is an order below the 1e-12 bar
    assert relerr(out, ref) < 1e-13


def test_oracle_target_subset(golden):
    r, v = golden["dense/r"], golden["dense/v"]
    a, eta = golden["dense/a_eta"]
    full = oracle.mobility_product("UF", "single_wall", r, v, eta, a)
    part = oracle.mobility_product("UF", "single_wall", r, v, eta, a, targets=(17, 60))
    assert np.array_equal(full[3 * 17:3 * 60], part)


@pytest.mark.parametrize("cloud", ["dense", "dilute", "c1"])
def test_vectorised_timing_variant_matches_reference_numba(golden, cloud):
    r, v = golden[f"{cloud}/r"], golden[f"{cloud}/v"]
    a, eta = golden[f"{cloud}/a_eta"]
    out = oracle.uf_wall_fast(r, v, eta, a)
    assert relerr(out, golden[f"{cloud}/open/single_wall/UF"]) < 1e-13
    part = oracle.uf_wall_fast(r, v, eta, a, targets=(5, 40))
    assert np.array_equal(part, out[15:120])


# ---- source-target kernels with per-blob radii (mobility_numba.py:1480, :1941) ----
def _st():
    import os
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "st_golden.npz"))


def _st_cases():
    return [str(c) for c in _st()["cases"]]


@pytest.mark.parametrize("case", _st_cases())
def test_oracle_source_target_matches_reference_numba(case):
    g = _st()
    name, lname, wall = case.split("/")
    src, trg, f = g[f"{name}/source"], g[f"{name}/target"], g[f"{name}/force"]
    rs, rt, eta, L = g[f"{name}/radius_source"], g[f"{name}/radius_target"], float(g[f"{name}/eta"]), g[f"{name}/L_{lname}"]
    if wall == "radii_single_wall":
        out = oracle.source_target("single_wall", src, src, f, rs, rs, eta, L)
    else:
        out = oracle.source_target(wall, src, trg, f, rs, rt, eta, L)
    assert relerr(out, g[case]) < 1e-13
