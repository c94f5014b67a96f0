"""More invariants on both backends (CPU oracle in the `not gpu` suite, CUDA path in the `gpu` suite), same harness as
tests/test_invariants.py: the a -> 0 limit against Blake's point-force image system, the fused clamp + damping as
B M(r_eff) B also below the wall plane, and the columns of the product against the reference's DENSE matrices."""
import os

import numpy as np
import pytest

from test_invariants import BLOCKS, dense, product  # noqa: F401  (product: the backend fixture)

HERE = os.path.dirname(os.path.abspath(__file__))


def _blake(x, y, eta):
    """Velocity at x from a point force at y above a no-slip wall z = 0 (Blake 1971): stokeslet, image stokeslet,
    Stokes doublet and potential dipole, written out from G = S(r) - S(R) + 2h D_jk d/dR_k [h R_i/R^3 - S_i3(R)],
    D = diag(1, 1, -1).  Independent of the Swan-Brady expressions the kernels evaluate."""
    h = y[2]
    r, R = x - y, x - np.array([y[0], y[1], -h])
    rn, Rn = np.linalg.norm(r), np.linalg.norm(R)
    G = np.eye(3) / rn + np.outer(r, r) / rn ** 3 - np.eye(3) / Rn - np.outer(R, R) / Rn ** 3
    sign = (1.0, 1.0, -1.0)
    for j in range(3):
        for i in range(3):
            t = h * ((i == j) / Rn ** 3 - 3 * R[i] * R[j] / Rn ** 5) + (i == 2) * R[j] / Rn ** 3 \
                - ((i == j) * R[2] + R[i] * (j == 2)) / Rn ** 3 + 3 * R[i] * R[j] * R[2] / Rn ** 5
            G[i, j] += 2 * h * sign[j] * t
    return G / (8 * np.pi * eta)


def test_finite_size_limit_is_the_point_force_image_system(product):
    """mobility_test.py:70-82 (finite-size mobility -> singular image system as a -> 0), against Blake's tensor and,
    without the wall, the Oseen tensor; the O(a^2 / r^2) corrections are ~1e-13 here"""
    rng = np.random.default_rng(16)
    n, a, eta = 4, 1e-6, 1.3
    r = rng.normal(5.0, 1.0, (n, 3))
    mw = dense(product, "UF", "single_wall", r, eta, a)
    mf = dense(product, "UF", "no_wall", r, eta, a)
    for i in range(n):
        for j in range(n):
            if i == j:
                continue
            g = _blake(r[i], r[j], eta)
            assert np.abs(mw[3 * i:3 * i + 3, 3 * j:3 * j + 3] - g).max() < 1e-10 * np.abs(g).max()
            d = r[i] - r[j]
            dn = np.linalg.norm(d)
            oseen = (np.eye(3) / dn + np.outer(d, d) / dn ** 3) / (8 * np.pi * eta)
            assert np.abs(mf[3 * i:3 * i + 3, 3 * j:3 * j + 3] - oseen).max() < 1e-10 * np.abs(oseen).max()


@pytest.mark.parametrize("block", BLOCKS)
def test_clamp_and_damping_are_B_M_B_also_below_the_wall_plane(product, block):
    """mobility.py:52-84, 1152-1162: the wall forms compute B M(r_eff) (B v) with z_eff = max(z, a) and B_ii = z_i / a
    for z_i < a -- also for z <= 0, where B <= 0 and the reference still 'works' (SURVEY 8b).  The fused clamp and
    damping of the product must equal the explicit three steps on already-clamped positions."""
    rng = np.random.default_rng(17)
    n, a, eta = 200, 0.13, 7.0
    r = 5 * a * rng.random((n, 3))
    r[:, 2] -= 1.5 * a                                   # ~30 % of the blobs have z < 0, ~50 % z < a
    assert (r[:, 2] < 0).sum() > 20
    v = rng.standard_normal((n, 3))
    B = np.where(r[:, 2] < a, r[:, 2] / a, 1.0)
    r_eff = r.copy()
    r_eff[:, 2] = np.maximum(r[:, 2], a)
    fused = product(block, "single_wall", r, v, eta, a)
    steps = np.repeat(B, 3) * product(block, "single_wall", r_eff, v * B[:, None], eta, a)
    assert np.linalg.norm(fused - steps) <= 1e-13 * np.linalg.norm(steps)


@pytest.mark.parametrize("wall", ["single_wall", "no_wall"])
def test_columns_of_the_product_are_the_reference_dense_matrix(product, wall):
    """mobility.single_wall_fluid_mobility / rotne_prager_tensor (mobility.py:967-1116, an independent code path of the
    reference: dense loops, not the numba kernels) on a cloud with blobs below z = a: tests/golden/dense_golden.npz"""
    g = np.load(os.path.join(HERE, "golden", "dense_golden.npz"))
    eta, a = g["eta_a"]
    m = dense(product, "UF", wall, g["r"], eta, a)
    assert np.linalg.norm(m - g[wall]) <= 1e-12 * np.linalg.norm(g[wall])
