"""The drop-in, executed: the UNMODIFIED reference driver (multi_bodies/multi_bodies.py, imported from the staged copy
oracle/_ref with the two harness stubs of SURVEY 8c) runs its own mobility problem -- set_mobility_vector_prod ->
linear_operator_rigid (multi_bodies.py:233-287, 424-471) -> build_block_diagonal_preconditioner (:752) -> SciPy GMRES
driven as general_application_utils.gmres does -- once with its own implementations ('python' = dense NumPy,
'numba') and once with the B200 CUDA product selected by string after integration.register().  This is the snippet
of INTEGRATION.md section 2, as a test."""
import numpy as np
import pytest

import bench

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mbod():
    from oracle import make_ref
    if not make_ref.available():
        pytest.skip("oracle/_ref not staged (run __graft_entry__.build() in the build container)")
    m = make_ref.import_multi_bodies()
    from rigidmultiblobswall_b200 import integration
    integration.register(m)
    return m


def relerr(x, ref):
    return float(np.linalg.norm(np.asarray(x) - np.asarray(ref)) / np.linalg.norm(ref))


@pytest.mark.parametrize("n_bodies", [20, 100])
def test_reference_mobility_problem_with_the_b200_product_selected_by_string(mbod, n_bodies):
    from oracle import ref_driver
    from rigidmultiblobswall_b200 import mobility as mb
    ref_cfg, loc, quat, FT, a, eta, tol = bench.make_c2(n_bodies)
    # one body so close to the wall that its lowest blob sits at z = 0.4 a: the clamp + damping path runs
    from rigidmultiblobswall_b200 import rigid
    z3 = rigid.RigidBodies(ref_cfg, loc, quat).get_r_vectors().reshape(-1, 3)[45:60, 2].min().item()
    loc[3, 2] -= z3 - 0.4 * a
    assert mbod.set_mobility_vector_prod("b200") is mb.single_wall_mobility_trans_times_force_b200
    out = {}
    for impl in ("python", "numba", "b200"):
        out[impl] = ref_driver.mobility_problem(ref_cfg, loc, quat, FT, a, eta, tol, implementation=impl, mbod=mbod)
        assert out[impl]["info"] == 0
    nb3 = 3 * out["b200"]["n_blobs"]
    assert (out["b200"]["r_vectors"][:, 2] < a).any()
    for ref in ("python", "numba"):
        assert out["b200"]["iterations"] == out[ref]["iterations"]
        assert relerr(out["b200"]["solution"][nb3:], out[ref]["solution"][nb3:]) < 1e-8      # body velocities
        assert relerr(out["b200"]["solution"][:nb3], out[ref]["solution"][:nb3]) < 1e-7      # blob forces
    assert out["b200"]["residual"] < 10 * tol


def test_one_operator_application_matches_to_rounding(mbod):
    """linear_operator_rigid itself (no solver in between): 1e-12"""
    from oracle import ref_driver
    ref_cfg, loc, quat, FT, a, eta, tol = bench.make_c2(30)
    bodies = ref_driver.build_bodies(mbod, ref_cfg, loc, quat, a)
    Nblobs = sum(b.Nblobs for b in bodies)
    r_vectors = mbod.get_blobs_r_vectors(bodies, Nblobs)
    x = np.random.default_rng(0).standard_normal(3 * Nblobs + 6 * len(bodies))
    res = {}
    for impl in ("numba", "b200", "b200_no_wall", "numba_no_wall"):
        mbod.mobility_vector_prod = mbod.set_mobility_vector_prod(impl)
        res[impl] = mbod.linear_operator_rigid(x, bodies, [], r_vectors, eta, a, periodic_length=np.zeros(3))
    assert relerr(res["b200"], res["numba"]) < 1e-12
    assert relerr(res["b200_no_wall"], res["numba_no_wall"]) < 1e-12


def test_free_surface_and_radii_strings(mbod):
    """the reference's other selector strings (multi_bodies.py:258-287) have B200 twins"""
    from oracle import ref_driver
    ref_cfg, loc, quat, FT, a, eta, tol = bench.make_c2(12)
    bodies = ref_driver.build_bodies(mbod, ref_cfg, loc, quat, a)
    rng = np.random.default_rng(3)
    for b in bodies:                                             # body.py: blobs_radius, one entry per blob
        b.blobs_radius = a * (0.7 + 0.6 * rng.random(b.Nblobs))
    Nblobs = sum(b.Nblobs for b in bodies)
    r_vectors = mbod.get_blobs_r_vectors(bodies, Nblobs)
    f = rng.standard_normal(3 * Nblobs)
    L = np.array([0.0, 0.0, 0.0])
    for ours, theirs in (("b200_free_surface", "numba_free_surface"), ("radii_b200", "radii_numba"),
                         ("radii_b200_no_wall", "radii_numba_no_wall"),
                         ("radii_b200_free_surface", "radii_numba_free_surface")):
        u = mbod.set_mobility_vector_prod(ours, bodies=bodies)(r_vectors, f, eta, a, periodic_length=L)
        v = mbod.set_mobility_vector_prod(theirs, bodies=bodies)(r_vectors, f, eta, a, periodic_length=L)
        assert relerr(np.asarray(u).reshape(-1), np.asarray(v).reshape(-1)) < 1e-12, ours
    # strings the reference already knows are untouched
    assert mbod.set_mobility_vector_prod("numba") is mbod.mb.single_wall_mobility_trans_times_force_numba


def test_mixed_size_bodies_use_the_bucketed_preconditioner(mbod):
    """Bodies of 12, 15 and 162 blobs in one suspension: the reference factorises each body at its own size
    (multi_bodies.py:766-785); the device preconditioner buckets them by size instead of padding all to 162 blobs."""
    import os
    from oracle import ref_driver
    from rigidmultiblobswall_b200 import rigid
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "rigid_golden.npz"))
    kinds = [g["struct/boomerang_N_15"]] * 9 + [g["struct/shell_N_162"]] * 1 + [g["struct/shell_N_12"]] * 6
    rng = np.random.default_rng(8)
    order = rng.permutation(len(kinds))
    cfgs = [kinds[k] for k in order]
    nb = len(cfgs)
    loc = np.array([[6.0 * (k % 4), 6.0 * (k // 4), 3.0 + rng.random()] for k in range(nb)])
    quat = rng.standard_normal((nb, 4))
    quat /= np.linalg.norm(quat, axis=1)[:, None]
    a, eta, tol = 0.25, 1.0, 1e-8
    FT = rng.standard_normal(6 * nb)
    ref = ref_driver.mobility_problem(cfgs, loc, quat, FT, a, eta, tol, implementation="numba", mbod=mbod)
    bodies = rigid.RigidBodies(cfgs, loc, quat)
    r = bodies.get_r_vectors()
    PC = rigid.block_diagonal_preconditioner(bodies, r, eta, a)
    assert isinstance(PC, rigid.GroupedBlockDiagonalPreconditioner) and len(PC.groups) == 2
    padded = rigid.BlockDiagonalPreconditioner(bodies, r, eta, a)
    assert PC.padded_bytes() < 0.25 * 8 * nb * padded.m * padded.m
    x = torch.from_numpy(rng.standard_normal(3 * bodies.n_blobs + 6 * nb)).cuda()
    PC_ref = mbod.build_block_diagonal_preconditioner(ref["bodies"], [], ref["r_vectors"], ref["n_blobs"], eta, a,
                                                      step=0, update_PC=1)
    y_ref = PC_ref(x.cpu().numpy())
    assert relerr(PC(x).cpu().numpy(), y_ref) < 1e-9
    assert relerr(padded(x).cpu().numpy(), y_ref) < 1e-9
    w = x[:3 * bodies.n_blobs]
    assert relerr(PC.M_inv_mult(w).cpu().numpy(), padded.M_inv_mult(w).cpu().numpy()) < 1e-10
    assert relerr(PC.P_inv_mult(w).cpu().numpy(), padded.P_inv_mult(w).cpu().numpy()) < 1e-10
    mult = lambda v: v * 2.0
    assert relerr(PC.mobility_pc(w, mult).cpu().numpy(), padded.mobility_pc(w, mult).cpu().numpy()) < 1e-10
    vel, lam, its, info = rigid.solve_mobility_problem(bodies, FT, eta, a, tolerance=tol)
    assert info == 0 and abs(its - ref["iterations"]) <= 1
    assert relerr(vel.reshape(-1).cpu().numpy(), ref["solution"][3 * bodies.n_blobs:]) < 1e-6
