"""GPU tests of the sharded resistance solve on one rank and of the dense-mobility drop-ins
(multi_bodies.set_mobility_blobs 'b200' / 'b200_no_wall'); fixtures shared with tests/test_schemes_gpu.py."""
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")
from test_schemes_gpu import gold, relerr, setup  # noqa: E402,F401

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def test_sharded_resistance_single_rank(gold, setup):
    """parallel.ShardedRigidSystem.solve_resistance on one rank against the reference's dense answer (the world-2
    plumbing is covered on CPU by tests/test_parallel_gloo.py)"""
    from rigidmultiblobswall_b200 import parallel
    rg, g = gold
    _, _, _, eta, a = setup
    system = parallel.ShardedRigidSystem(rg["struct/boomerang_N_15"], rg["rigid/location"], rg["rigid/quaternion"],
                                         eta, a, group=False)
    slip = torch.from_numpy(g["slip"]).cuda().reshape(-1)
    force, lam, its, info = system.solve_resistance(system.local_slice(g["velocity"].reshape(12, 6)), tolerance=1e-10,
                                                    slip_loc=slip)
    assert info == 0 and tuple(force.shape) == (12, 6)
    assert relerr(lam, g["resistance/slip/force_blobs"]) < 1e-8
    assert relerr(force, g["resistance/slip/force"]) < 1e-8


def test_dense_mobility_drop_ins(gold):
    """single_wall_fluid_mobility_b200 / rotne_prager_tensor_b200 (multi_bodies.set_mobility_blobs 'b200',
    'b200_no_wall') against the reference's dense matrices, NumPy and torch inputs, and against the product"""
    from rigidmultiblobswall_b200 import mobility as mb
    rg, _ = gold
    g = np.load(os.path.join(HERE, "golden", "dense_golden.npz"))
    eta, a = g["eta_a"]
    Mw = mb.single_wall_fluid_mobility_b200(g["r"], eta, a)
    Mf = mb.rotne_prager_tensor_b200(g["r"], eta, a)
    assert isinstance(Mw, np.ndarray) and Mw.shape == (120, 120)
    assert relerr(Mw, g["single_wall"]) < 1e-12 and relerr(Mf, g["no_wall"]) < 1e-12
    rt = torch.from_numpy(g["r"]).cuda()
    Mt = mb.single_wall_fluid_mobility_b200(rt, eta, a)
    assert Mt.is_cuda and relerr(Mt, g["single_wall"]) < 1e-12
    f = torch.randn(120, dtype=torch.float64, device="cuda", generator=torch.Generator(device="cuda").manual_seed(7))
    assert relerr(Mt @ f, mb.single_wall_mobility_trans_times_force_b200(rt, f, eta, a).cpu().numpy()) < 1e-13
    # one body of the rigid golden set: the same matrix the preconditioner build assembles
    a2, eta2, _ = rg["rigid/a_eta_tol"]
    assert relerr(mb.single_wall_fluid_mobility_b200(rg["rigid/r_vectors"][:15], eta2, a2), rg["rigid/M_body0"]) < 1e-12
    assert mb.rotne_prager_tensor_b200(np.zeros((0, 3)), eta, a).shape == (0, 0)
