"""Config 3's caller, executed: the reference's own QuaternionIntegratorRollers.compute_deterministic_velocity_and_torque
(quaternion_integrator_rollers.py:905-979, from the staged copy oracle/_ref) with (a) its numba block table, (b) the
B200 block table it finds under the PyCUDA names after integration.register(alias_pycuda=True) (:59-91), and (c) the
device-resident form rollers.compute_deterministic_velocity_and_torque."""
import numpy as np
import pytest

import bench

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def relerr(x, ref):
    x = x.detach().cpu().numpy() if hasattr(x, "detach") else np.asarray(x)
    return float(np.linalg.norm(x.reshape(-1) - np.asarray(ref).reshape(-1)) / np.linalg.norm(ref))


@pytest.fixture(scope="module")
def mbod():
    from oracle import make_ref
    if not make_ref.available():
        pytest.skip("oracle/_ref not staged")
    m = make_ref.import_multi_bodies()
    from rigidmultiblobswall_b200 import integration
    integration.register(m, alias_pycuda=True)
    return m


@pytest.mark.parametrize("domain,periodic", [("single_wall", True), ("single_wall", False), ("no_wall", True)])
def test_deterministic_velocity_and_torque(mbod, domain, periodic):
    from oracle import ref_driver
    from rigidmultiblobswall_b200 import mobility as mb, rollers
    n = 3000
    r, f, _, a, eta, L = bench.make_c3(n)
    if not periodic:
        L = np.zeros(3)
    omega, tol = np.array([0.0, 62.8, 0.0]), 1e-8
    ref = ref_driver.rollers_integrator(mbod, r, f, omega, eta, a, L, tol, implementation="numba", domain=domain)
    v_ref, T_ref = ref.compute_deterministic_velocity_and_torque()
    # (b) the reference's own method on the B200 blocks (NumPy boundary, four separate block products)
    ours = ref_driver.rollers_integrator(mbod, r, f, omega, eta, a, L, tol, implementation="pycuda", domain=domain)
    assert ours.mobility_rot_times_torque is getattr(mb, "%s_mobility_rot_times_torque_b200" % domain)
    v_b, T_b = ours.compute_deterministic_velocity_and_torque()
    assert ours.det_iterations_count == ref.det_iterations_count
    assert relerr(T_b, T_ref) < 1e-7 and relerr(v_b, v_ref) < 1e-7
    # (c) device resident
    v, T, its, info = rollers.compute_deterministic_velocity_and_torque(r, f, omega, eta, a, periodic_length=L,
                                                                        tolerance=tol, domain=domain)
    assert info == 0 and abs(its - ref.det_iterations_count) <= 1
    assert relerr(T, T_ref) < 1e-6 and relerr(v, v_ref) < 1e-6
    # residual of the device solve with the REFERENCE's numba blocks
    res = omega.repeat(n).reshape(n, 3).T.reshape(-1) if False else np.tile(omega, n)
    res = res - ref.mobility_rot_times_force(r, f.reshape(-1), eta, a, periodic_length=L) \
        - ref.mobility_rot_times_torque(r, T.cpu().numpy(), eta, a, periodic_length=L)
    assert np.linalg.norm(res) / np.linalg.norm(np.tile(omega, n)) < 10 * tol
    # free kinematics (:969-971): given torques, one fused pass
    v2, T2, its2, _ = rollers.compute_deterministic_velocity_and_torque(r, f, omega, eta, a, periodic_length=L,
                                                                       domain=domain, torque=T_ref)
    assert its2 == 0 and relerr(v2, v_ref) < 1e-7


@pytest.mark.parametrize("free_kinematics", [False, True])
def test_stochastic_velocity(mbod, free_kinematics):
    """compute_stochastic_velocity (:1082-1200) with the same Gaussian draws as the reference takes from np.random:
    Lanczos on the 6N grand mobility, random-finite-difference divergences, GMRES for the stochastic torque."""
    from oracle import ref_driver
    from rigidmultiblobswall_b200 import rollers
    n, dt, kT, tol = 1000, 0.01, 0.0041, 1e-8
    r, f, _, a, eta, L = bench.make_c3(n)
    ref = ref_driver.rollers_integrator(mbod, r, f, np.zeros(3), eta, a, L, tol, implementation="numba")
    ref.kT, ref.rf_delta = kT, 1e-3
    ref.free_kinematics = "True" if free_kinematics else "False"
    np.random.seed(77)
    v_ref = ref.compute_stochastic_velocity(dt)
    np.random.seed(77)
    z, dx = np.random.randn(6 * n), np.random.randn(3 * n)          # the reference's two draws, in its order
    v, it_l, its = rollers.compute_stochastic_velocity(r, dt, kT, eta, a, periodic_length=L, tolerance=tol, rf_delta=1e-3,
                                                       free_kinematics=free_kinematics, z=z, dx_stoch=dx)
    assert abs(it_l - ref.stoch_iterations_count) <= 1
    assert free_kinematics or abs(its - ref.det_iterations_count) <= 1
    # Lanczos at tolerance 1e-8 and a finite difference of two O(1) products divided by rf_delta a = 1e-3
    assert relerr(v, v_ref) < 1e-5
    # the grand mobility itself against the reference's four numba blocks: rounding
    x = np.random.randn(6 * n)
    got = rollers.grand_mobility_product(torch.from_numpy(r).cuda().reshape(-1), torch.from_numpy(x).cuda(), eta, a, L)
    want = np.concatenate([
        ref.mobility_trans_times_force(r, x[:3 * n], eta, a, periodic_length=L)
        + ref.mobility_trans_times_torque(r, x[3 * n:], eta, a, periodic_length=L),
        ref.mobility_rot_times_force(r, x[:3 * n], eta, a, periodic_length=L)
        + ref.mobility_rot_times_torque(r, x[3 * n:], eta, a, periodic_length=L)])
    assert relerr(got, want) < 1e-12
