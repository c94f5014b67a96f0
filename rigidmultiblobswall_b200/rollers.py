"""Device-resident deterministic step of the microroller integrator (BASELINE config 3's caller).

Mirrors QuaternionIntegratorRollers.compute_deterministic_velocity_and_torque
(/root/reference/quaternion_integrator/quaternion_integrator_rollers.py:905-979): rollers (single blobs) rotate with
a prescribed angular velocity and feel forces; solve

    M_rr T = omega - M_rt f          GMRES, no preconditioner, RHS scaled to unit norm (:944-965)
    v      = M_tt f + M_tr T         (:977-979)

The reference calls four NumPy-boundary block products per GMRES iteration (the block table of :59-91); here positions,
vectors and the Krylov basis stay on the GPU, M_rr is the WT product, M_rt the WF product, and the last line is ONE
fused UF f + UT T pass (the PyCUDA-only kernel the reference has commented out at :1116).
"""
import numpy as np
import torch

from . import mobility as _mb
from . import solvers

_WALL = {"single_wall": _mb.SINGLE_WALL, "no_wall": _mb.NO_WALL}


def _dev(x, dev):
    if isinstance(x, torch.Tensor):
        return x.to(device=dev, dtype=torch.float64).reshape(-1).contiguous()
    return torch.from_numpy(np.ascontiguousarray(np.asarray(x, dtype=np.float64).reshape(-1))).to(dev)


def compute_deterministic_velocity_and_torque(r_vectors, force, omega, eta, a, periodic_length=None, tolerance=1e-8,
                                              x0=None, domain="single_wall", torque=None, restart=20, maxiter=1000,
                                              callback=None, device=None):
    """Returns (velocity [3N], torque [3N], iterations, info) as device tensors.

    r_vectors, force: (N,3) / flat, NumPy or torch.  omega: 3 numbers (every roller, get_omega_one_roller) or 3N.
    torque: given torques (`free_kinematics == 'True'`, :969-971): no solve, only the velocity.
    x0: first guess for the SCALED unknown, as the reference keeps it (deterministic_torque_previous_step, :962).
    restart = 20 is SciPy's default, which the reference's call inherits; maxiter counts restart cycles."""
    if domain not in _WALL:
        raise ValueError("domain must be 'single_wall' or 'no_wall' (the in_plane table has no rotation blocks)")
    wall = _WALL[domain]
    dev = device or (r_vectors.device if isinstance(r_vectors, torch.Tensor) else _mb._device())
    r, f = _dev(r_vectors, dev), _dev(force, dev)
    n = r.numel() // 3
    L = periodic_length
    its, info = 0, 0
    if torque is None:
        om = _dev(omega, dev)
        if om.numel() == 3:
            om = om.repeat(n)
        rhs = om - _mb.mobility_product(_mb.WF, wall, r, f, eta, a, periodic_length=L)
        nrm = float(torch.linalg.norm(rhs))
        if nrm > 0:
            rhs = rhs / nrm
        A = lambda t: _mb.mobility_product(_mb.WT, wall, r, t, eta, a, periodic_length=L)
        g = None if x0 is None else _dev(x0, dev)
        T, info, its = solvers.gmres(A, rhs, x0=g, tol=tolerance, restart=restart, maxiter=maxiter, callback=callback)
        scaled = T
        if nrm > 0:
            T = T * nrm
    else:
        T = _dev(torque, dev)
        scaled = T
    if bool(torch.any(T != 0)):
        v = _mb.mobility_product(_mb.UF_UT, wall, r, f, eta, a, vec_b=T, periodic_length=L)
    else:
        v = _mb.mobility_product(_mb.UF, wall, r, f, eta, a, periodic_length=L)
    compute_deterministic_velocity_and_torque.last_scaled_solution = scaled      # what the reference keeps as x0
    return v, T, its, info


def grand_mobility_product(r_vectors, force_torque, eta, a, periodic_length=None, domain="single_wall"):
    """N [f; t] = [M_tt f + M_tr t ; M_rt f + M_rr t] for single-blob rollers (the closure `grand_mobility_matrix` of
    quaternion_integrator_rollers.py:1114-1121, four block products per call there): the translational half is ONE
    fused UF f + UT t pass, the rotational half the WF and WT products; device tensors in and out."""
    wall = _WALL[domain]
    half = force_torque.numel() // 2
    f, t = force_torque[:half], force_torque[half:]
    v = _mb.mobility_product(_mb.UF_UT, wall, r_vectors, f, eta, a, vec_b=t, periodic_length=periodic_length)
    w = _mb.mobility_product(_mb.WF, wall, r_vectors, f, eta, a, periodic_length=periodic_length)
    w = w + _mb.mobility_product(_mb.WT, wall, r_vectors, t, eta, a, periodic_length=periodic_length)
    return torch.cat([v, w])


def compute_stochastic_velocity(r_vectors, dt, kT, eta, a, periodic_length=None, tolerance=1e-8, rf_delta=1e-3,
                                domain="single_wall", free_kinematics=False, z=None, dx_stoch=None, generator=None,
                                restart=20, maxiter=1000, device=None):
    """Device-resident QuaternionIntegratorRollers.compute_stochastic_velocity (quaternion_integrator_rollers.py:
    1082-1200): noise = sqrt(2 kT / dt) N^{1/2} z by Lanczos on the 6N grand mobility (:1130-1136), the divergence
    terms div_t(M_rt), div_t(M_tt) by a centred random finite difference with displacement rf_delta a dx (:1139-1160,
    skipped without a wall or at kT = 0), then, for prescribed angular velocities, the torque from
    M_rr T = -noise_r - kT div_t(M_rt) / (rf_delta a) (GMRES, unit-norm RHS, :1165-1190), and
    v_stoch = M_tr T + noise_t + kT div_t(M_tt) / (rf_delta a) (:1197-1199).
    z (6N) and dx_stoch (3N) are the two Gaussian draws the reference takes from np.random, in that order; given here
    explicitly (or drawn on the device) so that a run can be reproduced.
    Returns (v_stoch [3N], lanczos_iterations, gmres_iterations)."""
    if domain not in _WALL:
        raise ValueError("domain must be 'single_wall' or 'no_wall'")
    wall = _WALL[domain]
    dev = device or (r_vectors.device if isinstance(r_vectors, torch.Tensor) else _mb._device())
    r = _dev(r_vectors, dev)
    n = r.numel() // 3
    L = periodic_length
    z = torch.randn(6 * n, dtype=torch.float64, device=dev, generator=generator) if z is None else _dev(z, dev)
    noise, it_l = solvers.lanczos(factor=float(np.sqrt(2.0 * kT / dt)), tolerance=tolerance, dim=6 * n, z=z,
                                  mobility_mult=lambda x: grand_mobility_product(r, x, eta, a, L, domain))
    noise_t, noise_r = noise[:3 * n], noise[3 * n:]
    if kT > 0.0 and domain != "no_wall":
        dx = (torch.randn(3 * n, dtype=torch.float64, device=dev, generator=generator) if dx_stoch is None
              else _dev(dx_stoch, dev))
        rp = r + dx * (rf_delta * a * 0.5)
        rm = rp - dx * (rf_delta * a)                          # the reference displaces back from the displaced array
        div_rt = (_mb.mobility_product(_mb.WF, wall, rp, dx, eta, a, periodic_length=L)
                  - _mb.mobility_product(_mb.WF, wall, rm, dx, eta, a, periodic_length=L))
        div_tt = (_mb.mobility_product(_mb.UF, wall, rp, dx, eta, a, periodic_length=L)
                  - _mb.mobility_product(_mb.UF, wall, rm, dx, eta, a, periodic_length=L))
        r = rm + dx * (rf_delta * a * 0.5)                     # :1160 -- the positions the solve below uses
    else:
        div_rt = torch.zeros(3 * n, dtype=torch.float64, device=dev)
        div_tt = torch.zeros(3 * n, dtype=torch.float64, device=dev)
    scale = kT / (rf_delta * a)
    its = 0
    if not free_kinematics:
        rhs = -noise_r - div_rt * scale
        nrm = float(torch.linalg.norm(rhs))
        if nrm > 0:
            rhs = rhs / nrm
        A = lambda t: _mb.mobility_product(_mb.WT, wall, r, t, eta, a, periodic_length=L)
        T, info, its = solvers.gmres(A, rhs, tol=tolerance, restart=restart, maxiter=maxiter)
        if info != 0:
            raise RuntimeError("compute_stochastic_velocity: GMRES did not converge (%d iterations)" % its)
        if nrm > 0:
            T = T * nrm
        v = _mb.mobility_product(_mb.UT, wall, r, T, eta, a, periodic_length=L)
    else:
        v = torch.zeros(3 * n, dtype=torch.float64, device=dev)
    return v + noise_t + scale * div_tt, it_l, its
