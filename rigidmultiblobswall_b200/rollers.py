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
