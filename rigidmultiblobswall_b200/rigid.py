"""Rigid-body operators on the device: K, K^T, the saddle-point operator of
the mobility problem and the block-diagonal preconditioner.

Mirrors (same names, same argument meaning) the module-level functions of
/root/reference/multi_bodies/multi_bodies.py that run inside every GMRES
iteration as interpreted loops over bodies:

    K_matrix_vector_prod              multi_bodies.py:328-347
    K_matrix_T_vector_prod            multi_bodies.py:350-371
    linear_operator_rigid             multi_bodies.py:424-471  (free and prescribed-kinematics bodies; no articulated constraints)
    build_block_diagonal_preconditioner  multi_bodies.py:752-903 (same)
    solve_mobility_problem / solve_resistance_problem / body_mobility
                                      the `mobility`, `resistance` and `body_mobility` schemes of
                                      multi_bodies_utilities.py:449-540, :549-574, :576-600
    get_blobs_r_vectors               multi_bodies.py:196-208 / body.py:64-80

The `bodies` argument is a RigidBodies object (the device-resident form of
the reference's list of body.Body); vectors are torch CUDA float64 tensors.
"""
import ctypes

import numpy as np
import torch

from . import _lib, mobility


def _rlib():
    lib = _lib.load()
    if not getattr(lib, "_rigid_bound", False):
        vp, cl, cd, ci = ctypes.c_void_p, ctypes.c_long, ctypes.c_double, ctypes.c_int
        lib.rmb_rigid_last_error.restype = ctypes.c_char_p
        lib.rmb_blobs_from_bodies.argtypes = [cl, vp, vp, vp, vp, vp, vp]
        lib.rmb_K_apply.argtypes = [cl, vp, vp, vp, vp, cd, cd, vp, vp]
        lib.rmb_KT_apply.argtypes = [cl, vp, vp, vp, vp, cd, vp, vp]
        lib.rmb_body_mobility_blocks.argtypes = [cl, vp, ci, vp, cd, cd, ci, vp, vp]
        lib._rigid_bound = True
    return lib


def _check(rc):
    if rc != 0:
        raise _lib.RmbError("librmb_mobility (rigid): error %d: %s" % (rc, _rlib().rmb_rigid_last_error().decode()))


def _stream(dev):
    return torch.cuda.current_stream(dev).cuda_stream


class RigidBodies:
    """Device-resident suspension: for every body its location, quaternion
    (s, p1, p2, p3) and reference configuration (body/body.py:17-40).  Body b
    owns the contiguous blob range [body_offset[b], body_offset[b+1])."""

    def __init__(self, reference_configurations, locations, quaternions, device=None, prescribed_velocity=None):
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.device = dev
        if isinstance(reference_configurations, np.ndarray) and reference_configurations.ndim == 2:
            reference_configurations = [reference_configurations] * len(locations)
        counts = np.array([len(c) for c in reference_configurations], dtype=np.int64)
        self.n_bodies = len(counts)
        self.n_blobs = int(counts.sum())
        self.nb_max = int(counts.max()) if self.n_bodies else 0
        self.uniform = bool(self.n_bodies and (counts == counts[0]).all())
        off = np.zeros(self.n_bodies + 1, dtype=np.int64)
        np.cumsum(counts, out=off[1:])
        self.counts = counts
        self.body_offset_host = off
        self.body_offset = torch.from_numpy(off).to(dev)
        self.blob_body = torch.from_numpy(np.repeat(np.arange(self.n_bodies, dtype=np.int32), counts)).to(dev)
        ref = np.concatenate([np.asarray(c, dtype=np.float64).reshape(-1, 3) for c in reference_configurations]) \
            if self.n_bodies else np.zeros((0, 3))
        self.reference_configuration = torch.from_numpy(np.ascontiguousarray(ref)).to(dev)
        self.location = torch.as_tensor(np.asarray(locations, dtype=np.float64).reshape(-1, 3)).to(dev).contiguous()
        self.quaternion = torch.as_tensor(np.asarray(quaternions, dtype=np.float64).reshape(-1, 4)).to(dev).contiguous()
        self.set_prescribed(prescribed_velocity)

    def set_prescribed(self, prescribed_velocity):
        """Bodies with prescribed kinematics (body.py:58-59, :150): prescribed_velocity is None (all bodies free) or
        a sequence with one entry per body, None for a free body, else its 6 prescribed components (U, omega)."""
        nb = self.n_bodies
        mask = np.zeros(nb, dtype=bool)
        vel = np.zeros((nb, 6))
        if prescribed_velocity is not None:
            if len(prescribed_velocity) != nb:
                raise ValueError("prescribed_velocity needs one entry (None or 6 numbers) per body")
            for k, v in enumerate(prescribed_velocity):
                if v is not None:
                    mask[k] = True
                    vel[k] = np.asarray(v, dtype=np.float64).reshape(6)
        self.any_prescribed = bool(mask.any())
        self.prescribed = torch.from_numpy(mask).to(self.device)
        self.prescribed_velocity = torch.from_numpy(vel).to(self.device)

    @classmethod
    def from_bodies(cls, bodies, device=None):
        """From the reference's own list of body.Body objects."""
        presc = [b.calc_prescribed_velocity() if getattr(b, "prescribed_kinematics", False) is True else None
                 for b in bodies]
        return cls([b.reference_configuration for b in bodies], [b.location for b in bodies],
                   [b.orientation.entries for b in bodies], device=device,
                   prescribed_velocity=presc if any(p is not None for p in presc) else None)

    def get_r_vectors(self):
        """multi_bodies.get_blobs_r_vectors (multi_bodies.py:196-208): flat 3*Nblobs tensor."""
        r = torch.empty(3 * self.n_blobs, dtype=torch.float64, device=self.device)
        _check(_rlib().rmb_blobs_from_bodies(self.n_blobs, self.blob_body.data_ptr(),
                                             self.reference_configuration.data_ptr(), self.location.data_ptr(),
                                             self.quaternion.data_ptr(), r.data_ptr(), _stream(self.device)))
        return r


def get_blobs_r_vectors(bodies, Nblobs=None):
    return bodies.get_r_vectors()


def K_matrix_vector_prod(bodies, vector, Nblobs=None, K_bodies=None, r_vectors=None, out=None, alpha=1.0, beta=0.0):
    """K * vector: (U_b, omega_b) per body -> blob velocities U_b + omega_b x (r_i - q_b); returns (Nblobs, 3).
    out/alpha/beta give the fused update out = beta*out + alpha*K*vector."""
    r = bodies.get_r_vectors() if r_vectors is None else r_vectors
    U = vector.reshape(-1).contiguous()
    if out is None:
        out = torch.empty(3 * bodies.n_blobs, dtype=torch.float64, device=bodies.device)
        beta = 0.0
    _check(_rlib().rmb_K_apply(bodies.n_blobs, bodies.blob_body.data_ptr(), r.data_ptr(), bodies.location.data_ptr(),
                               U.data_ptr(), float(alpha), float(beta), out.data_ptr(), _stream(bodies.device)))
    return out.view(-1, 3)


def K_matrix_T_vector_prod(bodies, vector, Nblobs=None, K_bodies=None, r_vectors=None, alpha=1.0):
    """K^T * vector: blob forces -> (force, torque) per body; returns (2*Nbodies, 3) like the reference."""
    r = bodies.get_r_vectors() if r_vectors is None else r_vectors
    lam = vector.reshape(-1).contiguous()
    out = torch.empty(6 * bodies.n_bodies, dtype=torch.float64, device=bodies.device)
    _check(_rlib().rmb_KT_apply(bodies.n_bodies, bodies.body_offset.data_ptr(), r.data_ptr(),
                                bodies.location.data_ptr(), lam.data_ptr(), float(alpha), out.data_ptr(),
                                _stream(bodies.device)))
    return out.view(-1, 3)


def linear_operator_rigid(vector, bodies, r_vectors, eta, a, mobility_vector_prod=None, *args, **kwargs):
    """Action of  |  M   -K | on [lambda; U]  (multi_bodies.py:424-471, no articulated constraints).
                  | -K^T  0 |
    For a body with prescribed kinematics (:458-464) the unknown in its U slot is the force-torque it needs: its
    K U term drops out of the blob rows and its body rows read -K^T lambda + F.
    mobility_vector_prod defaults to the single-wall B200 product."""
    mvp = mobility.single_wall_mobility_trans_times_force_b200 if mobility_vector_prod is None else mobility_vector_prod
    n3 = 3 * bodies.n_blobs
    res = torch.empty_like(vector)
    lam, U = vector[:n3], vector[n3:]
    res[:n3] = mvp(r_vectors, lam, eta, a, *args, **kwargs)
    if bodies.any_prescribed:
        pm = bodies.prescribed.unsqueeze(1)
        U2 = U.reshape(-1, 6)
        K_matrix_vector_prod(bodies, torch.where(pm, torch.zeros_like(U2), U2), r_vectors=r_vectors, out=res[:n3],
                             alpha=-1.0, beta=1.0)
        ktl = K_matrix_T_vector_prod(bodies, lam, r_vectors=r_vectors, alpha=-1.0).reshape(-1, 6)
        res[n3:] = torch.where(pm, ktl + U2, ktl).reshape(-1)
        return res
    K_matrix_vector_prod(bodies, U, r_vectors=r_vectors, out=res[:n3], alpha=-1.0, beta=1.0)
    res[n3:] = K_matrix_T_vector_prod(bodies, lam, r_vectors=r_vectors, alpha=-1.0).reshape(-1)
    return res


class BlockDiagonalPreconditioner:
    """Block-diagonal preconditioner of the mobility problem on the device.

    build (multi_bodies.py:766-785), per body b:
        M_b  dense blob mobility of the body alone (rmb_body_mobility_blocks)
        M_b^-1  via batched Cholesky (torch.linalg.cholesky + cholesky_inverse: plain library calls)
        N_b = pinv(K_b^T M_b^-1 K_b)   (6x6)
    apply (multi_bodies.py:825-862), per body:
        Lambda~ = M_b^-1 slip;  U = N_b (-F - K^T Lambda~);  Lambda = M_b^-1 (slip + K U)
        (prescribed kinematics: Lambda = Lambda~, and the body slot returns the force K^T Lambda)
    Bodies with fewer blobs than nb_max are identity-padded (the padding rows
    never couple to real ones)."""

    def __init__(self, bodies, r_vectors, eta, a, wall=True):
        self.bodies = bodies
        dev = bodies.device
        nb, m = bodies.n_bodies, 3 * bodies.nb_max
        self.m = m
        M = torch.empty(nb, m, m, dtype=torch.float64, device=dev)
        r = r_vectors.reshape(-1).contiguous()
        lib = _rlib()
        for b0 in range(0, nb, 65535):
            b1 = min(nb, b0 + 65535)
            _check(lib.rmb_body_mobility_blocks(b1 - b0, bodies.body_offset[b0:].data_ptr(), bodies.nb_max,
                                                r.data_ptr(), float(eta), float(a), int(bool(wall)),
                                                M[b0:].data_ptr(), _stream(dev)))
        self.M_blobs = M
        self.chol = torch.linalg.cholesky(M)
        self.M_inv = torch.cholesky_inverse(self.chol)
        # padded gather/scatter maps between the flat blob vector and [nb, 3 nb_max]
        if bodies.uniform:
            self.pad_index = None
        else:
            idx = torch.full((nb, m), 3 * bodies.n_blobs, dtype=torch.int64, device=dev)   # -> appended zero
            off, cnt = bodies.body_offset_host, bodies.counts
            rows = np.repeat(np.arange(nb), 3 * cnt)
            cols = np.concatenate([np.arange(3 * c) for c in cnt])
            idx[torch.from_numpy(rows).to(dev), torch.from_numpy(cols).to(dev)] = torch.arange(
                3 * bodies.n_blobs, device=dev)
            self.pad_index = idx
            self.valid = idx < 3 * bodies.n_blobs
        # dense K per body [nb, 3 nb_max, 6] (columns = K applied to unit rigid motions)
        K = torch.zeros(nb, m, 6, dtype=torch.float64, device=dev)
        for c in range(6):
            U = torch.zeros(nb, 6, dtype=torch.float64, device=dev)
            U[:, c] = 1.0
            K[:, :, c] = self._pad(K_matrix_vector_prod(bodies, U, r_vectors=r).reshape(-1))
        self.K = K
        self.N = torch.linalg.pinv(K.transpose(1, 2) @ self.M_inv @ K, hermitian=True)

    def _pad(self, flat):
        if self.pad_index is None:
            return flat.reshape(self.bodies.n_bodies, self.m)
        ext = torch.cat([flat, flat.new_zeros(1)])
        return ext[self.pad_index]

    def _unpad(self, padded):
        if self.pad_index is None:
            return padded.reshape(-1)
        return padded[self.valid]

    def __call__(self, vector):
        nb, n3 = self.bodies.n_bodies, 3 * self.bodies.n_blobs
        slip = self._pad(vector[:n3]).unsqueeze(2)                 # [nb, m, 1]
        F = vector[n3:].reshape(nb, 6, 1)
        lam_t = self.M_inv @ slip
        KT_lam_t = self.K.transpose(1, 2) @ lam_t
        U = self.N @ (-F - KT_lam_t)
        lam = self.M_inv @ (slip + self.K @ U)
        if self.bodies.any_prescribed:          # multi_bodies.py:850-862: Lambda = M_b^-1 (slip + K U), F = K^T Lambda
            pm = self.bodies.prescribed.reshape(nb, 1, 1)
            lam = torch.where(pm, lam_t, lam)
            U = torch.where(pm, KT_lam_t, U)
        out = torch.empty_like(vector)
        out[:n3] = self._unpad(lam.squeeze(2))
        out[n3:] = U.reshape(-1)
        return out


    # ---- stochastic preconditioner (multi_bodies.py:480-618): M_b = Lc Lc^T, P = Lc^-T, P^-1 = Lc^T ... the
    # reference stores P = inv(U) with M = U^T U (U = Lc^T): mobility_pc = P^T M P, P_inv_mult = U^T = Lc.
    def _ensure_stoch(self):
        if getattr(self, "L_inv", None) is None:
            eye = torch.eye(self.m, dtype=torch.float64, device=self.bodies.device).expand(self.bodies.n_bodies, -1, -1)
            self.L_inv = torch.linalg.solve_triangular(self.chol, eye, upper=False)

    def _P_mult(self, w):
        """P w = Lc^-T w"""
        self._ensure_stoch()
        return self._unpad((self.L_inv.transpose(1, 2) @ self._pad(w.reshape(-1)).unsqueeze(2)).squeeze(2))

    def _PT_mult(self, w):
        """P^T w = Lc^-1 w"""
        self._ensure_stoch()
        return self._unpad((self.L_inv @ self._pad(w.reshape(-1)).unsqueeze(2)).squeeze(2))

    def mobility_pc(self, w, mobility_mult):
        """y = P^T M P w with the block-diagonal P = Lc^-T; mobility_mult(v) applies the full blob mobility."""
        return self._PT_mult(mobility_mult(self._P_mult(w)).reshape(-1))

    def P_inv_mult(self, w):
        """y = P^-1 w = Lc w (maps the preconditioned Lanczos noise back, stochastic_forcing.py:256)."""
        return self._unpad((self.chol @ self._pad(w.reshape(-1)).unsqueeze(2)).squeeze(2))

    def M_inv_mult(self, w):
        """y = blockdiag(M_b^-1) w: the blob-level preconditioner of the resistance problem M lambda = rhs."""
        return self._unpad((self.M_inv @ self._pad(w.reshape(-1)).unsqueeze(2)).squeeze(2))


class GroupedBlockDiagonalPreconditioner:
    """The same preconditioner for suspensions whose bodies differ widely in size.  BlockDiagonalPreconditioner pads
    every body to the largest one ([n_bodies, 3 nb_max, 3 nb_max] per factor): one 2,000-blob body among 1,000
    15-blob bodies would cost 288 GB per factor and O(n_bodies nb_max^3) Cholesky work, where the reference
    factorises each body at its own size (multi_bodies.py:766-785).  Here the bodies are bucketed by blob count
    (a bucket spans at most a factor 1.25 in size), each bucket is one batched BlockDiagonalPreconditioner over its
    own sub-suspension, and vectors are gathered / scattered by index."""

    SPREAD = 1.25

    def __init__(self, bodies, r_vectors, eta, a, wall=True):
        self.bodies = bodies
        dev = bodies.device
        counts, off = bodies.counts, bodies.body_offset_host
        ref = bodies.reference_configuration.cpu().numpy()
        loc, quat = bodies.location.cpu().numpy(), bodies.quaternion.cpu().numpy()
        presc = None
        if bodies.any_prescribed:
            pm, pv = bodies.prescribed.cpu().numpy(), bodies.prescribed_velocity.cpu().numpy()
            presc = [pv[k] if pm[k] else None for k in range(bodies.n_bodies)]
        r3 = r_vectors.reshape(-1, 3)
        order = np.argsort(counts, kind="stable")
        self.groups = []
        k = 0
        while k < len(order):
            c0 = counts[order[k]]
            j = k
            while j < len(order) and counts[order[j]] <= self.SPREAD * c0:
                j += 1
            idx = np.sort(order[k:j])
            blob_idx = np.concatenate([np.arange(off[b], off[b + 1]) for b in idx])
            sub = RigidBodies([ref[off[b]:off[b + 1]] for b in idx], loc[idx], quat[idx], device=dev,
                              prescribed_velocity=None if presc is None else [presc[b] for b in idx])
            bi = torch.from_numpy(blob_idx).to(dev)
            comp = (3 * bi.unsqueeze(1) + torch.arange(3, device=dev)).reshape(-1)      # blob-vector components
            bodyc = (6 * torch.from_numpy(idx).to(dev).unsqueeze(1) + torch.arange(6, device=dev)).reshape(-1)
            pc = BlockDiagonalPreconditioner(sub, r3[bi].reshape(-1).contiguous(), eta, a, wall=wall)
            self.groups.append((pc, comp, bodyc))
            k = j

    def padded_bytes(self):
        return sum(8 * pc.bodies.n_bodies * pc.m * pc.m for pc, _, _ in self.groups)

    def __call__(self, vector):
        n3 = 3 * self.bodies.n_blobs
        out = torch.empty_like(vector)
        for pc, comp, bodyc in self.groups:
            y = pc(torch.cat([vector[comp], vector[n3 + bodyc]]))
            out[comp] = y[:comp.numel()]
            out[n3 + bodyc] = y[comp.numel():]
        return out

    def _blobwise(self, w, name, *extra):
        w = w.reshape(-1)
        out = torch.empty_like(w)
        for pc, comp, _ in self.groups:
            out[comp] = getattr(pc, name)(w[comp], *extra).reshape(-1)
        return out

    def mobility_pc(self, w, mobility_mult):
        """y = P^T M P w: the block-diagonal P^T and P act bucket by bucket, M on the whole blob vector"""
        v = self._blobwise(w, "_P_mult")
        return self._blobwise(mobility_mult(v).reshape(-1), "_PT_mult")

    def P_inv_mult(self, w):
        return self._blobwise(w, "P_inv_mult")

    def M_inv_mult(self, w):
        return self._blobwise(w, "M_inv_mult")


def block_diagonal_preconditioner(bodies, r_vectors, eta, a, wall=True):
    """BlockDiagonalPreconditioner, or its bucketed form when padding every body to the largest one would more than
    double the storage (sum over bodies of (3 nb_max)^2 against (3 nb_b)^2)."""
    c = bodies.counts.astype(np.float64)
    if bodies.uniform or bodies.n_bodies * float(bodies.nb_max) ** 2 <= 2.0 * float((c * c).sum()):
        return BlockDiagonalPreconditioner(bodies, r_vectors, eta, a, wall=wall)
    return GroupedBlockDiagonalPreconditioner(bodies, r_vectors, eta, a, wall=wall)


def build_block_diagonal_preconditioners_det_stoch(bodies, r_vectors, Nblobs, eta, a, *args, **kwargs):
    """multi_bodies.build_block_diagonal_preconditioners_det_stoch (multi_bodies.py:480): returns
    (block_diagonal_preconditioner, mobility_pc, P_inv_mult).  kwargs: mobility_mult (callable applying the blob
    mobility; default single-wall B200 product with kwargs['periodic_length']), wall."""
    PC = block_diagonal_preconditioner(bodies, r_vectors, eta, a, wall=kwargs.get("wall", True))
    mult = kwargs.get("mobility_mult")
    if mult is None:
        L = kwargs.get("periodic_length")
        fn = (mobility.single_wall_mobility_trans_times_force_b200 if kwargs.get("wall", True)
              else mobility.no_wall_mobility_trans_times_force_b200)
        mult = lambda v: fn(r_vectors, v, eta, a, periodic_length=L)
    return PC, (lambda w: PC.mobility_pc(w, mult)), PC.P_inv_mult


def build_block_diagonal_preconditioner(bodies, r_vectors, Nblobs, eta, a, *args, **kwargs):
    """Same role as multi_bodies.build_block_diagonal_preconditioner (multi_bodies.py:752): returns the callable."""
    return block_diagonal_preconditioner(bodies, r_vectors, eta, a, wall=kwargs.get("wall", True))


def solve_mobility_problem(bodies, force_torque, eta, a, tolerance=1e-8, periodic_length=None, slip=None,
                           PC=None, restart=60, maxiter=1000, callback=None, mobility_vector_prod=None, wall=True,
                           x0=None):
    """The `mobility` scheme of multi_bodies_utilities.py:420-540 (free rigid bodies) on the device:
    RHS = [slip ; -force_torque] scaled to unit norm, right-preconditioned GMRES(60) on
    linear_operator_rigid with the block-diagonal preconditioner, solution rescaled.
    Bodies with prescribed kinematics (RigidBodies.set_prescribed) keep their velocity; the force-torque they need is
    K^T lambda (K_matrix_T_vector_prod) restricted to them.
    x0: optional first guess [lambda ; U] (e.g. the previous time step's solution, as the reference's integrators
    pass it, quaternion_integrator_multi_bodies.py:1525).
    Returns (velocity [Nbodies, 6], lambda_blobs [Nblobs, 3], iterations, info)."""
    from . import solvers
    dev = bodies.device
    n3 = 3 * bodies.n_blobs
    r = bodies.get_r_vectors()
    FT = torch.as_tensor(force_torque, dtype=torch.float64).to(dev).reshape(-1)
    rhs = torch.zeros(n3 + 6 * bodies.n_bodies, dtype=torch.float64, device=dev)
    if slip is not None:
        rhs[:n3] = torch.as_tensor(slip, dtype=torch.float64).to(dev).reshape(-1)
    rhs[n3:] = -FT
    if bodies.any_prescribed:                  # multi_bodies_utilities.py:481-489: RHS += K U_prescribed, F rows = 0
        K_matrix_vector_prod(bodies, bodies.prescribed_velocity, r_vectors=r, out=rhs[:n3], alpha=1.0, beta=1.0)
        rhs[n3:] = torch.where(bodies.prescribed.unsqueeze(1), torch.zeros_like(FT.reshape(-1, 6)),
                               -FT.reshape(-1, 6)).reshape(-1)
    nrm = torch.linalg.norm(rhs)
    if float(nrm) > 0:
        rhs = rhs / nrm
    if PC is None:
        PC = block_diagonal_preconditioner(bodies, r, eta, a, wall=wall)
    if mobility_vector_prod is None:
        mobility_vector_prod = (mobility.single_wall_mobility_trans_times_force_b200 if wall
                                else mobility.no_wall_mobility_trans_times_force_b200)
    A = lambda v: linear_operator_rigid(v, bodies, r, eta, a, mobility_vector_prod, periodic_length=periodic_length)
    if x0 is not None:
        x0 = torch.as_tensor(x0, dtype=torch.float64).to(dev).reshape(-1)
        x0 = x0 / nrm if float(nrm) > 0 else x0
    x, info, its = solvers.gmres(A, rhs, x0=x0, tol=tolerance, restart=restart, maxiter=maxiter, M=PC, callback=callback)
    if float(nrm) > 0:
        x = x * nrm
    velocity = x[n3:].reshape(-1, 6)
    if bodies.any_prescribed:                  # :524-527: those bodies move as prescribed (their slot held the force)
        velocity = torch.where(bodies.prescribed.unsqueeze(1), bodies.prescribed_velocity, velocity)
    return velocity, x[:n3].reshape(-1, 3), its, info


def solve_resistance_problem(bodies, velocity, eta, a, tolerance=1e-8, periodic_length=None, slip=None,
                             PC=None, restart=60, maxiter=1000, callback=None, mobility_vector_prod=None, wall=True):
    """The `resistance` scheme of multi_bodies_utilities.py:549-574 on the device, matrix-free:
    blob forces lambda from  M lambda = slip + K velocity,  body force-torques = K^T lambda.
    The reference forms the dense blob mobility and calls LAPACK (O(N^3)); here M is applied by the pair kernel
    inside GMRES(60), right-preconditioned with the bodies' own M_b^-1 (the blocks the mobility problem's
    preconditioner already holds), RHS scaled to unit norm as in the mobility scheme.
    velocity: (U_b, omega_b) per body, any shape with 6*Nbodies entries.
    Returns (force [Nbodies, 6], lambda_blobs [Nblobs, 3], iterations, info)."""
    from . import solvers
    dev = bodies.device
    r = bodies.get_r_vectors()
    U = torch.as_tensor(velocity, dtype=torch.float64).to(dev).reshape(-1)
    rhs = K_matrix_vector_prod(bodies, U, r_vectors=r).reshape(-1)
    if slip is not None:
        rhs = rhs + torch.as_tensor(slip, dtype=torch.float64).to(dev).reshape(-1)
    nrm = torch.linalg.norm(rhs)
    if float(nrm) > 0:
        rhs = rhs / nrm
    if PC is None:
        PC = block_diagonal_preconditioner(bodies, r, eta, a, wall=wall)
    if mobility_vector_prod is None:
        mobility_vector_prod = (mobility.single_wall_mobility_trans_times_force_b200 if wall
                                else mobility.no_wall_mobility_trans_times_force_b200)
    A = lambda v: mobility_vector_prod(r, v, eta, a, periodic_length=periodic_length)
    lam, info, its = solvers.gmres(A, rhs, tol=tolerance, restart=restart, maxiter=maxiter, M=PC.M_inv_mult,
                                   callback=callback)
    if float(nrm) > 0:
        lam = lam * nrm
    force = K_matrix_T_vector_prod(bodies, lam, r_vectors=r).reshape(-1, 6)
    return force, lam.reshape(-1, 3), its, info


def body_mobility(bodies, eta, a, tolerance=1e-10, periodic_length=None, PC=None, mobility_vector_prod=None, wall=True):
    """The `body_mobility` scheme of multi_bodies_utilities.py:576-600: the dense (6 Nbodies)^2 grand mobility
    N = (K^T M^-1 K)^-1 of the suspension.  The reference inverts the dense blob mobility; here column c is the
    velocity answer of the mobility problem for the unit force-torque e_c (6 Nbodies preconditioned GMRES solves
    sharing one preconditioner), so it is meant for the small suspensions that scheme is used on.
    Returns N as a [6 Nbodies, 6 Nbodies] device tensor (row/column order: body-major, (U, omega) / (F, T))."""
    n6 = 6 * bodies.n_bodies
    if PC is None:
        PC = block_diagonal_preconditioner(bodies, bodies.get_r_vectors(), eta, a, wall=wall)
    N = torch.empty(n6, n6, dtype=torch.float64, device=bodies.device)
    e = torch.zeros(n6, dtype=torch.float64, device=bodies.device)
    for c in range(n6):
        e.zero_()
        e[c] = 1.0
        vel, _, _, info = solve_mobility_problem(bodies, e, eta, a, tolerance=tolerance, periodic_length=periodic_length,
                                                 PC=PC, mobility_vector_prod=mobility_vector_prod, wall=wall)
        if info != 0:
            raise RuntimeError("body_mobility: GMRES did not converge for column %d" % c)
        N[:, c] = vel.reshape(-1)
    return N
