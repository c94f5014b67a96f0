"""Multi-GPU sharding of the mobility product: one process per GPU.

Targets are split in contiguous index ranges, one per rank -- cut on body
boundaries when a body_offset table is given, so that K, K^T and the
block-diagonal preconditioner stay rank-local.  Sources are replicated:
positions once per configuration, the (B-scaled inside the kernel) vector by an
all-gather of the ranks' slices at every product.  There is nothing to match
in the reference: its Mf path is single process ("Distributed
communication backend: none").

The collective plumbing is torch.distributed (NCCL on GPUs, gloo in the CPU
tests); `local_product` is injectable so the host logic is testable without a
GPU.
"""
import numpy as np
import torch
import torch.distributed as dist

from . import mobility as _mb


def shard_ranges(n_items, world, boundaries=None):
    """[(offset, count)] * world covering [0, n_items).  boundaries: sorted cut candidates (e.g. the
    body_offset table in blobs); each cut is moved to the nearest candidate."""
    cuts = [(n_items * k) // world for k in range(world + 1)]
    if boundaries is not None:
        b = np.asarray(boundaries, dtype=np.int64)
        for k in range(1, world):
            j = int(np.searchsorted(b, cuts[k]))
            cand = [int(b[i]) for i in (j - 1, j) if 0 <= i < len(b)]
            cuts[k] = min(cand, key=lambda c: abs(c - cuts[k]))
        for i in range(1, len(cuts)):                 # keep the cuts monotone (empty shards are legal)
            cuts[i] = max(cuts[i], cuts[i - 1])
    return [(cuts[k], cuts[k + 1] - cuts[k]) for k in range(world)]


class ShardedVector:
    """All-gather of per-rank slices with unequal lengths (padded to the longest slice)."""

    def __init__(self, ranges, width, device, dtype=torch.float64, group=None):
        self.ranges, self.width, self.group = ranges, width, group
        self.world = len(ranges)
        self.max_cnt = max(c for _, c in ranges)
        self.equal = all(c == self.max_cnt for _, c in ranges)
        n = sum(c for _, c in ranges)
        self.full = torch.empty(n * width, dtype=dtype, device=device)
        if not self.equal:
            self._padded = torch.empty(self.world, self.max_cnt * width, dtype=dtype, device=device)
            self._mine = torch.zeros(self.max_cnt * width, dtype=dtype, device=device)

    def gather(self, local):
        local = local.reshape(-1)
        if self.world == 1:
            self.full.copy_(local)
        elif self.equal:
            dist.all_gather_into_tensor(self.full, local.contiguous(), group=self.group)
        else:
            self._mine[:local.numel()] = local
            dist.all_gather_into_tensor(self._padded.view(-1), self._mine, group=self.group)
            for k, (o, c) in enumerate(self.ranges):
                self.full[o * self.width:(o + c) * self.width] = self._padded[k, :c * self.width]
        return self.full


class ShardedMobility:
    """u_local = rows [offset, offset+count) of M(r_all) * vec_all, with vec_all gathered from the ranks.

    r_vectors: the FULL (replicated) position array on this rank's device.
    __call__(vec_local) -> u_local; signature-compatible with `mobility_mult` of the Lanczos solver and with
    the blob part of linear_operator_rigid."""

    def __init__(self, r_vectors, eta, a, block=_mb.UF, wall=_mb.SINGLE_WALL, periodic_length=None,
                 precision=_mb.FP64, group=None, ranges=None, boundaries=None, local_product=None, symmetric=None,
                 deterministic=False):
        single = group is False or not dist.is_initialized()      # group=False: single process on purpose
        self.group = None if single else (dist.group.WORLD if group is None else group)
        self.world = 1 if single else dist.get_world_size(group)
        self.rank = 0 if single else dist.get_rank(group)
        self.r = r_vectors.reshape(-1)
        n = self.r.numel() // 3
        self.ranges = ranges if ranges is not None else shard_ranges(n, self.world, boundaries)
        self.offset, self.count = self.ranges[self.rank]
        self.vec = ShardedVector(self.ranges, 3, self.r.device, group=self.group)
        self.eta, self.a, self.block, self.wall, self.L, self.precision = eta, a, block, wall, periodic_length, precision
        # deterministic=True: never the pair-symmetric kernel (row kernel: fixed summation order for a given sharding,
        # bit-reproducible run to run; different shardings agree to rounding)
        if deterministic:
            symmetric = False
        self._product = local_product or (lambda r, v, targets: _mb.mobility_product(
            self.block, self.wall, r, v, self.eta, self.a, periodic_length=self.L, precision=self.precision,
            targets=targets, flags=_mb.DETERMINISTIC if deterministic else 0))

        # Pair-symmetric schedule (UF / UT / WF / WT, wall or no wall; fp64, not periodic in z): the unordered pairs are dealt to the
        # ranks row group by row group, every rank accumulates raw sums for ALL blobs, the sums are all-reduced
        # (the one extra data-path collective: 24 N bytes) and each rank finishes its own target range.
        open_z = periodic_length is None or not bool(np.asarray(
            periodic_length.cpu() if isinstance(periodic_length, torch.Tensor) else periodic_length).reshape(-1)[2] > 0)
        eligible = block in (_mb.UF, _mb.UT, _mb.WF, _mb.WT) and wall in (_mb.SINGLE_WALL, _mb.NO_WALL)
        self.symmetric = (symmetric is not False and local_product is None and eligible and
                          precision == _mb.FP64 and open_z and
                          (symmetric is True or n >= _mb._lib.SYMMETRIC_MIN_N))
        self._acc = None

    def __call__(self, vec_local):
        full = self.vec.gather(vec_local)
        if not self.symmetric:
            return self._product(self.r, full, (self.offset, self.count))
        self._acc = _mb.symmetric_partial(self.r, full, self.eta, self.a, rank=self.rank, world=self.world,
                                          acc=self._acc, wall=self.wall, periodic_length=self.L, block=self.block)
        if self.world > 1:
            dist.all_reduce(self._acc, group=self.group)
        return _mb.symmetric_finish(self.r, full, self._acc, self.eta, self.a, targets=(self.offset, self.count),
                                    wall=self.wall, block=self.block)


def _global_norm(x, group):
    s = torch.dot(x, x).reshape(1)
    if group is not None and dist.is_initialized():
        dist.all_reduce(s, group=group)
    return float(torch.sqrt(s[0]))


class ShardedRigidSystem:
    """Rigid multiblob suspension sharded BY BODY over the ranks of `group`.

    Every rank holds the (tiny) body table and the full blob position array, replicated; it owns the bodies
    whose blobs fall in its contiguous blob range (cut on body boundaries), and for those bodies the slices
    lambda_loc (3 per blob) and U_loc (6 per body) of every Krylov vector, K, K^T and the block-diagonal
    preconditioner.  The only data-path collective is the all-gather of lambda slices inside the blob mobility
    product; inner products of GMRES / Lanczos are all-reduced by the solvers.

    Local vector layout: [lambda_loc (3 n_blobs_loc) ; U_loc (6 n_bodies_loc)] -- the rank-local rows of the
    reference's [lambda ; U] unknown (multi_bodies.py:424-471).

    backend: object providing RigidBodies, K_matrix_vector_prod, K_matrix_T_vector_prod and
    BlockDiagonalPreconditioner (default: rigidmultiblobswall_b200.rigid, CUDA); injectable for the CPU tests.
    """

    def __init__(self, reference_configurations, locations, quaternions, eta, a, group=None, wall=True,
                 periodic_length=None, precision=_mb.FP64, backend=None, local_product=None):
        if backend is None:
            from . import rigid as backend
        single = group is False or not dist.is_initialized()      # group=False: single process on purpose
        self.backend, self.eta, self.a, self.wall, self.L = backend, eta, a, wall, periodic_length
        self.group = None if single else (dist.group.WORLD if group is None else group)
        self.world = 1 if single else dist.get_world_size(group)
        self.rank = 0 if single else dist.get_rank(group)
        locations = np.asarray(locations, dtype=np.float64).reshape(-1, 3)
        quaternions = np.asarray(quaternions, dtype=np.float64).reshape(-1, 4)
        if isinstance(reference_configurations, np.ndarray) and reference_configurations.ndim == 2:
            reference_configurations = [reference_configurations] * len(locations)
        allb = backend.RigidBodies(reference_configurations, locations, quaternions)
        self.n_bodies_total, self.n_blobs_total = allb.n_bodies, allb.n_blobs
        self.r_full = allb.get_r_vectors().reshape(-1)
        off = np.asarray(allb.body_offset_host)
        self.ranges = shard_ranges(allb.n_blobs, self.world, boundaries=off)
        o, c = self.ranges[self.rank]
        self.b0 = int(np.searchsorted(off, o))
        self.b1 = int(np.searchsorted(off, o + c))
        self.body_ranges = [(int(np.searchsorted(off, oo)), int(np.searchsorted(off, oo + cc)))
                            for oo, cc in self.ranges]
        self.bodies = backend.RigidBodies(reference_configurations[self.b0:self.b1], locations[self.b0:self.b1],
                                          quaternions[self.b0:self.b1])
        self.r_loc = self.r_full[3 * o:3 * (o + c)]
        self.n3 = 3 * c
        self.mobility = ShardedMobility(self.r_full, eta, a, block=_mb.UF,
                                        wall=_mb.SINGLE_WALL if wall else _mb.NO_WALL,
                                        periodic_length=periodic_length, precision=precision,
                                        group=False if single else group, ranges=self.ranges,
                                        local_product=local_product)
        self.PC = None

    # -- operators on rank-local vectors ------------------------------------------------------------------
    def linear_operator(self, x):
        """rank-local rows of [[M, -K], [-K^T, 0]] [lambda; U]"""
        lam, U = x[:self.n3], x[self.n3:]
        res = torch.empty_like(x)
        res[:self.n3] = self.mobility(lam).reshape(-1)
        self.backend.K_matrix_vector_prod(self.bodies, U, r_vectors=self.r_loc, out=res[:self.n3], alpha=-1.0, beta=1.0)
        res[self.n3:] = self.backend.K_matrix_T_vector_prod(self.bodies, lam, r_vectors=self.r_loc, alpha=-1.0).reshape(-1)
        return res

    def build_preconditioner(self):
        make = getattr(self.backend, "block_diagonal_preconditioner", self.backend.BlockDiagonalPreconditioner)
        self.PC = make(self.bodies, self.r_loc, self.eta, self.a, wall=self.wall)
        return self.PC

    # -- solves -------------------------------------------------------------------------------------------
    def local_slice(self, per_body_full, width=6):
        """rows of this rank's bodies out of a replicated [n_bodies_total, width] array"""
        x = torch.as_tensor(per_body_full, dtype=torch.float64).reshape(self.n_bodies_total, width)
        return x[self.b0:self.b1].to(self.r_full.device).reshape(-1)

    def solve_mobility(self, force_torque_loc, tolerance=1e-8, slip_loc=None, restart=60, maxiter=1000, callback=None):
        """The mobility problem (multi_bodies_utilities.py:420-540) on sharded vectors: RHS = [slip; -FT] scaled to
        unit (global) norm, right-preconditioned GMRES(60).  Returns (U_loc [nb_loc, 6], lambda_loc [n_loc, 3],
        iterations, info)."""
        from . import solvers
        if self.PC is None:
            self.build_preconditioner()
        dev = self.r_full.device
        rhs = torch.zeros(self.n3 + 6 * self.bodies.n_bodies, dtype=torch.float64, device=dev)
        if slip_loc is not None:
            rhs[:self.n3] = slip_loc.reshape(-1)
        rhs[self.n3:] = -torch.as_tensor(force_torque_loc, dtype=torch.float64).to(dev).reshape(-1)
        grp = self.group if self.world > 1 else None
        nrm = _global_norm(rhs, grp)
        if nrm > 0:
            rhs = rhs / nrm
        x, info, its = solvers.gmres(self.linear_operator, rhs, tol=tolerance, restart=restart, maxiter=maxiter,
                                     M=self.PC, callback=callback, group=grp)
        if nrm > 0:
            x = x * nrm
        return x[self.n3:].reshape(-1, 6), x[:self.n3].reshape(-1, 3), its, info

    def solve_resistance(self, velocity_loc, tolerance=1e-8, slip_loc=None, restart=60, maxiter=1000, callback=None):
        """The resistance problem (multi_bodies_utilities.py:549-574) on sharded vectors, matrix-free:
        M lambda = slip + K U with the local bodies' M_b^-1 as right preconditioner, F = K^T lambda.
        velocity_loc: (U, omega) of this rank's bodies.  Returns (force_loc [nb_loc, 6], lambda_loc [n_loc, 3],
        iterations, info)."""
        from . import solvers
        if self.PC is None:
            self.build_preconditioner()
        dev = self.r_full.device
        U = torch.as_tensor(velocity_loc, dtype=torch.float64).to(dev).reshape(-1)
        rhs = self.backend.K_matrix_vector_prod(self.bodies, U, r_vectors=self.r_loc).reshape(-1).clone()
        if slip_loc is not None:
            rhs += torch.as_tensor(slip_loc, dtype=torch.float64).to(dev).reshape(-1)
        grp = self.group if self.world > 1 else None
        nrm = _global_norm(rhs, grp)
        if nrm > 0:
            rhs = rhs / nrm
        lam, info, its = solvers.gmres(lambda v: self.mobility(v).reshape(-1), rhs, tol=tolerance, restart=restart,
                                       maxiter=maxiter, M=self.PC.M_inv_mult, callback=callback, group=grp)
        if nrm > 0:
            lam = lam * nrm
        force = self.backend.K_matrix_T_vector_prod(self.bodies, lam, r_vectors=self.r_loc).reshape(-1, 6)
        return force, lam.reshape(-1, 3), its, info

    def stochastic_noise(self, z_loc, factor=1.0, tolerance=1e-6, max_iter=1000):
        """factor * M^{1/2} z on sharded blob vectors by preconditioned Lanczos: the operator is P^T M P with
        the block-diagonal P of the local bodies (multi_bodies.py:591-615) and the result is mapped back with
        P^-1 (stochastic_forcing.py:256).  Returns (noise_loc, iterations)."""
        from . import solvers
        if self.PC is None:
            self.build_preconditioner()
        grp = self.group if self.world > 1 else None
        return solvers.lanczos(factor=factor, tolerance=tolerance, max_iter=max_iter,
                               mobility_mult=lambda w: self.PC.mobility_pc(w, self.mobility),
                               L_mult=self.PC.P_inv_mult, z=z_loc, group=grp)
