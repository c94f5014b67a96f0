"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: target sharding, the padded all-gather of
unequal slices, the sharded product re-assembling to the full one, and GMRES / Lanczos with all-reduced inner
products.  The local pair kernel is injected (the CPU oracle) -- only the plumbing is under test here."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from rigidmultiblobswall_b200 import parallel, solvers


def test_shard_ranges():
    assert parallel.shard_ranges(10, 3) == [(0, 3), (3, 3), (6, 4)]
    assert parallel.shard_ranges(1000000, 8)[7] == (875000, 125000)
    body_offset = np.arange(0, 15 * 7 + 1, 15)            # 7 bodies of 15 blobs
    r = parallel.shard_ranges(105, 2, boundaries=body_offset)
    assert r == [(0, 45), (45, 60)] or r == [(0, 60), (60, 45)]
    assert all(o % 15 == 0 for o, _ in r)
    r = parallel.shard_ranges(30, 4, boundaries=[0, 15, 30])   # more ranks than bodies: empty shards are legal
    assert sum(c for _, c in r) == 30 and all(o % 15 == 0 for o, _ in r)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _spawn(worker, world, out, attempts=3):
    """mp.spawn with a fresh rendezvous port per attempt: the probed port can be taken by another process between
    the probe and the bind (seen once on a loaded box); a real failure fails every attempt"""
    for k in range(attempts):
        try:
            out.clear()
            mp.spawn(worker, args=(world, _free_port(), out), nprocs=world, join=True)
            return
        except Exception:
            if k == attempts - 1:
                raise


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oracle
        rng = np.random.default_rng(0)                     # same data on every rank
        n, a, eta = 151, 0.3, 1.1                          # odd n: unequal shards
        r = 6.0 * rng.random((n, 3)); r[:, 2] += 0.2
        f = rng.standard_normal((n, 3))
        prod = lambda rr, v, targets: torch.from_numpy(oracle.mobility_product(
            "UF", "single_wall", rr.numpy().reshape(-1, 3), v.numpy(), eta, a,
            targets=(targets[0], targets[0] + targets[1])))
        sm = parallel.ShardedMobility(torch.from_numpy(r), eta, a, local_product=prod)
        off, cnt = sm.offset, sm.count
        assert not sm.vec.equal
        u_loc = sm(torch.from_numpy(f[off:off + cnt].copy()))
        full = oracle.mobility_product("UF", "single_wall", r, f, eta, a)
        e1 = float(np.abs(u_loc.numpy() - full[3 * off:3 * (off + cnt)]).max())
        # GMRES on the sharded SPD operator M: solve M lam = u, every rank holds its slice
        b_loc = torch.from_numpy(full[3 * off:3 * (off + cnt)].copy())
        x_loc, info, its = solvers.gmres(sm, b_loc, tol=1e-10, restart=40, maxiter=20, group=dist.group.WORLD)
        e2 = float(np.abs(x_loc.numpy() - f.reshape(-1)[3 * off:3 * (off + cnt)]).max())
        # Lanczos M^{1/2} z with sharded vectors == single-process Lanczos
        z = rng.standard_normal(3 * n)
        noise_loc, it_l = solvers.lanczos(factor=1.0, tolerance=1e-9, mobility_mult=sm,
                                          z=torch.from_numpy(z[3 * off:3 * (off + cnt)].copy()), group=dist.group.WORLD)
        one = lambda v: torch.from_numpy(oracle.mobility_product("UF", "single_wall", r, v.numpy(), eta, a))
        noise_ref, it_r = solvers.lanczos(factor=1.0, tolerance=1e-9, mobility_mult=one, z=torch.from_numpy(z))
        e3 = float((noise_loc - noise_ref[3 * off:3 * (off + cnt)]).abs().max() / noise_ref.abs().max())
        out[rank] = (e1, info, e2, abs(it_l - it_r), e3)
    finally:
        dist.destroy_process_group()


def test_sharded_product_gmres_lanczos_world2():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    _spawn(_worker, world, out)
    assert len(out) == world
    for rank in range(world):
        e1, info, e2, dit, e3 = out[rank]
        assert e1 == 0.0                  # same oracle rows, bit-identical
        assert info == 0 and e2 < 1e-6    # M is ill-conditioned-ish; residual tol 1e-10
        assert dit <= 1 and e3 < 1e-7


# ---- ShardedRigidSystem: sharding by body, operator assembly, sharded GMRES + preconditioned Lanczos ------
class _CpuRigidBackend:
    """NumPy/oracle stand-in for rigidmultiblobswall_b200.rigid (test only): same call shapes."""

    class RigidBodies:
        def __init__(self, cfgs, loc, quat):
            self.loc = np.asarray(loc, dtype=np.float64).reshape(-1, 3)
            self.quat = np.asarray(quat, dtype=np.float64).reshape(-1, 4)
            self.cfgs = [np.asarray(c, dtype=np.float64) for c in cfgs]
            self.counts = np.array([len(c) for c in self.cfgs], dtype=np.int64)
            self.n_bodies, self.n_blobs = len(self.cfgs), int(self.counts.sum())
            self.body_offset_host = np.concatenate([[0], np.cumsum(self.counts)]).astype(np.int64)
            self.device = torch.device("cpu")

        def get_r_vectors(self):
            out = []
            for c, x, q in zip(self.cfgs, self.loc, self.quat):
                s, p = q[0], q[1:]
                P = np.array([[0, -p[2], p[1]], [p[2], 0, -p[0]], [-p[1], p[0], 0]])
                R = 2 * (np.outer(p, p) + (s * s - 0.5) * np.eye(3) + s * P)
                out.append(c @ R.T + x)
            return torch.from_numpy(np.concatenate(out).reshape(-1)) if out else torch.zeros(0, dtype=torch.float64)

    @staticmethod
    def _rel(bodies, r_vectors):
        r = r_vectors.numpy().reshape(-1, 3)
        return r - np.repeat(bodies.loc, bodies.counts, axis=0)

    @classmethod
    def K_matrix_vector_prod(cls, bodies, U, r_vectors=None, out=None, alpha=1.0, beta=0.0):
        U = U.numpy().reshape(-1, 6)
        Ub = np.repeat(U, bodies.counts, axis=0)
        v = Ub[:, :3] + np.cross(Ub[:, 3:], cls._rel(bodies, r_vectors))
        res = torch.from_numpy(alpha * v.reshape(-1))
        if out is None:
            return res.view(-1, 3)
        out.mul_(beta).add_(res)
        return out.view(-1, 3)

    @classmethod
    def K_matrix_T_vector_prod(cls, bodies, lam, r_vectors=None, alpha=1.0):
        lam = lam.numpy().reshape(-1, 3)
        tq = np.cross(cls._rel(bodies, r_vectors), lam)
        idx = np.repeat(np.arange(bodies.n_bodies), bodies.counts)
        F = np.zeros((bodies.n_bodies, 6))
        np.add.at(F[:, :3], idx, lam)
        np.add.at(F[:, 3:], idx, tq)
        return torch.from_numpy(alpha * F).reshape(-1, 3)

    class BlockDiagonalPreconditioner:
        def __init__(self, bodies, r_vectors, eta, a, wall=True):
            import oracle
            self.bodies = bodies
            r = r_vectors.numpy().reshape(-1, 3)
            self.blocks = []
            for b in range(bodies.n_bodies):
                sl = slice(bodies.body_offset_host[b], bodies.body_offset_host[b + 1])
                rb, m = r[sl], 3 * int(bodies.counts[b])
                M = np.stack([oracle.mobility_product("UF", "single_wall", rb, e, eta, a) for e in np.eye(m)], axis=1)
                K = np.zeros((m, 6))
                for i, x in enumerate(rb - bodies.loc[b]):
                    K[3 * i:3 * i + 3, :3] = np.eye(3)
                    K[3 * i:3 * i + 3, 3:] = -np.array([[0, -x[2], x[1]], [x[2], 0, -x[0]], [-x[1], x[0], 0]])
                Minv = np.linalg.inv(M)
                Lc = np.linalg.cholesky(M)
                self.blocks.append((Minv, K, np.linalg.pinv(K.T @ Minv @ K), Lc, np.linalg.inv(Lc)))

        def _split(self, v):
            o = 3 * self.bodies.body_offset_host
            return [v[o[b]:o[b + 1]] for b in range(self.bodies.n_bodies)]

        def __call__(self, vector):
            v = vector.numpy()
            n3 = 3 * self.bodies.n_blobs
            lam_out, U_out = [], []
            for (Minv, K, N, _, _), slip, F in zip(self.blocks, self._split(v[:n3]), v[n3:].reshape(-1, 6)):
                lt = Minv @ slip
                U = N @ (-F - K.T @ lt)
                lam_out.append(Minv @ (slip + K @ U)); U_out.append(U)
            return torch.from_numpy(np.concatenate(lam_out + U_out))

        def M_inv_mult(self, w):
            return torch.from_numpy(np.concatenate([Minv @ x for (Minv, _, _, _, _), x in zip(self.blocks, self._split(w.numpy()))]))

        def mobility_pc(self, w, mobility_mult):
            v = np.concatenate([Li.T @ x for (_, _, _, _, Li), x in zip(self.blocks, self._split(w.numpy()))])
            mv = mobility_mult(torch.from_numpy(v)).numpy().reshape(-1)
            return torch.from_numpy(np.concatenate([Li @ x for (_, _, _, _, Li), x in zip(self.blocks, self._split(mv))]))

        def P_inv_mult(self, w):
            return torch.from_numpy(np.concatenate([Lc @ x for (_, _, _, Lc, _), x in zip(self.blocks, self._split(w.numpy()))]))


def _rigid_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oracle
        rng = np.random.default_rng(3)
        nb, a, eta = 7, 0.2, 1.3                                   # 7 bodies: unequal body shards
        cfgs = [0.45 * rng.standard_normal((4 + (b % 3), 3)) for b in range(nb)]   # 4..6 blobs per body
        loc = np.stack([2.5 * np.arange(nb), np.zeros(nb), 2.0 + rng.random(nb)], axis=1)
        quat = rng.standard_normal((nb, 4)); quat /= np.linalg.norm(quat, axis=1)[:, None]
        FT = rng.standard_normal((nb, 6))
        prod = lambda rr, v, targets: torch.from_numpy(oracle.mobility_product(
            "UF", "single_wall", rr.numpy().reshape(-1, 3), v.numpy(), eta, a, targets=(targets[0], targets[0] + targets[1])))
        mk = lambda grp: parallel.ShardedRigidSystem(cfgs, loc, quat, eta, a, group=grp, backend=_CpuRigidBackend,
                                                     local_product=prod)
        sys_sh = mk(dist.group.WORLD)
        assert sum(b1 - b0 for b0, b1 in sys_sh.body_ranges) == nb
        U_loc, lam_loc, its, info = sys_sh.solve_mobility(sys_sh.local_slice(FT), tolerance=1e-10)
        z = rng.standard_normal(3 * sys_sh.n_blobs_total)
        o, c = sys_sh.ranges[rank]
        noise_loc, it_l = sys_sh.stochastic_noise(torch.from_numpy(z[3 * o:3 * (o + c)].copy()), tolerance=1e-9)
        # single-process reference with the same backend (group=False: single process)
        one = mk(False)
        U1, lam1, its1, info1 = one.solve_mobility(torch.from_numpy(FT.reshape(-1)), tolerance=1e-10)
        noise1, it1 = one.stochastic_noise(torch.from_numpy(z.copy()), tolerance=1e-9)
        eU = float((U_loc - U1[sys_sh.b0:sys_sh.b1]).abs().max() / U1.abs().max())
        eL = float((lam_loc.reshape(-1) - lam1.reshape(-1)[3 * o:3 * (o + c)]).abs().max() / lam1.abs().max())
        eN = float((noise_loc - noise1[3 * o:3 * (o + c)]).abs().max() / noise1.abs().max())
        # resistance problem: sharded against single process, and against a dense solve of the assembled system
        vel = rng.standard_normal((nb, 6))
        F_loc, lamr_loc, its_r, info_r = sys_sh.solve_resistance(sys_sh.local_slice(vel), tolerance=1e-10)
        F1, lamr1, its_r1, info_r1 = one.solve_resistance(torch.from_numpy(vel.reshape(-1)), tolerance=1e-10)
        r_all = one.r_full.numpy().reshape(-1, 3)
        Mdense = np.stack([oracle.mobility_product("UF", "single_wall", r_all, e, eta, a) for e in np.eye(r_all.size)], axis=1)
        KU = _CpuRigidBackend.K_matrix_vector_prod(one.bodies, torch.from_numpy(vel.reshape(-1)), r_vectors=one.r_full)
        lam_dense = np.linalg.solve(Mdense, KU.numpy().reshape(-1))
        eF = float((F_loc - F1[sys_sh.b0:sys_sh.b1]).abs().max() / F1.abs().max())
        eD = float(np.abs(lamr1.numpy().reshape(-1) - lam_dense).max() / np.abs(lam_dense).max())
        out[rank] = (info, info1, abs(its - its1), eU, eL, abs(it_l - it1), eN, info_r, info_r1, abs(its_r - its_r1), eF, eD)
    finally:
        dist.destroy_process_group()


def test_sharded_rigid_system_world2():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    _spawn(_rigid_worker, world, out)
    assert len(out) == world
    for rank in range(world):
        info, info1, dits, eU, eL, dl, eN, info_r, info_r1, dits_r, eF, eD = out[rank]
        assert info == 0 and info1 == 0 and dits <= 1
        assert eU < 1e-8 and eL < 1e-7
        assert dl <= 1 and eN < 1e-7
        assert info_r == 0 and info_r1 == 0 and dits_r <= 1 and eF < 1e-7 and eD < 1e-7
