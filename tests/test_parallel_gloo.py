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
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    assert len(out) == world
    for rank in range(world):
        e1, info, e2, dit, e3 = out[rank]
        assert e1 == 0.0                  # same oracle rows, bit-identical
        assert info == 0 and e2 < 1e-6    # M is ill-conditioned-ish; residual tol 1e-10
        assert dit <= 1 and e3 < 1e-7
