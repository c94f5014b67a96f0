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
                 precision=_mb.FP64, group=None, ranges=None, boundaries=None, local_product=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.r = r_vectors.reshape(-1)
        n = self.r.numel() // 3
        self.ranges = ranges if ranges is not None else shard_ranges(n, self.world, boundaries)
        self.offset, self.count = self.ranges[self.rank]
        self.vec = ShardedVector(self.ranges, 3, self.r.device, group=group)
        self.eta, self.a, self.block, self.wall, self.L, self.precision = eta, a, block, wall, periodic_length, precision
        self._product = local_product or (lambda r, v, targets: _mb.mobility_product(
            self.block, self.wall, r, v, self.eta, self.a, periodic_length=self.L, precision=self.precision,
            targets=targets))

    def __call__(self, vec_local):
        full = self.vec.gather(vec_local)
        return self._product(self.r, full, (self.offset, self.count))
