"""Batch sharding over the GPUs of one node (one process per GPU, torch.distributed).

The NMPC problems are independent, so the batch is cut into contiguous index ranges -- rank g of
G owns ``[g*B/G, (g+1)*B/G)`` for the whole closed-loop run (its warm-start store and last-valid
commands never migrate) -- and the ONLY communication is the gather of the per-problem results
``(u0[4], status)`` plus a small statistics vector.  With NCCL this is an all-gather over
NVLink 5 / NVSwitch of 36 B per problem; the same code runs on ``gloo`` for CPU tests.

The reference has no distributed code at all; this module exists because
``north_star`` shards the batch over the 8 GPUs of a box.
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np


def shard_bounds(rank: int, world: int, batch: int) -> Tuple[int, int]:
    """Contiguous, balanced partition: the first ``batch % world`` ranks get one extra problem."""
    if world < 1 or not 0 <= rank < world or batch < 0:
        raise ValueError('invalid shard request')
    base, extra = divmod(batch, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_sizes(world: int, batch: int):
    return [shard_bounds(r, world, batch)[1] - shard_bounds(r, world, batch)[0]
            for r in range(world)]


class ShardedBatchNMPC:
    """Drives one ``BatchedPositionNMPC`` (or any object with the same ``solve`` / ``command``)
    per rank on that rank's slice of a global batch and gathers the results on every rank.

    ``make_solver(local_batch)`` builds the rank-local solver; it is only called when the local
    slice is non-empty.
    """

    def __init__(self, make_solver, global_batch: int, group=None) -> None:
        import torch.distributed as dist
        self._dist = dist
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.global_batch = int(global_batch)
        self.lo, self.hi = shard_bounds(self.rank, self.world, self.global_batch)
        self.local_batch = self.hi - self.lo
        self.solver = make_solver(self.local_batch) if self.local_batch > 0 else None
        self.sizes = shard_sizes(self.world, self.global_batch)

    def local_slice(self, arr):
        """The rows of a global (host) array owned by this rank."""
        return arr[self.lo:self.hi]

    def solve_local(self, x0, yref, yref_e):
        if self.solver is None:
            return None, None
        return self.solver.solve(x0, yref, yref_e)

    def gather(self, u0, status, device: Optional[str] = None):
        """All-gather ``(u0[b,4], status[b])`` of every rank -> global ``(u0[B,4], status[B])``.
        Accepts numpy arrays or torch tensors; ragged shards are padded to the largest one."""
        import torch
        if self.world == 1:
            return (u0, status)
        if device is None:
            device = 'cuda' if (hasattr(u0, 'is_cuda') and u0.is_cuda) else 'cpu'
        m = max(self.sizes)
        pack = torch.zeros((m, 5), dtype=torch.float64, device=device)
        if self.local_batch:
            pack[:self.local_batch, :4] = torch.as_tensor(u0, dtype=torch.float64, device=device)
            pack[:self.local_batch, 4] = torch.as_tensor(status, device=device).to(torch.float64)
        out = torch.empty((self.world * m, 5), dtype=torch.float64, device=device)
        self._dist.all_gather_into_tensor(out, pack, group=self.group)
        out = out.view(self.world, m, 5)
        parts = [out[r, :n] for r, n in enumerate(self.sizes)]
        full = torch.cat(parts, dim=0)
        return full[:, :4].contiguous(), full[:, 4].to(torch.int32)

    def reduce_stats(self, counters, device: str = 'cpu'):
        """Sum a small vector of counters (status histogram, iteration sums ...) over ranks."""
        import torch
        t = torch.as_tensor(np.asarray(counters, dtype=np.float64), device=device).clone()
        if self.world > 1:
            self._dist.all_reduce(t, op=self._dist.ReduceOp.SUM, group=self.group)
        return t.cpu().numpy()
