"""Trajectory sharding across GPUs.

Trajectories are independent, so rank r of G owns a contiguous block of global trajectory ids and
builds its BatchContext with `traj_offset` = first id: the Philox counter uses the GLOBAL id, hence
every trajectory sees the same draws whatever G is.  No state ever crosses GPUs.  The only exchange
is one all-reduce (NCCL over NVLink on the box, gloo in CPU tests) of the packed statistics
{n_alive, sum, sum^2, sum_im} of the observables taken at a sampling point.
"""
from __future__ import annotations


def shard_range(ntraj_total: int, rank: int, world: int):
    """-> (offset, count): contiguous blocks, sizes differ by at most one, multiples of 32 where possible."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank/world")
    words = (ntraj_total + 31) // 32
    base, extra = divmod(words, world)
    w0 = rank * base + min(rank, extra)
    w1 = w0 + base + (1 if rank < extra else 0)
    lo, hi = min(w0 * 32, ntraj_total), min(w1 * 32, ntraj_total)
    return lo, hi - lo


class StatsBuffer:
    """A device (or CPU, for gloo tests) tensor of K x 4 doubles that observable() calls fill without a
    host sync; `allreduce()` sums it over ranks in ONE collective; `means()` reads it back."""

    def __init__(self, k: int, device):
        import torch
        self.t = torch.zeros(k, 4, dtype=torch.float64, device=device)
        self.k = k

    def slot_ptr(self, i: int) -> int:
        if not 0 <= i < self.k:     # the device kernels write through this pointer: never hand out a row past the end
            raise IndexError(f"statistics row {i} requested from a StatsBuffer of {self.k} rows")
        return self.t.data_ptr() + i * 32

    def allreduce(self):
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(self.t, op=dist.ReduceOp.SUM)
        return self

    def host(self):
        return self.t.cpu().numpy()

    def means(self):
        h = self.host()
        n = h[:, 0]
        with_alive = n > 0
        m = h[:, 1].copy()
        m[with_alive] /= n[with_alive]
        m[~with_alive] = float("nan")
        return m
