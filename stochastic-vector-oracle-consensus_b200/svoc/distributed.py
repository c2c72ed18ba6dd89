"""Trial-level data parallelism: rounds are independent (contract.cairo:442-503 recomputes everything from
the stored vectors), so rank r of G owns one contiguous range of trials, generator counters are keyed by
the GLOBAL trial index, and the only exchange is a SUM all-reduce of the two accumulator buffers
(NCCL over NVLink on GPUs; gloo in the CPU tests)."""
from __future__ import annotations

from typing import Tuple

import torch


def shard_range(n_trials: int, rank: int, world_size: int) -> Tuple[int, int]:
    """[lo, hi) of the trials rank `rank` owns: contiguous, sizes differ by at most one, union = all trials."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError("bad rank / world size")
    return n_trials * rank // world_size, n_trials * (rank + 1) // world_size


def allreduce_stats(counts: torch.Tensor, sums: torch.Tensor, group=None, out=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """SUM the (int64 counts, float64 sums) accumulator pair over all ranks.

    The local accumulators are left untouched: they are copied into `out` (a (counts, sums) pair of the same
    shapes, allocated once by the caller and reused every step; fresh tensors when omitted) and reduced there.
    The two dtypes cannot share one SUM buffer without losing the exactness of the 64-bit counters (per-oracle
    sums of wsad risks exceed 2^53 on long runs), so on NCCL both reductions are issued as ONE coalesced group
    (a single ncclGroupStart/End, one launch); other backends (gloo in the CPU tests) get two calls."""
    import torch.distributed as dist
    if out is None:
        out = (torch.empty_like(counts), torch.empty_like(sums))
    c, s = out
    c.copy_(counts)
    s.copy_(sums)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        coalesce = getattr(dist, "_coalescing_manager", None)
        if coalesce is not None and c.is_cuda and dist.get_backend(group) == "nccl":
            with coalesce(group=group, device=c.device, async_ops=False):
                dist.all_reduce(c, op=dist.ReduceOp.SUM, group=group)
                dist.all_reduce(s, op=dist.ReduceOp.SUM, group=group)
        else:
            dist.all_reduce(c, op=dist.ReduceOp.SUM, group=group)
            dist.all_reduce(s, op=dist.ReduceOp.SUM, group=group)
    return c, s
