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


def allreduce_stats(counts: torch.Tensor, sums: torch.Tensor, group=None, out=None):
    """SUM the (int64 counts, float64 sums) accumulator pair over all ranks with ONE all-reduce.

    The two dtypes cannot share a SUM buffer as they are (torch's coalescing manager refuses mixed types too), and
    the 64-bit counters must stay exact (per-oracle sums of wsad risks exceed 2^53 on long runs), so every counter
    travels as two float64 words — its high and low 32 bits, each below 2^32 — next to the float64 sums: the sum over
    any realistic number of ranks of such halves is exact in float64, and the halves are recombined in int64.
    The local accumulators are left untouched.  `out` = the (counts, sums, buffer) triple a previous call returned
    (reused, no allocation per step); returns that triple."""
    import torch.distributed as dist
    n, m = counts.numel(), sums.numel()
    if out is None:
        out = (torch.empty_like(counts), torch.empty_like(sums), torch.empty(2 * n + m, dtype=torch.float64, device=counts.device))
    c, s, buf = out
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        buf[:n] = counts >> 32
        buf[n:2 * n] = counts & 0xFFFFFFFF
        buf[2 * n:] = sums
        dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
        torch.add(buf[:n].to(torch.int64) << 32, buf[n:2 * n].to(torch.int64), out=c)
        s.copy_(buf[2 * n:])
    else:
        c.copy_(counts)
        s.copy_(sums)
    return out
