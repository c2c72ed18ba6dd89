"""Sequential-commit replay of whole epochs in one batch.

On chain an epoch is N successive `update_prediction` transactions; after each of them the contract replaces one
stored row (update_a_single_oracle, contract.cairo:331-343) and recomputes the whole consensus from the stored
vectors (:442-503).  The state after commit k of an epoch is therefore "rows of the oracles that have already
committed = new, the others = previous epoch": N independent rounds, which this module lays out as one
[E*N, N, D] batch for `svoc_consensus_fixed` instead of N launches per epoch.  A commit whose consensus panics
reverts (contract.cairo panics abort the transaction), so its row stays old for the commits after it; those
commits are recomputed until no new revert appears (at most N passes; one pass when nothing reverts).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch

from . import engine
from .engine import ConsensusResult


@dataclass
class EpochReplay:
    result: ConsensusResult        # E*N rounds: round e*N + k = the contract's state after the k-th commit of epoch e
    reverted: torch.Tensor         # [E, N] bool: the k-th commit of epoch e panicked and left the state untouched
    revert_status: torch.Tensor    # [E, N] int32: SVOC_ST_* code of that panic (0 where the commit went through)
    order: torch.Tensor            # [E, N] int64: which oracle commits k-th
    passes: int                    # launches it took

    def after_commit(self, e: int, k: int) -> dict:
        """Getter values (contract.cairo:605-659) after commit k of epoch e: those of the last commit that did
        not revert."""
        n = self.order.shape[1]
        while k >= 0 and bool(self.reverted[e, k]):
            k -= 1
        if k < 0:
            return {}
        i = e * n + k
        r = self.result
        return {"consensus": r.consensus[i], "rel1": r.rel1[i], "rel2": r.rel2[i], "skew": r.skew[i] if r.skew is not None else None,
                "kurt": r.kurt[i] if r.kurt is not None else None, "reliable": r.reliable[i]}


def replay_epochs(prev: torch.Tensor, new: torch.Tensor, n_failing: int, constrained: bool = True, max_spread: int = 0,
                  order: Optional[torch.Tensor] = None, want_moments: bool = True) -> EpochReplay:
    """prev, new: [E, N, D] wsad integers on a CUDA device (stored vectors before the epoch / the vectors committed
    during it).  order[e, k] = index of the oracle that commits k-th (default 0..N-1).  All oracles are taken as
    already enabled (the activation gate of contract.cairo:447-449 only matters for the very first epoch: there,
    pass the state after the N-th commit as `prev`)."""
    if prev.shape != new.shape or prev.dim() != 3:
        raise ValueError("prev and new must both be [E, N, D]")
    E, N, D = prev.shape
    dev = prev.device
    if order is None:
        order = torch.arange(N, device=dev).expand(E, N).contiguous()
    order = order.to(dev).long()
    # when[e, i] = position at which oracle i commits
    when = torch.empty_like(order)
    when.scatter_(1, order, torch.arange(N, device=dev).expand(E, N).contiguous())
    reverted = torch.zeros((E, N), dtype=torch.bool, device=dev)          # indexed by commit position k
    revert_status = torch.zeros((E, N), dtype=torch.int32, device=dev)
    k_idx = torch.arange(N, device=dev).view(1, N, 1)                     # commit position of the state
    passes = 0
    res = None
    while True:
        # row i of state k is new iff oracle i committed at a position <= k and that commit did not revert
        rev_of_oracle = torch.gather(reverted, 1, when)                   # [E, N] by oracle
        is_new = (when.view(E, 1, N) <= k_idx) & ~rev_of_oracle.view(E, 1, N)     # [E, K, N]
        batch = torch.where(is_new.unsqueeze(-1), new.unsqueeze(1), prev.unsqueeze(1)).reshape(E * N, N, D).contiguous()
        res = engine.consensus_fixed(batch, n_failing, constrained=constrained, max_spread=max_spread,
                                     want_moments=want_moments)
        passes += 1
        bad = (res.status.view(E, N) != 0)
        # statuses up to and including the first not-yet-known revert of an epoch are final
        fresh = bad & ~reverted
        if not bool(fresh.any()):
            break
        first = torch.where(fresh.any(dim=1), fresh.float().argmax(dim=1), torch.full((E,), N, device=dev))
        hit = fresh & (torch.arange(N, device=dev).view(1, N) == first.view(E, 1))
        reverted |= hit
        revert_status = torch.where(hit, res.status.view(E, N).to(torch.int32), revert_status)
        if passes > N:
            raise RuntimeError("replay did not converge")
    # the final batch no longer contains the rows of reverted commits, so at those positions it shows the state
    # the revert left behind; the panic that caused it is in revert_status
    return EpochReplay(result=res, reverted=reverted, revert_status=revert_status, order=order, passes=passes)
