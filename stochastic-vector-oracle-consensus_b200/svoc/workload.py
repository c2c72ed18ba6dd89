"""Synthetic workloads and running statistics on the device.

`generate` is the on-device counterpart of the notebook generators (generate_beta_oracles /
generate_2d_beta_oracles, ipynb beta :37-68; generate_2d_gaussian_oracles,
gaussian_distribution_for_tests.ipynb:37-52).  `RunningStats` keeps the per-oracle divergence
accumulators and reliability histograms that a replacement decision needs, in two buffers
that multi-GPU runs combine with one SUM all-reduce each.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional, Sequence

import torch

from . import _lib
from .engine import _DTYPES, ConsensusResult, make_config
from ._lib import SvocGenConfig, SvocOutputs


@dataclass
class Generated:
    pred: torch.Tensor           # [T, N, D]
    honest_words: torch.Tensor   # [T, ceil(N/32)] int32, bit = 1 for honest oracles
    n_oracles: int

    @property
    def honest(self) -> torch.Tensor:
        """[T, N] bool — the `validity` array the notebook generators return."""
        idx = torch.arange(self.n_oracles, device=self.honest_words.device)
        return ((self.honest_words[:, idx // 32] >> (idx % 32)) & 1).bool()


def _seq(v, n=8):
    v = list(v) if isinstance(v, (Sequence, torch.Tensor)) and not isinstance(v, str) else [v]
    if not 1 <= len(v) <= n:
        raise ValueError("between 1 and 8 law parameters are supported (dimension d uses parameter d % len)")
    return v


def _gen_config(n_oracles, dimension, n_failing, a, b, law, failing, shift, seed, trial_offset, dtype, failing_prob=0.0) -> SvocGenConfig:
    av, bv = _seq(a), _seq(b)
    if len(av) != len(bv):
        raise ValueError("a and b need the same length")
    cfg = SvocGenConfig(n_oracles=int(n_oracles), dimension=int(dimension), n_failing=int(n_failing),
                        law={"beta": _lib.LAW_BETA, "gaussian": _lib.LAW_GAUSSIAN, "kumaraswamy": _lib.LAW_KUMARASWAMY,
                             "arctan_gaussian": _lib.LAW_ARCTAN_GAUSSIAN}[law],
                        failing={"uniform": _lib.FAIL_UNIFORM, "colluding": _lib.FAIL_COLLUDING}[failing],
                        dtype=_DTYPES[dtype], n_params=len(av), reserved=0, shift=float(shift),
                        seed=int(seed) & (2**64 - 1), trial_offset=int(trial_offset), failing_prob=float(failing_prob))
    for k, (x, y) in enumerate(zip(av, bv)):
        cfg.a[k], cfg.b[k] = float(x), float(y)
    return cfg


def generate(n_rounds, n_oracles, dimension, n_failing, a=10.0, b=10.0, law="beta", failing="uniform", shift=0.15,
             seed=0, trial_offset=0, dtype=torch.float32, device=None, failing_prob=0.0) -> Generated:
    """Draw `n_rounds` rounds of oracle predictions on the device.

    failing_prob > 0 switches from "exactly n_failing failing oracles" to "every oracle fails independently
    with that probability" (the failing-fraction sweep of BASELINE config 2).

    law="beta": honest ~ beta(a[d], b[d]), failing ~ uniform(0, 1); law="gaussian": honest ~
    normal(mu=a[d], sigma=b[d]), failing ~ uniform(mu, 5 sigma).  failing="colluding" makes every
    failing oracle submit the same vector (mode + shift, clipped to [0, 1] for beta).  Integer
    dtypes store int(x * 1e6) (`to_wsad`).  Failing positions are a random subset per round.
    """
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    if device.type != "cuda":
        raise ValueError("svoc.generate runs on a CUDA device only")
    cfg = _gen_config(n_oracles, dimension, n_failing, a, b, law, failing, shift, seed, trial_offset, dtype, failing_prob)
    pred = torch.empty((n_rounds, n_oracles, dimension), dtype=dtype, device=device)
    words = torch.empty((n_rounds, (n_oracles + 31) // 32), dtype=torch.int32, device=device)
    if n_rounds:
        with torch.cuda.device(device):
            st = torch.cuda.current_stream().cuda_stream
            _lib.check(_lib.load().svoc_generate(C.byref(cfg), n_rounds, C.c_void_p(pred.data_ptr()),
                                                 C.c_void_p(words.data_ptr()), C.c_void_p(st)))
    return Generated(pred, words, n_oracles)


def montecarlo(n_rounds, n_oracles, dimension, n_failing, a=10.0, b=10.0, law="beta", failing="uniform", shift=0.15,
               seed=0, trial_offset=0, dtype=torch.float32, constrained=True, max_spread=0, centre=_lib.CENTRE_MEDIAN,
               centre2=_lib.CENTRE_DEFAULT, tie_rule=_lib.TIE_LOW_INDEX_KEPT, stats: "RunningStats" = None, device=None):
    """Fused generate -> consensus -> accumulate for `n_rounds` rounds (trials trial_offset ..): the
    reference's `benchmark` loop (ipynb beta :385-402) without materialising a single prediction.
    Returns the RunningStats it accumulated into (a new one unless `stats` is given)."""
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    if stats is None:
        stats = RunningStats(n_oracles, device=device)
    fixed = dtype in (torch.int32, torch.int64)
    cfg = make_config(n_oracles, dimension, n_failing, _DTYPES[dtype], constrained, max_spread, centre, centre2, tie_rule)
    gen = _gen_config(n_oracles, dimension, n_failing, a, b, law, failing, shift, seed, trial_offset, dtype)
    if n_rounds:
        with torch.cuda.device(device):
            st = torch.cuda.current_stream().cuda_stream
            _lib.check(_lib.load().svoc_montecarlo(C.byref(cfg), C.byref(gen), int(n_rounds), C.c_void_p(stats.counts.data_ptr()),
                                                   C.c_void_p(stats.sums.data_ptr()), C.c_void_p(st)))
    stats.fixed = fixed
    return stats


class RunningStats:
    """Accumulators over many rounds: status counts, rel1/rel2 histograms (256 bins over [0, 1]),
    per-oracle "not reliable" counts and per-oracle sum of quadratic risk."""

    def __init__(self, n_oracles: int, device=None, window: int = 0):
        """window = k > 0 also keeps the reliable flags of the LAST k accepted rounds (status 0) in a ring, for the
        "recently unreliable" rule of a replacement proposal; 0 keeps the running totals only."""
        self.n = int(n_oracles)
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.window = int(window)
        self._ring = None              # [k, ceil(N/32)] int32 reliable words of the last k accepted rounds, oldest first
        self._ring_len = 0
        self.counts = torch.zeros(_lib.STATS_COUNTS_BASE + 2 * self.n + _lib.MC_EXTRA, dtype=torch.int64, device=self.device)
        self.sums = torch.zeros(_lib.STATS_SUMS_BASE + self.n + _lib.MC_EXTRA, dtype=torch.float64, device=self.device)
        self._reduced = None

    def update(self, res: ConsensusResult) -> None:
        T = res.status.shape[0]
        if T == 0:
            return
        fixed = res.rel1.dtype == torch.int64
        dt = _lib.I64 if fixed else _DTYPES[res.rel1.dtype]
        cfg = make_config(self.n, res.consensus.shape[1], 0, dt)
        ptr = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None  # noqa: E731
        outs = SvocOutputs(essence1=None, qr=ptr(res.qr), rel1=ptr(res.rel1), reliable=ptr(res.reliable_words),
                           consensus=ptr(res.consensus), rel2=ptr(res.rel2), skew=None, kurt=None, status=ptr(res.status))
        with torch.cuda.device(self.device):
            st = torch.cuda.current_stream().cuda_stream
            _lib.check(_lib.load().svoc_accumulate_stats(C.byref(cfg), T, C.byref(outs), C.c_void_p(self.counts.data_ptr()),
                                                         C.c_void_p(self.sums.data_ptr()), C.c_void_p(st)))
        if self.window > 0:
            self._push_window(res)

    def _push_window(self, res: ConsensusResult) -> None:
        """Append the accepted rounds of `res` (status 0, in batch order) to the ring and keep its last k entries.
        Only the tail of the batch can survive, so at most k rounds are gathered."""
        k = self.window
        ok = torch.nonzero(res.status == 0).flatten()
        if ok.numel() == 0:
            return
        new = res.reliable_words.index_select(0, ok[-k:])
        old = self._ring[: self._ring_len] if self._ring is not None else new[:0]
        both = torch.cat([old, new], dim=0)[-k:]
        if self._ring is None:
            self._ring = torch.zeros((k, new.shape[1]), dtype=new.dtype, device=new.device)
        self._ring_len = both.shape[0]
        self._ring[: self._ring_len] = both

    def window_unreliable_counts(self) -> torch.Tensor:
        """[N] int64: in how many of the last k accepted rounds each oracle was NOT reliable (OracleInfo.reliable
        false, contract.cairo:345-363)."""
        out = torch.zeros(self.n, dtype=torch.int64, device=self.device)
        if self._ring is None or self._ring_len == 0:
            return out
        idx = torch.arange(self.n, device=self.device)
        bits = (self._ring[: self._ring_len][:, idx // 32] >> (idx % 32)) & 1
        return self._ring_len - bits.sum(dim=0).to(torch.int64)

    def window_rounds(self) -> int:
        return self._ring_len

    def allreduce(self, group=None):
        """SUM the local accumulators over all ranks (NCCL over NVLink on GPUs); returns (counts, sums)
        of the whole job and leaves the local accumulators untouched."""
        from .distributed import allreduce_stats
        # the reduced copies are allocated once and reused by every later call
        self._reduced = allreduce_stats(self.counts, self.sums, group, out=self._reduced)
        return self._reduced[0], self._reduced[1]

    def summary(self, reduced: bool = False) -> dict:
        c, s = self._reduced[:2] if (reduced and self._reduced is not None) else (self.counts, self.sums)
        c, s = c.cpu(), s.cpu()
        B, H, n = _lib.STATS_COUNTS_BASE, _lib.STATS_HIST_BINS, self.n
        ok = int(c[1])
        mc = int(c[B + 2 * n + 1])
        extra = {}
        if mc:   # filled by svoc.montecarlo only
            extra = {"identification_success_pct": 100.0 * int(c[B + 2 * n]) / mc,
                     "true_consensus_reliability_pct": 100.0 * (1 - 2 * float(s[_lib.STATS_SUMS_BASE + n]) / mc)}
        return {
            **extra,
            "rounds": int(c[0]), "rounds_ok": ok,
            "status_counts": {_lib.STATUS_NAMES[k]: int(c[2 + k]) for k in range(13) if int(c[2 + k])},
            "mean_rel1": float(s[0]) / ok if ok else None, "mean_rel2": float(s[1]) / ok if ok else None,
            "unreliable_fraction_max": float(c[B:B + n].max()) / ok if ok else None,
            "unreliable_fraction_min": float(c[B:B + n].min()) / ok if ok else None,
        }

    def histograms(self, reduced: bool = False):
        c = (self._reduced[0] if (reduced and self._reduced is not None) else self.counts).cpu()
        H = _lib.STATS_HIST_BINS
        return c[16:16 + H].clone(), c[16 + H:16 + 2 * H].clone()

    def unreliable_counts(self) -> torch.Tensor:
        B = _lib.STATS_COUNTS_BASE
        return self.counts[B:B + self.n]

    def qr_sums(self, fixed: bool) -> torch.Tensor:
        B = _lib.STATS_COUNTS_BASE
        return self.counts[B + self.n:B + 2 * self.n] if fixed else self.sums[_lib.STATS_SUMS_BASE:_lib.STATS_SUMS_BASE + self.n]

    def replacement_candidates(self, min_fraction: float = 0.5, recent: bool = False) -> list:
        """Oracles flagged unreliable in at least `min_fraction` of the accepted rounds — of all rounds seen, or
        (recent=True) of the last-k window — most often flagged first: the indices an admin would put in
        `update_proposition` (contract.cairo:661-717; documentation/README.md:152-175 leaves the rule to the admins)."""
        if recent:
            total, counts = self.window_rounds(), self.window_unreliable_counts()
        else:
            total, counts = int(self.counts[1]), self.unreliable_counts()
        if total == 0:
            return []
        frac = counts.double() / total
        hit = torch.nonzero(frac >= min_fraction).flatten()
        order = torch.argsort(frac[hit], descending=True, stable=True)
        return hit[order].tolist()
