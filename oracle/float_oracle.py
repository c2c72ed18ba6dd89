"""CPU oracle — floating-point forms of the consensus round (numpy, float64).

TEST INFRASTRUCTURE ONLY.  Nothing in the product path may import this module;
only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg do.

Two forms, each citing the reference lines it restates (paths relative to
/root/reference; "nb" = contract/drafts/beta_kumaraswamy_algorithm_demo copy.ipynb,
raw JSON line numbers):

* the NOTEBOOK form  (nb:342-376, client/oracle_scheduler.py:94-134): true
  np.median centre, L2 deviation, drop the F largest, median of the rest;
* the CONTRACT form  (contract/src/contract.cairo:370-503 evaluated in floats):
  the contract's "smooth median" centre, squared-L2 quadratic risk, the
  (value asc, index desc) selection order, both reliabilities, skew/kurtosis.

Pinned against the reference itself: tests/golden/make_golden.py exec's the
notebook's own definition cells in the build container and stores their
outputs on seeded inputs in tests/golden/notebook_golden.npz; the printed
known answers 0.762 / 0.807 (nb cells 14-15) are checked as well.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import numpy as np

CENTRE_SMOOTH_MEDIAN_CAIRO = 0
CENTRE_MEDIAN = 1
CENTRE_MEAN = 2

TIE_HIGH_INDEX_KEPT = 0   # contract: sort.cairo:96-102 => (value asc, index desc)
TIE_LOW_INDEX_KEPT = 1    # numpy stable argsort: among ties the higher index is dropped first


# --------------------------------------------------------------------------- notebook form
def rank_array(values):
    """nb:342-352 == client/oracle_scheduler.py:94-104.  Stable argsort is used so
    that tied values have a defined order (lower index first)."""
    s = np.argsort(np.asarray(values), kind="stable")
    n = s.size
    rev = np.zeros(n, dtype=np.int64)
    for from_idx, to_idx in enumerate(s):
        rev[to_idx] = n - from_idx - 1
    return rev / (n - 1), rev


def identify_failing_oracles(pred, n_failing):
    """nb:354-367."""
    p = np.asarray(pred, dtype=np.float64)
    if p.ndim == 1:
        p = p[:, None]
    med = np.median(p, axis=0)
    dev = np.sqrt(((p - med) ** 2).sum(axis=1))
    _, ranks = rank_array(dev)
    return ranks >= n_failing


def restricted_median(pred, activated):
    """nb:369-371."""
    p = np.asarray(pred, dtype=np.float64)
    return np.median(p[np.asarray(activated, dtype=bool)], axis=0)


def expected_reliability(arr):
    """nb:260-267 — np.median without axis: scalar median of the flattened array."""
    a = np.asarray(arr, dtype=np.float64)
    return 1.0 - 2.0 * np.mean(np.abs(a - np.median(a)))


def predictions_to_values(pred):
    """client/oracle_scheduler.py:106-111 (numeric part): mean, median, ranks."""
    p = np.asarray(pred, dtype=np.float64)
    mean = p.mean(axis=0)
    med = np.median(p, axis=0)
    dev = np.sqrt(((p - med) ** 2).sum(axis=1))
    nr, r = rank_array(dev)
    return mean, med, nr, r


# --------------------------------------------------------------------------- contract form
@dataclass
class FloatRound:
    essence1: np.ndarray
    qr: np.ndarray
    rel1: float
    reliable: np.ndarray
    consensus: np.ndarray
    rel2: float
    skew: Optional[np.ndarray]
    kurt: Optional[np.ndarray]
    cut_gap: float  # qr distance between the last kept and the first dropped oracle


def _centre(x: np.ndarray, kind: int) -> np.ndarray:
    n = x.shape[0]
    if kind == CENTRE_MEAN:
        return x.mean(axis=0)                       # gaussian_algorithm_demo.ipynb:191-204
    s = np.sort(x, axis=0)
    if kind == CENTRE_MEDIAN:
        return np.median(x, axis=0)                 # nb:355
    mid = n // 2                                    # math.cairo:113-126 (3-value branch dead)
    return (s[mid - 1] + s[mid]) / 2.0


def _reliability(avg_qr, dim, constrained, max_spread):
    if constrained:
        return 1.0 - 2.0 * np.sqrt(avg_qr / dim)    # contract.cairo:436-439
    return 1.0 - min(max_spread, np.sqrt(avg_qr)) / max_spread  # contract.cairo:365-368


def consensus_round(pred, n_failing, constrained=True, max_spread=0.0,
                    centre=CENTRE_SMOOTH_MEDIAN_CAIRO, centre2=None,
                    tie=TIE_HIGH_INDEX_KEPT, want_moments=True) -> FloatRound:
    """contract.cairo:442-503 / :370-434 with real arithmetic.  `centre2` defaults
    to the contract's choice: same centre when constrained, mean otherwise."""
    x = np.asarray(pred, dtype=np.float64)
    n, dim = x.shape
    r = n - n_failing
    if centre2 is None:
        centre2 = centre if constrained else CENTRE_MEAN
    e1 = _centre(x, centre)
    qr = ((x - e1) ** 2).sum(axis=1)                               # math.cairo:175-238
    rel1 = _reliability(qr.mean(), dim, constrained, max_spread)
    idx = np.arange(n)
    if tie == TIE_HIGH_INDEX_KEPT:
        order = np.lexsort((-idx, qr))                              # sort.cairo:96-102
    else:
        order = np.lexsort((idx, qr))
    reliable = np.zeros(n, dtype=bool)
    reliable[order[:r]] = True                                      # contract.cairo:345-363
    gap = float(qr[order[r]] - qr[order[r - 1]]) if 0 < r < n else np.inf
    kept = x[reliable]
    cons = _centre(kept, centre2)
    rel2 = _reliability(qr[reliable].mean(), dim, constrained, max_spread)  # vs FIRST-pass essence
    skew = kurt = None
    if want_moments:
        m = kept.mean(axis=0)
        var = ((kept - m) ** 2).mean(axis=0)                        # math.cairo:208-222
        with np.errstate(divide="ignore", invalid="ignore"):
            z = (kept - m) / np.sqrt(var)
            rr = float(r)
            skew = (z ** 3).sum(axis=0) * rr / ((rr - 1) * (rr - 2))        # math.cairo:320-338
            k = (z ** 4).sum(axis=0)
            kurt = (k * rr * (rr + 1) / (rr - 1) - 3.0 * (rr - 1) ** 2) / ((rr - 2) * (rr - 3))
    return FloatRound(e1, qr, float(rel1), reliable, cons, float(rel2), skew, kurt, gap)


def consensus_batch(pred, n_failing, **kw):
    """Vectorised over a leading trial axis: [T][N][D] -> dict of stacked arrays."""
    out = [consensus_round(p, n_failing, **kw) for p in pred]
    return {
        "essence1": np.stack([o.essence1 for o in out]),
        "qr": np.stack([o.qr for o in out]),
        "rel1": np.array([o.rel1 for o in out]),
        "reliable": np.stack([o.reliable for o in out]),
        "consensus": np.stack([o.consensus for o in out]),
        "rel2": np.array([o.rel2 for o in out]),
        "skew": None if out[0].skew is None else np.stack([o.skew for o in out]),
        "kurt": None if out[0].kurt is None else np.stack([o.kurt for o in out]),
        "cut_gap": np.array([o.cut_gap for o in out]),
    }
