"""Name-compatible mirror of the reference notebook's functions
(contract/drafts/beta_kumaraswamy_algorithm_demo copy.ipynb, raw JSON line numbers cited per function).

Same names, argument meaning and return types as the notebook; the arithmetic runs in libsvoc on
the current CUDA device.  Generators use the library's Philox stream instead of numpy's global
MT19937 state, so individual draws differ from the reference while their distribution is the same
(`seed()` re-seeds the stream).  `benchmark` evaluates all K trials in one batch instead of a
Python loop — that loop is the one this engine replaces.
"""
from __future__ import annotations

from itertools import compress
from typing import List, Tuple

import numpy as np
import torch

from . import engine, workload
from ._lib import CENTRE_MEDIAN, TIE_LOW_INDEX_KEPT

_state = {"seed": 0, "trial": 0}


def seed(s: int) -> None:
    """Counterpart of np.random.seed for the generators below."""
    _state["seed"], _state["trial"] = int(s), 0


def _next_trials(k: int) -> int:
    t = _state["trial"]
    _state["trial"] += k
    return t


def _dev():
    if not torch.cuda.is_available():
        raise engine._lib.SvocLibraryError("svoc.notebook needs a CUDA device (there is no CPU fallback)")
    return torch.device("cuda", torch.cuda.current_device())


# ------------------------------------------------------------------ generators (ipynb :37-81)
def generate_beta_oracles(N_oracles, N_failing_oracles, a: float, b: float):
    """ipynb :37-47 -> (list of float, list of bool)."""
    g = workload.generate(1, N_oracles, 1, N_failing_oracles, a=a, b=b, seed=_state["seed"],
                          trial_offset=_next_trials(1), dtype=torch.float64, device=_dev())
    return g.pred[0, :, 0].tolist(), g.honest[0].tolist()


def generate_2d_beta_oracles(N_oracles, N_failing_oracles, a: Tuple[float], b: Tuple[float]):
    """ipynb :49-68 -> (ndarray [N, 2], ndarray [N] bool)."""
    g = workload.generate(1, N_oracles, 2, N_failing_oracles, a=list(a), b=list(b), seed=_state["seed"],
                          trial_offset=_next_trials(1), dtype=torch.float64, device=_dev())
    return g.pred[0].cpu().numpy(), g.honest[0].cpu().numpy()


def kumaraswamy_mode(a: float, b: float) -> float:
    """ipynb :77-78."""
    return ((a - 1) / (a * b - 1)) ** (1 / a)


def beta_mode(a: float, b: float) -> float:
    """ipynb :80-81."""
    return (a - 1) / (a + b - 2)


def to_wsad(x: float) -> int:
    """ipynb :250-251."""
    return int(x * 1e6)


# ------------------------------------------------------------------ reliability (ipynb :260-267)
def reliability(std: float) -> float:
    return 1 - 2 * std


def expected_reliability(arr) -> float:
    a = torch.as_tensor(np.asarray(arr, dtype=np.float64)).reshape(1, -1).to(_dev())
    return float(engine.expected_reliability(a)[0])


# ------------------------------------------------------------------ estimator (ipynb :342-376)
def rank_array(list, normalized=True):  # noqa: A002 - the reference's parameter name
    """ipynb :342-352: (rank / (n - 1), rank), largest value -> rank 0."""
    v = torch.as_tensor(np.asarray(list, dtype=np.float64)).reshape(1, -1).to(_dev())
    r = engine.rank_array(v)[0].cpu().numpy().astype(np.int64)
    return r / (r.size - 1), r


def _as_batch(oracles_predictions) -> torch.Tensor:
    p = np.asarray(oracles_predictions, dtype=np.float64)
    if p.ndim == 1:
        p = p[:, None]
    return torch.from_numpy(np.ascontiguousarray(p))[None].to(_dev())


def identify_failing_oracles(oracles_predictions, N_failing_oracles):
    """ipynb :354-367 -> list of bool (True = kept)."""
    x = _as_batch(oracles_predictions)
    r = engine.consensus_float(x, N_failing_oracles, centre=CENTRE_MEDIAN, tie_rule=TIE_LOW_INDEX_KEPT,
                               want_qr=False, want_essence=False)
    return r.reliable[0].tolist()


def restricted_median(predictions, activated):
    """ipynb :369-371: component-wise median of the activated predictions."""
    kept = list(compress(predictions, activated))
    one_d = np.asarray(kept).ndim == 1
    x = _as_batch(kept)
    med = engine.consensus_float(x, 0, centre=CENTRE_MEDIAN, want_qr=False).essence1[0].cpu().numpy()
    return float(med[0]) if one_d else med


def verify(left, right):
    """ipynb :373-376."""
    for x, y in zip(left, right):
        if x != y:
            return False
    return True


# ------------------------------------------------------------------ Monte-Carlo benchmark (ipynb :385-408)
def benchmark_values(a: float, N_oracles, N_failing_oracles, K: int = 100):
    """The two statistics `benchmark` prints: (identification success %, true consensus reliability %)."""
    dev = _dev()
    g = workload.generate(K, N_oracles, 1, N_failing_oracles, a=a, b=a, seed=_state["seed"],
                          trial_offset=_next_trials(K), dtype=torch.float64, device=dev)
    r = engine.consensus_float(g.pred, N_failing_oracles, centre=CENTRE_MEDIAN, tie_rule=TIE_LOW_INDEX_KEPT,
                               want_qr=False, want_essence=False)
    identified = r.reliable
    success = (identified == g.honest).all(dim=1).double().mean().item()
    # restricted_median(oracles, true_oracles): median over the honest rows = first-pass centre of a round
    # whose failing rows are pushed to +inf and ignored through n_valid
    R = N_oracles - N_failing_oracles
    srt = torch.where(g.honest, g.pred[..., 0], torch.full_like(g.pred[..., 0], float("inf"))).sort(dim=1).values
    truth = 0.5 * (srt[:, (R - 1) // 2] + srt[:, R // 2])
    distance = (r.consensus[:, 0] - truth).abs().mean().item()
    return 100 * success, 100 * (1 - 2 * distance)


def benchmark(a: float, b, N_oracles, N_failing_oracles, K: int = 100):
    """ipynb :385-402 (the `b` argument is only printed there too: oracles are beta(a, a))."""
    success_rate, rel = benchmark_values(a, N_oracles, N_failing_oracles, K)
    s = "{:<40}".format(f"a={a} | b=({b[0]:0.5f}, {b[1]:0.4f})")
    print(f"{s}| identification success: {success_rate:0.2f} % | true consensus reliability : {rel:0.02f} % ")
    return success_rate, rel


def launch_benchmark(N_oracles, N_failing_oracles, K: int = 300):
    """ipynb :404-408."""
    for a in [10, 20, 30, 100]:
        print("---")
        for b in [(15, 30), (a, a), [a, a ** 3], (a ** 3, -a ** 3), (a ** 4, a ** -4)]:
            benchmark(a, b, N_oracles, N_failing_oracles, K)
