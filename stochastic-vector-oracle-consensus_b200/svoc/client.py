"""Name-compatible mirror of the client-side float preview and wire helpers of the reference
(client/oracle_scheduler.py:73-134, client/contract.py:35-53, client/common.py:7-31).

The numeric part (mean, median, deviation ranks) runs in libsvoc; the NLP pipeline, the sqlite
window, eel and the Starknet RPC of the client are out of scope.
"""
from __future__ import annotations

import random
from typing import List

import numpy as np
import torch

from . import engine
from ._lib import CENTRE_MEAN, CENTRE_MEDIAN, TIE_LOW_INDEX_KEPT
from .notebook import _dev, rank_array  # noqa: F401  (identical copy in the client, oracle_scheduler.py:94-104)

# client/common.py:7-31
N_ORACLES = 7
N_FAILING_ORACLES = 2
PREDICTION_WINDOW = 50
BOOTSTRAPING_SUBSET = 10
LABELS_KEYS = ["optimism", "anger", "annoyance", "excitement", "nervousness", "remorse"]
DIMENSION = len(LABELS_KEYS)

# client/contract.py:35-36
UPPER_BOUND_FELT252 = 3618502788666131213697322783095070105623107215331596699973092056135872020481
UPPER_BOUND__I128 = (2 ** 127) - 1


def fwsad_to_float(x):
    """client/contract.py:41-45: felt252-encoded wsad -> float."""
    return float((x - UPPER_BOUND_FELT252) if x > UPPER_BOUND__I128 else x) * 1e-6


def float_to_fwsad(x):
    """client/contract.py:48-53: float -> felt252-encoded wsad (truncation toward zero, negatives as v + P)."""
    as_wsad = int(x * 1e6)
    return as_wsad + UPPER_BOUND_FELT252 if as_wsad < 0 else as_wsad


def felt_to_i128(x: int) -> int:
    """contract/src/signed_decimal.cairo:47-49 / tests/test_math.cairo:95-96: P - 1 -> -1."""
    return x - UPPER_BOUND_FELT252 if x > UPPER_BOUND__I128 else x


def i128_to_felt(x: int) -> int:
    """contract/src/math.cairo:22-35 (`as_felt`)."""
    return x + UPPER_BOUND_FELT252 if x < 0 else x


def to_hex(x: int) -> str:
    return f"0x{x:0x}"


def from_hex(x: str) -> int:
    return int(x, 16)


def gen_oracles_predictions(sentiment_analysis: List[np.ndarray]) -> List[np.ndarray]:
    """client/oracle_scheduler.py:73-92: the first N_FAILING_ORACLES oracles are uniform noise, the others
    a bootstrap mean of BOOTSTRAPING_SUBSET rows of the sentiment window; then shuffled.  Host-side data
    preparation of a 7 x 6 matrix (not part of the consensus hot path)."""
    values = []
    for i in range(N_ORACLES):
        if i < N_FAILING_ORACLES:
            values.append(np.random.uniform(0, 1, DIMENSION))
        else:
            values.append(np.mean(random.sample(list(sentiment_analysis), BOOTSTRAPING_SUBSET), axis=0))
    np.random.shuffle(values)
    return values


def predictions_to_eel_values(oracles_predictions: List[np.ndarray]):
    """client/oracle_scheduler.py:106-134 -> (component, mean, median, ranks).  An oracle is shown as
    failing when ranks[i] < N_FAILING_ORACLES (:146)."""
    p = np.asarray(oracles_predictions, dtype=np.float64)
    n, dim = p.shape
    x = torch.from_numpy(np.ascontiguousarray(p))[None].to(_dev())
    med_run = engine.consensus_float(x, 0, centre=CENTRE_MEDIAN, tie_rule=TIE_LOW_INDEX_KEPT, want_rank=True)
    mean = engine.consensus_float(x, 0, centre=CENTRE_MEAN, want_qr=False).essence1[0].cpu().numpy()
    median = med_run.essence1[0].cpu().numpy()
    ranks = (n - 1) - med_run.rank[0].cpu().numpy().astype(np.int64)   # largest deviation -> rank 0
    normalized_ranks = ranks / (n - 1)
    component = []
    for i in range(0, dim, 2):
        two = i + 1 < dim
        names = [LABELS_KEYS[i] if i < len(LABELS_KEYS) else f"dim{i}",
                 (LABELS_KEYS[i + 1] if i + 1 < len(LABELS_KEYS) else f"dim{i + 1}") if two else "None"]
        data = [{"x": p[j][i], "y": p[j][i + 1] if two else 0, "score": normalized_ranks[j]} for j in range(n)]
        component.append({"columnNames": names, "data": data})
    return component, mean, median, ranks
