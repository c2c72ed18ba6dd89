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


# ---- wire format ---------------------------------------------------------------
# A deployment exchanges felt252 words: a wsad value v is sent as v (v >= 0) or v + P (v < 0); a FeltVector is a
# Cairo Array<felt252>, serialised as its length followed by its elements.  The batched forms below work on
# whole arrays so that engine inputs / outputs can be compared with, or fed to, a real contract.

def float_to_wsad_array(x) -> np.ndarray:
    """`int(x * 1e6)` of client/contract.py:48-53 and ipynb beta :250-251 on a whole array: float64 product,
    truncation toward zero, as int64.  Accepts numpy arrays and torch tensors (CPU or CUDA; returns the same kind)."""
    if isinstance(x, torch.Tensor):
        return torch.trunc(x.to(torch.float64) * 1e6).to(torch.int64)
    return np.trunc(np.asarray(x, dtype=np.float64) * 1e6).astype(np.int64)


def wsad_to_float_array(v):
    """`float(v) * 1e-6` of client/contract.py:41-45 on a whole array of (already signed) wsad integers."""
    if isinstance(v, torch.Tensor):
        return v.to(torch.float64) * 1e-6
    return np.asarray(v, dtype=np.int64).astype(np.float64) * 1e-6


def wsad_array_to_felts(v) -> list:
    """Signed wsad integers (any shape, numpy or torch) -> nested lists of felt252 Python ints (`as_felt`,
    contract/src/math.cairo:22-35: negatives become v + P)."""
    a = v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)
    conv = np.frompyfunc(lambda t: i128_to_felt(int(t)), 1, 1)
    return conv(a.astype(object)).tolist()


def felts_to_wsad_array(felts) -> np.ndarray:
    """Nested lists of felt252 ints (or 0x.. strings) -> int64 array of signed wsad values (`as_wsad`,
    math.cairo:37-50; raises if a value does not fit 64 bits, which no constrained prediction does)."""
    def one(t):
        t = from_hex(t) if isinstance(t, str) else int(t)
        s = felt_to_i128(t)
        if not -(1 << 63) <= s < (1 << 63):
            raise OverflowError(f"wsad value {s} does not fit int64")
        return s
    conv = np.frompyfunc(one, 1, 1)
    return conv(np.asarray(felts, dtype=object)).astype(np.int64)


def update_prediction_calldata(prediction) -> List[int]:
    """Calldata of `update_prediction(prediction: FeltVector)` (contract.cairo:588) for one oracle: the array
    length, then one felt per component.  `prediction` holds floats (client side, oracle_scheduler.py:150-166)."""
    felts = [float_to_fwsad(float(x)) for x in prediction]
    return [len(felts)] + felts


def constructor_calldata(admins, enable_oracle_replacement, required_majority, n_failing_oracles, constrained,
                         unconstrained_max_spread, dimension, oracles) -> List[int]:
    """Constructor calldata in the order of contract/README.md:40-66 / contract.cairo:235-247: admins span, five
    scalars (the spread as a felt-encoded wsad), the dimension, oracles span."""
    as_int = lambda a: from_hex(a) if isinstance(a, str) else int(a)     # noqa: E731
    return ([len(admins)] + [as_int(a) for a in admins]
            + [int(bool(enable_oracle_replacement)), int(required_majority), int(n_failing_oracles), int(bool(constrained)),
               i128_to_felt(int(unconstrained_max_spread)), int(dimension)]
            + [len(oracles)] + [as_int(o) for o in oracles])


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
