"""CPU oracle — fixed-point ("wsad", 1e6-scaled signed integer) consensus round.

TEST INFRASTRUCTURE ONLY.  Nothing in the product path may import this module;
only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg do.

A plain Python big-int restatement of the reference contract's arithmetic.
Every function cites the reference lines it follows (paths relative to
/root/reference).  The Cairo sources cannot be executed on this box (no scarb /
cairo VM), so parity is pinned by the reference's own unit-test expectations
(contract/tests/test_math.cairo:12-18, :36), by the three-digit comments of
contract/tests/test_contract.cairo:285-288 / :188, and by agreement between
this file, oracle/svoc_ref.c and the CUDA kernels.

Python ints are unbounded; Cairo's i128 panics on overflow.  `_chk` reproduces
that panic so the oracle reports SVOC_ST_OVERFLOW where the contract reverts.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional, Sequence

WSAD = 1_000_000          # contract/src/signed_decimal.cairo:82
HALF_WSAD = 500_000       # contract/src/signed_decimal.cairo:83
I128_MAX = (1 << 127) - 1
I128_MIN = -(1 << 127)
MAX_SQRT_ITERATIONS = 50  # contract/src/math.cairo:271

# Per-round status codes = first panic met in the contract's evaluation order
# (contract/src/contract.cairo:442-503 / :370-434).  Shared with include/svoc.h.
ST_OK = 0
ST_INTERVAL_INPUT = 1     # math.cairo:298-310 via contract.cairo:591-593
ST_SQRT_REL1 = 2          # math.cairo:271-292 sqrt(1) -> wsad_div(.,0)
ST_MAX_SPREAD_ZERO = 3    # contract.cairo:365-368 wsad_div(., 0)
ST_REL1_RANGE = 4         # math.cairo:294-296 after contract.cairo:467 / :397
ST_SECOND_PASS_SHAPE = 5  # smooth_median on < 2 rows (math.cairo:117), empty average
ST_SQRT_REL2 = 6
ST_REL2_RANGE = 7         # contract.cairo:488 / :419
ST_SKEW_PANIC = 8         # math.cairo:320-338 (var in {0,1} or R <= 2)
ST_KURT_PANIC = 9         # math.cairo:340-363 (R <= 3)
ST_OVERFLOW = 10          # i128 overflow anywhere
ST_FIRST_PASS_SHAPE = 11  # N < 2: sorted.at(mid-1) underflows (math.cairo:117)


class CairoPanic(Exception):
    def __init__(self, code: int):
        super().__init__(f"panic {code}")
        self.code = code


def _chk(v: int) -> int:
    """Cairo i128 arithmetic panics on overflow (corelib)."""
    if v > I128_MAX or v < I128_MIN:
        raise CairoPanic(ST_OVERFLOW)
    return v


def tdiv(a: int, b: int) -> int:
    """I128Div::div — signed_decimal.cairo:52-63: divide magnitudes, apply sign
    (0 counts as positive, :14-16) => truncation toward zero; b == 0 panics."""
    if b == 0:
        raise ZeroDivisionError
    q = abs(a) // abs(b)
    return q if (a >= 0) == (b >= 0) else -q


def wsad_mul(a: int, b: int) -> int:
    """signed_decimal.cairo:110-112."""
    return tdiv(_chk(_chk(a * b) + HALF_WSAD), WSAD)


def wsad_div(a: int, b: int) -> int:
    """signed_decimal.cairo:114-116."""
    return tdiv(_chk(_chk(a * WSAD) + tdiv(b, 2)), b)


def sqrt(value: int) -> int:
    """math.cairo:271-292 — Newton iteration in wsad, 50-iteration cap.
    sqrt(1) divides by zero (g = 0)."""
    if value == 0:
        return 0
    g = tdiv(value, 2)
    g2 = _chk(g + WSAD)
    i = 0
    while True:
        if g == g2 or i == MAX_SQRT_ITERATIONS:
            return g
        n = wsad_div(value, g)
        g2 = g
        g = tdiv(_chk(g + n), 2)
        i += 1


def smooth_median(values: Sequence[int]) -> int:
    """math.cairo:113-126.  The three-value branch is guarded by
    `(len & 2) == 1`, which never holds, so the result is always
    (sorted[mid-1] + sorted[mid]) / 2 with mid = len / 2 (odd lengths too)."""
    n = len(values)
    s = sorted(values)  # alexandria MergeSort: plain ascending value sort
    mid = n // 2
    if mid - 1 < 0:
        raise IndexError  # usize underflow / out-of-bounds panic
    a, b = s[mid - 1], s[mid]
    if (n & 2) == 1:  # dead code, kept to mirror the source
        return tdiv(_chk(_chk(a + b) + s[mid + 1]), 3)
    return tdiv(_chk(a + b), 2)


def nd_array_split(rows: Sequence[Sequence[int]]) -> List[List[int]]:
    """math.cairo:55-85 (transpose); panics on an empty array (`array.at(0)`)."""
    if len(rows) == 0:
        raise IndexError
    d = len(rows[0])
    return [[r[k] for r in rows] for k in range(d)]


def nd_smooth_median(rows) -> List[int]:
    """math.cairo:152-165."""
    return [smooth_median(col) for col in nd_array_split(rows)]


def quadratic_deviation(a: int, b: int) -> int:
    """math.cairo:170-173."""
    x = _chk(a - b)
    return wsad_mul(x, x)


def nd_quadratic_deviation(a, b) -> int:
    """math.cairo:175-189 — rounding per component, then the sum."""
    r = 0
    for k in range(len(a)):
        r = _chk(r + quadratic_deviation(a[k], b[k]))
    return r


def nd_quadratic_risk(rows, center) -> List[int]:
    """math.cairo:225-238."""
    return [nd_quadratic_deviation(r, center) for r in rows]


def average(values: Sequence[int]) -> int:
    """math.cairo:240-254."""
    s = 0
    for v in values:
        s = _chk(s + v)
    return tdiv(s, len(values))


def nd_average(rows) -> List[int]:
    """math.cairo:256-269."""
    return [average(col) for col in nd_array_split(rows)]


def nd_component_wise_variance(rows, center) -> List[int]:
    """math.cairo:208-222 with quadratic_risk :193-206."""
    cols = nd_array_split(rows)
    return [average([quadratic_deviation(v, center[k]) for v in cols[k]])
            for k in range(len(cols))]


def skewness(values, mean: int, variance: int) -> int:
    """math.cairo:320-338."""
    n = len(values)
    sd = sqrt(variance)
    skew = 0
    for v in values:
        diff = wsad_div(_chk(v - mean), sd)
        skew = _chk(skew + wsad_mul(wsad_mul(diff, diff), diff))
    return tdiv(_chk(skew * n), _chk((n - 1) * (n - 2)))


def kurtosis(values, mean: int, variance: int) -> int:
    """math.cairo:340-363 (safe_wsad_mul == wsad_mul, signed_decimal.cairo:95-96)."""
    n = len(values)
    sd = sqrt(variance)
    kurt = 0
    for v in values:
        diff = wsad_div(_chk(v - mean), sd)
        d2 = wsad_mul(diff, diff)
        kurt = _chk(kurt + wsad_mul(d2, d2))
    term1 = tdiv(_chk(_chk(kurt * n) * (n + 1)), n - 1)
    term2 = _chk(_chk(3 * WSAD * (n - 1)) * (n - 1))
    return tdiv(_chk(term1 - term2), _chk((n - 2) * (n - 3)))


def indexed_merge_sort(values: Sequence[int]):
    """sort.cairo:13-103 — top-down merge sort of (index, value) pairs; the merge
    takes the left head only when `left < right` strictly (:96-102), so equal
    values come out right-half first: total order (value asc, index DESC)."""
    def rec(arr):
        n = len(arr)
        if n <= 1:
            return list(arr)
        m = n // 2
        l, r = rec(arr[:m]), rec(arr[m:])
        out, i, j = [], 0, 0
        while len(out) < len(l) + len(r):
            if i == len(l):
                out.append(r[j]); j += 1
            elif j == len(r):
                out.append(l[i]); i += 1
            elif l[i][1] < r[j][1]:
                out.append(l[i]); i += 1
            else:
                out.append(r[j]); j += 1
        return out
    return rec([(i, v) for i, v in enumerate(values)])


@dataclass
class RoundResult:
    status: int = ST_OK
    essence1: Optional[List[int]] = None      # first-pass essence
    qr: Optional[List[int]] = None            # per-oracle quadratic risk (first pass)
    rel1: Optional[int] = None
    reliable: Optional[List[bool]] = None
    consensus: Optional[List[int]] = None
    rel2: Optional[int] = None
    skew: Optional[List[int]] = None
    kurt: Optional[List[int]] = None
    order: Optional[List[int]] = field(default=None)  # oracle indices, most reliable first


def _reliability(avg_qr: int, dim: int, constrained: bool, max_spread: int,
                 st_sqrt: int) -> int:
    try:
        if constrained:
            # contract.cairo:436-439
            return _chk(WSAD - _chk(sqrt(tdiv(avg_qr, dim)) * 2))
        # contract.cairo:365-368 (+ min, math.cairo:312-318)
        sd = sqrt(avg_qr)
    except ZeroDivisionError:
        raise CairoPanic(st_sqrt)
    m = sd if max_spread > sd else max_spread
    try:
        return _chk(WSAD - wsad_div(m, max_spread))
    except ZeroDivisionError:
        raise CairoPanic(ST_MAX_SPREAD_ZERO)


def consensus_round(pred: Sequence[Sequence[int]], n_failing: int, constrained: bool,
                    max_spread: int = 0, want_moments: bool = True) -> RoundResult:
    """One full update with all N rows present — contract.cairo:442-503
    (constrained) / :370-434 (unconstrained), in the contract's evaluation order.
    Fields computed before the first panic are kept; later ones stay None."""
    res = RoundResult()
    rows = [[int(v) for v in r] for r in pred]
    n = len(rows)
    dim = len(rows[0])
    try:
        # contract.cairo:591-593 -> math.cairo:298-310
        if constrained:
            for r in rows:
                for v in r:
                    if not (0 <= v <= WSAD):
                        raise CairoPanic(ST_INTERVAL_INPUT)
        for r in rows:
            for v in r:
                _chk(v)
        # ---- first pass (contract.cairo:455-470 / :383-400)
        try:
            res.essence1 = nd_smooth_median(rows)
        except IndexError:
            raise CairoPanic(ST_FIRST_PASS_SHAPE)
        res.qr = nd_quadratic_risk(rows, res.essence1)
        rel1 = _reliability(average(res.qr), dim, constrained, max_spread, ST_SQRT_REL1)
        if not (0 <= rel1 <= WSAD):
            res.rel1 = rel1
            raise CairoPanic(ST_REL1_RANGE)
        res.rel1 = rel1
        ordered = indexed_merge_sort(res.qr)
        res.order = [i for i, _ in ordered]
        # contract.cairo:345-363 (usize subtraction panics when F > N)
        threshold = n - n_failing
        if threshold < 0:
            raise CairoPanic(ST_SECOND_PASS_SHAPE)
        reliable = [False] * n
        for pos, (i, _) in enumerate(ordered):
            reliable[i] = pos < threshold
        res.reliable = reliable
        # ---- second pass (contract.cairo:472-489 / :402-420)
        kept = [rows[i] for i in range(n) if reliable[i]]  # contract.cairo:310-329
        try:
            if constrained:
                res.consensus = nd_smooth_median(kept)
            else:
                res.consensus = nd_average(kept)
        except (IndexError, ZeroDivisionError):
            raise CairoPanic(ST_SECOND_PASS_SHAPE)
        qr2 = nd_quadratic_risk(kept, res.essence1)   # against the FIRST-pass essence
        rel2 = _reliability(average(qr2), dim, constrained, max_spread, ST_SQRT_REL2)
        if not (0 <= rel2 <= WSAD):
            res.rel2 = rel2
            raise CairoPanic(ST_REL2_RANGE)
        res.rel2 = rel2
        if not want_moments:
            return res
        # ---- skewness / kurtosis (contract.cairo:491-500 / :422-431)
        means = nd_average(kept)
        variances = nd_component_wise_variance(kept, means)
        cols = nd_array_split(kept)
        try:
            res.skew = [skewness(cols[k], means[k], variances[k]) for k in range(dim)]
        except ZeroDivisionError:
            raise CairoPanic(ST_SKEW_PANIC)
        try:
            res.kurt = [kurtosis(cols[k], means[k], variances[k]) for k in range(dim)]
        except ZeroDivisionError:
            raise CairoPanic(ST_KURT_PANIC)
    except CairoPanic as p:
        res.status = p.code
    return res


# ---- obsolete reliability (contract/obsolete/src/contract_nd.cairo:418, :437):
# 1 - 2*sqrt(avg qr) without the /dim.  Only used as a second numeric anchor
# (the "0.798" comment of contract/tests/test_contract.cairo:188).
def obsolete_constrained_reliability(avg_qr: int) -> int:
    return WSAD - sqrt(avg_qr) * 2


# ---- wire format (client/contract.py:35-53, contract/tests/test_math.cairo:95-96)
UPPER_BOUND_FELT252 = 3618502788666131213697322783095070105623107215331596699973092056135872020481
UPPER_BOUND_I128 = (1 << 127) - 1


def float_to_fwsad(x: float) -> int:
    w = int(x * 1e6)
    return w + UPPER_BOUND_FELT252 if w < 0 else w


def fwsad_to_float(x: int) -> float:
    return float((x - UPPER_BOUND_FELT252) if x > UPPER_BOUND_I128 else x) * 1e-6
