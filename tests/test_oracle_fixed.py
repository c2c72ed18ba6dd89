"""CPU: the fixed-point oracles against the reference's own vectors and against each other."""
import numpy as np

from helpers import W, gen_fixed
from oracle import cref
from oracle import fixed_oracle as fo


def test_reference_unit_tests(contract_vectors):
    u = contract_vectors["unit"]
    # contract/tests/test_math.cairo:12-18 (exact assert)
    got = fo.indexed_merge_sort(u["indexed_merge_sort"]["input"])
    assert [list(p) for p in got] == u["indexed_merge_sort"]["expected"]
    # contract/tests/test_math.cairo:36 (exact) and :25-33 (comment)
    for v, want in u["sqrt"]["cases"]:
        assert fo.sqrt(v) == want
    # SURVEY probe: ties come out higher-index first
    assert [i for i, _ in fo.indexed_merge_sort([5, 5, 5, 1, 5])] == [3, 4, 2, 1, 0]
    # felt252 encoding of negatives, test_math.cairo:95-96
    assert fo.fwsad_to_float(fo.UPPER_BOUND_FELT252 - 1) == -1e-6
    assert fo.float_to_fwsad(-1e-6) == fo.UPPER_BOUND_FELT252 - 1


def test_contract_vectors_and_anchors(contract_vectors):
    for v in contract_vectors["vectors"]:
        r = fo.consensus_round(v["pred"], v["n_failing"], v["constrained"], v["max_spread"])
        e = v["expected"]
        assert r.status == e["status"]
        assert r.essence1 == e["essence1"] and r.qr == e["qr"] and r.rel1 == e["rel1"]
        assert r.reliable == e["reliable"] and r.consensus == e["consensus"] and r.rel2 == e["rel2"]
        assert r.skew == e["skew"] and r.kurt == e["kurt"]
        a = v["anchors"]
        # the reference prints wsad values truncated to 3 digits (utils.cairo:283-297)
        trunc3 = lambda w: int(w / 1000) / 1000  # noqa: E731
        if "rel1_3digits" in a:
            assert trunc3(r.rel1) == a["rel1_3digits"] and trunc3(r.rel2) == a["rel2_3digits"]
            assert trunc3(r.consensus[0]) == a["consensus_3digits"][0]
            assert abs(r.consensus[1] / W - a["consensus_3digits"][1]) < 1e-3
        if "obsolete_rel2_3digits" in a:
            kept = [q for q, k in zip(r.qr, r.reliable) if k]
            ob = fo.obsolete_constrained_reliability(fo.average(kept))
            assert ob == e["obsolete_rel2"] and trunc3(ob) == a["obsolete_rel2_3digits"]
            assert fo.obsolete_constrained_reliability(fo.average(r.qr)) == e["obsolete_rel1"]


def _compare(x, F, constrained, ms):
    o = cref.fixed_batch(x, F, constrained, ms, True, nthreads=2)
    seen = set()
    for t in range(x.shape[0]):
        r = fo.consensus_round(x[t].tolist(), F, constrained, ms, True)
        seen.add(r.status)
        assert r.status == o["status"][t]
        for name in ("essence1", "qr", "reliable", "consensus", "skew", "kurt"):
            v = getattr(r, name)
            if v is not None:
                assert list(v) == list(o[name][t]), (name, t)
        for name in ("rel1", "rel2"):
            v = getattr(r, name)
            if v is not None:
                assert v == o[name][t]
    return seen


def test_c_port_equals_python_oracle():
    rng = np.random.default_rng(2024)
    seen = set()
    for (N, D, F, kind, T) in [(7, 2, 2, "beta", 200), (7, 2, 2, "coarse", 200), (7, 2, 2, "tiny", 200), (7, 2, 2, "extreme", 200),
                               (7, 2, 2, "outside", 100), (5, 3, 2, "coarse", 100), (5, 3, 3, "beta", 50), (3, 2, 2, "beta", 20),
                               (2, 2, 2, "beta", 20), (20, 1, 15, "beta", 50), (32, 8, 6, "colluders", 30), (64, 16, 12, "beta", 20)]:
        seen |= _compare(gen_fixed(kind, rng, T, N, D), F, True, 0)
    for ms in (10 * W, 0, 3 * W, -5 * W):
        seen |= _compare(gen_fixed("wide", rng, 60, 7, 2), 2, False, ms)
    seen |= _compare(gen_fixed("tiny", rng, 60, 7, 2), 2, False, 10 * W)
    # every reachable panic class is exercised
    assert {0, 1, 3, 4, 5, 8, 9} <= seen, seen  # 7 is unreachable: the kept rows have the smallest qr, so rel2 >= rel1


def test_sqrt_of_one_paths():
    # avg_qr / D == 1 -> sqrt(1) -> division by zero in rel1 (status 2); same in rel2 (status 6)
    x = np.full((1, 4, 1), 500000, dtype=np.int64)
    x[0, 0, 0] += 2000  # qr = [4, 0, 0, 0] -> avg 1
    r = fo.consensus_round(x[0].tolist(), 0, True)
    assert r.status == fo.ST_SQRT_REL1
    assert cref.fixed_batch(x, 0, True)["status"][0] == fo.ST_SQRT_REL1
    x = np.full((1, 5, 1), 500000, dtype=np.int64)
    x[0, 0, 0] += 2000   # qr 4 (kept), x[1] far (dropped): first avg large, second avg = 4/4 = 1
    x[0, 1, 0] += 300000
    r = fo.consensus_round(x[0].tolist(), 1, True)
    assert r.status == fo.ST_SQRT_REL2, r
    assert cref.fixed_batch(x, 1, True)["status"][0] == fo.ST_SQRT_REL2
