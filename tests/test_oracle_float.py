"""CPU: the float oracles against the reference notebook's own outputs (golden fixtures)."""
import numpy as np

from oracle import cref
from oracle import float_oracle as flo
from helpers import ranks_equal_up_to_ties


def test_notebook_known_answers(golden):
    # ipynb beta cells 14-15 print 0.762 and 0.807
    for k in ("kat1d", "kat2d"):
        v = flo.expected_reliability(golden[k + "_wsad"] * 1e-6)
        assert f"{v:0.3f}" == f"{float(golden[k + '_printed']):0.3f}"
        assert abs(v - float(golden[k + "_expected_reliability"])) < 1e-15


def test_restatement_matches_reference_outputs(golden):
    for (n, d, f, t) in golden["shapes"]:
        key = f"n{n}_d{d}_f{f}"
        pred = golden[key + "_pred"]
        for k in range(t):
            m = flo.identify_failing_oracles(pred[k], f)
            assert np.array_equal(m, golden[key + "_mask"][k]), (key, k)
            np.testing.assert_allclose(flo.restricted_median(pred[k], m), golden[key + "_consensus"][k], rtol=0, atol=0)
            med = np.median(pred[k], axis=0)
            dev = np.sqrt(((pred[k] - med) ** 2).sum(axis=1))
            nr, r = flo.rank_array(dev)
            # for N > 16 numpy's default argsort is not stable, so exactly tied deviations (the two
            # middle oracles of an even-N median) may swap ranks in the reference: compare up to ties
            assert ranks_equal_up_to_ties(r, golden[key + "_ranks"][k], dev), (key, k)
            arr = pred[k][:, 0] if d == 1 else pred[k]
            assert abs(flo.expected_reliability(arr) - golden[key + "_expected_reliability"][k]) < 1e-14


def test_contract_form_notebook_mode_equals_reference(golden):
    """consensus_round(centre=MEDIAN, tie=LOW_INDEX_KEPT) is the notebook estimator."""
    for (n, d, f, t) in golden["shapes"]:
        key = f"n{n}_d{d}_f{f}"
        pred = golden[key + "_pred"]
        o = cref.float_batch(pred, int(f), centre=1, tie=1, want_moments=False, nthreads=2)
        assert np.array_equal(o["reliable"], golden[key + "_mask"]), key
        np.testing.assert_allclose(o["consensus"], golden[key + "_consensus"], rtol=1e-15)
        for k in range(min(t, 4)):
            r = flo.consensus_round(pred[k], int(f), centre=flo.CENTRE_MEDIAN, tie=flo.TIE_LOW_INDEX_KEPT, want_moments=False)
            assert np.array_equal(r.reliable, golden[key + "_mask"][k])
            np.testing.assert_allclose(r.consensus, golden[key + "_consensus"][k], rtol=1e-15)


def test_c_port_equals_numpy_oracle():
    rng = np.random.default_rng(5)
    for (N, D, F) in [(7, 2, 2), (20, 2, 4), (32, 8, 6), (64, 16, 12)]:
        x = rng.beta(10, 15, (12, N, D))
        for kw in (dict(constrained=True, centre=0, tie=0), dict(constrained=False, max_spread=10.0, centre=0, tie=0),
                   dict(constrained=True, centre=1, tie=1), dict(constrained=True, centre=2, tie=0)):
            o = cref.float_batch(x, F, want_moments=True, nthreads=1, **kw)
            p = flo.consensus_batch(x, F, want_moments=True, **kw)
            for k in ("essence1", "qr", "rel1", "consensus", "rel2", "skew", "kurt"):
                np.testing.assert_allclose(o[k], p[k], rtol=1e-10, atol=1e-12, err_msg=k)
            assert np.array_equal(o["reliable"], p["reliable"])


def test_fixed_and_float_forms_agree_to_wsad_precision(contract_vectors):
    """The float restatement of the contract tracks the fixed-point one to ~1e-6 on the reference vectors."""
    from oracle import fixed_oracle as fo
    for v in contract_vectors["vectors"]:
        x = np.array(v["pred"], dtype=np.float64) * 1e-6
        r = flo.consensus_round(x, v["n_failing"], constrained=v["constrained"], max_spread=v["max_spread"] * 1e-6)
        e = v["expected"]
        np.testing.assert_allclose(r.consensus, np.array(e["consensus"]) * 1e-6, atol=2e-6)
        assert abs(r.rel1 - e["rel1"] * 1e-6) < 1e-4 and abs(r.rel2 - e["rel2"] * 1e-6) < 1e-4
        assert r.reliable.tolist() == e["reliable"]
