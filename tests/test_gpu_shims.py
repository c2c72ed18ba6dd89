"""GPU: the name-compatible mirrors of the reference's Python functions and of the contract ABI."""
import numpy as np
import pytest
import torch

from helpers import ranks_equal_up_to_ties

pytestmark = pytest.mark.gpu


def test_notebook_functions_on_reference_outputs(golden):
    """identify_failing_oracles / restricted_median / rank_array / expected_reliability called the way the
    notebook calls them (2-D array, list of arrays, list of floats), against the reference's own outputs."""
    from svoc import notebook as nb
    for (n, d, f, t) in golden["shapes"]:
        if n > 64:
            continue
        key = f"n{n}_d{d}_f{f}"
        pred = golden[key + "_pred"]
        for k in range(min(int(t), 6)):
            arg = pred[k][:, 0].tolist() if d == 1 else (pred[k] if k % 2 else [row for row in pred[k]])
            m = nb.identify_failing_oracles(arg, int(f))
            assert m == golden[key + "_mask"][k].tolist(), (key, k)
            c = nb.restricted_median(arg, m)
            np.testing.assert_allclose(np.atleast_1d(c), golden[key + "_consensus"][k], rtol=1e-12)
            assert abs(nb.expected_reliability(arg) - golden[key + "_expected_reliability"][k]) < 1e-12
            med = np.median(pred[k], axis=0)
            dev = [np.linalg.norm(row - med) for row in pred[k]]
            nr, r = nb.rank_array(dev)
            assert ranks_equal_up_to_ties(r, golden[key + "_ranks"][k], np.asarray(dev))
            np.testing.assert_allclose(nr, r / (n - 1))
    assert nb.verify([True, False], [True, False]) and not nb.verify([True, False], [True, True])
    assert nb.to_wsad(0.4926) == 492600 and abs(nb.beta_mode(10, 15) - 9 / 23) < 1e-15
    assert f"{nb.expected_reliability(golden['kat2d_wsad'] * 1e-6):0.3f}" == "0.807"


def test_notebook_generators_and_benchmark(golden, capsys):
    from svoc import notebook as nb
    nb.seed(123)
    vals, truth = nb.generate_beta_oracles(7, 2, 10, 10)
    assert len(vals) == 7 and sum(truth) == 5 and all(0 <= v <= 1 for v in vals)
    v2, t2 = nb.generate_2d_beta_oracles(20, 4, [10, 10], [15, 30])
    assert v2.shape == (20, 2) and t2.sum() == 16 and t2.dtype == bool
    nb.seed(123)
    again, _ = nb.generate_beta_oracles(7, 2, 10, 10)
    assert again == vals
    # benchmark(): same printed format, statistics inside the reference's own K=20000 numbers
    nb.seed(5)
    succ, rel = nb.benchmark(10, (10, 10), 7, 2, 100000)
    out = capsys.readouterr().out
    assert out.startswith("a=10 | b=(10.00000, 10.0000)") and "identification success" in out
    ref = {(int(a), int(b), int(c)): (s, r) for a, b, c, _, s, r in golden["benchmark_stats"]}[(7, 2, 10)]
    assert abs(succ - ref[0]) < 1.5 and abs(rel - ref[1]) < 0.3


def test_predictions_to_eel_values_matches_reference_math():
    from oracle import float_oracle as flo
    from svoc import client
    rng = np.random.default_rng(1)
    preds = [rng.random(6) for _ in range(7)]
    component, mean, median, ranks = client.predictions_to_eel_values(preds)
    m0, med0, nr0, r0 = flo.predictions_to_values(preds)
    np.testing.assert_allclose(mean, m0, rtol=1e-12)
    np.testing.assert_allclose(median, med0, rtol=1e-12)
    assert ranks.tolist() == r0.tolist()
    assert len(component) == 3 and component[0]["columnNames"] == ["optimism", "anger"]
    assert component[2]["data"][3]["x"] == preds[3][4] and abs(component[0]["data"][2]["score"] - nr0[2]) < 1e-15
    flagged = [i for i in range(7) if ranks[i] < client.N_FAILING_ORACLES]
    assert len(flagged) == 2


def _deploy(constrained, dimension, max_spread=0):
    from svoc.contract import OracleConsensusNDS
    admins = ["Akashi", "Ozu", "Higuchi"]
    oracles = [f"oracle_{i:02d}" for i in range(7)]
    return OracleConsensusNDS(admins, True, 2, 2, constrained, max_spread, dimension, oracles), admins, oracles


def test_contract_mirror_replays_the_reference_integration_tests(contract_vectors):
    """contract/tests/test_contract.cairo:116-215, 218-313, 315-396: deploy, check the zero state, commit the
    seven vectors one by one, read the getters (felt-encoded), run the replacement vote."""
    from svoc.client import fwsad_to_float, i128_to_felt
    from svoc.contract import ContractPanic
    for v in contract_vectors["vectors"]:
        dim = len(v["pred"][0])
        c, admins, oracles = _deploy(v["constrained"], dim, v["max_spread"])
        c.set_caller(admins[0])
        assert c.consensus_active() is False
        assert c.get_consensus_value() == [0] * dim
        assert c.get_first_pass_consensus_reliability() == 0 and c.get_second_pass_consensus_reliability() == 0
        assert c.get_skewness() == [0] * dim and c.get_kurtosis() == [0] * dim
        for i, row in enumerate(v["pred"]):
            c.set_caller(c.get_oracle_list()[i])
            c.update_prediction(row)
            assert c.consensus_active() == (i == 6)
        e = v["expected"]
        assert c.get_consensus_value() == [i128_to_felt(x) for x in e["consensus"]]
        assert c.get_first_pass_consensus_reliability() == e["rel1"]
        assert c.get_second_pass_consensus_reliability() == e["rel2"]
        assert c.get_skewness() == [i128_to_felt(x) for x in e["skew"]]
        assert [fwsad_to_float(x) for x in c.get_kurtosis()] == [x * 1e-6 for x in e["kurt"]]
        c.set_caller(admins[0])
        assert [row[3] for row in c.get_oracle_value_list()] == e["reliable"]
        # replacement vote, test_contract.cairo:195-213
        c.update_proposition((6, "oracle_XX"))
        assert c.get_oracle_list()[6] == "oracle_06"
        c.vote_for_a_proposition(0, True)
        assert c.get_oracle_list()[6] == "oracle_06"
        c.set_caller(admins[1])
        c.vote_for_a_proposition(0, True)
        assert c.get_oracle_list()[6] == "oracle_XX"
        assert c.get_replacement_propositions() == [None, None, None]
        # a later commit recomputes the consensus from the stored vectors (sequential-commit replay)
        c.set_caller("oracle_00")
        before = c.get_consensus_value()
        c.update_prediction(v["pred"][1])
        assert c.get_consensus_value() != before or v["pred"][0] == v["pred"][1]
        with pytest.raises(ContractPanic):
            c.set_caller("nobody")
            c.update_prediction(v["pred"][0])


def test_contract_mirror_reverts_on_panic():
    from svoc.contract import ContractPanic
    c, admins, oracles = _deploy(True, 2)
    for i in range(6):
        c.set_caller(oracles[i])
        c.update_prediction([500000, 500000])
        assert not c.consensus_active()
    c.set_caller(oracles[6])
    with pytest.raises(ContractPanic) as ei:
        c.update_prediction([500000, 500000])      # all equal: variance 0 -> skewness divides by zero
    assert ei.value.status == 8
    assert not c.consensus_active() and c.n_active_oracles == 6     # the 7th transaction reverted
    with pytest.raises(ContractPanic) as ei:
        c.update_prediction([1_000_001, 0])                          # interval error
    assert ei.value.status == 1
    with pytest.raises(ContractPanic):
        c.update_prediction([1, 2, 3])                               # wrong dimension
    assert c.get_consensus_value() == [0, 0]


def test_integration_md_ctypes_stub_runs(golden):
    """The numpy-only ctypes stub printed in INTEGRATION.md is executed as written."""
    import os
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    text = open(os.path.join(root, "INTEGRATION.md")).read()
    code = re.search(r"```python\n(.*?)```", text, flags=re.S).group(1)
    ns = {}
    cwd = os.getcwd()
    os.chdir(root)
    try:
        exec(code, ns)
        key = "n20_d2_f4"
        mask, cons = ns["identify_and_consensus"](golden[key + "_pred"], 4)
    finally:
        os.chdir(cwd)
    assert np.array_equal(mask, golden[key + "_mask"])
    np.testing.assert_allclose(cons, golden[key + "_consensus"], rtol=1e-12)


@pytest.mark.parametrize("kind", ["beta", "tiny"])
def test_batched_epoch_replay_matches_sequential_contract(kind):
    """an epoch of N update_prediction calls replayed as one batch gives, after every commit,
    what the contract mirror computes one transaction at a time — including reverted commits ('tiny' rounds hit
    the sqrt(1) / division-by-zero panics of math.cairo:271-292, 320-338)."""
    import numpy as np
    import svoc
    from svoc.client import i128_to_felt
    from svoc.contract import ContractPanic, OracleConsensusNDS
    from helpers import gen_fixed
    rng = np.random.default_rng(11)
    E, N, D, F = 6, 7, 3, 2
    prev = gen_fixed("beta", rng, E, N, D)
    new = gen_fixed(kind, rng, E, N, D)
    order = np.stack([rng.permutation(N) for _ in range(E)])
    rp = svoc.replay_epochs(torch.from_numpy(prev).cuda(), torch.from_numpy(new).cuda(), F, constrained=True,
                            order=torch.from_numpy(order).cuda())
    n_reverts = 0
    for e in range(E):
        c = OracleConsensusNDS(["adm"], False, 1, F, True, 0, D, [f"o{i}" for i in range(N)])
        for i in range(N):                                   # bring the mirror to the state before the epoch
            try:
                c.update_prediction([i128_to_felt(int(v)) for v in prev[e, i]], caller=f"o{i}")
            except ContractPanic:
                pytest.skip("the previous-epoch state itself panics")
        for k in range(N):
            i = int(order[e, k])
            try:
                c.update_prediction([i128_to_felt(int(v)) for v in new[e, i]], caller=f"o{i}")
                assert not bool(rp.reverted[e, k])
            except ContractPanic as ex:
                n_reverts += 1
                assert bool(rp.reverted[e, k]) and int(rp.revert_status[e, k]) == ex.status
            got = rp.after_commit(e, k)
            if got:
                assert int(got["rel1"]) == c.reliability_first_pass and int(got["rel2"]) == c.reliability_second_pass
                assert got["consensus"].tolist() == c.consensus_value
                assert got["reliable"].tolist() == c.reliable
    if kind == "tiny":
        assert n_reverts > 0 and rp.passes > 1
    else:
        assert rp.passes == 1


@pytest.mark.parametrize("kind", ["beta", "tiny", "coarse"])
def test_batched_epoch_replay_matches_the_bigint_oracle(kind):
    """f2 against oracle/ (not against the repo's own contract mirror): the stored vectors are walked one commit
    at a time in plain Python, every candidate state goes through oracle/fixed_oracle.consensus_round (big-int
    restatement of contract.cairo:442-503), a panicking commit is dropped (the transaction reverts,
    contract.cairo:331-343 never ran) — and the batched replay must report the same reverts, panic codes and
    getter values after every commit."""
    import numpy as np
    import svoc
    from helpers import gen_fixed
    from oracle import fixed_oracle as fo
    rng = np.random.default_rng(23)
    E, N, D, F = 8, 7, 3, 2
    prev = gen_fixed("beta", rng, E, N, D)
    new = gen_fixed(kind, rng, E, N, D)
    if kind == "coarse":                                   # a few commits outside [0, W]: interval panics
        new[::2, 3, 1] = -5
    order = np.stack([rng.permutation(N) for _ in range(E)])
    rp = svoc.replay_epochs(torch.from_numpy(prev).cuda(), torch.from_numpy(new).cuda(), F, constrained=True,
                            order=torch.from_numpy(order).cuda())
    n_reverts = 0
    for e in range(E):
        state = [list(map(int, row)) for row in prev[e]]
        last = None
        for k in range(N):
            i = int(order[e, k])
            cand = [list(r) for r in state]
            cand[i] = list(map(int, new[e, i]))
            o = fo.consensus_round(cand, F, True, 0, True)
            if o.status != 0:
                n_reverts += 1
                assert bool(rp.reverted[e, k]) and int(rp.revert_status[e, k]) == o.status, (e, k, o.status)
            else:
                assert not bool(rp.reverted[e, k]), (e, k)
                state, last = cand, o
            got = rp.after_commit(e, k)
            if last is None:
                assert got == {}
                continue
            assert int(got["rel1"]) == last.rel1 and int(got["rel2"]) == last.rel2
            assert got["consensus"].tolist() == last.consensus and got["reliable"].tolist() == last.reliable
            assert got["skew"].tolist() == last.skew and got["kurt"].tolist() == last.kurt
    if kind != "beta":
        assert n_reverts > 0
