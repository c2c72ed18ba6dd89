"""GPU: the on-device generators (a27) and the running accumulators."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_positions_follow_the_pinned_philox_stream():
    """Failing positions = the F smallest Philox position keys; checked against the KAT-pinned Python Philox."""
    import svoc
    from philox_ref import failing_flags
    seed, off = 0x1234_5678_9ABC_DEF0, 1_000_000_007
    g = svoc.generate(64, 7, 2, 2, a=10, b=15, seed=seed, trial_offset=off)
    honest = g.honest.cpu().numpy()
    for t in range(64):
        assert (~honest[t]).tolist() == failing_flags(seed, off + t, 7, 2), t
    g = svoc.generate(4, 256, 3, 51, seed=seed)
    honest = g.honest.cpu().numpy()
    for t in range(4):
        assert (~honest[t]).tolist() == failing_flags(seed, t, 256, 51)


def test_sharding_invariance_and_determinism():
    import svoc
    a = svoc.generate(1000, 32, 8, 6, a=[10, 20], b=[15, 30], seed=42)
    b0 = svoc.generate(400, 32, 8, 6, a=[10, 20], b=[15, 30], seed=42, trial_offset=0)
    b1 = svoc.generate(600, 32, 8, 6, a=[10, 20], b=[15, 30], seed=42, trial_offset=400)
    assert torch.equal(a.pred, torch.cat([b0.pred, b1.pred])) and torch.equal(a.honest_words, torch.cat([b0.honest_words, b1.honest_words]))
    c = svoc.generate(1000, 32, 8, 6, a=[10, 20], b=[15, 30], seed=43)
    assert not torch.equal(a.pred, c.pred)


@pytest.mark.parametrize("a,b", [(10.0, 15.0), (10.0, 30.0), (100.0, 100.0), (2.0, 5.0), (0.5, 0.7)])
def test_beta_distribution(a, b):
    import svoc
    from scipy import stats
    g = svoc.generate(20000, 20, 2, 4, a=a, b=b, seed=7)
    x = g.pred.cpu().numpy().astype(np.float64)
    h = g.honest.cpu().numpy()
    assert np.all(h.sum(axis=1) == 16)
    honest, failing = x[h].ravel(), x[~h].ravel()
    m, v = stats.beta.stats(a, b, moments="mv")
    assert abs(honest.mean() - m) < 4 * np.sqrt(v / honest.size) + 1e-4
    assert abs(honest.var() - v) / v < 0.02
    assert stats.kstest(honest[:20000], stats.beta(a, b).cdf).pvalue > 1e-3
    assert stats.kstest(failing[:20000], "uniform").pvalue > 1e-3
    # failing positions are uniform over the oracles (np.random.shuffle in the reference)
    frac = (~h).mean(axis=0)
    assert np.all(np.abs(frac - 0.2) < 0.02)


def test_gaussian_law_and_colluders_and_wsad():
    import svoc
    from scipy import stats
    g = svoc.generate(5000, 20, 2, 4, a=[20.0, 12.0], b=[3.0, 2.0], law="gaussian", seed=3, dtype=torch.float64)
    x, h = g.pred.cpu().numpy(), g.honest.cpu().numpy()
    for d, (mu, sg) in enumerate([(20.0, 3.0), (12.0, 2.0)]):
        hv = x[..., d][h]
        assert abs(hv.mean() - mu) < 0.05 and abs(hv.std() - sg) < 0.05
        assert stats.kstest((hv[:20000] - mu) / sg, "norm").pvalue > 1e-3
        fv = x[..., d][~h]       # np.random.uniform(mu, 5 sigma), gaussian_distribution_for_tests.ipynb:41-44
        lo, hi = min(mu, 5 * sg), max(mu, 5 * sg)
        assert fv.min() >= lo - 1e-6 and fv.max() <= hi + 1e-6
    c = svoc.generate(100, 16, 4, 5, a=10, b=15, failing="colluding", shift=0.15, seed=1)
    x, h = c.pred.cpu().numpy(), c.honest.cpu().numpy()
    mode = (10 - 1) / (10 + 15 - 2)
    assert np.allclose(x[~h], mode + 0.15, atol=1e-6)
    # integer outputs are to_wsad(x) = int(x * 1e6) of the same draws
    f = svoc.generate(500, 7, 2, 2, a=10, b=15, seed=5, dtype=torch.float32).pred
    i = svoc.generate(500, 7, 2, 2, a=10, b=15, seed=5, dtype=torch.int32).pred
    assert torch.equal(i, (f * torch.tensor(1e6, dtype=torch.float32, device=f.device)).to(torch.int32))
    assert int(i.min()) >= 0 and int(i.max()) <= 1_000_000


def test_montecarlo_statistics_match_the_reference_benchmark(golden):
    """benchmark() of the reference (ipynb beta :385-402) re-run at K = 8000..20000 by make_golden.py:
    identification success and 'true consensus reliability' of the on-device pipeline must agree
    within Monte-Carlo noise."""
    import svoc
    for (n, f, a, k, succ, rel) in golden["benchmark_stats"]:
        n, f = int(n), int(f)
        T = 200_000
        g = svoc.generate(T, n, 1, f, a=a, b=a, seed=int(1000 * n + f + a))
        r = svoc.consensus_float(g.pred, f, centre=svoc.CENTRE_MEDIAN, tie_rule=svoc.TIE_LOW_INDEX_KEPT, want_qr=False)
        ident = r.reliable
        success = (ident == g.honest).all(dim=1).float().mean().item() * 100
        # truth median: median over the honest rows
        big = torch.where(g.honest, g.pred[..., 0], torch.full_like(g.pred[..., 0], float("inf")))
        srt = big.sort(dim=1).values
        R = n - f
        truth_med = 0.5 * (srt[:, (R - 1) // 2] + srt[:, R // 2])
        dist = (r.consensus[:, 0] - truth_med).abs().mean().item()
        reliability = 100 * (1 - 2 * dist)
        p = succ / 100
        tol = 4 * 100 * np.sqrt(max(p * (1 - p), 1e-4) / k) + 0.3
        assert abs(success - succ) < tol, (n, f, a, success, succ)
        assert abs(reliability - rel) < 0.35, (n, f, a, reliability, rel)


def test_running_stats_match_numpy():
    import svoc
    g = svoc.generate(30000, 16, 4, 3, a=10, b=15, seed=11, dtype=torch.int32)
    x = g.pred.clone()
    x[::7, 0, 0] = 2_000_000          # some rounds panic with INTERVAL_INPUT
    r = svoc.consensus_fixed(x, 3, want_moments=True)
    st = svoc.RunningStats(16)
    st.update(r)
    st.update(r)
    torch.cuda.synchronize()
    c = st.counts.cpu().numpy()
    status = r.status.cpu().numpy()
    ok = status == 0
    assert c[0] == 2 * len(status) and c[1] == 2 * ok.sum()
    for code in range(13):
        assert c[2 + code] == 2 * (status == code).sum()
    rel1 = r.rel1.cpu().numpy()[ok] * 1e-6
    rel2 = r.rel2.cpu().numpy()[ok] * 1e-6
    h1 = np.bincount(np.clip((rel1 * 256).astype(int), 0, 255), minlength=256)
    h2 = np.bincount(np.clip((rel2 * 256).astype(int), 0, 255), minlength=256)
    assert np.array_equal(c[16:272], 2 * h1) and np.array_equal(c[272:528], 2 * h2)
    unrel = (~r.reliable.cpu().numpy())[ok].sum(axis=0)
    assert np.array_equal(c[528:544], 2 * unrel)
    qsum = r.qr.cpu().numpy()[ok].sum(axis=0)
    assert np.array_equal(c[544:560], 2 * qsum)
    s = st.sums.cpu().numpy()
    np.testing.assert_allclose(s[0], 2 * rel1.sum(), rtol=1e-12)
    np.testing.assert_allclose(s[3], 2 * (rel2 ** 2).sum(), rtol=1e-12)
    # float form: qr sums land in the float buffer
    rf = svoc.consensus_float(g.pred.float() * 1e-6, 3)
    sf = svoc.RunningStats(16)
    sf.update(rf)
    np.testing.assert_allclose(sf.qr_sums(False).cpu().numpy(), rf.qr.double().sum(dim=0).cpu().numpy(), rtol=1e-9)
    assert sf.summary()["rounds"] == 30000


def test_kumaraswamy_and_arctan_gaussian_laws():
    """The two remaining laws of the drafts: Kumaraswamy (only its mode formula is in the reference, ipynb beta
    :77-78) and the arctan-normalised gaussian of gaussian_algorithm_demo.ipynb:46-64."""
    import svoc
    from scipy import stats
    a, b = 3.0, 5.0
    g = svoc.generate(20000, 10, 2, 2, a=a, b=b, law="kumaraswamy", seed=4)
    x, h = g.pred.cpu().numpy().astype(np.float64), g.honest.cpu().numpy()
    hv = x[h].ravel()[:40000]
    cdf = lambda t: 1 - (1 - np.clip(t, 0, 1) ** a) ** b          # noqa: E731  Kumaraswamy CDF
    assert stats.kstest(hv, cdf).pvalue > 1e-3
    assert stats.kstest(x[~h].ravel()[:20000], "uniform").pvalue > 1e-3
    c = svoc.generate(50, 10, 2, 3, a=a, b=b, law="kumaraswamy", failing="colluding", shift=0.1, seed=4)
    mode = ((a - 1) / (a * b - 1)) ** (1 / a)
    assert np.allclose(c.pred.cpu().numpy()[~c.honest.cpu().numpy()], mode + 0.1, atol=1e-6)
    e, sigma = 0.9, 5.0
    g = svoc.generate(20000, 10, 1, 2, a=e, b=sigma, law="arctan_gaussian", seed=6)
    x, h = g.pred.cpu().numpy().astype(np.float64), g.honest.cpu().numpy()
    z = (np.tan(np.pi * (x[h].ravel()[:40000] - 0.5)) - np.tan(np.pi * (e - 0.5))) / sigma     # denormalize, standardise
    assert stats.kstest(z, "norm").pvalue > 1e-3
    assert x.min() > 0 and x.max() < 1


def test_bernoulli_failing_mode():
    """failing_prob > 0: every oracle fails independently with that probability (the failing-fraction sweep of
    BASELINE config 2) — the number of failing oracles per round is Binomial(N, p), positions are uniform, failing
    rows are uniform(0, 1), honest rows keep the law; failing_prob = 0 keeps the exact-count mode."""
    import svoc
    from scipy import stats
    T, N, D = 200_000, 7, 2
    for p in (0.1, 0.4):
        g = svoc.generate(T, N, D, 2, a=10.0, b=15.0, seed=21, failing_prob=p)
        h = g.honest.cpu().numpy()
        k = (~h).sum(axis=1)
        expect = stats.binom.pmf(np.arange(N + 1), N, p) * T
        obs = np.bincount(k, minlength=N + 1)
        assert stats.chisquare(obs[expect > 5], expect[expect > 5] * obs[expect > 5].sum() / expect[expect > 5].sum()).pvalue > 1e-4
        per_pos = (~h).mean(axis=0)
        assert np.all(np.abs(per_pos - p) < 5 * np.sqrt(p * (1 - p) / T))
        x = g.pred.cpu().numpy().astype(np.float64)
        assert stats.kstest(x[~h].ravel()[:50000], "uniform").pvalue > 1e-4
        assert stats.kstest(x[h][:, 0][:50000], stats.beta(10, 15).cdf).pvalue > 1e-4
    g0 = svoc.generate(1000, N, D, 2, a=10.0, b=15.0, seed=21)
    assert torch.all((~g0.honest).sum(dim=1) == 2)
    with pytest.raises(RuntimeError):
        svoc.generate(10, N, D, 2, failing_prob=1.5)


def test_replacement_candidates_and_last_k_window_match_numpy():
    """running per-oracle "not reliable" counts, the last-k window and the proposal rule, against a
    numpy restatement on the same per-round flags (OracleInfo.reliable, contract.cairo:345-363)."""
    import svoc
    N, D, F, k = 16, 4, 3, 500
    g = svoc.generate(6000, N, D, F, a=10, b=15, seed=31, dtype=torch.int32)
    x = g.pred.clone()
    x[:, 5, :] = torch.randint(0, 1_000_000, (6000, D), device=x.device, dtype=torch.int32)    # oracle 5 always noisy
    x[4000:, 9, :] = 990_000                                                                     # oracle 9 breaks late
    x[::13, 0, 0] = 2_000_000                                                                    # some rounds revert
    st = svoc.RunningStats(N, window=k)
    chunks = [x[:2500], x[2500:2501], x[2501:5200], x[5200:]]        # uneven batches, one of a single round
    masks, status = [], []
    for c in chunks:
        r = svoc.consensus_fixed(c, F, want_moments=False)
        st.update(r)
        masks.append(r.reliable.cpu().numpy())
        status.append(r.status.cpu().numpy())
    masks, status = np.concatenate(masks), np.concatenate(status)
    ok = status == 0
    unrel = ~masks[ok]
    assert np.array_equal(st.unreliable_counts().cpu().numpy(), unrel.sum(axis=0))
    assert st.window_rounds() == k
    assert np.array_equal(st.window_unreliable_counts().cpu().numpy(), unrel[-k:].sum(axis=0))
    frac = unrel.mean(axis=0)
    want = [i for i in np.argsort(-frac, kind="stable") if frac[i] >= 0.5]
    assert st.replacement_candidates(0.5) == want and want[0] == 5 and 9 not in want
    frac_k = unrel[-k:].mean(axis=0)
    want_k = [i for i in np.argsort(-frac_k, kind="stable") if frac_k[i] >= 0.5]
    assert st.replacement_candidates(0.5, recent=True) == want_k and 9 in want_k and 5 in want_k
    small = svoc.RunningStats(N, window=k)
    small.update(svoc.consensus_fixed(x[:40], F, want_moments=False))     # fewer accepted rounds than the window holds
    assert small.window_rounds() == int((status[:40] == 0).sum())


def test_rounds_to_proposal_to_vote_to_replacement():
    """End to end over f1: consensus rounds -> RunningStats -> replacement candidate -> update_proposition -> votes
    of the admins -> the oracle's address is swapped and the vote state reset (contract.cairo:547-580, 661-738)."""
    import svoc
    from svoc.contract import OracleConsensusNDS
    N, D, F = 7, 2, 2
    g = svoc.generate(4000, N, D, F, a=10, b=15, seed=77, dtype=torch.int32)
    x = g.pred.clone()
    x[:, 3, :] = torch.randint(0, 1_000_000, (4000, D), device=x.device, dtype=torch.int32)     # oracle 3 is dead: pure noise
    st = svoc.RunningStats(N, window=1000)
    st.update(svoc.consensus_fixed(x, F, want_moments=False))
    cands = st.replacement_candidates(0.5, recent=True)
    assert cands and cands[0] == 3
    admins, oracles = ["Akashi", "Ozu", "Higuchi"], [f"oracle_{i:02d}" for i in range(N)]
    c = OracleConsensusNDS(admins, True, 2, F, True, 0, D, oracles)
    c.set_caller(admins[0])
    c.update_proposition((cands[0], "oracle_new"))           # the proposer votes for itself
    assert c.get_oracle_list()[3] == "oracle_03"
    c.set_caller(admins[1])
    c.vote_for_a_proposition(0, True)                        # second vote: the required majority of 2 is reached
    assert c.get_oracle_list()[3] == "oracle_new"
    assert c.get_replacement_propositions() == [None] * 3 and not any(any(v) for v in c.vote_matrix)
    # the replaced oracle commits under its new address
    c.set_caller("oracle_new")
    c.update_prediction([400000, 300000])
    assert c.n_active_oracles == 1


@pytest.mark.parametrize("n,d,f", [(7, 2, 2), (33, 5, 6), (20, 3, 4), (256, 128, 51), (70, 1000, 9), (300, 300, 40)])
def test_generator_places_every_component_where_its_flags_say(n, d, f):
    """fill_components (generate_device.cuh) decodes (round, oracle, dim) from the element index with multiply-high
    divisions inside their exact range and plain divisions outside it (70 x 1000 and 300 x 300 leave it): a wrong
    decode would put the colluders' constant rows, or the per-dimension parameters, in the wrong place."""
    import svoc
    a = [10.0 + 3.0 * (k % 4) for k in range(min(d, 4))]
    b = [15.0 + 2.0 * (k % 4) for k in range(min(d, 4))]
    rounds = 6
    g = svoc.generate(rounds, n, d, f, a=a, b=b, failing="colluding", shift=0.2, seed=11)
    x, h = g.pred.cpu().numpy().astype(np.float64), g.honest.cpu().numpy()
    assert x.shape == (rounds, n, d) and np.all(h.sum(axis=1) == n - f)
    centre = np.array([min(max((a[k % len(a)] - 1) / (a[k % len(a)] + b[k % len(a)] - 2) + 0.2, 0.0), 1.0) for k in range(d)])
    assert np.allclose(x[~h], np.broadcast_to(centre, x[~h].shape), atol=1e-6)          # colluders: mode + shift per dimension
    hon = x[h].reshape(-1, d)
    assert not np.any(np.isclose(hon, centre, atol=1e-9).all(axis=1))
    # honest components follow THEIR dimension's beta law: the mean per dimension separates the four parameter pairs
    if hon.shape[0] * (d // len(a)) >= 2000:
        for k in range(len(a)):
            m = hon[:, k::len(a)].mean()
            assert abs(m - a[k] / (a[k] + b[k])) < 0.01, (k, m)
    # and the same draws come out of a call that starts later (another decomposition into blocks)
    g2 = svoc.generate(rounds - 2, n, d, f, a=a, b=b, failing="colluding", shift=0.2, seed=11, trial_offset=2)
    assert torch.equal(g.pred[2:], g2.pred)
