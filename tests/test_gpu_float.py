"""GPU parity, float form: libsvoc vs the reference notebook's own outputs (golden fixtures)
and vs the float64 CPU oracle.  Tolerances: 1e-6 relative in fp64, 1e-4 in fp32 (north star)."""
import numpy as np
import pytest
import torch

from helpers import masks_equal_up_to_cut_ties, ranks_equal_up_to_ties, seed_of

pytestmark = pytest.mark.gpu

RTOL = {torch.float64: 1e-6, torch.float32: 1e-4}


def _np(r):
    d = {k: (getattr(r, k).cpu().numpy() if getattr(r, k) is not None else None)
         for k in ("consensus", "rel1", "rel2", "status", "essence1", "qr", "skew", "kurt")}
    d["reliable"] = r.reliable.cpu().numpy()
    return d


def test_notebook_golden(golden):
    """identify_failing_oracles + restricted_median of the reference, on its own seeded draws."""
    import svoc
    for (n, d, f, t) in golden["shapes"]:
        key = f"n{n}_d{d}_f{f}"
        pred = golden[key + "_pred"]
        for dt in (torch.float64, torch.float32):
            x = torch.from_numpy(pred).to(dt).cuda()
            r = _np(svoc.consensus_float(x, int(f), centre=svoc.CENTRE_MEDIAN, tie_rule=svoc.TIE_LOW_INDEX_KEPT))
            # the reference's mask: exact, except that oracles whose deviation sits within rounding of the
            # cut may swap sides in fp32
            safe = masks_equal_up_to_cut_ties(r["reliable"], golden[key + "_mask"], r["qr"].astype(np.float64), n, f, 10 * RTOL[dt])
            # measured: no round differs in fp64; in fp32 (inputs rounded to fp32) a few per cent of these tiny batches
            print(f"{key} {dt}: {1 - safe.mean():.4f} of the rounds excluded")
            assert safe.mean() >= (0.999 if dt == torch.float64 else 0.95)
            np.testing.assert_allclose(r["consensus"][safe], golden[key + "_consensus"][safe], rtol=RTOL[dt], atol=1e-7)
            # rank_array (ipynb :342-352): rank 0 = largest deviation
            ranks = (n - 1) - np.argsort(np.argsort(r["qr"], axis=1, kind="stable"), axis=1, kind="stable")
            if dt == torch.float64:
                for k in range(int(t)):
                    # compare up to (near-)ties: sqrt in the reference can merge or split 1-ulp neighbours
                    dev = np.round(np.sqrt(r["qr"][k]), 12)
                    assert ranks_equal_up_to_ties(ranks[k], golden[key + "_ranks"][k], dev), (key, k)


def test_expected_reliability_kats(golden):
    """The two printed known answers of the notebook (cells 14-15): 0.762 and 0.807."""
    import svoc
    for k in ("kat1d", "kat2d"):
        a = torch.from_numpy(golden[k + "_wsad"] * 1e-6).reshape(1, -1).cuda()
        v = float(svoc.expected_reliability(a)[0])
        assert f"{v:0.3f}" == f"{float(golden[k + '_printed']):0.3f}"
        assert abs(v - float(golden[k + "_expected_reliability"])) < 1e-12
    for (n, d, f, t) in golden["shapes"]:
        key = f"n{n}_d{d}_f{f}"
        a = torch.from_numpy(golden[key + "_pred"]).cuda()
        v = svoc.expected_reliability(a).cpu().numpy()
        np.testing.assert_allclose(v, golden[key + "_expected_reliability"], rtol=1e-12, atol=1e-12)


SHAPES = [(7, 2, 2), (7, 6, 2), (7, 1, 2), (20, 2, 4), (32, 8, 6), (33, 5, 6), (64, 16, 12), (100, 3, 20), (256, 128, 51)]


@pytest.mark.parametrize("N,D,F", SHAPES)
@pytest.mark.parametrize("dt", [torch.float64, torch.float32])
@pytest.mark.parametrize("mode", ["contract", "unconstrained", "notebook", "mean"])
def test_float_vs_oracle(N, D, F, dt, mode):
    import svoc
    from oracle import cref
    rng = np.random.default_rng(seed_of(N, D, F, mode))
    T = 300 if N * D <= 1024 else (30 if N * D <= 8192 else 4)
    if mode == "unconstrained":
        x = rng.normal(20.0, 3.0, (T, N, D))
    else:
        x = rng.beta(10, 15, (T, N, D))
        bad = rng.random((T, N)) < 0.2
        x = np.where(bad[:, :, None], rng.random((T, N, D)), x)
    x = x.astype(np.float32).astype(np.float64)   # identical inputs for both precisions
    kw = {"contract": dict(constrained=True, centre=0, tie=0),
          "unconstrained": dict(constrained=False, max_spread=10.0, centre=0, tie=0),
          "notebook": dict(constrained=True, centre=1, tie=1),
          "mean": dict(constrained=True, centre=2, tie=0)}[mode]
    ora = cref.float_batch(x, F, want_moments=True, nthreads=0, **kw)
    r = _np(svoc.consensus_float(torch.from_numpy(x).to(dt).cuda(), F, constrained=kw["constrained"],
                                 max_spread=kw.get("max_spread", 0.0), centre=kw["centre"], tie_rule=kw["tie"],
                                 want_moments=True))
    rt = RTOL[dt]
    np.testing.assert_allclose(r["essence1"], ora["essence1"], rtol=rt, atol=1e-7)
    np.testing.assert_allclose(r["qr"], ora["qr"], rtol=10 * rt, atol=1e-9)
    np.testing.assert_allclose(r["rel1"], ora["rel1"], rtol=rt, atol=10 * rt)
    same = masks_equal_up_to_cut_ties(r["reliable"], ora["reliable"], ora["qr"], N, F, 20 * rt)
    # identical inputs in both precisions: fp64 agrees on every round; fp32 may flip an oracle whose risk is within
    # rounding of the cut (measured: none on these batches; test_gpu_bulk_parity.py prints the bulk fractions)
    print(f"{N}x{D} {dt} {mode}: {1 - same.mean():.4f} of the rounds excluded")
    assert same.mean() >= (0.999 if dt == torch.float64 else 0.97)
    safe = same
    np.testing.assert_allclose(r["consensus"][safe], ora["consensus"][safe], rtol=rt, atol=1e-7)
    np.testing.assert_allclose(r["rel2"][safe], ora["rel2"][safe], rtol=rt, atol=10 * rt)
    mt = 1e-6 if dt == torch.float64 else 2e-2   # z-scores amplify fp32 rounding
    np.testing.assert_allclose(r["skew"][safe], ora["skew"][safe], rtol=mt, atol=mt)
    np.testing.assert_allclose(r["kurt"][safe], ora["kurt"][safe], rtol=mt, atol=10 * mt)


@pytest.mark.parametrize("N,D", [(7, 2), (20, 3), (64, 16), (100, 5)])
def test_single_kept_oracle_is_valid_for_median_and_mean(N, D):
    """F = N - 1: restricted_median over ONE kept oracle is that oracle's vector (ipynb beta :369-371); only the
    contract's smooth median (sorted[mid - 1], math.cairo:117) needs two rows and reports status 5."""
    import svoc
    from oracle import cref
    rng = np.random.default_rng(seed_of("single", N, D))
    x = rng.beta(10, 15, (50, N, D))
    for centre, tie in ((1, 1), (2, 0)):
        ora = cref.float_batch(x, N - 1, centre=centre, tie=tie, want_moments=False)
        r = _np(svoc.consensus_float(torch.from_numpy(x).cuda(), N - 1, centre=centre, tie_rule=tie))
        assert np.all(r["status"] == 0)
        assert np.array_equal(r["reliable"], ora["reliable"]) and np.all(r["reliable"].sum(axis=1) == 1)
        np.testing.assert_allclose(r["consensus"], ora["consensus"], rtol=1e-12)
        np.testing.assert_allclose(r["consensus"], x[r["reliable"]], rtol=1e-12)
        np.testing.assert_allclose(r["rel2"], ora["rel2"], rtol=1e-9, atol=1e-9)
    r = _np(svoc.consensus_float(torch.from_numpy(x).cuda(), N - 1, centre=0))
    assert np.all(r["status"] == 5)


def test_ties_follow_the_rule():
    """Identical rows (colluders) tie exactly; the contract keeps the higher index, numpy the lower."""
    import svoc
    x = np.array([[[0.5, 0.5], [0.4, 0.4], [0.6, 0.6], [0.4, 0.4], [0.6, 0.6], [0.5, 0.5], [0.9, 0.1]]])
    t = torch.from_numpy(x).cuda()
    hi = svoc.consensus_float(t, 2, tie_rule=svoc.TIE_HIGH_INDEX_KEPT).reliable[0].tolist()
    lo = svoc.consensus_float(t, 2, tie_rule=svoc.TIE_LOW_INDEX_KEPT).reliable[0].tolist()
    # essence (0.5, 0.45): rows 2 and 4 tie for the second-largest qr; row 6 is the largest.
    assert hi == [True, True, False, True, True, True, False]
    assert lo == [True, True, True, True, False, True, False]


def test_full_size_properties_c4():
    """BASELINE config 4 shape (N=256, D=128, fp32, colluders): translation equivariance of the
    consensus, invariance of qr, and exact reliable count, on a batch far larger than the oracle runs."""
    import svoc
    g = torch.Generator(device="cuda").manual_seed(9)
    T, N, D, F = 2048, 256, 128, 51
    x = torch.rand((T, N, D), generator=g, device="cuda") * 0.25 + 0.25
    x[:, :F, :] = x[:, :1, :] + 0.125            # F colluders share one vector
    r = svoc.consensus_float(x, F)
    assert torch.all(r.reliable.sum(dim=1) == N - F)
    assert torch.all(~r.reliable[:, :F])          # the colluders are the outliers here
    shift = 0.25                                  # exact in binary: results must shift exactly
    r2 = svoc.consensus_float(x + shift, F)
    assert torch.equal(r.reliable_words, r2.reliable_words)
    torch.testing.assert_close(r2.consensus, r.consensus + shift, rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(r2.qr, r.qr, rtol=1e-4, atol=1e-7)
    kept = x[:, F:, :]
    torch.testing.assert_close(r.consensus, 0.5 * (kept.sort(dim=1).values[:, (N - F) // 2 - 1] + kept.sort(dim=1).values[:, (N - F) // 2]))


@pytest.mark.parametrize("N,D,F", [(256, 128, 51), (100, 3, 20), (65, 7, 13)])
def test_rank_output_and_threshold_search_agree(N, D, F):
    """The wide variant finds the reliable set either by ranking every row (when ranks are requested) or
    by a histogram threshold search: both must give the same flags, and the ranks must order qr."""
    import svoc
    g = svoc.generate(64, N, D, F, a=10, b=15, failing="colluding", shift=0.02, seed=3)   # colluders tie exactly
    for tie in (svoc.TIE_HIGH_INDEX_KEPT, svoc.TIE_LOW_INDEX_KEPT):
        a = svoc.consensus_float(g.pred, F, tie_rule=tie, want_rank=True)
        b = svoc.consensus_float(g.pred, F, tie_rule=tie, want_rank=False)
        assert torch.equal(a.reliable_words, b.reliable_words) and torch.equal(a.consensus, b.consensus)
        rank = a.rank.long()
        assert torch.equal(rank.sort(dim=1).values, torch.arange(N, device=rank.device).expand(64, N))
        assert torch.equal(a.reliable, rank < N - F)
        q_sorted = torch.gather(a.qr, 1, rank.argsort(dim=1))
        assert torch.all(q_sorted[:, 1:] >= q_sorted[:, :-1])
        # among equal qr the tie rule orders by index
        idx = rank.argsort(dim=1)
        same = q_sorted[:, 1:] == q_sorted[:, :-1]
        step = idx[:, 1:] - idx[:, :-1]
        assert torch.all(step[same] < 0) if tie == svoc.TIE_HIGH_INDEX_KEPT else torch.all(step[same] > 0)
        assert same.any()


def _selection_inputs(kind, T, N, D, rng):
    """Rounds built to reach every branch of the single-warp row selection of the wide variant."""
    if kind == "cluster":        # 90 % of the rows within 1e-4 of each other, the rest far away: the cut bucket of
        x = 0.5 + 1e-4 * rng.standard_normal((T, N, D))   # the first level holds > 32 distinct risks -> second level
        far = rng.random((T, N)) < 0.1
        x = np.where(far[:, :, None], rng.random((T, N, D)), x)
    elif kind == "ties":         # groups of identical rows of random sizes: exact ties of every size at the cut
        base = rng.beta(10, 15, (T, 8, D))
        x = base[:, rng.integers(0, 8, N), :]
    elif kind == "two_values":   # every risk is one of two values
        a, b = rng.random((T, 1, D)), rng.random((T, 1, D))
        x = np.where(rng.random((T, N, 1)) < 0.5, a, b)
    elif kind == "constant":     # all risks equal: the index order alone decides
        x = np.broadcast_to(rng.random((T, 1, D)), (T, N, D)).copy()
    elif kind == "spread":       # risks over many orders of magnitude
        x = 0.5 + rng.standard_normal((T, N, D)) * (10.0 ** rng.uniform(-6, -1, (T, N, 1)))
    else:
        raise ValueError(kind)
    return np.clip(x, 0.0, 1.0)


@pytest.mark.parametrize("kind", ["cluster", "ties", "two_values", "constant", "spread"])
@pytest.mark.parametrize("N,D", [(256, 128), (130, 40), (512, 8)])
def test_row_selection_matches_all_pairs_ranking(kind, N, D):
    """want_rank=True ranks every row by all-pairs counting; want_rank=False runs the bisection-based selection.
    Same flags, consensus and reliabilities for both tie rules, both dtypes and cuts from 0 to N-2 dropped rows."""
    import svoc
    rng = np.random.default_rng(seed_of(kind, N, D))
    T = 24
    x = _selection_inputs(kind, T, N, D, rng)
    idx = torch.arange(N, device="cuda")
    for F in (0, 1, N // 5, N // 2, N - 33, N - 2):
        for tie in (svoc.TIE_HIGH_INDEX_KEPT, svoc.TIE_LOW_INDEX_KEPT):
            t = torch.from_numpy(x.astype(np.float32)).cuda()
            a = svoc.consensus_float(t, F, tie_rule=tie, want_rank=True)
            b = svoc.consensus_float(t, F, tie_rule=tie, want_rank=False)
            # each result against its OWN risks (the two calls may run different kernels, whose fp32 risk sums differ in
            # the last bit: 'two_values' and 'constant' rounds tie exactly in real arithmetic, so rounding alone orders them)
            for res in (a, b):
                q = res.qr
                before = (q[:, None, :] < q[:, :, None]) | ((q[:, None, :] == q[:, :, None]) & (
                    (idx[None, None, :] > idx[None, :, None]) if tie == svoc.TIE_HIGH_INDEX_KEPT else (idx[None, None, :] < idx[None, :, None])))
                assert torch.equal(res.reliable, before.sum(dim=2) < N - F), (kind, F, tie, "f32")
            assert torch.equal(a.status, b.status)
            if torch.equal(a.qr, b.qr):
                assert torch.equal(a.reliable_words, b.reliable_words) and torch.equal(a.consensus, b.consensus)
                torch.testing.assert_close(a.rel2, b.rel2, rtol=1e-5, atol=1e-6)
            assert torch.all(b.reliable.sum(dim=1) == N - F)
            # fixed-point form on the same rounds: risks are 64-bit integers, selected through their float images
            ti = torch.from_numpy((x * 1e6).astype(np.int32)).cuda()
            ai = svoc.consensus_fixed(ti, F, constrained=True, tie_rule=tie, want_rank=True)
            bi = svoc.consensus_fixed(ti, F, constrained=True, tie_rule=tie, want_rank=False)
            assert torch.equal(ai.reliable_words, bi.reliable_words), (kind, F, tie, "i32")
            assert torch.equal(ai.consensus, bi.consensus) and torch.equal(ai.rel2, bi.rel2) and torch.equal(ai.status, bi.status)


def _equal_image_rounds(T, N, D, rng):
    x = np.full((T, N, D), 900000, dtype=np.int64)
    x[:, ::2, :] = 100000                          # centre 500000: every risk is about 128 * 0.16 wsad = 2.05e7 > 2^24
    x[:, :, 0] += rng.integers(0, 3, (T, N))       # risks that differ by one or two units
    return x


def test_row_selection_with_equal_float_images_of_integer_risks():
    """Fixed-point risks above 2^24 that differ by a few units share a float image: more than 32 such rows at the
    cut cannot be settled on the images, and the kernel must fall back to exact all-pairs ranking."""
    import svoc
    T, N, D, F = 16, 256, 128, 100
    x = _equal_image_rounds(T, N, D, np.random.default_rng(5))
    t = torch.from_numpy(x.astype(np.int32)).cuda()
    for tie in (svoc.TIE_HIGH_INDEX_KEPT, svoc.TIE_LOW_INDEX_KEPT):
        a = svoc.consensus_fixed(t, F, constrained=True, tie_rule=tie, want_rank=True)
        b = svoc.consensus_fixed(t, F, constrained=True, tie_rule=tie, want_rank=False)
        assert torch.equal(a.reliable_words, b.reliable_words)
        assert torch.equal(a.consensus, b.consensus) and torch.equal(a.rel2, b.rel2)
        q0 = a.qr[0]
        assert int(q0.min()) > (1 << 24) and len(torch.unique(q0)) > len(torch.unique(q0.float()))


@pytest.mark.parametrize("N,D", [(64, 16), (50, 8), (40, 32), (256, 128), (20, 4)])
def test_identical_rows_tie_exactly_in_every_kernel(N, D):
    """Colluders submit identical vectors: their risks must be BIT-identical (every row summed in the same order),
    so that a cut inside the group is decided by the index rule alone — the contract keeps the higher indices
    (sort.cairo:96-102), numpy's stable argsort the lower ones — exactly as the float64 oracle does."""
    import svoc
    from oracle import cref
    rng = np.random.default_rng(seed_of("identical", N, D))
    T, G = 200, N // 3
    x = rng.beta(10, 15, (T, N, D))
    pos = np.stack([rng.permutation(N)[:G] for _ in range(T)])
    shared = np.clip(rng.beta(10, 15, (T, 1, D)) + 0.3, 0, 1)      # far from the honest rows: the group has the largest risk
    for t in range(T):
        x[t, pos[t]] = shared[t]
    x = x.astype(np.float32)
    F = G // 2                                          # the cut falls inside the group of identical rows
    for tie in (0, 1):
        ora = cref.float_batch(x.astype(np.float64), F, tie=tie, want_moments=False, nthreads=0)
        r = svoc.consensus_float(torch.from_numpy(x).cuda(), F, tie_rule=tie)
        q = r.qr.cpu().numpy()
        for t in range(0, T, 17):
            assert len(np.unique(q[t, pos[t]])) == 1
        grp = ora["reliable"][np.arange(T)[:, None], pos]
        inside = ~np.all(grp, axis=1) & np.any(grp, axis=1)
        assert inside.mean() > 0.5                      # the test is about cuts inside the tie group
        same = (r.reliable.cpu().numpy() == ora["reliable"]).all(axis=1)
        assert same[inside].mean() >= 0.99, (N, D, tie, same[inside].mean())
