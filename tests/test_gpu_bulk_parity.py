"""GPU bulk parity on the BENCHMARK configurations: tens of thousands of rounds per shape, drawn by the bench
generator itself, through the C ABI and against oracle/svoc_ref.c — with the option set bench.py times
(no qr / essence / moments outputs) and with every output on.

The short oracle comparisons of test_gpu_fixed.py / test_gpu_float.py see 4-6 rounds of the 256x128 shape; the
kernels have data-dependent slow paths (crowded windows, rescue scans, re-keyed selections) that only a bulk
run reaches.  Reference lines: contract.cairo:442-503 (round), sort.cairo:96-102 (tie order), ipynb beta
:354-371 (float form).  Tolerances: bit-exact for the fixed form; 1e-4 (fp32) / 1e-6 (fp64) relative for the
float form, reliable masks exact except for oracles whose risk lies within rounding of the cut."""
import numpy as np
import pytest
import torch

from helpers import (W, assert_fixed_equal, gen_fixed, masks_equal_up_to_cut_ties, result_to_numpy, seed_of)

pytestmark = pytest.mark.gpu

RTOL = {torch.float64: 1e-6, torch.float32: 1e-4}


def _float_compare(r, ora, N, F, dt, label, min_same, full):
    rt = RTOL[dt]
    assert np.all(r["status"] == 0), f"{label}: status"
    np.testing.assert_allclose(r["rel1"], ora["rel1"], rtol=rt, atol=10 * rt, err_msg=f"{label}: rel1")
    same = masks_equal_up_to_cut_ties(r["reliable"], ora["reliable"], ora["qr"], N, F, 20 * rt)
    excluded = 1.0 - same.mean()
    print(f"{label}: {excluded:.5f} of the rounds excluded (mask differs within rounding of the cut)")
    assert same.mean() >= min_same, f"{label}: only {same.mean():.4f} of the masks identical"
    np.testing.assert_allclose(r["consensus"][same], ora["consensus"][same], rtol=rt, atol=1e-7, err_msg=f"{label}: consensus")
    np.testing.assert_allclose(r["rel2"][same], ora["rel2"][same], rtol=rt, atol=10 * rt, err_msg=f"{label}: rel2")
    if full:
        np.testing.assert_allclose(r["essence1"], ora["essence1"], rtol=rt, atol=1e-7, err_msg=f"{label}: essence1")
        np.testing.assert_allclose(r["qr"], ora["qr"], rtol=10 * rt, atol=1e-9, err_msg=f"{label}: qr")


def _words_to_bool(words, N):
    idx = np.arange(N)
    return ((words[:, idx // 32].astype(np.int64) >> (idx % 32)) & 1).astype(bool)


def _np_result(r, N):
    d = {k: (getattr(r, k).cpu().numpy() if getattr(r, k) is not None else None)
         for k in ("consensus", "rel1", "rel2", "status", "essence1", "qr")}
    d["reliable"] = _words_to_bool(r.reliable_words.cpu().numpy(), N)
    return d


# ------------------------------------------------------------------------------------------ config 4
@pytest.mark.parametrize("F_used,T,min_same", [(51, 20000, 0.999), (40, 4000, 0.999), (60, 4000, 0.995)])
def test_config4_bulk_vs_oracle(F_used, T, min_same):
    """BASELINE config 4 (256 x 128 fp32, 51 colluders at mode + 0.15), rounds from svoc.generate exactly as
    bench.py draws them.  F_used = 51: the cut separates honest rows from colluders; 40: the cut falls INSIDE
    the 51-way tie of identical colluder rows (the tie rule decides, exactly, in both precisions); 60: the cut
    falls among the honest rows, a contested cut with near-continuous risks."""
    import svoc
    from oracle import cref
    N, D = 256, 128
    g = svoc.generate(T, N, D, 51, a=10.0, b=15.0, failing="colluding", shift=0.15, seed=1234)
    x = g.pred
    xd = x.double().cpu().numpy()
    for tie in ((0, 1) if F_used == 40 else (0,)):
        ora = cref.float_batch(xd, F_used, want_moments=False, tie=tie, nthreads=0)
        # the option set of the benchmark's timed loop
        r = _np_result(svoc.consensus_float(x, F_used, tie_rule=tie, want_moments=False, want_qr=False, want_essence=False), N)
        _float_compare(r, ora, N, F_used, torch.float32, f"c4 F={F_used} tie={tie} bench options", min_same, False)
        # every output on
        r = _np_result(svoc.consensus_float(x, F_used, tie_rule=tie, want_moments=False, want_qr=True, want_essence=True), N)
        _float_compare(r, ora, N, F_used, torch.float32, f"c4 F={F_used} tie={tie} all outputs", min_same, True)
    if F_used == 51:
        # 20 % colluders at mode + 0.15 pull the first-pass median towards themselves: their common risk
        # (~1.6) sits inside the upper tail of the honest risks (1.3 +- 0.15), so the cut usually falls INSIDE
        # the 51-way tie and the tie rule picks which colluders stay — the benchmark's own rounds are contested
        honest = g.honest.cpu().numpy()
        kept_colluders = (r["reliable"] & ~honest).sum(axis=1)
        print(f"c4: rounds with colluders kept by the tie rule: {(kept_colluders > 0).mean():.3f}, mean {kept_colluders.mean():.2f}")
        assert (kept_colluders > 0).mean() > 0.5


def test_config4_bulk_fixed_point_vs_oracle():
    """The same 256 x 128 colluding rounds as wsad integers through the fixed form (int32), bit-exact."""
    import svoc
    from oracle import cref
    T, N, D, F = 6000, 256, 128, 51
    g = svoc.generate(T, N, D, F, a=10.0, b=15.0, failing="colluding", shift=0.15, seed=77, dtype=torch.int32)
    x = g.pred.cpu().numpy().astype(np.int64)
    ora = cref.fixed_batch(x, F, True, 0, False, nthreads=0)
    r = result_to_numpy(svoc.consensus_fixed(g.pred, F, want_moments=False))
    assert_fixed_equal(r, ora, T, want_moments=False, label="c4 fixed all outputs")
    r = svoc.consensus_fixed(g.pred, F, want_moments=False, want_qr=False, want_essence=False)
    st = r.status.cpu().numpy()
    assert np.array_equal(st, ora["status"])
    assert np.array_equal(r.consensus.cpu().numpy(), ora["consensus"]) and np.array_equal(r.rel2.cpu().numpy(), ora["rel2"])
    assert np.array_equal(r.rel1.cpu().numpy(), ora["rel1"]) and np.array_equal(r.reliable.cpu().numpy(), ora["reliable"])


def test_config4_shape_in_float64_and_int64():
    """N > 64 in the 64-bit element types — the reference's own float precision (np.median on float64, ipynb beta
    :354-371; the 1e-6 tolerance is stated for it) and unconstrained int64 contract values (contract.cairo:370-434) —
    run the wide2 kernel too: 10 000 rounds of the config-4 shape against the oracle."""
    import svoc
    from oracle import cref
    T, N, D, F = 10_000, 256, 128, 51
    assert svoc.kernel_variant(N, D, torch.float64) == "wide2"
    g = svoc.generate(T, N, D, F, a=10.0, b=15.0, failing="colluding", shift=0.15, seed=99, dtype=torch.float64)
    xd = g.pred.cpu().numpy()
    for centre, tie in ((1, 1), (0, 0)):
        ora = cref.float_batch(xd, F, centre=centre, tie=tie, want_moments=False, nthreads=0)
        r = _np_result(svoc.consensus_float(g.pred, F, centre=centre, tie_rule=tie, want_moments=False), N)
        _float_compare(r, ora, N, F, torch.float64, f"c4 shape fp64 centre={centre}", 0.9999, True)
    gi = svoc.generate(3000, N, D, F, a=10.0, b=15.0, failing="colluding", shift=0.15, seed=98, dtype=torch.int64)
    xi = gi.pred.cpu().numpy()
    ora = cref.fixed_batch(xi, F, True, 0, True, nthreads=0)
    r = result_to_numpy(svoc.consensus_fixed(gi.pred, F, want_moments=True))
    assert_fixed_equal(r, ora, 3000, label="c4 shape int64")
    wide = (xi.astype(np.int64) - 500_000) * 37 + 20_000_000          # unconstrained values around 20 +- 18 (wsad)
    ora = cref.fixed_batch(wide, F, False, 10 * W, False, nthreads=0)
    r = result_to_numpy(svoc.consensus_fixed(torch.from_numpy(wide).cuda(), F, constrained=False, max_spread=10 * W, want_moments=False))
    assert_fixed_equal(r, ora, 3000, want_moments=False, label="c4 shape int64 unconstrained")


# ------------------------------------------------------------------------------------------ config 3
def _c3_rounds(T, N, D, F, seed):
    """bench.py's config-3 draw (int32 wsad, beta(10,15) honest, F uniform failing) with blocks of rounds replaced
    by forced ties at the medians and at the cut, colluders, variance-0/1 panics, reliability-range panics and
    out-of-interval inputs ('include forced-tie and panic-triggering rounds')."""
    import svoc
    g = svoc.generate(T, N, D, F, a=10.0, b=15.0, failing="uniform", seed=seed, dtype=torch.int32)
    x = g.pred.cpu().numpy().astype(np.int64)
    rng = np.random.default_rng(seed)
    kinds = [("coarse", 0.06), ("colluders", 0.06), ("tiny", 0.03), ("extreme", 0.02), ("outside", 0.02)]
    pos = 0
    for kind, frac in kinds:
        n = int(T * frac)
        x[pos:pos + n] = gen_fixed(kind, rng, n, N, D)
        pos += n
    perm = rng.permutation(T)
    return x[perm]


@pytest.mark.parametrize("N,D,F,T", [(64, 16, 12, 100_000), (32, 8, 6, 100_000), (7, 2, 2, 100_000)])
def test_fixed_bulk_vs_oracle(N, D, F, T):
    """BASELINE configs 3, 5 (shape) and 2 in the fixed form: 1e5 rounds each, bit-exact on every field, with
    every output on (incl. skew / kurt) and with the benchmark's option set."""
    import svoc
    from oracle import cref
    x = _c3_rounds(T, N, D, F, seed_of("bulk", N, D))
    ora = cref.fixed_batch(x, F, True, 0, True, nthreads=0)
    seen = set(np.unique(ora["status"]).tolist())
    assert {0, 1, 8}.issubset(seen), seen          # ok, interval panic, variance panic all present
    for dt in (torch.int32, torch.int64):
        t = torch.from_numpy(x.astype(np.int32 if dt == torch.int32 else np.int64)).cuda()
        r = result_to_numpy(svoc.consensus_fixed(t, F, want_moments=True))
        assert_fixed_equal(r, ora, T, label=f"bulk {N}x{D} {dt} all outputs")
    # the benchmark's option set: statuses 8 / 9 (moments) cannot appear, every earlier panic must
    ora0 = cref.fixed_batch(x, F, True, 0, False, nthreads=0)
    t = torch.from_numpy(x.astype(np.int32)).cuda()
    r = svoc.consensus_fixed(t, F, want_moments=False, want_qr=False, want_essence=False)
    st = r.status.cpu().numpy()
    assert np.array_equal(st, ora0["status"])
    done = np.isin(st, [0])
    assert np.array_equal(r.consensus.cpu().numpy()[done], ora0["consensus"][done])
    assert np.array_equal(r.rel1.cpu().numpy()[done], ora0["rel1"][done]) and np.array_equal(r.rel2.cpu().numpy()[done], ora0["rel2"][done])
    assert np.array_equal(r.reliable.cpu().numpy()[done], ora0["reliable"][done])


@pytest.mark.parametrize("N,D,F,T", [(32, 8, 6, 100_000), (7, 2, 2, 100_000), (64, 16, 12, 30_000)])
@pytest.mark.parametrize("dt", [torch.float32, torch.float64])
def test_float_bulk_vs_oracle(N, D, F, T, dt):
    """Float form at the config-5 / config-2 / config-3 shapes, rounds from the bench generator, notebook
    semantics (true median, lower index kept) and contract semantics."""
    import svoc
    from oracle import cref
    g = svoc.generate(T, N, D, F, a=10.0, b=15.0, failing="uniform", seed=seed_of("fbulk", N, D), dtype=torch.float32)
    xd = g.pred.double().cpu().numpy()
    for centre, tie in ((1, 1), (0, 0)):
        ora = cref.float_batch(xd, F, centre=centre, tie=tie, want_moments=False, nthreads=0)
        x = g.pred.to(dt)
        r = _np_result(svoc.consensus_float(x, F, centre=centre, tie_rule=tie, want_moments=False, want_qr=False, want_essence=False), N)
        # measured: fp64 differs on no round; fp32 on a few per thousand (risks within 1e-6 of each other at the cut)
        _float_compare(r, ora, N, F, dt, f"{N}x{D} {dt} centre={centre} bench options",
                       0.9999 if dt == torch.float64 else 0.999, False)
        r = _np_result(svoc.consensus_float(x, F, centre=centre, tie_rule=tie, want_moments=False), N)
        _float_compare(r, ora, N, F, dt, f"{N}x{D} {dt} centre={centre} all outputs",
                       0.9999 if dt == torch.float64 else 0.999, True)


# ------------------------------------------------------------------------------------------ rare statuses
def test_shape_panics_and_sqrt_of_one_on_the_gpu():
    """Status codes the random generators never reach, compared with the oracle on the GPU:
    5 (fewer than two reliable rows in constrained mode: sorted.at(mid - 1), math.cairo:117; and a negative
    reliable count, usize underflow at contract.cairo:345-346), 11 (N < 2), 2 and 6 (sqrt(1) divides by zero,
    math.cairo:271-292) from explicit vectors."""
    import svoc
    from oracle import cref
    rng = np.random.default_rng(5)

    def both(x, F, constrained=True, ms=0):
        ora = cref.fixed_batch(x, F, constrained, ms, True, nthreads=1)
        for dt in (torch.int32, torch.int64):
            t = torch.from_numpy(x.astype(np.int32 if dt == torch.int32 else np.int64)).cuda()
            r = result_to_numpy(svoc.consensus_fixed(t, F, constrained=constrained, max_spread=ms, want_moments=True))
            assert_fixed_equal(r, ora, x.shape[0], label=f"N={x.shape[1]} F={F}")
        return ora["status"]

    for N, D in ((7, 2), (33, 5), (64, 16), (100, 3), (256, 128)):
        x = gen_fixed("beta", rng, 8, N, D)
        assert np.all(both(x, N - 1) == 5)                  # R = 1: second-pass smooth median of one row
        assert np.all(both(x, N) == 5)                      # R = 0
        assert np.all(both(x, N + 3) == 5)                  # R < 0
        st = both(gen_fixed("wide", rng, 8, N, D), N, constrained=False, ms=10 * W)   # unconstrained, R = 0
        assert np.all(st == 5)
    x = gen_fixed("beta", rng, 16, 1, 3)
    assert np.all(both(x, 0) == 11)                         # N = 1
    # sqrt(1): avg_qr / D == 1 in the first pass; in the second pass only
    x = np.full((1, 4, 1), 500000, dtype=np.int64)
    x[0, 0, 0] += 2000
    assert both(x, 0)[0] == 2
    x = np.full((1, 5, 1), 500000, dtype=np.int64)
    x[0, 0, 0] += 2000
    x[0, 1, 0] += 300000
    assert both(x, 1)[0] == 6
    # the same two panics at wide shapes: N = 100 / 256 rows, one of them sqrt(N) * 1000 off the centre -> avg qr = 1
    for N in (100, 256):
        x = np.full((1, N, 1), 500000, dtype=np.int64)
        x[0, 0, 0] += 1000 * int(round(np.sqrt(N)))
        assert both(x, 0)[0] == 2
        x = np.full((1, N + 1, 1), 500000, dtype=np.int64)
        x[0, 0, 0] += 1000 * int(round(np.sqrt(N)))          # kept: qr = N over N kept rows -> avg 1
        x[0, 1, 0] += 300000                                  # dropped
        assert both(x, 1)[0] == 6


# ------------------------------------------------------------------------------------------ kernel variants
MID_SHAPES = [(64, 16, 12), (64, 32, 10), (40, 32, 8), (50, 8, 10), (33, 5, 6), (64, 1, 3), (48, 12, 9), (64, 7, 20)]


@pytest.mark.parametrize("N,D,F", MID_SHAPES)
@pytest.mark.parametrize("kind", ["beta", "coarse", "colluders", "tiny", "extreme", "outside"])
def test_mid_kernel_every_geometry_vs_oracle(N, D, F, kind):
    """The warp-per-round kernel (32 < N <= 64) in all its geometries — 1, 2 or 4 lanes per column, D equal to or
    below the lane grid, N = 64 or less — without moments (with them the small<64> kernel runs): bit-exact in fixed
    point for every input family, fp32 within tolerance, both against oracle/svoc_ref.c."""
    import svoc
    from oracle import cref
    assert svoc.kernel_variant(N, D, torch.int32) == "mid"
    rng = np.random.default_rng(seed_of("mid", N, D, kind))
    T = 300
    x = gen_fixed(kind, rng, T, N, D)
    ora = cref.fixed_batch(x, F, True, 0, False, nthreads=0)
    r = result_to_numpy(svoc.consensus_fixed(torch.from_numpy(x.astype(np.int32)).cuda(), F, want_moments=False))
    assert_fixed_equal(r, ora, T, want_moments=False, label=f"mid {N}x{D} {kind}")
    if kind == "beta":
        xf = (x * 1e-6).astype(np.float32)
        for centre, tie in ((1, 1), (0, 0)):
            oraf = cref.float_batch(xf.astype(np.float64), F, centre=centre, tie=tie, want_moments=False, nthreads=0)
            rf = _np_result(svoc.consensus_float(torch.from_numpy(xf).cuda(), F, centre=centre, tie_rule=tie), N)
            _float_compare(rf, oraf, N, F, torch.float32, f"mid {N}x{D} fp32 centre={centre}", 0.98, True)


@pytest.mark.parametrize("N,F", [(256, 51), (200, 40), (130, 26), (65, 13)])
@pytest.mark.parametrize("kind", ["beta", "coarse", "colluders", "tiny", "extreme", "outside"])
def test_wide3_kernel_partial_rounds_vs_oracle(N, F, kind):
    """The cluster kernel (D = 128) with full and partial rounds (N < 256: the landing box reads past the round and
    the extra rows are masked), every input family, fixed point bit-exact; fp32 on the beta family."""
    import svoc
    from oracle import cref
    D = 128
    assert svoc.kernel_variant(N, D, torch.int32) == "wide3"
    rng = np.random.default_rng(seed_of("w3", N, kind))
    T = 40
    x = gen_fixed(kind, rng, T, N, D)
    ora = cref.fixed_batch(x, F, True, 0, False, nthreads=0)
    r = result_to_numpy(svoc.consensus_fixed(torch.from_numpy(x.astype(np.int32)).cuda(), F, want_moments=False))
    assert_fixed_equal(r, ora, T, want_moments=False, label=f"wide3 {N}x{D} {kind}")
    if kind == "beta":
        xf = (x * 1e-6).astype(np.float32)
        oraf = cref.float_batch(xf.astype(np.float64), F, want_moments=False, nthreads=0)
        rf = _np_result(svoc.consensus_float(torch.from_numpy(xf).cuda(), F), N)
        _float_compare(rf, oraf, N, F, torch.float32, f"wide3 {N}x{D} fp32", 0.95, True)
