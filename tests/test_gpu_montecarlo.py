"""GPU: the fused generate -> consensus -> accumulate kernel against the materialised pipeline."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,D,F,dtype", [(7, 1, 2, torch.float64), (7, 2, 2, torch.float32), (32, 8, 6, torch.float32),
                                         (20, 1, 15, torch.float64), (64, 16, 12, torch.int32), (100, 3, 20, torch.int32)])  # N > 64 runs different kernel variants: only the fixed form is order-free
def test_fused_equals_materialised(N, D, F, dtype):
    import svoc
    T, seed, off = 20000 if N <= 64 else 2000, 77, 12345
    fixed = dtype == torch.int32
    kw = dict(a=[10, 12], b=[15, 30], seed=seed, trial_offset=off, dtype=dtype)
    ckw = dict(centre=svoc.CENTRE_SMOOTH_MEDIAN_CAIRO, tie_rule=svoc.TIE_HIGH_INDEX_KEPT) if fixed else \
        dict(centre=svoc.CENTRE_MEDIAN, tie_rule=svoc.TIE_LOW_INDEX_KEPT)
    fused = svoc.montecarlo(T, N, D, F, **kw, **ckw)
    g = svoc.generate(T, N, D, F, **kw)
    r = (svoc.consensus_fixed if fixed else svoc.consensus_float)(g.pred, F, want_moments=False, **ckw)
    ref = svoc.RunningStats(N)
    ref.update(r)
    torch.cuda.synchronize()
    B = 528
    fc, rc = fused.counts.cpu().numpy(), ref.counts.cpu().numpy()
    assert np.array_equal(fc[:B + N], rc[:B + N])          # rounds, status counts, both histograms, unreliable counts
    np.testing.assert_allclose(fused.sums.cpu().numpy()[:4], ref.sums.cpu().numpy()[:4], rtol=1e-9)
    ok = (r.status == 0)
    success = ((r.reliable == g.honest).all(dim=1) & ok).sum().item()
    assert fc[B + 2 * N] == success and fc[B + 2 * N + 1] == ok.sum().item()
    # distance to the centre of the honest rows
    x = g.pred.double() * (1e-6 if fixed else 1.0)
    big = torch.where(g.honest[:, :, None], x, torch.full_like(x, float("inf"))).sort(dim=1).values
    R = N - F
    lo = (R // 2 - 1) if fixed else (R - 1) // 2
    truth = 0.5 * (big[:, lo] + big[:, R // 2])
    if fixed:   # integer division of the contract: (a + b) / 2 truncated, on wsad integers
        bi = torch.where(g.honest[:, :, None], g.pred.long(), torch.full_like(g.pred.long(), 2**40)).sort(dim=1).values
        truth = torch.div(bi[:, lo] + bi[:, R // 2], 2, rounding_mode="trunc").double() * 1e-6
    cons = r.consensus.double() * (1e-6 if fixed else 1.0)
    dist = ((cons - truth) ** 2).sum(dim=1).sqrt()[ok].sum().item()
    np.testing.assert_allclose(fused.sums.cpu().numpy()[4 + N], dist, rtol=2e-3 if dtype == torch.float32 else 1e-9)  # fp32 medians round at 1 ulp, distances are ~1e-4


def test_fused_is_independent_of_sharding():
    import svoc
    whole = svoc.montecarlo(30000, 32, 8, 6, a=10, b=15, seed=5)
    parts = svoc.montecarlo(10000, 32, 8, 6, a=10, b=15, seed=5, trial_offset=0)
    svoc.montecarlo(20000, 32, 8, 6, a=10, b=15, seed=5, trial_offset=10000, stats=parts)
    assert torch.equal(whole.counts, parts.counts)
    torch.testing.assert_close(whole.sums, parts.sums, rtol=1e-9, atol=1e-9)
    s = whole.summary()
    assert s["rounds"] == 30000 and 0 <= s["identification_success_pct"] <= 100


def test_fused_reproduces_the_reference_benchmark_table(golden):
    """Same check as for the materialised path, at 2M rounds per line in one launch each."""
    import svoc
    for (n, f, a, k, succ, rel) in golden["benchmark_stats"]:
        s = svoc.montecarlo(2_000_000, int(n), 1, int(f), a=a, b=a, seed=int(31 * n + f + a), dtype=torch.float64).summary()
        p = succ / 100
        tol = 4 * 100 * np.sqrt(max(p * (1 - p), 1e-4) / k) + 0.25
        assert abs(s["identification_success_pct"] - succ) < tol, (n, f, a, s, succ)
        assert abs(s["true_consensus_reliability_pct"] - rel) < 0.3, (n, f, a, s, rel)
