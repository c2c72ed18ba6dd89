"""GPU parity, fixed-point form: libsvoc (through the C ABI) vs the CPU oracles, bit-exact."""
import numpy as np
import pytest
import torch

from helpers import W, assert_fixed_equal, gen_fixed, masks_equal_up_to_cut_ties, result_to_numpy, seed_of

pytestmark = pytest.mark.gpu


def _run(x, F, constrained, ms, dtype, want_moments=True):
    import svoc
    t = torch.from_numpy(x.astype(np.int32 if dtype == torch.int32 else np.int64)).cuda()
    r = svoc.consensus_fixed(t, F, constrained=constrained, max_spread=ms, want_moments=want_moments)
    torch.cuda.synchronize()
    return result_to_numpy(r)


def test_contract_vectors(contract_vectors):
    for v in contract_vectors["vectors"]:
        x = np.array([v["pred"]], dtype=np.int64)
        for dt in (torch.int32, torch.int64):
            r = _run(x, v["n_failing"], v["constrained"], v["max_spread"], dt)
            e = v["expected"]
            assert r["status"][0] == e["status"]
            for k in ("essence1", "qr", "consensus", "skew", "kurt"):
                assert r[k][0].tolist() == e[k], (v["name"], k)
            assert int(r["rel1"][0]) == e["rel1"] and int(r["rel2"][0]) == e["rel2"]
            assert r["reliable"][0].tolist() == e["reliable"]


SHAPES = [(7, 2, 2), (7, 6, 2), (7, 1, 2), (5, 3, 2), (5, 3, 3), (4, 2, 0), (2, 1, 0), (3, 2, 1), (20, 1, 15), (16, 4, 3),
          (32, 8, 6), (33, 5, 6), (64, 16, 12), (100, 3, 20), (65, 7, 13), (256, 128, 51), (130, 40, 26)]


@pytest.mark.parametrize("N,D,F", SHAPES)
@pytest.mark.parametrize("kind", ["beta", "coarse", "colluders", "tiny", "extreme", "outside"])
def test_constrained_vs_c_oracle(N, D, F, kind):
    from oracle import cref
    rng = np.random.default_rng(seed_of(N, D, F, kind))
    T = 600 if N * D <= 1024 else (40 if N * D <= 8192 else 6)
    x = gen_fixed(kind, rng, T, N, D)
    ora = cref.fixed_batch(x, F, True, 0, True, nthreads=0)
    for dt in (torch.int32, torch.int64):
        r = _run(x, F, True, 0, dt)
        assert_fixed_equal(r, ora, T, label=f"{N}x{D} F={F} {kind} {dt}")


@pytest.mark.parametrize("N,D,F", [(7, 2, 2), (20, 2, 4), (64, 16, 12), (100, 3, 20)])
@pytest.mark.parametrize("kind,ms", [("wide", 10 * W), ("wide", 0), ("wide", 3 * W), ("wide", -5 * W), ("huge", 10 * W), ("beta", W)])
def test_unconstrained_vs_c_oracle(N, D, F, kind, ms):
    from oracle import cref
    rng = np.random.default_rng(seed_of(N, D, F, kind, ms))
    T = 300 if N * D <= 1024 else 20
    x = gen_fixed(kind, rng, T, N, D)
    ora = cref.fixed_batch(x, F, False, ms, True, nthreads=0)
    r = _run(x, F, False, ms, torch.int64)
    if kind == "huge":   # beyond +-2^30 the engine reports ENGINE_RANGE instead of computing
        big = np.abs(x).reshape(T, -1).max(axis=1) > (1 << 30)
        assert np.all(r["status"][big] == 12)
        keep = ~big
        r = {k: (v[keep] if v is not None else None) for k, v in r.items()}
        ora = {k: v[keep] for k, v in ora.items()}
    assert_fixed_equal(r, ora, len(r["status"]), label=f"unconstrained {N}x{D} {kind} ms={ms}")


def test_python_bigint_oracle_small():
    """The slow big-int oracle on a few hundred rounds (the C port is the bulk checker)."""
    from oracle import fixed_oracle as fo
    rng = np.random.default_rng(7)
    for (N, D, F, kind) in [(7, 2, 2, "beta"), (7, 2, 2, "coarse"), (7, 2, 2, "tiny"), (32, 8, 6, "beta"), (64, 16, 12, "colluders")]:
        T = 64 if N <= 8 else 8
        x = gen_fixed(kind, rng, T, N, D)
        r = _run(x, F, True, 0, torch.int64)
        for t in range(T):
            o = fo.consensus_round(x[t].tolist(), F, True, 0, True)
            assert r["status"][t] == o.status
            if o.essence1 is not None:
                assert r["essence1"][t].tolist() == o.essence1 and r["qr"][t].tolist() == o.qr
            if o.consensus is not None:
                assert r["consensus"][t].tolist() == o.consensus and r["reliable"][t].tolist() == o.reliable
            if o.rel2 is not None and o.status in (0, 8, 9):
                assert int(r["rel1"][t]) == o.rel1 and int(r["rel2"][t]) == o.rel2
            if o.kurt is not None:
                assert r["skew"][t].tolist() == o.skew and r["kurt"][t].tolist() == o.kurt


def test_empty_and_ragged_batches():
    import svoc
    r = svoc.consensus_fixed(torch.empty((0, 7, 2), dtype=torch.int64, device="cuda"), 2)
    assert r.consensus.shape == (0, 2) and r.status.numel() == 0
    # a batch that does not fill the last CTA and a misaligned view (offset pointer, no 16-byte alignment)
    from oracle import cref
    rng = np.random.default_rng(3)
    x = gen_fixed("beta", rng, 131, 7, 2).astype(np.int32)
    big = torch.from_numpy(np.concatenate([np.zeros(1, np.int32), x.reshape(-1)])).cuda()
    view = big[1:].view(131, 7, 2)
    assert view.data_ptr() % 16 != 0
    r = result_to_numpy(svoc.consensus_fixed(view, 2))
    assert_fixed_equal(r, cref.fixed_batch(x.astype(np.int64), 2, True, 0, True), 131, label="misaligned")


def test_host_entry_point_matches_device():
    import svoc
    rng = np.random.default_rng(11)
    x = gen_fixed("beta", rng, 5000, 32, 8).astype(np.int32)
    a = result_to_numpy(svoc.consensus_fixed(torch.from_numpy(x).cuda(), 6))
    b = svoc.consensus_fixed_host(torch.from_numpy(x).pin_memory(), 6)
    b = {k: (getattr(b, k).numpy() if getattr(b, k) is not None else None) for k in ("consensus", "rel1", "rel2", "status", "essence1", "qr", "skew", "kurt")} | {"reliable": b.reliable.numpy()}
    for k in a:
        assert np.array_equal(a[k], b[k]), k


def test_full_size_properties_c3():
    """BASELINE config 3 shape (N=64, D=16, fixed) at 2^18 rounds: size-independent properties —
    permutation of oracles permutes qr/flags and leaves consensus and reliabilities unchanged;
    the reliable count is exactly N - F; consensus lies within the kept rows' range."""
    import svoc
    g = torch.Generator(device="cuda").manual_seed(5)
    T, N, D, F = 1 << 18, 64, 16, 12
    x = (torch.rand((T, N, D), generator=g, device="cuda") * 1e6).to(torch.int32)
    r = svoc.consensus_fixed(x, F, want_moments=False)
    perm = torch.randperm(N, generator=g, device="cuda")
    rp = svoc.consensus_fixed(x[:, perm, :].contiguous(), F, want_moments=False)
    assert torch.equal(r.essence1, rp.essence1) and torch.equal(r.rel1, rp.rel1)
    assert torch.equal(r.qr[:, perm], rp.qr)
    rel = r.reliable
    assert torch.all(rel.sum(dim=1) == N - F)
    # ties at the cut are broken by index, so compare flags only where the cut is strict
    qs = torch.sort(r.qr, dim=1).values
    strict = qs[:, N - F - 1] < qs[:, N - F]
    assert strict.float().mean() > 0.99
    assert torch.equal(rel[:, perm][strict], rp.reliable[strict])
    assert torch.equal(r.consensus[strict], rp.consensus[strict]) and torch.equal(r.rel2[strict], rp.rel2[strict])
    big = torch.where(rel[:, :, None], x.long(), torch.full_like(x, -1).long()).max(dim=1).values
    small = torch.where(rel[:, :, None], x.long(), torch.full_like(x, 2 * W).long()).min(dim=1).values
    assert torch.all((r.consensus >= small) & (r.consensus <= big))
    assert torch.all(r.status == 0)


def test_obsolete_reliability_anchor(contract_vectors):
    """The "0.798" comment of contract/tests/test_contract.cairo:188 is the OBSOLETE contract's second-pass
    reliability (1 - 2*sqrt(avg qr), no division by the dimension, contract/obsolete/src/contract_nd.cairo:437)."""
    import svoc
    v = contract_vectors["vectors"][0]
    x = torch.tensor([v["pred"]], dtype=torch.int64, device="cuda")
    r = svoc.consensus_fixed(x, v["n_failing"], want_moments=False, obsolete_reliability=True)
    assert int(r.rel1[0]) == v["expected"]["obsolete_rel1"] and int(r.rel2[0]) == v["expected"]["obsolete_rel2"]
    assert int(int(r.rel2[0]) / 1000) / 1000 == v["anchors"]["obsolete_rel2_3digits"]
    assert r.consensus[0].tolist() == v["expected"]["consensus"]       # everything else is unchanged


@pytest.mark.parametrize("N,D,F", [(100, 3, 20), (256, 128, 51), (130, 40, 26)])
def test_misaligned_wide_rounds_take_the_staged_fallback(N, D, F):
    """Rounds that are not 16-byte aligned cannot use the global-tile kernel: the staged `wide` variant runs."""
    import svoc
    from oracle import cref
    rng = np.random.default_rng(seed_of(N, D, F, "misaligned"))
    T = 12 if N * D > 8192 else 60
    x = gen_fixed("colluders", rng, T, N, D).astype(np.int32)
    big = torch.from_numpy(np.concatenate([np.zeros(1, np.int32), x.reshape(-1)])).cuda()
    view = big[1:].view(T, N, D)
    assert view.data_ptr() % 16 != 0
    r = result_to_numpy(svoc.consensus_fixed(view, F))
    assert_fixed_equal(r, cref.fixed_batch(x.astype(np.int64), F, True, 0, True), T, label="misaligned wide")
    rf = svoc.consensus_float(view.float() * 1e-6, F)     # aligned copy: the default wide kernel
    rm = svoc.consensus_float((big.float() * 1e-6)[1:].view(T, N, D), F)
    # the two float kernels sum qr in different orders: flags may differ only for rows within rounding of the cut
    same = masks_equal_up_to_cut_ties(rf.reliable.cpu().numpy(), rm.reliable.cpu().numpy(),
                                      rf.qr.double().cpu().numpy(), N, F, 1e-4)
    assert same.mean() > 0.7
    sel = torch.from_numpy(same).cuda()
    torch.testing.assert_close(rf.consensus[sel], rm.consensus[sel], rtol=1e-6, atol=1e-7)
