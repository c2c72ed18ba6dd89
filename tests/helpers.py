"""Shared input generators and comparison helpers for the parity tests."""
import numpy as np

W = 1_000_000


def seed_of(*a):
    import zlib
    return zlib.crc32(repr(a).encode())


def gen_fixed(kind, rng, T, N, D):
    if kind == "beta":      # honest beta(10,15) + a few uniform outliers, like ipynb beta :37-68
        x = rng.beta(10, 15, (T, N, D))
        bad = rng.random((T, N)) < 0.2
        x = np.where(bad[:, :, None], rng.random((T, N, D)), x)
        return (x * 1e6).astype(np.int64)
    if kind == "coarse":    # few distinct values: ties at the medians and at the selection cut
        return (rng.integers(0, 4, (T, N, D)) * 250000).astype(np.int64)
    if kind == "colluders":  # F-ish identical rows (ties in qr), beta elsewhere
        x = (rng.beta(10, 15, (T, N, D)) * 1e6).astype(np.int64)
        k = max(1, N // 4)
        x[:, :k, :] = x[:, :1, :]
        return x
    if kind == "tiny":      # variance 0/1/2: sqrt(1) and division-by-zero panics
        return (500000 + rng.integers(-2, 3, (T, N, D))).astype(np.int64)
    if kind == "extreme":   # 0 / 1e6 only: reliability leaves [0, 1]
        return (rng.integers(0, 2, (T, N, D)) * W).astype(np.int64)
    if kind == "outside":   # a few components outside [0, W]
        x = (rng.beta(10, 15, (T, N, D)) * 1e6).astype(np.int64)
        hit = rng.random((T, N, D)) < 0.02
        return np.where(hit, np.where(rng.random((T, N, D)) < 0.5, -1, W + 1), x)
    if kind == "wide":      # unconstrained gaussian, like gaussian_distribution_for_tests.ipynb:37-52
        return rng.normal(20e6, 3e6, (T, N, D)).astype(np.int64)
    if kind == "huge":      # beyond the engine's +-2^30 range
        x = rng.normal(20e6, 3e6, (T, N, D)).astype(np.int64)
        x[::3, 0, 0] = (1 << 30) + 5
        return x
    raise ValueError(kind)


def assert_fixed_equal(res, ora, T, want_moments=True, label=""):
    """res: svoc ConsensusResult moved to numpy dict; ora: oracle/cref.fixed_batch dict.
    Bit-exact on every field the contract has written before its first panic."""
    st = res["status"]
    assert np.array_equal(st, ora["status"]), f"{label}: status differs at {np.nonzero(st != ora['status'])[0][:5]}"
    # which fields the contract has produced before each panic code
    valid = {
        "essence1": [0, 2, 3, 4, 5, 6, 7, 8, 9], "qr": [0, 2, 3, 4, 5, 6, 7, 8, 9],
        "rel1": [0, 4, 5, 6, 7, 8, 9], "reliable": [0, 6, 7, 8, 9], "consensus": [0, 6, 7, 8, 9],
        "rel2": [0, 7, 8, 9],
    }
    for name, codes in valid.items():
        sel = np.isin(st, codes)
        a, b = res[name][sel], ora[name][sel]
        assert np.array_equal(a, b), f"{label}: {name} differs ({np.sum(a != b)} elements)"
    if want_moments:
        ok = st == 0
        assert np.array_equal(res["skew"][ok | (st == 9)], ora["skew"][ok | (st == 9)]), f"{label}: skew"
        assert np.array_equal(res["kurt"][ok], ora["kurt"][ok]), f"{label}: kurt"


def result_to_numpy(r):
    d = {k: (getattr(r, k).cpu().numpy() if getattr(r, k) is not None else None)
         for k in ("consensus", "rel1", "rel2", "status", "essence1", "qr", "skew", "kurt")}
    d["reliable"] = r.reliable.cpu().numpy()
    return d


def ranks_equal_up_to_ties(r_a, r_b, dev):
    """rank vectors agree except for permutations among exactly equal deviations."""
    r_a, r_b, dev = np.asarray(r_a), np.asarray(r_b), np.asarray(dev)
    if np.array_equal(r_a, r_b):
        return True
    diff = np.nonzero(r_a != r_b)[0]
    for i in diff:
        j = np.nonzero(r_b == r_a[i])[0][0]     # who holds my rank in the other vector
        if dev[i] != dev[j]:
            return False
    return True


def masks_equal_up_to_cut_ties(mask_a, mask_b, qr, N, F, rtol):
    """Asserts that two reliable masks differ only on oracles whose qr lies within `rtol` (relative) of the
    selection cut, and returns the per-round bool vector 'masks identical'."""
    qs = np.sort(qr, axis=1)
    cut = 0.5 * (qs[:, N - F - 1] + qs[:, N - F]) if 0 < F < N else qs[:, -1]
    diff = mask_a != mask_b
    near = np.abs(qr - cut[:, None]) <= rtol * np.abs(cut[:, None]) + 1e-300
    assert not np.any(diff & ~near), "reliable masks differ away from the selection cut"
    return ~diff.any(axis=1)
