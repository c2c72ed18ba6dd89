"""Pure-Python Philox4x32-10 (Salmon et al., SC'11; Random123 reference vectors) — test-side
restatement used to pin the bit stream of csrc/philox.cuh."""
M0, M1, W0, W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85
MASK = 0xFFFFFFFF


def philox4x32_10(c, k):
    c0, c1, c2, c3 = c
    k0, k1 = k
    for _ in range(10):
        p0, p1 = M0 * c0, M1 * c2
        hi0, lo0, hi1, lo1 = p0 >> 32, p0 & MASK, p1 >> 32, p1 & MASK
        c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
        k0, k1 = (k0 + W0) & MASK, (k1 + W1) & MASK
    return c0, c1, c2, c3


def position_key(seed, trial, oracle):
    """csrc/svoc_generate.cu position_key(): counter (trial_lo, trial_hi, 0xFFFF0000 + oracle, 0)."""
    return philox4x32_10((trial & MASK, (trial >> 32) & MASK, (0xFFFF0000 + oracle) & MASK, 0),
                         (seed & MASK, (seed >> 32) & MASK))[0]


def failing_flags(seed, trial, n, f):
    keys = [position_key(seed, trial, i) for i in range(n)]
    order = sorted(range(n), key=lambda i: (keys[i], i))
    flags = [False] * n
    for i in order[:f]:
        flags[i] = True
    return flags
