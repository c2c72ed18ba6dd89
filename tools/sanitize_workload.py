"""CI-size calls of every kernel family, for compute-sanitizer (tools/sanitize.sh): small<8/32/64>, mid, wide
(misaligned rounds), wide2 (moments / rank / mean centre), wide3, the fused Monte-Carlo kernel, the generator and
the accumulators.  Small batches: the sanitizer slows kernels down by two to three orders of magnitude."""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "stochastic-vector-oracle-consensus_b200"))
import svoc  # noqa: E402

seen = {}
for (N, D, F, T) in [(7, 2, 2, 64), (32, 8, 6, 48), (64, 16, 12, 40), (50, 8, 10, 24), (256, 128, 51, 6), (200, 128, 40, 5), (130, 40, 26, 6)]:
    g = svoc.generate(T, N, D, F, a=10.0, b=15.0, failing="colluding", shift=0.15, seed=3)
    gi = svoc.generate(T, N, D, F, a=10.0, b=15.0, seed=4, dtype=torch.int32)
    for kw in (dict(), dict(want_moments=True), dict(want_rank=True), dict(centre=svoc.CENTRE_MEAN)):
        r = svoc.consensus_float(g.pred, F, **kw)
        assert int(r.status.abs().sum()) >= 0
    ri = svoc.consensus_fixed(gi.pred, F, want_moments=False)
    rm = svoc.consensus_fixed(gi.pred, F, want_moments=True)
    st = svoc.RunningStats(N)
    st.update(ri)
    seen[(N, D)] = svoc.kernel_variant(N, D, torch.float32)
    if N > 64:          # misaligned rounds: the staged `wide` kernel
        big = torch.cat([torch.zeros(1, device="cuda"), g.pred.reshape(-1)])
        svoc.consensus_float(big[1:].view(T, N, D), F)
mc = svoc.montecarlo(2048, 32, 8, 6, a=10.0, b=15.0, seed=9)
mc7 = svoc.montecarlo(2048, 7, 2, 2, a=10.0, b=15.0, seed=9, dtype=torch.int32)
torch.cuda.synchronize()
print("sanitize workload done", seen, mc.summary()["rounds"], mc7.summary()["rounds"])
