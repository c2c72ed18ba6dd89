"""Times the 64-bit wide2 kernel on the config-4 shape (fp64 and int64) — kernel only, CUDA events."""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "stochastic-vector-oracle-consensus_b200"))
import svoc  # noqa: E402

T, N, D, F = 131072, 256, 128, 51
for dt in (torch.float64, torch.int64):
    g = svoc.generate(T, N, D, F, a=10.0, b=15.0, failing="colluding", shift=0.15, seed=7, dtype=dt)
    f = svoc.consensus_float if dt == torch.float64 else svoc.consensus_fixed
    for _ in range(2):
        f(g.pred, F, want_moments=False, want_qr=False, want_essence=False)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(3):
        f(g.pred, F, want_moments=False, want_qr=False, want_essence=False)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 3
    print(str(dt), svoc.kernel_variant(N, D, dt), f"{T / ms / 1e3:.2f} M rounds/s", f"{T * (N * D * 8 + 8 * D + 52) / ms / 1e6:.0f} GB/s", flush=True)
    del g
