"""Time svoc.generate on config-4 and config-3 shaped rounds (CUDA events); prints a checksum of the output."""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "stochastic-vector-oracle-consensus_b200"))
import svoc  # noqa: E402

tag = sys.argv[1] if len(sys.argv) > 1 else ""
for name, kw in (("c4", dict(n_rounds=65536, n_oracles=256, dimension=128, n_failing=51, a=(10.0,), b=(15.0,), failing="colluding", shift=0.25, dtype=torch.float32)),
                 ("c3", dict(n_rounds=1 << 20, n_oracles=64, dimension=16, n_failing=12, a=(10.0,), b=(15.0,), dtype=torch.int32))):
    try:
        out = svoc.generate(seed=7, **kw)
    except TypeError as e:
        print("signature:", e)
        raise
    pred = out[0] if isinstance(out, tuple) else out.pred
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for _ in range(3):
        out = svoc.generate(seed=7, **kw)
    ev[1].record()
    torch.cuda.synchronize()
    ms = ev[0].elapsed_time(ev[1]) / 3
    pred = out[0] if isinstance(out, tuple) else out.pred
    nbytes = pred.numel() * pred.element_size()
    print(tag, name, "generate %.2f ms  %.0f GB/s  checksum %.9g" % (ms, nbytes / ms / 1e6, pred.double().sum().item()))
