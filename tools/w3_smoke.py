"""Quick liveness / parity probe of the wide kernels on a handful of rounds (run under `timeout`):
    python tools/w3_smoke.py [rounds] [N]
compares the default dispatch (wide3 where eligible) with SVOC_VARIANT=wide2 run in a child process."""
import os
import subprocess
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "stochastic-vector-oracle-consensus_b200"))
import svoc  # noqa: E402

T = int(sys.argv[1]) if len(sys.argv) > 1 else 2
N = int(sys.argv[2]) if len(sys.argv) > 2 else 256
F = N // 5
print("variant", svoc.kernel_variant(N, 128, torch.float32), "rounds", T, flush=True)
g = svoc.generate(T, N, 128, F, a=10.0, b=15.0, failing="colluding", shift=0.15, seed=5)
r = svoc.consensus_float(g.pred, F)
torch.cuda.synchronize()
print("float done", r.consensus[0, :4].tolist(), r.rel1[:2].tolist(), r.status[:4].tolist(), flush=True)
gi = svoc.generate(T, N, 128, F, a=10.0, b=15.0, failing="colluding", shift=0.15, seed=5, dtype=torch.int32)
ri = svoc.consensus_fixed(gi.pred, F, want_moments=False)
torch.cuda.synchronize()
print("fixed done", ri.consensus[0, :4].tolist(), ri.rel1[:2].tolist(), ri.status[:4].tolist(), flush=True)
if os.environ.get("SVOC_VARIANT") != "wide2":
    torch.save({"c": r.consensus.cpu(), "m": r.reliable_words.cpu(), "ci": ri.consensus.cpu(), "mi": ri.reliable_words.cpu(),
                "q": r.qr.cpu(), "qi": ri.qr.cpu()}, "/tmp/w3_smoke.pt")
    env = dict(os.environ, SVOC_VARIANT="wide2")
    out = subprocess.run([sys.executable, __file__, str(T), str(N)], env=env, capture_output=True, text=True)
    print(out.stdout, out.stderr[-500:])
else:
    a = torch.load("/tmp/w3_smoke.pt")
    print("wide2 vs default: consensus equal", torch.equal(a["c"], r.consensus.cpu()), "mask equal", torch.equal(a["m"], r.reliable_words.cpu()),
          "fixed consensus equal", torch.equal(a["ci"], ri.consensus.cpu()), "fixed mask equal", torch.equal(a["mi"], ri.reliable_words.cpu()),
          "fixed qr equal", torch.equal(a["qi"], ri.qr.cpu()), "max |dqr|", float((a["q"] - r.qr.cpu()).abs().max()))
