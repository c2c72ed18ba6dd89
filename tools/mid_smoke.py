"""Liveness / parity probe of the `mid` kernel (32 < N <= 64): default dispatch vs SVOC_VARIANT=small in a child."""
import os
import subprocess
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "stochastic-vector-oracle-consensus_b200"))
import svoc  # noqa: E402

T = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
shapes = [(64, 16, 12), (64, 32, 12), (50, 8, 10), (33, 5, 6), (64, 1, 3)]
out = {}
for (N, D, F) in shapes:
    print("variant", svoc.kernel_variant(N, D, torch.int32), N, D, flush=True)
    gi = svoc.generate(T, N, D, F, a=10.0, b=15.0, seed=5, dtype=torch.int32)
    x = gi.pred.clone()
    x[::7, : N // 4, :] = x[::7, :1, :]          # some rounds with identical rows (ties at the cut)
    x[::11, 0, 0] = 2_000_000                   # some interval panics
    ri = svoc.consensus_fixed(x, F, want_moments=False)
    rf = svoc.consensus_float(x.float() * 1e-6, F, centre=svoc.CENTRE_MEDIAN, tie_rule=svoc.TIE_LOW_INDEX_KEPT)
    torch.cuda.synchronize()
    out[(N, D)] = {k: getattr(ri, k).cpu() for k in ("consensus", "rel1", "rel2", "status", "reliable_words", "qr", "essence1")}
    out[(N, D, "f")] = {k: getattr(rf, k).cpu() for k in ("consensus", "rel1", "rel2", "status", "reliable_words", "qr", "essence1")}
    print("done", ri.status[:12].tolist(), flush=True)
if os.environ.get("SVOC_VARIANT") != "small":
    torch.save(out, "/tmp/mid_smoke.pt")
    r = subprocess.run([sys.executable, __file__, str(T)], env=dict(os.environ, SVOC_VARIANT="small"), capture_output=True, text=True)
    print(r.stdout[-1500:], r.stderr[-800:])
else:
    a = torch.load("/tmp/mid_smoke.pt")
    for key, d in out.items():
        ok = d["status"] == 0
        res = {k: (torch.equal(a[key][k][ok], d[k][ok]) if len(key) == 2 else bool(torch.allclose(a[key][k][ok].double(), d[k][ok].double(), rtol=1e-5, atol=1e-7))) for k in d if k != "status"}
        print(key, "status equal", torch.equal(a[key]["status"], d["status"]), res)
