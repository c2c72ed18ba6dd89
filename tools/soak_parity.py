"""Bulk parity soak on the benchmark shapes: for a time budget per leg, draw chunks of rounds with the bench
generator (fresh seed per chunk), run the CUDA path through the C ABI and compare with oracle/svoc_ref.c on all
host threads.  Fixed-point legs are compared bit for bit on every output; float legs with the tolerances of
tests/test_gpu_bulk_parity.py.  Writes one summary line per leg.

    python tools/soak_parity.py [seconds per leg] > gpurun_out/soak.txt
"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "stochastic-vector-oracle-consensus_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)
import svoc  # noqa: E402
from helpers import assert_fixed_equal, result_to_numpy  # noqa: E402
from oracle import cref  # noqa: E402
from test_gpu_bulk_parity import _float_compare, _np_result  # noqa: E402

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0


def leg_fixed(name, N, D, F, chunk, failing, dtype=torch.int32):
    t0, rounds, chunks, variant = time.time(), 0, 0, None
    status = {}
    while time.time() - t0 < budget:
        seed = 0x5EED0000 + chunks
        g = svoc.generate(chunk, N, D, F, a=10.0, b=15.0, failing=failing, shift=0.15, seed=seed, dtype=dtype)
        ora = cref.fixed_batch(g.pred.cpu().numpy().astype(np.int64), F, True, 0, False, nthreads=0)
        r = result_to_numpy(svoc.consensus_fixed(g.pred, F, want_moments=False))
        assert_fixed_equal(r, ora, chunk, want_moments=False, label=f"{name} seed {seed:#x}")
        for k, v in zip(*np.unique(ora["status"], return_counts=True)):
            status[int(k)] = status.get(int(k), 0) + int(v)
        rounds += chunk
        chunks += 1
    print(f"{name}: {rounds} rounds in {chunks} chunks, every output bit-identical to the oracle; status counts {status}; {time.time() - t0:.0f} s", flush=True)


def leg_float(name, N, D, F, chunk, failing, min_same):
    t0, rounds, chunks = time.time(), 0, 0
    while time.time() - t0 < budget:
        seed = 0xF10A0000 + chunks
        g = svoc.generate(chunk, N, D, F, a=10.0, b=15.0, failing=failing, shift=0.15, seed=seed)
        ora = cref.float_batch(g.pred.double().cpu().numpy(), F, want_moments=False, tie=0, nthreads=0)
        r = _np_result(svoc.consensus_float(g.pred, F, want_moments=False, want_qr=True, want_essence=True), N)
        _float_compare(r, ora, N, F, torch.float32, f"{name} seed {seed:#x}", min_same, True)
        rounds += chunk
        chunks += 1
    print(f"{name}: {rounds} rounds in {chunks} chunks within the fp32 tolerances of the bulk tests; {time.time() - t0:.0f} s", flush=True)


print(f"# soak parity, {budget:.0f} s per leg, {cref.max_threads()} oracle threads, device {torch.cuda.get_device_name(0)}", flush=True)
leg_fixed("config 3 (64 x 16 int32, mid kernel)", 64, 16, 12, 400_000, "uniform")
leg_fixed("config 4 shape (256 x 128 int32, wide3 kernel, colluders)", 256, 128, 51, 8_000, "colluding")
leg_fixed("config 5 shape (32 x 8 int32, small<32> kernel)", 32, 8, 6, 1_000_000, "uniform")
leg_float("config 4 (256 x 128 fp32, wide3 kernel, colluders)", 256, 128, 51, 8_000, "colluding", 0.999)
leg_float("config 5 shape (32 x 8 fp32, small<32> kernel)", 32, 8, 6, 500_000, "uniform", 0.99)
