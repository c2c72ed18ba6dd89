"""GPU: svoc_consensus_float_host_noalloc, the host-buffer consensus call for callers that must not allocate
device memory (a plugin inside a process whose GPU memory belongs to someone else, code running after the
memory map is frozen, a latency-bound loop that cannot take cudaMalloc's driver locks).  It stages through a
fixed buffer built into the library instead of the cudaMalloc'ed cache of svoc_consensus_float_host, so it has a
size limit and its calls share that buffer.

Checked here, through the C ABI with plain numpy buffers: two calls in a row give what the allocating entry
point gives on the same inputs (the second call is smaller and of another shape, so anything left in the
shared buffer by the first would show), and a batch beyond the limit is refused with a message that names
the limit.  The file name sorts last on purpose: a run that stops at the first failure has then already run
every other test."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SCRATCH_BYTES = 16 << 20      # SVOC_NOALLOC_SCRATCH_BYTES in include/svoc.h


def _call(lib, L, fn, x, F, tie):
    """One call of a *_host entry point on numpy buffers; returns the outputs as a dict."""
    T, N, D = x.shape
    f32 = x.dtype == np.float32
    cfg = L.SvocConfig(n_oracles=N, dimension=D, n_failing=F, constrained=1, max_spread_wsad=0, max_spread=0.0,
                       centre=L.CENTRE_SMOOTH_MEDIAN_CAIRO, centre2=L.CENTRE_DEFAULT, tie_rule=tie,
                       dtype=L.F32 if f32 else L.F64, want_moments=0, flags=0)
    words = (N + 31) // 32
    o = {"essence1": np.full((T, D), np.nan, x.dtype), "qr": np.full((T, N), np.nan, x.dtype),
         "rel1": np.full(T, np.nan, x.dtype), "reliable": np.full((T, words), 0xdeadbeef, np.uint32),
         "consensus": np.full((T, D), np.nan, x.dtype), "rel2": np.full(T, np.nan, x.dtype),
         "status": np.full(T, -7, np.int32)}
    outs = L.SvocOutputs(essence1=o["essence1"].ctypes.data, qr=o["qr"].ctypes.data, rel1=o["rel1"].ctypes.data,
                         reliable=o["reliable"].ctypes.data, consensus=o["consensus"].ctypes.data,
                         rel2=o["rel2"].ctypes.data, skew=None, kurt=None, status=o["status"].ctypes.data, rank=None)
    rc = fn(C.byref(cfg), x.ctypes.data_as(C.c_void_p), T, C.byref(outs))
    return rc, o


def _rounds(rng, T, N, D, F, dtype):
    x = rng.beta(10, 15, (T, N, D))
    x[:, :F, :] = rng.random((T, F, D))
    return np.ascontiguousarray(x.astype(dtype))


def test_no_alloc_entry_point_matches_the_allocating_one_twice_in_a_row():
    from svoc import _lib as L
    lib = L.load()
    rng = np.random.default_rng(2024)
    # first call: the config-4 shape (wide kernel), 3.2 MB of the buffer; second: a small fp64 batch
    calls = [(_rounds(rng, 24, 256, 128, 51, np.float32), 51, L.TIE_HIGH_INDEX_KEPT, 1e-4),
             (_rounds(rng, 300, 7, 2, 2, np.float64), 2, L.TIE_LOW_INDEX_KEPT, 1e-6)]
    for x, F, tie, tol in calls:
        rc_a, a = _call(lib, L, lib.svoc_consensus_float_host_noalloc, x, F, tie)
        assert rc_a == 0, lib.svoc_last_error().decode()
        rc_b, b = _call(lib, L, lib.svoc_consensus_float_host, x, F, tie)
        assert rc_b == 0, lib.svoc_last_error().decode()
        assert np.array_equal(a["status"], b["status"]) and np.all(a["status"] == 0)
        assert np.array_equal(a["reliable"], b["reliable"])
        for k in ("essence1", "qr", "rel1", "consensus", "rel2"):
            np.testing.assert_allclose(a[k], b[k], rtol=tol, atol=tol, err_msg=k)
        N = x.shape[1]
        kept = int(np.unpackbits(a["reliable"][0].view(np.uint8)).sum())
        assert kept == N - F


def test_no_alloc_entry_point_refuses_a_batch_beyond_its_buffer():
    from svoc import _lib as L
    lib = L.load()
    T = SCRATCH_BYTES // (256 * 128 * 4) + 1          # the predictions alone exceed the buffer
    x = np.zeros((T, 256, 128), np.float32)
    rc, _ = _call(lib, L, lib.svoc_consensus_float_host_noalloc, x, 51, L.TIE_HIGH_INDEX_KEPT)
    msg = lib.svoc_last_error().decode()
    assert rc != 0 and "SVOC_NOALLOC_SCRATCH_BYTES" in msg and "rounds of this shape fit" in msg
    # the same call with a batch that fits still works afterwards
    rc, o = _call(lib, L, lib.svoc_consensus_float_host_noalloc, x[:4] + 0.5, 51, L.TIE_HIGH_INDEX_KEPT)
    assert rc == 0 and np.all(o["status"] == 0) and np.all(o["consensus"] == 0.5)
