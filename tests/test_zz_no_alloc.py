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


def _call(lib, L, fn, x, F, tie, moments=False, extra=False):
    """One call of a *_host entry point on numpy buffers; returns the outputs as a dict.  `extra` also passes
    skew / kurt / rank buffers (pre-filled with a marker), `moments` sets cfg.want_moments."""
    T, N, D = x.shape
    f32 = x.dtype == np.float32
    cfg = L.SvocConfig(n_oracles=N, dimension=D, n_failing=F, constrained=1, max_spread_wsad=0, max_spread=0.0,
                       centre=L.CENTRE_SMOOTH_MEDIAN_CAIRO, centre2=L.CENTRE_DEFAULT, tie_rule=tie,
                       dtype=L.F32 if f32 else L.F64, want_moments=int(moments), flags=0)
    words = (N + 31) // 32
    o = {"essence1": np.full((T, D), np.nan, x.dtype), "qr": np.full((T, N), np.nan, x.dtype),
         "rel1": np.full(T, np.nan, x.dtype), "reliable": np.full((T, words), 0xdeadbeef, np.uint32),
         "consensus": np.full((T, D), np.nan, x.dtype), "rel2": np.full(T, np.nan, x.dtype),
         "status": np.full(T, -7, np.int32)}
    if extra:
        o["skew"] = np.full((T, D), MARKER, x.dtype)
        o["kurt"] = np.full((T, D), MARKER, x.dtype)
        o["rank"] = np.full((T, N), -7, np.int32)
    outs = L.SvocOutputs(essence1=o["essence1"].ctypes.data, qr=o["qr"].ctypes.data, rel1=o["rel1"].ctypes.data,
                         reliable=o["reliable"].ctypes.data, consensus=o["consensus"].ctypes.data,
                         rel2=o["rel2"].ctypes.data, skew=o["skew"].ctypes.data if extra else None,
                         kurt=o["kurt"].ctypes.data if extra else None, status=o["status"].ctypes.data,
                         rank=o["rank"].ctypes.data if extra else None)
    rc = fn(C.byref(cfg), x.ctypes.data_as(C.c_void_p), T, C.byref(outs))
    return rc, o


MARKER = -12345.0


def _oracle(x, F, tie, moments=False):
    """The CPU oracle (oracle/svoc_ref.c, float64) on the same rounds, mask packed like the library's."""
    from oracle import cref
    return cref.float_batch(x.astype(np.float64), F, centre=0, tie=tie, want_moments=moments, nthreads=0)


def _mask_bool(words, N):
    idx = np.arange(N)
    return ((words[:, idx // 32].astype(np.int64) >> (idx % 32)) & 1).astype(bool)


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
        # and against the CPU oracle, not only against the library's other entry point
        ora = _oracle(x, F, tie)
        assert np.array_equal(_mask_bool(a["reliable"], N), ora["reliable"])
        for k in ("essence1", "rel1", "consensus", "rel2"):
            np.testing.assert_allclose(a[k], ora[k], rtol=tol, atol=10 * tol, err_msg="oracle " + k)
        np.testing.assert_allclose(a["qr"], ora["qr"], rtol=10 * tol, atol=1e-9)


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


def test_no_alloc_entry_point_allocates_no_device_memory():
    """The property the entry point exists for: after one warm call (library loaded, kernels resident, the
    buffer's address looked up) a further call — of a NEW, larger shape, with every optional output — leaves
    the device's free memory exactly where it was, while the allocating entry point grows its staging cache for
    the same batch.  The optional outputs (rank, skew, kurt) are checked against the oracle on the way."""
    import torch
    from svoc import _lib as L
    lib = L.load()
    rng = np.random.default_rng(7)
    small = _rounds(rng, 40, 7, 2, 2, np.float64)
    rc, _ = _call(lib, L, lib.svoc_consensus_float_host_noalloc, small, 2, L.TIE_HIGH_INDEX_KEPT)    # warm call
    assert rc == 0, lib.svoc_last_error().decode()
    T, N, D, F = 1500, 64, 16, 12       # 14.4 MB of the 16 MiB buffer with every output on
    x = _rounds(rng, T, N, D, F, np.float64)
    # kernels of this shape resident before measuring (module loading is not what is measured here)
    rc, _ = _call(lib, L, lib.svoc_consensus_float_host_noalloc, x[:8], F, L.TIE_HIGH_INDEX_KEPT, moments=True, extra=True)
    assert rc == 0
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    rc, a = _call(lib, L, lib.svoc_consensus_float_host_noalloc, x, F, L.TIE_HIGH_INDEX_KEPT, moments=True, extra=True)
    assert rc == 0, lib.svoc_last_error().decode()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, f"the no-alloc call changed the device's free memory by {free0 - free1} bytes"
    ora = _oracle(x, F, 0, moments=True)
    assert np.all(a["status"] == 0) and np.array_equal(_mask_bool(a["reliable"], N), ora["reliable"])
    np.testing.assert_allclose(a["consensus"], ora["consensus"], rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose(a["skew"], ora["skew"], rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(a["kurt"], ora["kurt"], rtol=1e-6, atol=1e-5)
    order = np.argsort(a["rank"], axis=1)                    # rank = position in the (qr asc, index desc) order
    q = np.take_along_axis(ora["qr"], order, axis=1)
    assert np.all(q[:, 1:] >= q[:, :-1] * (1 - 1e-9)) and np.array_equal(np.sort(a["rank"], axis=1), np.tile(np.arange(N), (T, 1)))
    # the allocating entry point on a batch beyond anything it has staged so far (two 32 MiB prediction chunks of a
    # one-dimensional shape with every per-oracle output on: ~85 MiB per staging buffer) grows its cache
    big = _rounds(rng, 150_000, 64, 1, 12, np.float64)
    rc, _ = _call(lib, L, lib.svoc_consensus_float_host, big, 12, L.TIE_HIGH_INDEX_KEPT, moments=True, extra=True)
    assert rc == 0, lib.svoc_last_error().decode()
    free2, _ = torch.cuda.mem_get_info()
    assert free2 < free1


def test_moment_outputs_without_want_moments_stay_untouched():
    """skew / kurt buffers passed with cfg.want_moments = 0: the kernels do not produce them, so neither host
    entry point may hand back staging bytes of an earlier call — the caller's buffers keep their contents."""
    from svoc import _lib as L
    lib = L.load()
    rng = np.random.default_rng(8)
    x = _rounds(rng, 200, 20, 3, 4, np.float64)
    for fn in (lib.svoc_consensus_float_host_noalloc, lib.svoc_consensus_float_host):
        rc, a = _call(lib, L, fn, x, 4, L.TIE_HIGH_INDEX_KEPT, moments=True, extra=True)     # leaves moments in staging
        assert rc == 0 and not np.any(a["skew"] == MARKER)
        rc, b = _call(lib, L, fn, x, 4, L.TIE_HIGH_INDEX_KEPT, moments=False, extra=True)
        assert rc == 0
        assert np.all(b["skew"] == MARKER) and np.all(b["kurt"] == MARKER)
        assert np.array_equal(a["consensus"], b["consensus"]) and np.array_equal(a["rank"], b["rank"])
