"""Torch-tensor front end of libsvoc: batched consensus rounds on the current CUDA device.

`consensus_fixed` / `consensus_float` take a [T, N, D] CUDA tensor and launch the fused
sm_100a kernel on torch's current stream; `*_host` variants take pinned/pageable CPU tensors
and go through the library's own chunked H2D -> kernel -> D2H pipeline.  PyTorch is only the
memory/stream plumbing: all arithmetic happens in libsvoc.so.  No CPU fallback exists.

Reference semantics: contract/src/contract.cairo:442-503 / :370-434 (fixed form) and the
notebook estimator ipynb beta :354-371 (float form); see include/svoc.h.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional

import torch

from . import _lib
from ._lib import (CENTRE_DEFAULT, CENTRE_MEAN, CENTRE_MEDIAN, CENTRE_SMOOTH_MEDIAN_CAIRO,  # noqa: F401
                   TIE_HIGH_INDEX_KEPT, TIE_LOW_INDEX_KEPT, SvocConfig, SvocOutputs)

WSAD = 1_000_000

_DTYPES = {torch.int32: _lib.I32, torch.int64: _lib.I64, torch.float32: _lib.F32, torch.float64: _lib.F64}


@dataclass
class ConsensusResult:
    """Outputs of T rounds.  Value tensors are int64 (fixed form) or the input dtype (float form)."""
    consensus: torch.Tensor                 # [T, D]  get_consensus_value()
    rel1: torch.Tensor                      # [T]     first-pass reliability
    rel2: torch.Tensor                      # [T]     second-pass reliability
    status: torch.Tensor                    # [T]     int32, 0 = ok, else the contract's first panic
    reliable_words: Optional[torch.Tensor]  # [T, ceil(N/32)] int32 bit mask
    essence1: Optional[torch.Tensor]        # [T, D]  first-pass essence
    qr: Optional[torch.Tensor]              # [T, N]  per-oracle quadratic risk
    skew: Optional[torch.Tensor]            # [T, D]
    kurt: Optional[torch.Tensor]            # [T, D]
    rank: Optional[torch.Tensor] = None     # [T, N] int32 position in the selection order (0 = smallest qr)
    n_oracles: int = 0

    @property
    def reliable(self) -> torch.Tensor:
        """[T, N] bool — OracleInfo.reliable of every oracle (contract.cairo:345-363)."""
        w = self.reliable_words
        idx = torch.arange(self.n_oracles, device=w.device)
        return ((w[:, idx // 32] >> (idx % 32)) & 1).bool()


def make_config(n, d, n_failing, dtype, constrained=True, max_spread=0, centre=CENTRE_SMOOTH_MEDIAN_CAIRO,
                centre2=CENTRE_DEFAULT, tie_rule=TIE_HIGH_INDEX_KEPT, want_moments=False, obsolete_reliability=False) -> SvocConfig:
    fixed = dtype in (_lib.I32, _lib.I64)
    return SvocConfig(n_oracles=int(n), dimension=int(d), n_failing=int(n_failing), constrained=int(bool(constrained)),
                      max_spread_wsad=int(max_spread) if fixed else 0, max_spread=0.0 if fixed else float(max_spread),
                      centre=int(centre), centre2=int(centre2), tie_rule=int(tie_rule), dtype=int(dtype),
                      want_moments=int(bool(want_moments)),
                      flags=_lib.FLAG_OBSOLETE_RELIABILITY if obsolete_reliability else 0)


def _alloc_outputs(T, N, D, val_dtype, device, want_moments, want_qr, want_essence, pin=False, want_rank=False):
    kw = dict(device=device)
    if pin:
        kw["pin_memory"] = True
    e = lambda *s, dt=val_dtype: torch.empty(*s, dtype=dt, **kw)  # noqa: E731
    res = ConsensusResult(
        consensus=e(T, D), rel1=e(T), rel2=e(T), status=e(T, dt=torch.int32),
        reliable_words=e(T, (N + 31) // 32, dt=torch.int32),
        essence1=e(T, D) if want_essence else None, qr=e(T, N) if want_qr else None,
        skew=e(T, D) if want_moments else None, kurt=e(T, D) if want_moments else None,
        rank=e(T, N, dt=torch.int32) if want_rank else None, n_oracles=N)
    ptr = lambda t: C.c_void_p(t.data_ptr()) if t is not None and t.numel() > 0 else None  # noqa: E731
    outs = SvocOutputs(essence1=ptr(res.essence1), qr=ptr(res.qr), rel1=ptr(res.rel1), reliable=ptr(res.reliable_words),
                       consensus=ptr(res.consensus), rel2=ptr(res.rel2), skew=ptr(res.skew), kurt=ptr(res.kurt),
                       status=ptr(res.status), rank=ptr(res.rank))
    return res, outs


def _run(pred: torch.Tensor, fixed: bool, host: bool, n_failing, constrained, max_spread, centre, centre2, tie_rule,
         want_moments, want_qr, want_essence, want_rank=False, obsolete_reliability=False) -> ConsensusResult:
    if pred.dim() != 3:
        raise ValueError("predictions must be [T, N, D]")
    if pred.dtype not in _DTYPES:
        raise TypeError(f"unsupported dtype {pred.dtype}")
    dt = _DTYPES[pred.dtype]
    if fixed != (dt in (_lib.I32, _lib.I64)):
        raise TypeError("fixed form takes int32/int64, float form takes float32/float64")
    if host == pred.is_cuda:
        raise ValueError("device entry points take CUDA tensors, *_host entry points take CPU tensors")
    pred = pred.contiguous()
    T, N, D = pred.shape
    lib = _lib.load()
    cfg = make_config(N, D, n_failing, dt, constrained, max_spread, centre, centre2, tie_rule, want_moments,
                      obsolete_reliability)
    val_dtype = torch.int64 if fixed else pred.dtype
    res, outs = _alloc_outputs(T, N, D, val_dtype, pred.device, want_moments, want_qr, want_essence,
                               pin=host and pred.is_pinned(), want_rank=want_rank)
    if T == 0:
        return res
    if host:
        fn = lib.svoc_consensus_fixed_host if fixed else lib.svoc_consensus_float_host
        _lib.check(fn(C.byref(cfg), C.c_void_p(pred.data_ptr()), T, C.byref(outs)))
    else:
        fn = lib.svoc_consensus_fixed if fixed else lib.svoc_consensus_float
        with torch.cuda.device(pred.device):
            stream = torch.cuda.current_stream().cuda_stream
            _lib.check(fn(C.byref(cfg), C.c_void_p(pred.data_ptr()), T, C.byref(outs), C.c_void_p(stream)))
    return res


def consensus_fixed(pred, n_failing, constrained=True, max_spread=0, centre=CENTRE_SMOOTH_MEDIAN_CAIRO,
                    centre2=CENTRE_DEFAULT, tie_rule=TIE_HIGH_INDEX_KEPT, want_moments=True, want_qr=True,
                    want_essence=True, want_rank=False, obsolete_reliability=False) -> ConsensusResult:
    """T rounds of update_{un,}constrained_consensus on wsad integers (CUDA tensor in, CUDA tensors out)."""
    return _run(pred, True, False, n_failing, constrained, max_spread, centre, centre2, tie_rule, want_moments,
                want_qr, want_essence, want_rank, obsolete_reliability)


def consensus_float(pred, n_failing, constrained=True, max_spread=0.0, centre=CENTRE_SMOOTH_MEDIAN_CAIRO,
                    centre2=CENTRE_DEFAULT, tie_rule=TIE_HIGH_INDEX_KEPT, want_moments=False, want_qr=True,
                    want_essence=True, want_rank=False) -> ConsensusResult:
    """T rounds in float32/float64 (CUDA tensor in, CUDA tensors out)."""
    return _run(pred, False, False, n_failing, constrained, max_spread, centre, centre2, tie_rule, want_moments,
                want_qr, want_essence, want_rank)


def consensus_fixed_host(pred, n_failing, **kw) -> ConsensusResult:
    """Same as consensus_fixed for a CPU tensor (pinned for full PCIe speed); outputs are CPU tensors."""
    kw = {**dict(constrained=True, max_spread=0, centre=CENTRE_SMOOTH_MEDIAN_CAIRO, centre2=CENTRE_DEFAULT,
                 tie_rule=TIE_HIGH_INDEX_KEPT, want_moments=True, want_qr=True, want_essence=True), **kw}
    return _run(pred, True, True, n_failing, **kw)


def consensus_float_host(pred, n_failing, **kw) -> ConsensusResult:
    kw = {**dict(constrained=True, max_spread=0.0, centre=CENTRE_SMOOTH_MEDIAN_CAIRO, centre2=CENTRE_DEFAULT,
                 tie_rule=TIE_HIGH_INDEX_KEPT, want_moments=False, want_qr=True, want_essence=True), **kw}
    return _run(pred, False, True, n_failing, **kw)


def expected_reliability(arrays: torch.Tensor) -> torch.Tensor:
    """[T, ...] float64 CUDA tensor -> [T]: 1 - 2*mean|A - median(A)| per flattened array (ipynb beta :260-267)."""
    if not arrays.is_cuda or arrays.dtype != torch.float64:
        raise TypeError("expected a float64 CUDA tensor")
    a = arrays.contiguous().view(arrays.shape[0], -1)
    out = torch.empty(a.shape[0], dtype=torch.float64, device=a.device)
    if a.shape[0] == 0:
        return out
    with torch.cuda.device(a.device):
        stream = torch.cuda.current_stream().cuda_stream
        _lib.check(_lib.load().svoc_expected_reliability(C.c_void_p(a.data_ptr()), a.shape[0], a.shape[1],
                                                         C.c_void_p(out.data_ptr()), C.c_void_p(stream)))
    return out


def rank_array(values: torch.Tensor) -> torch.Tensor:
    """[T, n] float64 CUDA tensor -> [T, n] int32 ranks, largest value = rank 0 (ipynb beta :342-352)."""
    if not values.is_cuda or values.dtype != torch.float64 or values.dim() != 2:
        raise TypeError("expected a [T, n] float64 CUDA tensor")
    v = values.contiguous()
    out = torch.empty(v.shape, dtype=torch.int32, device=v.device)
    if v.shape[0]:
        with torch.cuda.device(v.device):
            stream = torch.cuda.current_stream().cuda_stream
            _lib.check(_lib.load().svoc_rank_array(C.c_void_p(v.data_ptr()), v.shape[0], v.shape[1],
                                                   C.c_void_p(out.data_ptr()), C.c_void_p(stream)))
    return out


def kernel_variant(n, d, dtype=torch.float32) -> str:
    cfg = make_config(n, d, 0, _DTYPES[dtype])
    return _lib.load().svoc_kernel_variant(C.byref(cfg)).decode()
