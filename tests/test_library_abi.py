"""CPU: libsvoc.so loads and exports every symbol include/svoc.h declares; argument errors are
reported without touching a GPU.  No compute call is made here."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "svoc.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(svoc_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported_and_bound():
    from svoc import _lib
    lib = _lib.load()
    names = _declared()
    assert len(names) >= 9
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/svoc.h but not exported"
        assert n in _lib.SYMBOLS, f"{n} has no ctypes signature in svoc/_lib.py"
    assert lib.svoc_version() == 101


def test_struct_layout_matches_header():
    from svoc import _lib
    assert C.sizeof(_lib.SvocConfig) == 56
    assert C.sizeof(_lib.SvocOutputs) == 10 * 8
    assert _lib.SvocConfig.max_spread_wsad.offset == 16 and _lib.SvocConfig.centre.offset == 32
    assert C.sizeof(_lib.SvocGenConfig) == 32 + 128 + 32 and _lib.SvocGenConfig.failing_prob.offset == 184


def test_argument_errors_without_gpu():
    from svoc import _lib
    lib = _lib.load()
    outs = _lib.SvocOutputs()
    assert lib.svoc_consensus_fixed(None, None, 1, C.byref(outs), None) == -1
    assert b"null" in lib.svoc_last_error()
    cfg = _lib.SvocConfig(n_oracles=7, dimension=2, n_failing=2, constrained=1, dtype=_lib.F32)
    assert lib.svoc_consensus_fixed(C.byref(cfg), C.c_void_p(16), 1, C.byref(outs), None) == -1
    assert b"dtype" in lib.svoc_last_error()
    cfg.dtype = _lib.I32
    assert lib.svoc_consensus_fixed(C.byref(cfg), C.c_void_p(16), 1, C.byref(outs), None) == -1
    assert b"required" in lib.svoc_last_error()
    assert lib.svoc_consensus_fixed(C.byref(cfg), None, 0, C.byref(outs), None) == -1  # outputs still required
    assert lib.svoc_kernel_variant(C.byref(cfg)) == b"small8"
    cfg.n_oracles = 256
    assert lib.svoc_kernel_variant(C.byref(cfg)) in (b"generic", b"wide", b"wide2", b"wide3")


def test_product_path_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under the package may reference it."""
    pkg = os.path.join(ROOT, "stochastic-vector-oracle-consensus_b200")
    for dp, _, files in os.walk(pkg):
        if "build" in dp.split(os.sep):
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dp, f)).read()
                assert "oracle." not in text.replace("oracle.cairo", "") or "import oracle" not in text, f
                assert "from oracle" not in text and "import oracle" not in text and "svoc_ref" not in text, f


def test_torch_front_end_refuses_cpu_tensors():
    import torch
    import svoc
    with pytest.raises(ValueError):
        svoc.consensus_fixed(torch.zeros((1, 7, 2), dtype=torch.int64), 2)
    with pytest.raises(TypeError):
        svoc.consensus_fixed_host(torch.zeros((1, 7, 2), dtype=torch.float32), 2)
