"""ctypes binding of libsvoc.so (C ABI declared in include/svoc.h).

The library is the product: there is no Python or CPU fallback.  If the shared object is
missing, importing this module still works (so host-only helpers stay usable on a box
without the build), but every compute call raises SvocLibraryError.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SVOC_LIB") or os.path.join(_HERE, "libsvoc.so")   # SVOC_LIB: A/B builds while tuning

I32, I64, F32, F64 = 0, 1, 2, 3
CENTRE_SMOOTH_MEDIAN_CAIRO, CENTRE_MEDIAN, CENTRE_MEAN, CENTRE_DEFAULT = 0, 1, 2, -1
TIE_HIGH_INDEX_KEPT, TIE_LOW_INDEX_KEPT = 0, 1

STATUS_NAMES = {
    0: "OK", 1: "INTERVAL_INPUT", 2: "SQRT_REL1", 3: "MAX_SPREAD_ZERO", 4: "REL1_RANGE",
    5: "SECOND_PASS_SHAPE", 6: "SQRT_REL2", 7: "REL2_RANGE", 8: "SKEW_PANIC", 9: "KURT_PANIC",
    10: "OVERFLOW", 11: "FIRST_PASS_SHAPE", 12: "ENGINE_RANGE",
}


class SvocLibraryError(RuntimeError):
    pass


class SvocConfig(C.Structure):
    _fields_ = [
        ("n_oracles", C.c_int32), ("dimension", C.c_int32), ("n_failing", C.c_int32),
        ("constrained", C.c_int32), ("max_spread_wsad", C.c_int64), ("max_spread", C.c_double),
        ("centre", C.c_int32), ("centre2", C.c_int32), ("tie_rule", C.c_int32), ("dtype", C.c_int32),
        ("want_moments", C.c_int32), ("flags", C.c_int32),
    ]


class SvocOutputs(C.Structure):
    _fields_ = [
        ("essence1", C.c_void_p), ("qr", C.c_void_p), ("rel1", C.c_void_p), ("reliable", C.c_void_p),
        ("consensus", C.c_void_p), ("rel2", C.c_void_p), ("skew", C.c_void_p), ("kurt", C.c_void_p),
        ("status", C.c_void_p), ("rank", C.c_void_p),
    ]


class SvocGenConfig(C.Structure):
    _fields_ = [
        ("n_oracles", C.c_int32), ("dimension", C.c_int32), ("n_failing", C.c_int32), ("law", C.c_int32),
        ("failing", C.c_int32), ("dtype", C.c_int32), ("n_params", C.c_int32), ("reserved", C.c_int32),
        ("a", C.c_double * 8), ("b", C.c_double * 8), ("shift", C.c_double), ("seed", C.c_uint64),
        ("trial_offset", C.c_int64), ("failing_prob", C.c_double),
    ]


LAW_BETA, LAW_GAUSSIAN, LAW_KUMARASWAMY, LAW_ARCTAN_GAUSSIAN = 0, 1, 2, 3
FLAG_OBSOLETE_RELIABILITY = 1
FAIL_UNIFORM, FAIL_COLLUDING = 0, 1
STATS_COUNTS_BASE, STATS_SUMS_BASE, STATS_HIST_BINS, MC_EXTRA = 528, 4, 256, 2

# every symbol include/svoc.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "svoc_consensus_fixed": (C.c_int, [C.POINTER(SvocConfig), C.c_void_p, C.c_int64, C.POINTER(SvocOutputs), C.c_void_p]),
    "svoc_consensus_float": (C.c_int, [C.POINTER(SvocConfig), C.c_void_p, C.c_int64, C.POINTER(SvocOutputs), C.c_void_p]),
    "svoc_consensus_fixed_host": (C.c_int, [C.POINTER(SvocConfig), C.c_void_p, C.c_int64, C.POINTER(SvocOutputs)]),
    "svoc_consensus_float_host": (C.c_int, [C.POINTER(SvocConfig), C.c_void_p, C.c_int64, C.POINTER(SvocOutputs)]),
    "svoc_consensus_float_host_noalloc": (C.c_int, [C.POINTER(SvocConfig), C.c_void_p, C.c_int64, C.POINTER(SvocOutputs)]),
    "svoc_expected_reliability": (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p]),
    "svoc_rank_array": (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p]),
    "svoc_generate": (C.c_int, [C.POINTER(SvocGenConfig), C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "svoc_montecarlo": (C.c_int, [C.POINTER(SvocConfig), C.POINTER(SvocGenConfig), C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "svoc_accumulate_stats": (C.c_int, [C.POINTER(SvocConfig), C.c_int64, C.POINTER(SvocOutputs), C.c_void_p, C.c_void_p, C.c_void_p]),
    "svoc_last_error": (C.c_char_p, []),
    "svoc_version": (C.c_int, []),
    "svoc_launch_count": (C.c_int64, []),
    "svoc_kernel_variant": (C.c_char_p, [C.POINTER(SvocConfig)]),
}

_lib = None


def load():
    """dlopen libsvoc.so and type every entry point; raises when the build is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SvocLibraryError(
                f"{LIB_PATH} not found: build it with `python stochastic-vector-oracle-consensus_b200/build.py` "
                "(there is no CPU fallback)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc: int):
    if rc != 0:
        raise SvocLibraryError(f"libsvoc error {rc}: {load().svoc_last_error().decode()}")


def launch_count() -> int:
    return int(load().svoc_launch_count())
