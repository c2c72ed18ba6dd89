"""ctypes loader for oracle/libsvoc_ref.so (the C restatement, oracle/svoc_ref.c).

TEST INFRASTRUCTURE ONLY — used by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs, never by the product path.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libsvoc_ref.so")
_lib = None


def _cpu_stamp() -> str:
    """Identifies the host CPU: the library is built with -march=native, so one built on another machine
    (it travels with the tree) must be rebuilt before it is loaded."""
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name"):
                    flags = next((l for l in f if l.startswith("flags")), "")
                    import zlib
                    return f"{line.split(':', 1)[1].strip()} / {zlib.crc32(flags.encode()):08x}"
    except OSError:
        pass
    return "unknown"


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "svoc_ref.c")
    stamp_file = _SO + ".cpu"
    stamp = _cpu_stamp()
    try:
        same_cpu = open(stamp_file).read() == stamp
    except OSError:
        same_cpu = False
    if force or not same_cpu or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        env = dict(os.environ)
        env["CC"] = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc"
        subprocess.check_call(["make", "-B", "-C", _HERE, "libsvoc_ref.so"], env=env,
                              stdout=subprocess.DEVNULL)
        with open(stamp_file, "w") as f:
            f.write(stamp)
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.svoc_ref_max_threads.restype = C.c_int
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def max_threads() -> int:
    return int(lib().svoc_ref_max_threads())


def fixed_batch(pred, n_failing, constrained=True, max_spread=0, want_moments=True, nthreads=1):
    x = np.ascontiguousarray(pred, dtype=np.int64)
    T, N, D = x.shape
    o = {
        "essence1": np.zeros((T, D), np.int64), "qr": np.zeros((T, N), np.int64),
        "rel1": np.zeros(T, np.int64), "reliable": np.zeros((T, N), np.uint8),
        "consensus": np.zeros((T, D), np.int64), "rel2": np.zeros(T, np.int64),
        "skew": np.zeros((T, D), np.int64), "kurt": np.zeros((T, D), np.int64),
        "status": np.zeros(T, np.int32),
    }
    rc = lib().svoc_ref_fixed(_p(x), C.c_long(T), N, D, int(n_failing), int(bool(constrained)),
                              C.c_int64(int(max_spread)), int(bool(want_moments)),
                              _p(o["essence1"]), _p(o["qr"]), _p(o["rel1"]), _p(o["reliable"]),
                              _p(o["consensus"]), _p(o["rel2"]), _p(o["skew"]), _p(o["kurt"]),
                              _p(o["status"]), int(nthreads))
    if rc != 0:
        raise ValueError(f"svoc_ref_fixed rc={rc}")
    o["reliable"] = o["reliable"].astype(bool)
    return o


def float_batch(pred, n_failing, constrained=True, max_spread=0.0, centre=0, centre2=None, tie=0,
                want_moments=True, nthreads=1):
    x = np.ascontiguousarray(pred, dtype=np.float64)
    T, N, D = x.shape
    if centre2 is None:
        centre2 = centre if constrained else 2
    o = {
        "essence1": np.zeros((T, D)), "qr": np.zeros((T, N)), "rel1": np.zeros(T),
        "reliable": np.zeros((T, N), np.uint8), "consensus": np.zeros((T, D)), "rel2": np.zeros(T),
        "skew": np.zeros((T, D)), "kurt": np.zeros((T, D)),
    }
    rc = lib().svoc_ref_float(_p(x), C.c_long(T), N, D, int(n_failing), int(bool(constrained)),
                              C.c_double(float(max_spread)), int(centre), int(centre2), int(tie),
                              int(bool(want_moments)), _p(o["essence1"]), _p(o["qr"]), _p(o["rel1"]),
                              _p(o["reliable"]), _p(o["consensus"]), _p(o["rel2"]), _p(o["skew"]),
                              _p(o["kurt"]), int(nthreads))
    if rc != 0:
        raise ValueError(f"svoc_ref_float rc={rc}")
    o["reliable"] = o["reliable"].astype(bool)
    return o
