"""ctypes loader for the CPU restatement (oracle/probe_ref.c).

TEST INFRASTRUCTURE, NOT PRODUCT.  Only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import this
package.  Nothing under ``kind-gpu-sim_b200/`` imports, links or executes it.

PARITY UNPINNED against the reference: maryamtahhan/kind-gpu-sim has no
arithmetic (``/root/reference/pods/nvidia-gpu-test-pod.yaml:9`` is an ``echo``),
so this follows the spec in SURVEY.md section 8(d) and is pinned to the
published SplitMix64 vector and to ``tests/pyspec.py`` instead.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libprobe_ref.so")

SEED_B = 0x0123456789ABCDEF
SEED_C = 0xFEDCBA9876543210
SCALAR = 3.0

_u64 = ctypes.c_uint64
_dp = ctypes.POINTER(ctypes.c_double)
_up = ctypes.POINTER(ctypes.c_uint64)


def build(force: bool = False) -> str:
    """Compile oracle/libprobe_ref.so with the committed Makefile."""
    src_m = max(os.path.getmtime(os.path.join(_HERE, f)) for f in ("probe_ref.c", "probe_ref.h", "Makefile"))
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < src_m:
        subprocess.run(["make", "-C", _HERE, "-B", "libprobe_ref.so"], check=True, capture_output=True)
    return _LIB_PATH


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_LIB_PATH)
        L.ref_gen.restype = _u64
        L.ref_gen.argtypes = [_u64, _u64]
        L.ref_init_fill.restype = None
        L.ref_init_fill.argtypes = [_dp, _u64, _u64, _u64]
        L.ref_triad.restype = None
        L.ref_triad.argtypes = [_dp, _dp, _dp, ctypes.c_double, _u64]
        L.ref_triad_unfused.restype = None
        L.ref_triad_unfused.argtypes = [_dp, _dp, _dp, ctypes.c_double, _u64]
        L.ref_checksum_u64.restype = None
        L.ref_checksum_u64.argtypes = [_up, _u64, _u64, _up]
        L.ref_stream_lanes.restype = None
        L.ref_stream_lanes.argtypes = [_u64, _u64, _u64, _u64, ctypes.c_double, _up]
        L.ref_max_threads.restype = ctypes.c_int
        L.ref_set_threads.argtypes = [ctypes.c_int]
        _lib = L
    return _lib


def _d(a: np.ndarray):
    assert a.dtype == np.float64 and a.flags.c_contiguous
    return a.ctypes.data_as(_dp)


def _u(a: np.ndarray):
    assert a.dtype in (np.uint64, np.float64) and a.flags.c_contiguous
    return a.ctypes.data_as(_up)


def gen(i: int, seed: int) -> int:
    return int(lib().ref_gen(i & (2**64 - 1), seed & (2**64 - 1)))


def init_fill(n: int, first_index: int, seed: int) -> np.ndarray:
    out = np.empty(n, dtype=np.float64)
    lib().ref_init_fill(_d(out), n, first_index, seed)
    return out


def triad(b: np.ndarray, c: np.ndarray, s: float = SCALAR, out: np.ndarray | None = None) -> np.ndarray:
    a = np.empty_like(b) if out is None else out
    lib().ref_triad(_d(a), _d(b), _d(c), s, b.size)
    return a


def triad_unfused(b: np.ndarray, c: np.ndarray, s: float = SCALAR) -> np.ndarray:
    a = np.empty_like(b)
    lib().ref_triad_unfused(_d(a), _d(b), _d(c), s, b.size)
    return a


def checksum_u64(w: np.ndarray, first_index: int = 0) -> tuple[int, int]:
    """Lanes over the 64-bit words of ``w`` (uint64, or float64 bit patterns)."""
    lanes = (ctypes.c_uint64 * 2)()
    lib().ref_checksum_u64(_u(w), w.size, first_index & (2**64 - 1), lanes)
    return int(lanes[0]), int(lanes[1])


def stream_lanes(n: int, first_index: int = 0, seed_b: int = SEED_B, seed_c: int = SEED_C,
                 s: float = SCALAR) -> tuple[int, int]:
    lanes = (ctypes.c_uint64 * 2)()
    lib().ref_stream_lanes(n, first_index & (2**64 - 1), seed_b, seed_c, s, lanes)
    return int(lanes[0]), int(lanes[1])


def max_threads() -> int:
    return int(lib().ref_max_threads())


def set_threads(n: int) -> None:
    lib().ref_set_threads(int(n))
