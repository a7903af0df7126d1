"""ctypes binding of lib/libgpuprobe.so (include/gpu_probe.h) for tests/ and bench.py.

This is NOT the product's host side -- that is bash + the C binary
(kind-gpu-sim.sh -> pods/probe-*.yaml -> bin/gpu-probe).  It exists so the parity
tests and the bench can call the very same C ABI a maintainer would bind.

No fallback: if the shared library is missing or a call fails, this raises.
"""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libgpuprobe.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "gpu_probe.h")

SEED_B = 0x0123456789ABCDEF
SEED_C = 0xFEDCBA9876543210
SCALAR = 3.0

VARIANT_AUTO, VARIANT_LDG, VARIANT_TMA, VARIANT_LDG256, VARIANT_TMA_PACED = -1, 0, 1, 2, 3
VARIANTS = {"auto": VARIANT_AUTO, "ldg128": VARIANT_LDG, "tma": VARIANT_TMA, "ldg256": VARIANT_LDG256,
            "tma_paced": VARIANT_TMA_PACED}

EINVAL, EALIGN, ETOOBIG, ENOTREADY, ETIMEDOUT = -100, -101, -102, -103, -104
EXCHANGE_MAX_RANKS, EXCHANGE_HANDLE_BYTES = 16, 64
ALLREDUCE_AUTO, ALLREDUCE_PULL_PUSH, ALLREDUCE_PUSH, ALLREDUCE_MULTIMEM = 0, 1, 2, 3

_u64 = ctypes.c_uint64
_vp = ctypes.c_void_p


class ProbeError(RuntimeError):
    def __init__(self, rc: int, what: str):
        self.rc = rc
        super().__init__(f"{what}: rc={rc} ({error_string(rc)})")


_lib = None


def lib() -> ctypes.CDLL:
    """Load the CUDA library; raise (never fall back) when it is absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(
                f"{LIB_PATH} not built: run `make -C kind-gpu-sim_b200` (or __graft_entry__.build()); "
                "the probe has no CPU fallback")
        L = ctypes.CDLL(LIB_PATH)
        L.gpu_probe_abi_version.restype = ctypes.c_int
        L.gpu_probe_kernel_info.restype = ctypes.c_char_p
        L.gpu_probe_error_string.restype = ctypes.c_char_p
        L.gpu_probe_error_string.argtypes = [ctypes.c_int]
        L.gpu_probe_init_fill.argtypes = [_vp, _u64, _u64, _u64, _vp]
        L.gpu_probe_triad.argtypes = [_vp, _vp, _vp, ctypes.c_double, _u64, ctypes.c_int, _vp]
        L.gpu_probe_triad_calibrate.argtypes = [_vp, _vp, _vp, _u64, _vp]
        L.gpu_probe_triad_calibrate.restype = ctypes.c_int
        L.gpu_probe_checksum_workspace_bytes.restype = ctypes.c_size_t
        L.gpu_probe_checksum_u64.argtypes = [_vp, _u64, _u64, _vp, _vp, ctypes.c_int, _vp]
        L.gpu_probe_checksum_u64_noalloc.argtypes = [_vp, _u64, _u64, _vp, _vp]
        L.gpu_probe_checksum_u64_noalloc.restype = ctypes.c_int
        L.gpu_probe_triad_checksum.argtypes = [_vp, _vp, _vp, ctypes.c_double, _u64, _u64, _vp, _vp, ctypes.c_int, _vp]
        L.gpu_probe_p2p_copy.argtypes = [_vp, _vp, _u64, _vp]
        L.gpu_probe_pipeline_create.argtypes = [ctypes.POINTER(_vp), _u64, ctypes.c_int]
        L.gpu_probe_pipeline_run_host.argtypes = [_vp, _vp, _vp, _vp, ctypes.c_double, _u64, _u64,
                                                  ctypes.POINTER(_u64)]
        L.gpu_probe_pipeline_destroy.argtypes = [_vp]
        L.gpu_probe_exchange_create.argtypes = [ctypes.POINTER(_vp), ctypes.c_int, ctypes.c_int, _u64]
        L.gpu_probe_exchange_buffer.argtypes = [_vp]
        L.gpu_probe_exchange_buffer.restype = _vp
        L.gpu_probe_exchange_buffer_bytes.argtypes = [_vp]
        L.gpu_probe_exchange_buffer_bytes.restype = _u64
        L.gpu_probe_allreduce_f32.argtypes = [_vp, _u64, _u64, ctypes.c_int, ctypes.c_int, _vp]
        L.gpu_probe_allreduce_f32.restype = ctypes.c_int
        L.gpu_probe_exchange_connect_local.argtypes = [ctypes.POINTER(_vp), ctypes.c_int]
        L.gpu_probe_exchange_export.argtypes = [_vp, ctypes.c_char_p]
        L.gpu_probe_exchange_connect_ipc.argtypes = [_vp, ctypes.c_char_p]
        L.gpu_probe_exchange_set_timeout_ms.argtypes = [_vp, _u64]
        L.gpu_probe_exchange_status.argtypes = [_vp]
        L.gpu_probe_exchange_has_multicast.argtypes = [_vp]
        L.gpu_probe_exchange_has_multicast.restype = ctypes.c_int
        L.gpu_probe_exchange_destroy.argtypes = [_vp]
        L.gpu_probe_checksum_u64_exchange.argtypes = [_vp, _u64, _u64, _vp, _vp, ctypes.c_int, _vp, _vp]
        L.gpu_probe_triad_checksum_exchange.argtypes = [_vp, _vp, _vp, ctypes.c_double, _u64, _u64, _vp, _vp,
                                                        ctypes.c_int, _vp, _vp]
        L.gpu_probe_checksum_u64_exchange_group.argtypes = [ctypes.POINTER(_vp), ctypes.POINTER(_u64), ctypes.POINTER(_u64),
                                                            ctypes.POINTER(_vp), ctypes.POINTER(_vp), ctypes.POINTER(_vp),
                                                            ctypes.c_int, _vp]
        L.gpu_probe_checksum_u64_exchange_group.restype = ctypes.c_int
        L.gpu_probe_allreduce_f32_group.argtypes = [ctypes.POINTER(_vp), ctypes.c_int, _u64, _u64, ctypes.c_int,
                                                    ctypes.c_int, _vp]
        L.gpu_probe_allreduce_f32_group.restype = ctypes.c_int
        for name in ("gpu_probe_exchange_create", "gpu_probe_exchange_connect_local", "gpu_probe_exchange_export",
                     "gpu_probe_exchange_connect_ipc", "gpu_probe_exchange_set_timeout_ms", "gpu_probe_exchange_status",
                     "gpu_probe_exchange_destroy", "gpu_probe_checksum_u64_exchange",
                     "gpu_probe_triad_checksum_exchange"):
            getattr(L, name).restype = ctypes.c_int
        for name in ("gpu_probe_init_fill", "gpu_probe_triad", "gpu_probe_checksum_u64", "gpu_probe_triad_checksum",
                     "gpu_probe_p2p_copy", "gpu_probe_pipeline_create", "gpu_probe_pipeline_run_host",
                     "gpu_probe_pipeline_destroy"):
            getattr(L, name).restype = ctypes.c_int
        _lib = L
    return _lib


def error_string(rc: int) -> str:
    return lib().gpu_probe_error_string(rc).decode()


def kernel_info() -> str:
    return lib().gpu_probe_kernel_info().decode()


def triad_pace_ns() -> int:
    """AUTO's pacing on the current device: 0 never calibrated, -1 ring is as fast, -2 demoted by the drift guard, > 0 ns per refill."""
    L = lib()
    L.gpu_probe_triad_pace_ns.restype = ctypes.c_int
    return int(L.gpu_probe_triad_pace_ns())


def triad_calibrate(a: int, b: int, c: int, n: int, stream: int = 0) -> int:
    """[blocking] Calibrate AUTO's paced triad for the current device on the caller's buffers; returns the verdict."""
    _check(lib().gpu_probe_triad_calibrate(a, b, c, n, stream), "gpu_probe_triad_calibrate")
    return triad_pace_ns()


def triad_pace_report() -> "tuple[int, int]":
    """(duration in ns of the last finished paced AUTO launch, overruns in a row) on the current device."""
    out = (_u64 * 2)()
    L = lib()
    L.gpu_probe_triad_pace_report.restype = ctypes.c_int
    _check(L.gpu_probe_triad_pace_report(out), "gpu_probe_triad_pace_report")
    return int(out[0]), int(out[1])


def triad_recalibrate() -> None:
    L = lib()
    L.gpu_probe_triad_recalibrate.restype = ctypes.c_int
    _check(L.gpu_probe_triad_recalibrate(), "gpu_probe_triad_recalibrate")


def workspace_bytes() -> int:
    return int(lib().gpu_probe_checksum_workspace_bytes())


def _check(rc: int, what: str) -> None:
    if rc != 0:
        raise ProbeError(rc, what)


# The functions below take raw device addresses (ints) and a raw cudaStream_t (int).
# tests/ and bench.py get them from torch tensors (.data_ptr()) -- torch is only the
# allocator and the stream/event provider there.

def init_fill(dst: int, n: int, first_index: int, seed: int, stream: int = 0) -> None:
    _check(lib().gpu_probe_init_fill(dst, n, first_index & (2**64 - 1), seed, stream), "gpu_probe_init_fill")


def triad(a: int, b: int, c: int, s: float, n: int, variant: int = VARIANT_AUTO, stream: int = 0) -> None:
    _check(lib().gpu_probe_triad(a, b, c, s, n, variant, stream), "gpu_probe_triad")


def checksum_u64(w: int, n: int, first_index: int, out_lanes: int, workspace: int,
                 variant: int = VARIANT_AUTO, stream: int = 0) -> None:
    _check(lib().gpu_probe_checksum_u64(w, n, first_index & (2**64 - 1), out_lanes, workspace, variant, stream),
           "gpu_probe_checksum_u64")


NOALLOC_MAX_WORDS = 1 << 34


def checksum_u64_noalloc(w: int, n: int, first_index: int, out_lanes: int, stream: int = 0) -> None:
    """Checksum that uses the library's built-in device scratch: no workspace argument."""
    _check(lib().gpu_probe_checksum_u64_noalloc(w, n, first_index & (2**64 - 1), out_lanes, stream),
           "gpu_probe_checksum_u64_noalloc")


def triad_checksum(a: int, b: int, c: int, s: float, n: int, first_index: int, out_lanes: int, workspace: int,
                   variant: int = VARIANT_AUTO, stream: int = 0) -> None:
    _check(lib().gpu_probe_triad_checksum(a, b, c, s, n, first_index & (2**64 - 1), out_lanes, workspace, variant,
                                          stream), "gpu_probe_triad_checksum")


def p2p_copy(dst: int, src: int, nbytes: int, stream: int = 0) -> None:
    _check(lib().gpu_probe_p2p_copy(dst, src, nbytes, stream), "gpu_probe_p2p_copy")


class Pipeline:
    """Host-buffer end-to-end path (gpu_probe_pipeline_*)."""

    def __init__(self, chunk_elems: int = 0, depth: int = 0):
        self._h = _vp()
        _check(lib().gpu_probe_pipeline_create(ctypes.byref(self._h), chunk_elems, depth), "gpu_probe_pipeline_create")

    def run_host(self, b_host: int, c_host: int, a_host: int | None, s: float, n: int,
                 first_index: int = 0) -> tuple[int, int]:
        lanes = (_u64 * 2)()
        _check(lib().gpu_probe_pipeline_run_host(self._h, b_host, c_host, a_host or None, s, n,
                                                 first_index & (2**64 - 1), lanes), "gpu_probe_pipeline_run_host")
        return int(lanes[0]), int(lanes[1])

    def close(self) -> None:
        if self._h:
            _check(lib().gpu_probe_pipeline_destroy(self._h), "gpu_probe_pipeline_destroy")
            self._h = _vp()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


class Exchange:
    """One rank's mailbox for the fused checksum + lanes exchange (gpu_probe_exchange_*).
    Create with the rank's device current; then either Exchange.connect_local([...]) for
    ranks in one process, or export_handle() / connect_ipc(all_handles) across processes."""

    def __init__(self, rank: int, world: int, buffer_bytes: int = 0):
        self._h = _vp()
        self.rank, self.world = rank, world
        _check(lib().gpu_probe_exchange_create(ctypes.byref(self._h), rank, world, buffer_bytes),
               "gpu_probe_exchange_create")

    @property
    def buffer(self) -> int:
        """Device address of this rank's data buffer inside the segment (0 if none)."""
        return int(lib().gpu_probe_exchange_buffer(self._h) or 0)

    @property
    def buffer_bytes(self) -> int:
        return int(lib().gpu_probe_exchange_buffer_bytes(self._h))

    def allreduce_f32(self, offset_bytes: int, count: int, ctas: int = 0, stream: int = 0,
                      algo: int = 0) -> None:
        """algo: ALLREDUCE_AUTO / ALLREDUCE_PULL_PUSH / ALLREDUCE_PUSH."""
        _check(lib().gpu_probe_allreduce_f32(self._h, offset_bytes, count, algo, ctas, stream),
               "gpu_probe_allreduce_f32")

    @staticmethod
    def connect_local(xs: "list[Exchange]") -> None:
        arr = (_vp * len(xs))(*[x._h for x in xs])
        _check(lib().gpu_probe_exchange_connect_local(arr, len(xs)), "gpu_probe_exchange_connect_local")

    @staticmethod
    def checksum_u64_group(xs: "list[Exchange]", w: "list[int]", n: "list[int]", first_index: "list[int]",
                           out_lanes: "list[int]", workspace: "list[int]", stream: int = 0) -> None:
        """All ranks (exchanges on ONE device) in one launch: the single-GPU self-test form."""
        g = len(xs)
        _check(lib().gpu_probe_checksum_u64_exchange_group(
            (_vp * g)(*w), (_u64 * g)(*n), (_u64 * g)(*[f & (2**64 - 1) for f in first_index]), (_vp * g)(*out_lanes),
            (_vp * g)(*workspace), (_vp * g)(*[x._h for x in xs]), g, stream), "gpu_probe_checksum_u64_exchange_group")

    @staticmethod
    def allreduce_f32_group(xs: "list[Exchange]", offset_bytes: int, count: int, ctas: int = 0, stream: int = 0,
                            algo: int = 0) -> None:
        g = len(xs)
        _check(lib().gpu_probe_allreduce_f32_group((_vp * g)(*[x._h for x in xs]), g, offset_bytes, count, algo, ctas,
                                                   stream), "gpu_probe_allreduce_f32_group")

    def export_handle(self) -> bytes:
        buf = ctypes.create_string_buffer(EXCHANGE_HANDLE_BYTES)
        _check(lib().gpu_probe_exchange_export(self._h, buf), "gpu_probe_exchange_export")
        return buf.raw

    def connect_ipc(self, handles: bytes) -> None:
        if len(handles) != self.world * EXCHANGE_HANDLE_BYTES:
            raise ValueError("need world x 64 bytes of handles, in rank order")
        _check(lib().gpu_probe_exchange_connect_ipc(self._h, handles), "gpu_probe_exchange_connect_ipc")

    def set_timeout_ms(self, ms: int) -> None:
        _check(lib().gpu_probe_exchange_set_timeout_ms(self._h, ms), "gpu_probe_exchange_set_timeout_ms")

    def status(self) -> int:
        return int(lib().gpu_probe_exchange_status(self._h))

    @property
    def has_multicast(self) -> bool:
        return bool(lib().gpu_probe_exchange_has_multicast(self._h))

    def checksum_u64(self, w: int, n: int, first_index: int, out_lanes: int, workspace: int,
                     variant: int = VARIANT_AUTO, stream: int = 0) -> None:
        _check(lib().gpu_probe_checksum_u64_exchange(w, n, first_index & (2**64 - 1), out_lanes, workspace, variant,
                                                     self._h, stream), "gpu_probe_checksum_u64_exchange")

    def triad_checksum(self, a: int, b: int, c: int, s: float, n: int, first_index: int, out_lanes: int,
                       workspace: int, variant: int = VARIANT_AUTO, stream: int = 0) -> None:
        _check(lib().gpu_probe_triad_checksum_exchange(a, b, c, s, n, first_index & (2**64 - 1), out_lanes, workspace,
                                                       variant, self._h, stream), "gpu_probe_triad_checksum_exchange")

    def close(self) -> None:
        if self._h:
            _check(lib().gpu_probe_exchange_destroy(self._h), "gpu_probe_exchange_destroy")
            self._h = _vp()
