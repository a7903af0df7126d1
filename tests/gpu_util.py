"""Helpers for the -m gpu tests: torch is the allocator / stream provider only;
every compute call goes through the C ABI (gpuprobe -> lib/libgpuprobe.so)."""
from __future__ import annotations

import numpy as np
import torch

import gpuprobe

M64 = (1 << 64) - 1


def stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def f64(n: int, offset_elems: int = 0) -> torch.Tensor:
    """n doubles on the device; offset_elems=1 gives an 8- but not 16-byte-aligned view."""
    base = torch.empty(n + offset_elems + 4, dtype=torch.float64, device="cuda")
    return base[offset_elems:offset_elems + n]


def workspace() -> torch.Tensor:
    return torch.zeros(gpuprobe.workspace_bytes() // 8, dtype=torch.int64, device="cuda")


def lanes_buf() -> torch.Tensor:
    return torch.full((2,), -1, dtype=torch.int64, device="cuda")


def lanes_of(t: torch.Tensor) -> tuple[int, int]:
    v = t.cpu().numpy().view(np.uint64)
    return int(v[0]), int(v[1])


def to_u64(t: torch.Tensor) -> np.ndarray:
    return t.cpu().numpy().view(np.uint64)


def from_np(a: np.ndarray, offset_elems: int = 0) -> torch.Tensor:
    """Copy a float64/uint64 numpy array to the device as float64 storage."""
    t = f64(a.size, offset_elems)
    t.copy_(torch.from_numpy(a.view(np.float64)))
    return t


def gpu_init(n: int, first: int, seed: int, offset_elems: int = 0) -> torch.Tensor:
    t = f64(n, offset_elems)
    gpuprobe.init_fill(t.data_ptr(), n, first, seed, stream())
    return t


def gpu_checksum(t: torch.Tensor, first: int, variant: int, ws: torch.Tensor | None = None) -> tuple[int, int]:
    ws = workspace() if ws is None else ws
    out = lanes_buf()
    gpuprobe.checksum_u64(t.data_ptr(), t.numel(), first, out.data_ptr(), ws.data_ptr(), variant, stream())
    torch.cuda.synchronize()
    return lanes_of(out)


def hx(l) -> list[str]:
    return [f"0x{l[0]:016x}", f"0x{l[1]:016x}"]
