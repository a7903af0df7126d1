"""CPU-side checks of the C ABI: the library loads, exports every symbol that
include/gpu_probe.h declares, validates arguments, and FAILS LOUDLY (CUDA error,
never a CPU result) when asked to compute on a machine without a GPU."""
import ctypes
import os
import re
import subprocess

import pytest

import gpuprobe

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(gpuprobe.HEADER_PATH).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gpu_probe_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    L = gpuprobe.lib()
    names = declared_symbols()
    assert len(names) >= 12
    for name in names:
        assert hasattr(L, name), f"{name} declared in include/gpu_probe.h but not exported"
    # and nothing is exported under the prefix that the header does not declare
    out = subprocess.run(["nm", "-D", "--defined-only", gpuprobe.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = sorted(set(re.findall(r" T (gpu_probe_[a-z0-9_]+)", out)))
    assert exported == names


def test_abi_version_info_and_error_strings():
    L = gpuprobe.lib()
    header = open(gpuprobe.HEADER_PATH).read()
    assert L.gpu_probe_abi_version() == int(re.search(r"#define GPU_PROBE_ABI_VERSION (\d+)", header).group(1))
    info = gpuprobe.kernel_info()
    assert "sm_100a" in info and "cpu_fallback=none" in info
    assert gpuprobe.error_string(0) == "ok"
    for rc in (gpuprobe.EINVAL, gpuprobe.EALIGN, gpuprobe.ETOOBIG, gpuprobe.ENOTREADY, -5, -1234, 1, 700):
        assert gpuprobe.error_string(rc)
    assert gpuprobe.workspace_bytes() >= 16 * 1184 + 8  # room for 148 SMs x 8 CTAs of partial lanes + ticket


def test_argument_validation_needs_no_gpu():
    L = gpuprobe.lib()
    assert L.gpu_probe_init_fill(None, 0, 0, 0, None) == 0          # n == 0 is a no-op
    assert L.gpu_probe_init_fill(None, 8, 0, 0, None) == gpuprobe.EINVAL
    assert L.gpu_probe_init_fill(0x1004, 8, 0, 0, None) == gpuprobe.EALIGN
    assert L.gpu_probe_triad(None, None, None, 3.0, 0, 0, None) == 0
    assert L.gpu_probe_triad(None, None, None, 3.0, 4, 0, None) == gpuprobe.EINVAL
    assert L.gpu_probe_triad(0x1000, 0x2000, 0x3000, 3.0, 4, 7, None) == gpuprobe.EINVAL   # unknown variant
    assert L.gpu_probe_triad(0x1000, 0x2004, 0x3000, 3.0, 4, 0, None) == gpuprobe.EALIGN
    assert L.gpu_probe_triad_calibrate(None, 0x2000, 0x3000, 1 << 26, None) == gpuprobe.EINVAL
    assert L.gpu_probe_triad_calibrate(0x1004, 0x2000, 0x3000, 4, None) == gpuprobe.EALIGN
    assert L.gpu_probe_triad_calibrate(0x1000, 0x1000, 0x3000, 4, None) == gpuprobe.EINVAL   # a aliases b
    assert L.gpu_probe_triad_calibrate(0x1000, 0x2000, 0x1010, 4, None) == gpuprobe.EINVAL   # a overlaps c
    assert L.gpu_probe_triad_calibrate(0x1000, 0x2000, 0x3000, 4, None) == 0                 # too small to pace: no-op
    assert L.gpu_probe_checksum_u64(0x1000, 4, 0, None, None, 0, None) == gpuprobe.EINVAL
    assert L.gpu_probe_checksum_u64(0x1000, 4, 0, 0x2000, 0x3000, -2, None) == gpuprobe.EINVAL
    assert L.gpu_probe_triad_checksum(None, None, None, 3.0, 4, 0, 0x2000, 0x3000, 0, None) == gpuprobe.EINVAL
    assert L.gpu_probe_p2p_copy(None, None, 0, None) == 0
    assert L.gpu_probe_p2p_copy(None, 0x1000, 16, None) == gpuprobe.EINVAL
    assert L.gpu_probe_pipeline_create(None, 0, 0) == gpuprobe.EINVAL
    h = ctypes.c_void_p()
    assert L.gpu_probe_pipeline_create(ctypes.byref(h), 3, 4) == gpuprobe.EINVAL   # odd chunk
    assert L.gpu_probe_pipeline_create(ctypes.byref(h), 1 << 20, 1) == gpuprobe.EINVAL  # depth < 2
    assert L.gpu_probe_pipeline_destroy(None) == gpuprobe.ENOTREADY
    assert L.gpu_probe_exchange_create(None, 0, 1, 0) == gpuprobe.EINVAL
    assert L.gpu_probe_exchange_create(ctypes.byref(h), 0, 17, 0) == gpuprobe.EINVAL
    assert L.gpu_probe_exchange_create(ctypes.byref(h), 0, 2, 1 << 40) == gpuprobe.ETOOBIG
    assert L.gpu_probe_exchange_status(None) == gpuprobe.ENOTREADY
    assert L.gpu_probe_allreduce_f32(None, 0, 4, 0, 0, None) == gpuprobe.ENOTREADY
    assert L.gpu_probe_checksum_u64_exchange(0x1000, 4, 0, 0x2000, 0x3000, 0, None, None) == gpuprobe.ENOTREADY
    assert L.gpu_probe_exchange_buffer(None) is None and L.gpu_probe_exchange_buffer_bytes(None) == 0


def _no_gpu():
    try:
        import torch
        return not torch.cuda.is_available()
    except Exception:
        return True


@pytest.mark.skipif(not _no_gpu(), reason="only meaningful on a machine without a GPU")
def test_compute_without_gpu_fails_loudly():
    # no CPU fallback: a compute call on a GPU-less host must return a CUDA error
    rc = gpuprobe.lib().gpu_probe_init_fill(0x10000, 1024, 0, 1, None)
    assert rc > 0, rc
    with pytest.raises(gpuprobe.ProbeError):
        gpuprobe.init_fill(0x10000, 1024, 0, 1)
    h = ctypes.c_void_p()
    assert gpuprobe.lib().gpu_probe_pipeline_create(ctypes.byref(h), 0, 0) > 0


def test_product_sources_never_touch_the_oracle():
    # the shipped probe must not include, link or load anything under oracle/
    pkg = os.path.join(ROOT, "kind-gpu-sim_b200")
    offenders = []
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".cu", ".cuh", ".c", ".h", ".py", ".sh")) or f == "Makefile":
                text = open(os.path.join(dirpath, f), errors="replace").read()
                if re.search(r"oracle/|probe_ref|libprobe_ref|import oracle|from oracle", text):
                    offenders.append(os.path.join(dirpath, f))
    assert not offenders, offenders
    deps = subprocess.run(["ldd", gpuprobe.LIB_PATH], capture_output=True, text=True).stdout
    assert "probe_ref" not in deps
