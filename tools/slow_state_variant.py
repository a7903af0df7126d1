#!/usr/bin/env python
"""One two-process variant for tools/slow_state_bisect3.sh (torchrun, one rank per GPU).
usage: slow_state_variant.py <backend: gloo|nccl> <ingredients, comma separated>
  ingredients: kernels (probe init_fill/triad/checksum at 2^28), torchadd (a torch HBM kernel instead),
               allreduce (NCCL fp32 allreduce 64 MiB x 100 -- needs backend nccl or adds an NCCL subgroup),
               calibrate, nodestroy (exit without destroy_process_group), nvml (sample clocks with NVML)"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kind-gpu-sim_b200"))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

backend, ing = sys.argv[1], set(sys.argv[2].split(",")) if len(sys.argv) > 2 else set()
local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
if backend == "nccl":
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
else:
    dist.init_process_group("gloo")
n = 1 << 28
if "nvml" in ing:
    import pynvml
    pynvml.nvmlInit()
    h = pynvml.nvmlDeviceGetHandleByIndex(local)
    pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
if "kernels" in ing or "calibrate" in ing:
    import gpuprobe
    st = torch.cuda.current_stream().cuda_stream
    b = torch.empty(n, dtype=torch.float64, device="cuda")
    c = torch.empty_like(b)
    a = torch.empty_like(b)
    ws = torch.zeros(gpuprobe.workspace_bytes() // 8, dtype=torch.int64, device="cuda")
    out = torch.zeros(2, dtype=torch.int64, device="cuda")
    gpuprobe.init_fill(b.data_ptr(), n, 0, gpuprobe.SEED_B, st)
    gpuprobe.init_fill(c.data_ptr(), n, 0, gpuprobe.SEED_C, st)
    if "calibrate" in ing:
        gpuprobe.triad_calibrate(a.data_ptr(), b.data_ptr(), c.data_ptr(), n, st)
    for _ in range(20):
        gpuprobe.triad(a.data_ptr(), b.data_ptr(), c.data_ptr(), 3.0, n, gpuprobe.VARIANT_AUTO, st)
        gpuprobe.checksum_u64(a.data_ptr(), n, 0, out.data_ptr(), ws.data_ptr(), gpuprobe.VARIANT_AUTO, st)
    torch.cuda.synchronize()
if "torchadd" in ing:
    x = torch.ones(n, dtype=torch.float64, device="cuda")
    y = torch.ones_like(x)
    for _ in range(20):
        torch.add(x, y, out=x)
    torch.cuda.synchronize()
if "allreduce" in ing:
    grp = None if backend == "nccl" else dist.new_group(backend="nccl", device_id=torch.device("cuda", local))
    buf = torch.zeros(1 << 24, dtype=torch.float32, device="cuda")
    for _ in range(100):
        dist.all_reduce(buf, group=grp)
    torch.cuda.synchronize()
dist.barrier()
if "nodestroy" not in ing:
    dist.destroy_process_group()
