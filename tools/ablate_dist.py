#!/usr/bin/env python
"""ablate_dist.py -- what in a multi-rank run slows LOCAL HBM streaming by ~4 %?  (dev tool)
torchrun --nproc-per-node 2 tools/ablate_dist.py [gloo|nccl]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kind-gpu-sim_b200"))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import gpuprobe  # noqa: E402
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import pynvml  # noqa: E402
pynvml.nvmlInit()

backend = sys.argv[1] if len(sys.argv) > 1 else "nccl"
rank, local = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
world = int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
N = 1 << 29
st = torch.cuda.current_stream().cuda_stream
b = torch.empty(N, dtype=torch.float64, device="cuda")
c = torch.empty(N, dtype=torch.float64, device="cuda")
a = torch.empty(N, dtype=torch.float64, device="cuda")
gpuprobe.init_fill(b.data_ptr(), N, 0, gpuprobe.SEED_B, st)
gpuprobe.init_fill(c.data_ptr(), N, 0, gpuprobe.SEED_C, st)
torch.cuda.synchronize()


def measure(tag, sync_ranks=False):
    for _ in range(3):
        gpuprobe.triad(a.data_ptr(), b.data_ptr(), c.data_ptr(), 3.0, N, 1, st)
    torch.cuda.synchronize()
    if sync_ranks:
        dist.barrier()
    ts = []
    h = pynvml.nvmlDeviceGetHandleByIndex(local)
    smp, pw, mem = [], [], []
    evs = []
    for _ in range(60):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); gpuprobe.triad(a.data_ptr(), b.data_ptr(), c.data_ptr(), 3.0, N, 1, st); e1.record()
        evs.append((e0, e1))
        if len(evs) % 6 == 0:  # queue stays deep; sample while the GPU is busy
            smp.append(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
            mem.append(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_MEM))
            pw.append(pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0)
    torch.cuda.synchronize()
    ts = sorted(e0.elapsed_time(e1) for e0, e1 in evs)
    print(f"rank {rank} [{backend}] {tag}: triad ring med {ts[30]:.4f} ms = {24 * N / ts[30] / 1e6:.0f} GB/s  "
          f"sm_mhz {sorted(smp)[len(smp) // 2]} mem_mhz {sorted(mem)[len(mem) // 2]} power_w {sorted(pw)[len(pw) // 2]:.0f}", flush=True)


measure("before init_process_group (ranks not synchronised: may or may not overlap)")
if world > 1:
    if backend == "nccl":
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        dist.init_process_group("gloo")
    measure("after init_process_group, ranks running at the same time", sync_ranks=True)
    if backend == "nccl":
        t = torch.ones(4, device="cuda")
        dist.all_reduce(t)
        torch.cuda.synchronize()
        measure("after the first NCCL allreduce", sync_ranks=True)
    if backend == "gloo":  # our own cross-process segments (fd passing, peer mappings, multicast) without NCCL
        x = gpuprobe.Exchange(rank, world, 64 << 20)
        hs = [None] * world
        dist.all_gather_object(hs, x.export_handle())
        x.connect_ipc(b"".join(hs))
        dist.barrier()
        measure(f"after Exchange connect_ipc (multicast={x.has_multicast})", sync_ranks=True)
        x.allreduce_f32(0, 1 << 20, 0, st)
        torch.cuda.synchronize()
        measure("after one K4 allreduce", sync_ranks=True)
    # one rank at a time
    for r in range(world):
        dist.barrier()
        if r == rank:
            measure(f"alone (other ranks idle at a barrier), turn of rank {r}")
        dist.barrier()
    dist.destroy_process_group()
