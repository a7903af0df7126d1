#!/usr/bin/env python
"""ablate_peer.py -- does enabling peer access / mapping peer segments / multicast slow LOCAL HBM
streaming?  (dev tool; one process, 2 GPUs)"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kind-gpu-sim_b200"))
import torch  # noqa: E402

import gpuprobe  # noqa: E402

N = 1 << 29
torch.cuda.set_device(0)
st = torch.cuda.current_stream().cuda_stream
b = torch.empty(N, dtype=torch.float64, device="cuda:0")
c = torch.empty(N, dtype=torch.float64, device="cuda:0")
a = torch.empty(N, dtype=torch.float64, device="cuda:0")
gpuprobe.init_fill(b.data_ptr(), N, 0, gpuprobe.SEED_B, st)
gpuprobe.init_fill(c.data_ptr(), N, 0, gpuprobe.SEED_C, st)
ws = torch.zeros(gpuprobe.workspace_bytes() // 8, dtype=torch.int64, device="cuda:0")
out = torch.zeros(2, dtype=torch.int64, device="cuda:0")
torch.cuda.synchronize()


def measure(tag):
    res = {}
    for name, fn in (("triad_ring", lambda: gpuprobe.triad(a.data_ptr(), b.data_ptr(), c.data_ptr(), 3.0, N, 1, st)),
                     ("triad_paced_1040", lambda: gpuprobe.triad(a.data_ptr(), b.data_ptr(), c.data_ptr(), 3.0, N, 3, st)),
                     ("checksum", lambda: gpuprobe.checksum_u64(a.data_ptr(), N, 0, out.data_ptr(), ws.data_ptr(), 1, st)),
                     ("init_fill", lambda: gpuprobe.init_fill(a.data_ptr(), N, 0, 1, st))):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(15):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts.sort()
        by = {"triad_ring": 24, "triad_paced_1040": 24, "checksum": 8, "init_fill": 8}[name] * N
        res[name] = round(by / ts[len(ts) // 2] / 1e6, 1)
    print(tag, res, flush=True)


measure("baseline")
if torch.cuda.device_count() > 1:
    x = torch.empty(1 << 20, device="cuda:1")
    y = torch.empty(1 << 20, device="cuda:0")
    y.copy_(x)  # torch enables peer access 0 <-> 1 (cudaDeviceEnablePeerAccess)
    torch.cuda.synchronize()
    measure("after cudaDeviceEnablePeerAccess")
    xs = []
    for r in range(2):
        with torch.cuda.device(r):
            xs.append(gpuprobe.Exchange(r, 2, 64 << 20))
    torch.cuda.set_device(0)
    measure("after Exchange.create x2 (VMM allocations, not connected)")
    gpuprobe.Exchange.connect_local(xs)
    torch.cuda.set_device(0)
    print("multicast:", xs[0].has_multicast)
    measure("after connect_local (peer access on the segments + multicast)")
    for x_ in xs:
        x_.close()
    torch.cuda.set_device(0)
    measure("after closing the exchanges")
