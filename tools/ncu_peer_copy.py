#!/usr/bin/env python
"""Four 1 GiB peer copies GPU0 -> GPU1 through K3's TMA ring, for an ncu capture of the NVLink
counters (user bytes vs protocol bytes: how much of the 900 GB/s a stream of SM-issued stores can
carry).  One process, both GPUs visible; ncu profiles the kernel on GPU 0.
  ncu --metrics nvltx__bytes.sum,nvltx__bytes_data_user.sum,nvltx__bytes_data_protocol.sum,\\
nvlrx__bytes.sum,nvlrx__bytes_data_user.sum,nvlrx__bytes_data_protocol.sum,gpu__time_duration.sum \\
      -k regex:copy_tma --clock-control none python tools/ncu_peer_copy.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kind-gpu-sim_b200"))
import torch  # noqa: E402

import gpuprobe  # noqa: E402

n = 1 << 30
torch.cuda.set_device(0)
src = torch.empty(n, dtype=torch.uint8, device="cuda:0")
src.random_(0, 255)
dst = torch.empty(n, dtype=torch.uint8, device="cuda:1")
dst[:16].copy_(src[:16])  # torch enables peer access 0 -> 1
torch.cuda.synchronize()
st = torch.cuda.current_stream().cuda_stream
for _ in range(2):
    gpuprobe.p2p_copy(dst.data_ptr(), src.data_ptr(), n, st)   # push: local loads, stores cross NVLink
torch.cuda.synchronize()
for _ in range(2):
    gpuprobe.p2p_copy(src.data_ptr(), dst.data_ptr(), n, st)   # pull: loads cross NVLink, local stores
torch.cuda.synchronize()
print("done")
