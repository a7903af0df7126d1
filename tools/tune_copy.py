#!/usr/bin/env python
"""tune_copy.py -- K3 geometry sweep: LDG/STG copy vs TMA ring copy, local and to a peer. (dev tool)
Each configuration runs in a fresh process because the knob is read once per process."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = r'''
import os, sys
sys.path.insert(0, os.path.join(%r, "kind-gpu-sim_b200"))
import torch, gpuprobe
n = 1 << 30
src = torch.empty(n, dtype=torch.uint8, device="cuda:0"); src.random_(0, 255)
st = torch.cuda.current_stream().cuda_stream
targets = [("local", torch.empty(n, dtype=torch.uint8, device="cuda:0"))]
if torch.cuda.device_count() > 1:
    dst1 = torch.empty(n, dtype=torch.uint8, device="cuda:1")
    dst1[:16].copy_(src[:16])  # makes torch enable peer access 0 <-> 1
    torch.cuda.synchronize()
    targets.append(("peer", dst1))
for name, dst in targets:
    for _ in range(3): gpuprobe.p2p_copy(dst.data_ptr(), src.data_ptr(), n, st)
    torch.cuda.synchronize()
    ts = []
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); gpuprobe.p2p_copy(dst.data_ptr(), src.data_ptr(), n, st); e1.record()
        torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    ts.sort()
    ok = bool(torch.equal(dst.cpu()[:1 << 20], src.cpu()[:1 << 20]) and torch.equal(dst[-4096:].cpu(), src[-4096:].cpu()))
    mult = 2 if name == "local" else 1
    print(f"{os.environ.get('GPU_PROBE_TUNE_COPY', 'ldg')} {name} best_ms={ts[0]:.4f} med_ms={ts[5]:.4f} GBps={mult * n / ts[5] / 1e6:.1f} ok={ok}", flush=True)
''' % ROOT

configs = [None] + [f"{k}:{s}:{c}" for (k, s) in ((4, 4), (8, 4), (16, 4), (32, 4), (8, 8), (16, 2), (32, 2), (64, 2), (32, 3))
                    for c in (1, 2, 4, 8) if 32 <= k * s * c <= 200]
for cfg in configs:
    env = dict(os.environ)
    if cfg:
        env["GPU_PROBE_TUNE_COPY"] = cfg
    r = subprocess.run([sys.executable, "-c", WORKER], env=env, capture_output=True, text=True)
    sys.stdout.write(r.stdout)
    if r.returncode:
        sys.stdout.write(f"{cfg} FAILED: {r.stderr[-300:]}\n")
    sys.stdout.flush()
