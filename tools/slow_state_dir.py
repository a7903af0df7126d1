#!/usr/bin/env python
"""slow_state_dir.py -- WHICH NVLink traffic leaves a B200's HBM ~4 % slower?  (follow-up to slow_state_probe.py)

slow_state_probe.py showed: copy-engine peer copies and merely mapping peer memory do nothing; 10 s of
SM-driven allreduce (the probe's K4 or NCCL) costs the GPU 4 % of its HBM bandwidth for at least 90 s
after the process is gone.  Here each pair of fresh GPUs gets one kind of SM-driven traffic, in
growing doses, and BOTH ends are re-measured (unpaced ring triad, best of 8) after every dose:
  pair (0,1): GPU0 PUSHES into GPU1   -- TMA bulk stores (gpu_probe_p2p_copy), then LSU stores (init_fill)
  pair (2,3): GPU2 PULLS from GPU3    -- TMA bulk loads (checksum, TMA variant), then LSU loads (LDG variant)
`gpurun --gpus 4 -- python tools/slow_state_dir.py`"""
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kind-gpu-sim_b200"))
import torch  # noqa: E402

import gpuprobe  # noqa: E402

N = 1 << 29
G = min(torch.cuda.device_count(), 4)
rt = ctypes.CDLL("libcudart.so.12")
bufs = {}
for d in range(G):
    torch.cuda.set_device(d)
    st = torch.cuda.current_stream().cuda_stream
    b = torch.empty(N, dtype=torch.float64, device="cuda")
    c = torch.empty_like(b)
    a = torch.empty_like(b)
    gpuprobe.init_fill(b.data_ptr(), N, 0, gpuprobe.SEED_B, st)
    gpuprobe.init_fill(c.data_ptr(), N, 0, gpuprobe.SEED_C, st)
    ws = torch.zeros(gpuprobe.workspace_bytes() // 8, dtype=torch.int64, device="cuda")
    out = torch.zeros(2, dtype=torch.int64, device="cuda")
    bufs[d] = (a, b, c, ws, out)
    torch.cuda.synchronize()
for d in range(G):
    torch.cuda.set_device(d)
    for p in range(G):
        if p != d:
            rc = rt.cudaDeviceEnablePeerAccess(p, 0)
            assert rc in (0, 704), rc
    rt.cudaGetLastError()


def ring_ms(d):
    torch.cuda.set_device(d)
    a, b, c, _, _ = bufs[d]
    st = torch.cuda.current_stream().cuda_stream
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(8)]
    gpuprobe.triad(a.data_ptr(), b.data_ptr(), c.data_ptr(), 3.0, N, gpuprobe.VARIANT_TMA, st)
    for e0, e1 in ev:
        e0.record()
        gpuprobe.triad(a.data_ptr(), b.data_ptr(), c.data_ptr(), 3.0, N, gpuprobe.VARIANT_TMA, st)
        e1.record()
    torch.cuda.synchronize()
    return round(min(e0.elapsed_time(e1) for e0, e1 in ev), 4)


def report(tag):
    print(json.dumps({"after": tag, "ring_best_ms": {d: ring_ms(d) for d in range(G)}}), flush=True)


def dose(active, passive, kind, nbytes, reps):
    """`active`'s SMs move nbytes x reps over NVLink; `passive` only serves."""
    torch.cuda.set_device(active)
    st = torch.cuda.current_stream().cuda_stream
    a, b, c, ws, out = bufs[active]
    pa = bufs[passive][0]
    for _ in range(reps):
        if kind == "push_tma":      # local b -> peer a, UBLKCP both ways (>= 8 MiB), LDG/STG below
            gpuprobe.p2p_copy(pa.data_ptr(), b.data_ptr(), nbytes, st)
        elif kind == "push_lsu":    # st.global to peer memory
            gpuprobe.init_fill(pa.data_ptr(), nbytes // 8, 0, 1, st)
        elif kind == "pull_tma":    # bulk loads from peer memory
            gpuprobe.checksum_u64(pa.data_ptr(), nbytes // 8, 0, out.data_ptr(), ws.data_ptr(), gpuprobe.VARIANT_TMA, st)
        elif kind == "pull_lsu":    # ld.global from peer memory
            gpuprobe.checksum_u64(pa.data_ptr(), nbytes // 8, 0, out.data_ptr(), ws.data_ptr(), gpuprobe.VARIANT_LDG, st)
    torch.cuda.synchronize()


report("start")
report("start")
plans = [((0, 1), ["push_tma", "push_lsu"])]
if G >= 4:
    plans.append(((2, 3), ["pull_tma", "pull_lsu"]))
for (act, pas), kinds in plans:
    for kind in kinds:
        for nbytes, reps in ((1 << 20, 1), (1 << 30, 1), (1 << 30, 50), (1 << 30, 2000)):
            dose(act, pas, kind, nbytes, reps)
            report(f"gpu{act} {kind} gpu{pas}: {nbytes >> 20} MiB x {reps}")
