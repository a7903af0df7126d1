#!/usr/bin/env python
"""slow_state_algo.py -- third step: one-sided NVLink traffic does not cause the slow state
(slow_state_dir.py); which ingredient of a collective does?  Escalating doses on two fresh pairs,
both ends re-measured (unpaced ring triad, best of 8) after each:
  pair (0,1): both GPUs push at once (plain stores) -> both pull at once -> K4 PULL_PUSH, 1 KiB x 20000
              (flags, hardly any data) -> K4 PULL_PUSH 1 GiB x 1 / x 10 / x 100
  pair (2,3): K4 one-shot PUSH 2 MiB x 5000 (stores + flags only) -> K4 MULTIMEM 1 GiB x 50 (switch
              reduce) -> K4 PULL_PUSH 1 GiB x 100
`gpurun --gpus 4 -- python tools/slow_state_algo.py`"""
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
bufs, streams = {}, {}
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
    streams[d] = torch.cuda.Stream()
    torch.cuda.synchronize()
for d in range(G):
    torch.cuda.set_device(d)
    for p in range(G):
        if p != d:
            assert rt.cudaDeviceEnablePeerAccess(p, 0) in (0, 704)
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


def sync(pair):
    for d in pair:
        torch.cuda.set_device(d)
        torch.cuda.synchronize()


def both(pair, kind, reps):
    for _ in range(reps):
        for me, other in (pair, pair[::-1]):
            torch.cuda.set_device(me)
            a, b, c, ws, out = bufs[me]
            pa = bufs[other][0]
            st = streams[me].cuda_stream
            if kind == "push":
                gpuprobe.init_fill(pa.data_ptr(), 1 << 27, 0, 1, st)       # 1 GiB of plain stores into the peer
            else:
                gpuprobe.checksum_u64(pa.data_ptr(), 1 << 27, 0, out.data_ptr(), ws.data_ptr(), gpuprobe.VARIANT_LDG, st)
    sync(pair)


def k4(xs, pair, count, algo, reps):
    for _ in range(reps):
        for i, d in enumerate(pair):
            torch.cuda.set_device(d)
            xs[i].allreduce_f32(0, count, ctas=0, stream=streams[d].cuda_stream, algo=algo)
    sync(pair)
    assert all(x.status() == 0 for x in xs)


def exchanges(pair):
    xs = []
    for i, d in enumerate(pair):
        torch.cuda.set_device(d)
        xs.append(gpuprobe.Exchange(i, 2, 1 << 30))
    gpuprobe.Exchange.connect_local(xs)
    return xs


report("start")
report("start")
p = (0, 1)
both(p, "push", 200); report("gpus 0,1 push into each other at once, plain stores, 200 x 1 GiB each way")
for k in range(3):
    both(p, "pull", 200); report(f"gpus 0,1 pull from each other at once, plain loads, 200 x 1 GiB each way (round {k + 1})")
if G >= 4:
    p = (2, 3)
    ys = exchanges(p); report("gpus 2,3: segments created + connected (multicast=%d)" % ys[0].has_multicast)
    k4(ys, p, 256, gpuprobe.ALLREDUCE_PULL_PUSH, 20000); report("gpus 2,3: K4 PULL_PUSH 1 KiB x 20000 (flags, hardly any data)")
    k4(ys, p, 1 << 19, gpuprobe.ALLREDUCE_PUSH, 5000); report("gpus 2,3: K4 one-shot PUSH 2 MiB x 5000 (stores + flags)")
    if ys[0].has_multicast:
        k4(ys, p, 1 << 28, gpuprobe.ALLREDUCE_MULTIMEM, 50); report("gpus 2,3: K4 MULTIMEM 1 GiB x 50")
    for reps in (1, 10, 100):
        k4(ys, p, 1 << 28, gpuprobe.ALLREDUCE_PULL_PUSH, reps); report(f"gpus 2,3: K4 PULL_PUSH 1 GiB x {reps}")
