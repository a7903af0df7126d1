#!/usr/bin/env python
"""tune_allreduce.py -- grid-size sweep of K4 (gpu_probe_allreduce_f32) in one process. (dev tool)

    python tools/tune_allreduce.py [world]      # env GPU_PROBE_TUNE_MM_U = 1|2|4|8 picks the NVLS unroll
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kind-gpu-sim_b200"))

import torch  # noqa: E402

import gpuprobe  # noqa: E402

world = int(sys.argv[1]) if len(sys.argv) > 1 else torch.cuda.device_count()
MAXB = 1 << 30
xs, sts = [], []
for r in range(world):
    with torch.cuda.device(r):
        xs.append(gpuprobe.Exchange(r, world, MAXB))
        sts.append(torch.cuda.Stream())
gpuprobe.Exchange.connect_local(xs)
print(f"world={world} multicast={xs[0].has_multicast} U={os.environ.get('GPU_PROBE_TUNE_MM_U', 'default')}", flush=True)
for r in range(world):
    with torch.cuda.device(r):
        torch.cuda.synchronize()


def timed(size, algo, ctas, iters):
    ev = []
    for i in range(-3, iters):
        for r in range(world):
            with torch.cuda.device(r):
                if i == 0:
                    e = torch.cuda.Event(enable_timing=True)
                    e.record(sts[r])
                    ev.append(e)
                xs[r].allreduce_f32(0, size // 4, ctas, sts[r].cuda_stream, algo)
    ms = 0.0
    for r in range(world):
        with torch.cuda.device(r):
            e = torch.cuda.Event(enable_timing=True)
            e.record(sts[r])
            torch.cuda.synchronize()
            ms = max(ms, ev[r].elapsed_time(e) / iters)
    return ms


names = {1: "pull_push", 2: "push", 3: "multimem"}
algos = [1, 3] if xs[0].has_multicast else [1]
for size in (1 << 30, 64 << 20, 8 << 20):
    for algo in algos:
        for ctas in (16, 24, 32, 48, 64, 96, 128, 148):
            ms = timed(size, algo, ctas, 10 if size >= (1 << 28) else 20)
            bus = size / ms / 1e6 * 2 * (world - 1) / world
            print(f"size={size} algo={names[algo]} ctas={ctas} us={ms * 1e3:.1f} busbw={bus:.1f}", flush=True)
for size in (1 << 20, 1 << 21, 1 << 18, 1 << 14):
    for algo in [1, 2] + ([3] if xs[0].has_multicast else []):
        for ctas in (0, 4, 8, 16, 32, 64):
            try:
                ms = timed(size, algo, ctas, 50)
            except gpuprobe.ProbeError as e:
                print(f"size={size} algo={names[algo]} ctas={ctas} error rc={e.rc}")
                continue
            print(f"size={size} algo={names[algo]} ctas={ctas} us={ms * 1e3:.2f}", flush=True)
for x in xs:
    x.close()
