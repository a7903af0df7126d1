#!/usr/bin/env python
"""How far do paced AUTO launches sit from what their pace allows?  (drift-guard threshold)

For n = 2^26..2^29: calibrate, then time launches (device-side duration from
gpu_probe_triad_pace_report) in three settings: back to back, after a checksum (the bench's step),
and right after a kernel that leaves L2 full of dirty lines of a[] (torch fill_)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kind-gpu-sim_b200"))
import torch  # noqa: E402

import gpuprobe  # noqa: E402

st = torch.cuda.current_stream().cuda_stream
N = 1 << 29
b = torch.empty(N, dtype=torch.float64, device="cuda")
c = torch.empty_like(b)
a = torch.empty_like(b)
ws = torch.zeros(gpuprobe.workspace_bytes() // 8, dtype=torch.int64, device="cuda")
out = torch.zeros(2, dtype=torch.int64, device="cuda")
gpuprobe.init_fill(b.data_ptr(), N, 0, gpuprobe.SEED_B, st)
gpuprobe.init_fill(c.data_ptr(), N, 0, gpuprobe.SEED_C, st)
torch.cuda.synchronize()
rows = []
for log2n in (26, 27, 28, 29):
    n = 1 << log2n
    verdicts = [gpuprobe.triad_calibrate(a.data_ptr(), b.data_ptr(), c.data_ptr(), n, st) for _ in range(3)]
    pace = verdicts[-1]
    if pace <= 0:
        rows.append({"log2n": log2n, "verdicts": verdicts})
        continue
    tiles_per_cta = -(-(n // 1024) // 296)
    ideal = tiles_per_cta * pace
    for setting in ("back_to_back", "after_checksum", "after_fill"):
        ratios = []
        for k in range(8):
            if setting == "after_checksum":
                gpuprobe.checksum_u64(a.data_ptr(), n, 0, out.data_ptr(), ws.data_ptr(), gpuprobe.VARIANT_AUTO, st)
            elif setting == "after_fill":
                a[:n].fill_(-1.0)
            gpuprobe.triad(a.data_ptr(), b.data_ptr(), c.data_ptr(), gpuprobe.SCALAR, n, gpuprobe.VARIANT_AUTO, st)
            torch.cuda.synchronize()
            ns, over = gpuprobe.triad_pace_report()
            ratios.append(round(ns / ideal, 4))
            if gpuprobe.triad_pace_ns() <= 0:  # guard fired: calibrate again so the series continues
                gpuprobe.triad_calibrate(a.data_ptr(), b.data_ptr(), c.data_ptr(), n, st)
        rows.append({"log2n": log2n, "verdicts": verdicts, "setting": setting, "ideal_us": round(ideal / 1e3, 1),
                     "ratio_to_ideal": ratios})
for r in rows:
    print(json.dumps(r))
