#!/usr/bin/env python
"""paced_mode_probe.py -- does staggering the CTAs' refill phases, or a soft start, help the paced triad?
For each mode (0: plain ladder, 1: ladder, then ack-clocked fallback -- the shipped default): calibrate at 2^29, then time the AUTO triad
(device-side duration from gpu_probe_triad_pace_report and CUDA events) back to back, after a checksum, and
right after a writer kernel that leaves L2 full of dirty lines; the unpaced ring in the same three settings."""
import json
import os
import sys

os.environ["GPU_PROBE_TUNE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kind-gpu-sim_b200"))
import torch  # noqa: E402

import gpuprobe  # noqa: E402

st = torch.cuda.current_stream().cuda_stream
n = 1 << 29
b = torch.empty(n, dtype=torch.float64, device="cuda")
c = torch.empty_like(b)
a = torch.empty_like(b)
ws = torch.zeros(gpuprobe.workspace_bytes() // 8, dtype=torch.int64, device="cuda")
out = torch.zeros(2, dtype=torch.int64, device="cuda")
gpuprobe.init_fill(b.data_ptr(), n, 0, gpuprobe.SEED_B, st)
gpuprobe.init_fill(c.data_ptr(), n, 0, gpuprobe.SEED_C, st)
torch.cuda.synchronize()


def series(variant, setting, reps=8):
    ms = []
    for _ in range(reps):
        if setting == "after_checksum":
            gpuprobe.checksum_u64(a.data_ptr(), n, 0, out.data_ptr(), ws.data_ptr(), gpuprobe.VARIANT_AUTO, st)
        elif setting == "after_fill":
            a.fill_(-1.0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        gpuprobe.triad(a.data_ptr(), b.data_ptr(), c.data_ptr(), 3.0, n, variant, st)
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
        if variant == gpuprobe.VARIANT_AUTO and gpuprobe.triad_pace_ns() <= 0:
            gpuprobe.triad_calibrate(a.data_ptr(), b.data_ptr(), c.data_ptr(), n, st)
    ms.sort()
    return round(24 * n / ms[len(ms) // 2] / 1e6, 0), round(24 * n / ms[0] / 1e6, 0)


for setting in ("back_to_back", "after_checksum", "after_fill"):
    print(json.dumps({"kernel": "ring", "setting": setting, "GBps_median_best": series(gpuprobe.VARIANT_TMA, setting)}), flush=True)
for mode in (1, 0, 1):
    os.environ["GPU_PROBE_TUNE_PACED_MODE"] = str(mode)
    verdicts = [gpuprobe.triad_calibrate(a.data_ptr(), b.data_ptr(), c.data_ptr(), n, st) for _ in range(3)]
    row = {"kernel": "paced", "mode": mode, "verdicts": verdicts}
    for setting in ("back_to_back", "after_checksum", "after_fill"):
        gpuprobe.triad_calibrate(a.data_ptr(), b.data_ptr(), c.data_ptr(), n, st)
        row[setting] = series(gpuprobe.VARIANT_AUTO, setting)
    print(json.dumps(row), flush=True)
