#!/usr/bin/env python
"""slow_state_probe.py -- what leaves a B200 ~5 % short of HBM bandwidth for tens of seconds?

Round 1's scaling run showed GPUs that another multi-GPU process had used 30-45 s earlier
calibrating to the slow end of the pace ladder (1082 ns) and running the triad in 1.92 ms instead
of 1.81-1.85.  This tool keeps one measuring process on GPU 0 (buffers resident, ~100 ms of work
per sample), runs each suspect in a child process, and samples GPU 0 afterwards at
t = 0.5, 5, 15, 30 (60, 90 while still slow) seconds, logging NVML temperatures, clocks and power
beside every sample.  `gpurun --gpus 2 -- python tools/slow_state_probe.py > profile.jsonl`."""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kind-gpu-sim_b200"))
import pynvml  # noqa: E402
import torch  # noqa: E402

import gpuprobe  # noqa: E402

pynvml.nvmlInit()
H = pynvml.nvmlDeviceGetHandleByIndex(0)
torch.cuda.set_device(0)
N = 1 << 29
st = torch.cuda.current_stream().cuda_stream
b = torch.empty(N, dtype=torch.float64, device="cuda")
c = torch.empty_like(b)
a = torch.empty_like(b)
gpuprobe.init_fill(b.data_ptr(), N, 0, gpuprobe.SEED_B, st)
gpuprobe.init_fill(c.data_ptr(), N, 0, gpuprobe.SEED_C, st)
torch.cuda.synchronize()


def nvml():
    out = {}
    try:
        out["temp_gpu"] = pynvml.nvmlDeviceGetTemperature(H, pynvml.NVML_TEMPERATURE_GPU)
    except Exception as e:
        out["temp_gpu"] = str(e)
    try:  # HBM temperature: field value 82 (NVML_FI_DEV_MEMORY_TEMP)
        out["temp_mem"] = pynvml.nvmlDeviceGetFieldValues(H, [pynvml.NVML_FI_DEV_MEMORY_TEMP])[0].value.uiVal
    except Exception as e:
        out["temp_mem"] = str(e)
    for k, fn in (("sm_mhz", lambda: pynvml.nvmlDeviceGetClockInfo(H, pynvml.NVML_CLOCK_SM)),
                  ("mem_mhz", lambda: pynvml.nvmlDeviceGetClockInfo(H, pynvml.NVML_CLOCK_MEM)),
                  ("power_w", lambda: pynvml.nvmlDeviceGetPowerUsage(H) / 1e3),
                  ("reasons", lambda: hex(pynvml.nvmlDeviceGetCurrentClocksEventReasons(H)))):
        try:
            out[k] = fn()
        except Exception as e:
            out[k] = str(e)
    return out


def timed(variant, reps=8):
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    gpuprobe.triad(a.data_ptr(), b.data_ptr(), c.data_ptr(), 3.0, N, variant, st)
    for e0, e1 in ev:
        e0.record()
        gpuprobe.triad(a.data_ptr(), b.data_ptr(), c.data_ptr(), 3.0, N, variant, st)
        e1.record()
    torch.cuda.synchronize()
    ms = sorted(e0.elapsed_time(e1) for e0, e1 in ev)
    return round(ms[0], 4), round(ms[len(ms) // 2], 4)


def sample(tag, t_after):
    before = nvml()
    ring = timed(gpuprobe.VARIANT_TMA)
    verdict = gpuprobe.triad_calibrate(a.data_ptr(), b.data_ptr(), c.data_ptr(), N, st)
    paced = timed(gpuprobe.VARIANT_AUTO) if verdict > 0 else (None, None)
    row = {"after": tag, "t_s": round(t_after, 1), "ring_best_ms": ring[0], "ring_med_ms": ring[1], "verdict_ns": verdict,
           "paced_best_ms": paced[0], "paced_med_ms": paced[1],
           "ring_GBps": round(24 * N / ring[0] / 1e6, 0), "paced_GBps": round(24 * N / paced[0] / 1e6, 0) if paced[0] else None,
           "nvml": before}
    print(json.dumps(row), flush=True)
    return row


def run_suspect(name, argv, timeout=300):
    t0 = time.time()
    r = subprocess.run(argv, capture_output=True, text=True, timeout=timeout)
    print(json.dumps({"suspect": name, "rc": r.returncode, "ran_s": round(time.time() - t0, 1),
                      "tail": (r.stdout[-200:] + r.stderr[-300:]) if r.returncode else r.stdout[-120:]}), flush=True)
    return time.time()


py = sys.executable
child = os.path.join(ROOT, "tools", "slow_state_child.py")
torchrun = [py, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr=127.0.0.1"]
suspects = [
    ("idle_10s", [py, "-c", "import time; time.sleep(10)"]),
    ("heat_gpu1_15s", [py, child, "heat1", "15"]),
    ("heat_gpu0_15s", [py, child, "heat0", "15"]),
    ("connect_only", [py, child, "connect"]),
    ("memcpy_peer_10s", [py, child, "memcpy_peer", "10"]),
    ("k4_allreduce_10s", [py, child, "k4", "10"]),
    ("nccl_allreduce_10s", torchrun + ["--master-port=29561", child, "nccl", "10"]),
    ("bench_n2", torchrun + ["--master-port=29562", os.path.join(ROOT, "bench.py"), "--gpus", "2", "--steps", "20", "--warmup", "5"]),
]
only = [s for s in sys.argv[1:] if not s.startswith("-")]
base = sample("start", 0.0)
for _ in range(2):
    time.sleep(2)
    base = sample("start", 0.0)
for name, argv in suspects:
    if only and name not in only:
        continue
    t_end = run_suspect(name, argv)
    slow = False
    for t in (0.5, 5, 15, 30, 60, 90):
        if t > 30 and not slow:
            break
        while time.time() - t_end < t:
            time.sleep(0.05)
        row = sample(name, time.time() - t_end)
        slow = row["ring_best_ms"] > 1.02 * base["ring_best_ms"]
