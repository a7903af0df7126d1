#!/usr/bin/env python
"""One suspect of tools/slow_state_probe.py, run in its own process (or two, under torchrun)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kind-gpu-sim_b200"))
import torch  # noqa: E402

import gpuprobe  # noqa: E402

what = sys.argv[1]
secs = float(sys.argv[2]) if len(sys.argv) > 2 else 10.0


def triad_loop(dev, seconds):
    torch.cuda.set_device(dev)
    n = 1 << 29
    st = torch.cuda.current_stream().cuda_stream
    b = torch.empty(n, dtype=torch.float64, device="cuda")
    c = torch.empty_like(b)
    a = torch.empty_like(b)
    gpuprobe.init_fill(b.data_ptr(), n, 0, gpuprobe.SEED_B, st)
    gpuprobe.init_fill(c.data_ptr(), n, 0, gpuprobe.SEED_C, st)
    t0 = time.time()
    while time.time() - t0 < seconds:
        for _ in range(50):
            gpuprobe.triad(a.data_ptr(), b.data_ptr(), c.data_ptr(), 3.0, n, gpuprobe.VARIANT_TMA, st)
        torch.cuda.synchronize()


def exchanges(buffer_bytes):
    xs = []
    for d in (0, 1):
        torch.cuda.set_device(d)
        xs.append(gpuprobe.Exchange(d, 2, buffer_bytes))
    gpuprobe.Exchange.connect_local(xs)
    return xs


if what in ("heat0", "heat1"):
    triad_loop(int(what[-1]), secs)
elif what == "connect":          # segments, peer access, multicast bind -- and no kernel at all
    xs = exchanges(1 << 30)
    print("multicast", xs[0].has_multicast, flush=True)
    time.sleep(1.0)
    for x in xs:
        x.close()
elif what == "k4":               # the probe's own allreduce over NVLink for `secs`
    xs = exchanges(1 << 30)
    sts = []
    for d in (0, 1):
        torch.cuda.set_device(d)
        sts.append(torch.cuda.Stream())
    t0 = time.time()
    while time.time() - t0 < secs:
        for _ in range(20):
            for d in (0, 1):
                torch.cuda.set_device(d)
                xs[d].allreduce_f32(0, (1 << 30) // 4, 0, sts[d].cuda_stream)
        for d in (0, 1):
            torch.cuda.set_device(d)
            torch.cuda.synchronize()
    for x in xs:
        x.close()
elif what == "memcpy_peer":      # copy-engine traffic GPU0 -> GPU1 for `secs`
    src = torch.empty(1 << 30, dtype=torch.uint8, device="cuda:0")
    dst = torch.empty(1 << 30, dtype=torch.uint8, device="cuda:1")
    t0 = time.time()
    while time.time() - t0 < secs:
        for _ in range(20):
            dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize(0)
        torch.cuda.synchronize(1)
elif what == "nccl":             # under torchrun, 2 ranks: communicator, 1 GiB allreduces for `secs`, destroy
    import torch.distributed as dist
    rank = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
    buf = torch.zeros(1 << 28, dtype=torch.float32, device="cuda")
    t0 = time.time()
    while time.time() - t0 < secs:
        for _ in range(20):
            dist.all_reduce(buf)
        torch.cuda.synchronize()
    dist.destroy_process_group()
else:
    raise SystemExit(f"unknown suspect {what}")
