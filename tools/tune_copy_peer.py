#!/usr/bin/env python
"""tune_copy_peer.py -- K3 across NVLink: CTAs x asked rate of the TMA ring, uni- and bidirectional,
beside the copy engine.  One process; GPU_PROBE_TUNE=1 makes the library re-read its two peer knobs
on every call.  `gpurun --gpus 2 -- python tools/tune_copy_peer.py`"""
import os
import sys

os.environ["GPU_PROBE_TUNE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kind-gpu-sim_b200"))
import torch  # noqa: E402

import gpuprobe  # noqa: E402

n = 1 << 30
bufs, sts = {}, {}
for d in (0, 1):
    torch.cuda.set_device(d)
    src = torch.empty(n, dtype=torch.uint8, device="cuda")
    src.random_(0, 255)
    bufs[d] = (src, torch.empty(n, dtype=torch.uint8, device="cuda"))
    sts[d] = torch.cuda.Stream()
bufs[1][1][:16].copy_(bufs[0][0][:16])  # torch enables peer access both ways
bufs[0][1][:16].copy_(bufs[1][0][:16])
for d in (0, 1):
    torch.cuda.synchronize(d)


def run(kind, bidir, reps=6):
    """GPU0 pushes its src into GPU1's dst (and, bidir, GPU1 into GPU0's at the same time); ms per copy, max over the two."""
    dirs = [(0, 1)] + ([(1, 0)] if bidir else [])
    ev = {}
    for i in range(-2, reps):
        for me, other in dirs:
            torch.cuda.set_device(me)
            if i == 0:
                ev[me] = torch.cuda.Event(enable_timing=True)
                ev[me].record(sts[me])
            if kind == "sm":
                gpuprobe.p2p_copy(bufs[other][1].data_ptr(), bufs[me][0].data_ptr(), n, sts[me].cuda_stream)
            else:
                with torch.cuda.stream(sts[me]):
                    bufs[other][1].copy_(bufs[me][0], non_blocking=True)
    ms = 0.0
    for me, _ in dirs:
        torch.cuda.set_device(me)
        e = torch.cuda.Event(enable_timing=True)
        e.record(sts[me])
        sts[me].synchronize()
        ms = max(ms, ev[me].elapsed_time(e) / reps)
    return ms


for bidir in (False, True):
    ms = run("ce", bidir)
    print(f"copy_engine bidir={int(bidir)} GBps_per_dir={n / ms / 1e6:.1f}", flush=True)
for ctas in (0, 16, 32, 48, 64, 96, 148, 296):
    for gbs in (0, 720, 760, 800, 840, 900):
        os.environ["GPU_PROBE_COPY_PEER_CTAS"] = str(ctas)
        os.environ["GPU_PROBE_COPY_PEER_GBS"] = str(gbs)
        uni, bi = run("sm", False), run("sm", True)
        print(f"tma_ring ctas={ctas} asked_GBps={gbs} uni_GBps={n / uni / 1e6:.1f} bidir_GBps_per_dir={n / bi / 1e6:.1f}", flush=True)
ok = bool(torch.equal(bufs[1][1].cpu(), bufs[0][0].cpu()))
print("last copy verified:", ok)
