#!/usr/bin/env python
"""Times every kernel x variant at BASELINE size with CUDA events (3 warm-up + 20 timed
launches each; inputs 4 GiB/array >> L2).  Writes gpurun_out/sweep_variants.json.
Usage (GPU box): python tools/sweep_variants.py [log2n]"""
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kind-gpu-sim_b200"))
import torch  # noqa: E402

import gpuprobe  # noqa: E402


def timeit(fn, warm=3, iters=20):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for e0, e1 in evs:
        e0.record()
        fn()
        e1.record()
    torch.cuda.synchronize()
    ms = [e0.elapsed_time(e1) for e0, e1 in evs]
    return min(ms), statistics.median(ms)


def main():
    log2n = int(sys.argv[1]) if len(sys.argv) > 1 else 29
    n = 1 << log2n
    st = torch.cuda.current_stream().cuda_stream
    b = torch.empty(n, dtype=torch.float64, device="cuda")
    c = torch.empty(n, dtype=torch.float64, device="cuda")
    a = torch.empty(n, dtype=torch.float64, device="cuda")
    ws = torch.zeros(gpuprobe.workspace_bytes() // 8, dtype=torch.int64, device="cuda")
    out = torch.zeros(2, dtype=torch.int64, device="cuda")
    gpuprobe.init_fill(b.data_ptr(), n, 0, gpuprobe.SEED_B, st)
    gpuprobe.init_fill(c.data_ptr(), n, 0, gpuprobe.SEED_C, st)
    peak = 6532.5
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    rows = []

    def rec(kernel, variant, nbytes, fn):
        best, med = timeit(fn)
        r = {"kernel": kernel, "variant": variant, "bytes": nbytes, "best_ms": round(best, 4), "median_ms": round(med, 4),
             "GBps_best": round(nbytes / best / 1e6, 1), "GBps_median": round(nbytes / med / 1e6, 1),
             "frac_measured": round(nbytes / med / 1e6 / peak, 4), "frac_nominal_8000": round(nbytes / med / 1e6 / 8000, 4)}
        rows.append(r)
        print(json.dumps(r), flush=True)

    rec("init_fill", "stg128", 8 * n, lambda: gpuprobe.init_fill(a.data_ptr(), n, 0, 1, st))
    for name, v in (("ldg128", 0), ("tma", 1), ("ldg256", 2)):
        rec("triad", name, 24 * n, lambda v=v: gpuprobe.triad(a.data_ptr(), b.data_ptr(), c.data_ptr(), 3.0, n, v, st))
    for name, v in (("ldg128", 0), ("tma", 1), ("ldg256", 2)):
        rec("checksum_u64", name, 8 * n,
            lambda v=v: gpuprobe.checksum_u64(a.data_ptr(), n, 0, out.data_ptr(), ws.data_ptr(), v, st))
    for name, v in (("ldg128", 0), ("tma", 1), ("ldg256", 2)):
        rec("triad_checksum", name, 24 * n,
            lambda v=v: gpuprobe.triad_checksum(a.data_ptr(), b.data_ptr(), c.data_ptr(), 3.0, n, 0, out.data_ptr(),
                                                ws.data_ptr(), v, st))
    rec("copy(local)", "ldg128", 16 * n, lambda: gpuprobe.p2p_copy(a.data_ptr(), b.data_ptr(), 8 * n, st))
    rec("torch.copy_(ref)", "lib", 16 * n, lambda: a.copy_(b))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump({"log2n": log2n, "peak_measured": peak, "rows": rows},
              open(os.path.join(ROOT, "gpurun_out", "sweep_variants.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
