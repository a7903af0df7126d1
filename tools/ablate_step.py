#!/usr/bin/env python
"""ablate_step.py -- why is a kernel slower inside the bench step than alone?  (dev tool)

Times triad and checksum with CUDA events under different launch patterns and array
placements, all through the C ABI:  T only / C only / T,C alternating / T,T,C,C / with an
idle gap / arrays from torch vs one slab with padding."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "kind-gpu-sim_b200"))

import torch  # noqa: E402

import gpuprobe  # noqa: E402

N = 1 << 29
st = torch.cuda.current_stream().cuda_stream
ws = torch.zeros(gpuprobe.workspace_bytes() // 8, dtype=torch.int64, device="cuda")
lanes = torch.zeros(2, dtype=torch.int64, device="cuda")


def run(pattern, a, b, c, reps=30, variant=-1):
    def T():
        gpuprobe.triad(a, b, c, 3.0, N, variant, st)

    def C():
        gpuprobe.checksum_u64(a, N, 0, lanes.data_ptr(), ws.data_ptr(), variant, st)

    def F():
        gpuprobe.triad_checksum(a, b, c, 3.0, N, 0, lanes.data_ptr(), ws.data_ptr(), variant, st)

    ops = {"T": T, "C": C, "F": F}
    for _ in range(3):
        for ch in pattern:
            ops[ch]()
    torch.cuda.synchronize()
    evs = []
    for _ in range(reps):
        for ch in pattern:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            ops[ch]()
            e1.record()
            evs.append((ch, e0, e1))
    torch.cuda.synchronize()
    out = {}
    for ch in set(pattern):
        ts = sorted(e0.elapsed_time(e1) for k, e0, e1 in evs if k == ch)
        by = {"T": 24, "C": 8, "F": 24}[ch] * N
        out[ch] = {"med_ms": round(ts[len(ts) // 2], 4), "best_ms": round(ts[0], 4),
                   "GBps_med": round(by / ts[len(ts) // 2] / 1e6, 1)}
    t0, t1 = evs[0][1], evs[-1][2]
    out["total_ms_per_pattern"] = round(t0.elapsed_time(t1) / reps, 4)
    return out


def placements():
    # torch allocator, bench order (b, c, a)
    b = torch.empty(N, dtype=torch.float64, device="cuda")
    c = torch.empty(N, dtype=torch.float64, device="cuda")
    a = torch.empty(N, dtype=torch.float64, device="cuda")
    yield "torch b,c,a", a.data_ptr(), b.data_ptr(), c.data_ptr(), (a, b, c)
    del a, b, c
    torch.cuda.empty_cache()
    a = torch.empty(N, dtype=torch.float64, device="cuda")
    b = torch.empty(N, dtype=torch.float64, device="cuda")
    c = torch.empty(N, dtype=torch.float64, device="cuda")
    yield "torch a,b,c", a.data_ptr(), b.data_ptr(), c.data_ptr(), (a, b, c)
    del a, b, c
    torch.cuda.empty_cache()
    for pad in (0, 256, 4096, 65536, 1 << 20, (2 << 20) + 4096, (1 << 20) * 33 + 8192):
        slab = torch.empty(3 * N * 8 + 4 * pad + 4096, dtype=torch.uint8, device="cuda")
        base = (slab.data_ptr() + 255) & ~255
        pa, pb, pc = base, base + 8 * N + pad, base + 2 * (8 * N + pad)
        yield f"slab a,b,c pad={pad}", pa, pb, pc, (slab,)
        del slab
        torch.cuda.empty_cache()


def orders():
    """One slab; every order of (a, b, c) in address space x a few paddings; plus separate
    torch allocations in every allocation order."""
    import itertools
    for pad in (0, 4096 + 256, (1 << 20) + 8192 + 256, (2 << 20) * 7 + 4096):
        for perm in itertools.permutations("abc"):
            slab = torch.empty(3 * N * 8 + 4 * pad + 4096, dtype=torch.uint8, device="cuda")
            base = (slab.data_ptr() + 255) & ~255
            addr = {name: base + k * (8 * N + pad) for k, name in enumerate(perm)}
            yield f"slab order={''.join(perm)} pad={pad}", addr["a"], addr["b"], addr["c"], (slab,)
            del slab
            torch.cuda.empty_cache()
    for perm in itertools.permutations("abc"):
        t = {}
        for name in perm:
            t[name] = torch.empty(N, dtype=torch.float64, device="cuda")
        yield f"torch alloc order={''.join(perm)}", t["a"].data_ptr(), t["b"].data_ptr(), t["c"].data_ptr(), (t,)
        del t
        torch.cuda.empty_cache()


def main():
    print(json.dumps({"info": gpuprobe.kernel_info()}))
    if len(sys.argv) > 1 and sys.argv[1] == "orders":
        for name, a, b, c, keep in orders():
            gpuprobe.init_fill(b, N, 0, gpuprobe.SEED_B, st)
            gpuprobe.init_fill(c, N, 0, gpuprobe.SEED_C, st)
            torch.cuda.synchronize()
            r = run("T", a, b, c, reps=15)["T"]
            f = run("F", a, b, c, reps=15)["F"]
            l = run("T", a, b, c, reps=15, variant=0)["T"]
            print(json.dumps({"placement": name, "a": hex(a), "b": hex(b), "c": hex(c), "T_tma": r["GBps_med"],
                              "F_tma": f["GBps_med"], "T_ldg128": l["GBps_med"]}), flush=True)
        return
    for name, a, b, c, keep in placements():
        gpuprobe.init_fill(b, N, 0, gpuprobe.SEED_B, st)
        gpuprobe.init_fill(c, N, 0, gpuprobe.SEED_C, st)
        torch.cuda.synchronize()
        row = {"placement": name, "a": hex(a), "b": hex(b), "c": hex(c)}
        for pattern in ("T", "C", "TC", "TTCC", "F"):
            row[pattern] = run(pattern, a, b, c)
        print(json.dumps(row), flush=True)
        if name.startswith("torch b"):
            for v, vn in ((0, "ldg128"), (2, "ldg256")):
                print(json.dumps({"placement": name, "variant": vn, "TC": run("TC", a, b, c, variant=v),
                                  "T": run("T", a, b, c, variant=v)}), flush=True)


if __name__ == "__main__":
    main()
