#!/usr/bin/env python
"""sanitize_driver.py -- every single-GPU kernel of the probe at small, ragged sizes, each result
checked against the CPU oracle: the workload for compute-sanitizer (memcheck / racecheck /
synccheck / initcheck run it unchanged, about 100x slower than native).

    compute-sanitizer --tool racecheck python tools/sanitize_driver.py

Covers K0 (vector + scalar paths), K1 and K1+K2 in all five variants (LDG128, LDG256, TMA ring,
paced TMA ring incl. its PaceTrack reporting after a forced calibration verdict, scalar), K2 in
all variants + the no-allocation entry point, K3 (LDG, TMA ring, byte tail), the host-buffer
pipeline, and the single-GPU group forms of the lanes exchange and of K4 (pull_push + one-shot
push: all ranks of the protocol as rows of one launch).  racecheck tracks shared memory only and
does not follow system-scope traffic between GPUs, so the cross-GPU launches are not run here."""
import os
import sys

os.environ.setdefault("GPU_PROBE_TRIAD_FORCE_VERDICT_NS", "1040")  # calibration = bookkeeping only, no 27 timed launches

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "kind-gpu-sim_b200"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import gpuprobe  # noqa: E402
import oracle  # noqa: E402

st = torch.cuda.current_stream().cuda_stream
checks = 0


def lanes_of(t):
    v = t.cpu().numpy().view(np.uint64)
    return int(v[0]), int(v[1])


def dev(n, off=0):
    base = torch.empty(n + off + 4, dtype=torch.float64, device="cuda")
    return base[off:off + n]


def expect(got, want, what):
    global checks
    checks += 1
    if got != want:
        raise SystemExit(f"sanitize_driver: {what}: {got} != {want}")


ws = torch.zeros(gpuprobe.workspace_bytes() // 8, dtype=torch.int64, device="cuda")
out = torch.zeros(2, dtype=torch.int64, device="cuda")
# sizes: whole TMA tiles + ragged tails; (1 << 26) + 9 is there for AUTO's paced path only
for n, first, off in (((1 << 20) + 77, 1 << 34, 0), (4096 * 5 + 3, 7, 0), (3 * 1024 + 1, 0, 1), (1, 5, 0)):
    b, c, a = dev(n, off), dev(n, off), dev(n, off)
    gpuprobe.init_fill(b.data_ptr(), n, first, gpuprobe.SEED_B, st)
    gpuprobe.init_fill(c.data_ptr(), n, first, gpuprobe.SEED_C, st)
    torch.cuda.synchronize()
    hb = oracle.init_fill(n, first, oracle.SEED_B)
    expect(bool(np.array_equal(b.cpu().numpy(), hb)), True, f"init_fill n={n}")
    want = oracle.stream_lanes(n, first)
    for name, variant in gpuprobe.VARIANTS.items():
        a.zero_()
        gpuprobe.triad(a.data_ptr(), b.data_ptr(), c.data_ptr(), gpuprobe.SCALAR, n, variant, st)
        gpuprobe.checksum_u64(a.data_ptr(), n, first, out.data_ptr(), ws.data_ptr(), variant, st)
        torch.cuda.synchronize()
        expect(lanes_of(out), want, f"triad+checksum {name} n={n}")
        gpuprobe.triad_checksum(a.data_ptr(), b.data_ptr(), c.data_ptr(), gpuprobe.SCALAR, n, first, out.data_ptr(),
                                ws.data_ptr(), variant, st)
        torch.cuda.synchronize()
        expect(lanes_of(out), want, f"fused {name} n={n}")
    gpuprobe.checksum_u64_noalloc(a.data_ptr(), n, first, out.data_ptr(), st)
    gpuprobe.checksum_u64_noalloc(a.data_ptr(), n, first, out.data_ptr(), st)
    torch.cuda.synchronize()
    expect(lanes_of(out), want, f"noalloc n={n}")
    # K3: LDG path (small), byte tail, unaligned
    d = dev(n, off)
    gpuprobe.p2p_copy(d.data_ptr(), a.data_ptr(), 8 * n - 3 if n > 1 else 5, st)
    torch.cuda.synchronize()
    nb = 8 * n - 3 if n > 1 else 5
    expect(bool(np.array_equal(d.cpu().numpy().view(np.uint8)[:nb], a.cpu().numpy().view(np.uint8)[:nb])), True, f"copy n={n}")

# AUTO's paced path (n >= 2^26) with tracking: calibrate (forced verdict), launch, read the report
n = (1 << 26) + 9
b, c, a = dev(n), dev(n), dev(n)
gpuprobe.init_fill(b.data_ptr(), n, 0, gpuprobe.SEED_B, st)
gpuprobe.init_fill(c.data_ptr(), n, 0, gpuprobe.SEED_C, st)
expect(gpuprobe.triad_calibrate(a.data_ptr(), b.data_ptr(), c.data_ptr(), n, st), 1040, "forced verdict")
for _ in range(2):
    gpuprobe.triad(a.data_ptr(), b.data_ptr(), c.data_ptr(), gpuprobe.SCALAR, n, gpuprobe.VARIANT_AUTO, st)
gpuprobe.checksum_u64(a.data_ptr(), n, 0, out.data_ptr(), ws.data_ptr(), gpuprobe.VARIANT_AUTO, st)  # paced TMA checksum
torch.cuda.synchronize()
expect(lanes_of(out), oracle.stream_lanes(n, 0), "paced AUTO triad + paced checksum at 2^26+9")
expect(gpuprobe.triad_pace_report()[0] > 0, True, "pace report written by the last CTA")
# K3 TMA ring (>= 8 MiB, 16-byte aligned) with a ragged end
d = dev(n)
gpuprobe.p2p_copy(d.data_ptr(), a.data_ptr(), 8 * (1 << 21) + 40, st)
torch.cuda.synchronize()
expect(bool(torch.equal(d[: (1 << 21) + 5], a[: (1 << 21) + 5])), True, "copy_tma")
del b, c, d

# host-buffer pipeline, small chunks so several are in flight
ne = (1 << 18) + 2
hb = torch.from_numpy(oracle.init_fill(ne, 3, oracle.SEED_B)).pin_memory()
hc = torch.from_numpy(oracle.init_fill(ne, 3, oracle.SEED_C)).pin_memory()
pipe = gpuprobe.Pipeline(1 << 16, 3)
expect(pipe.run_host(hb.data_ptr(), hc.data_ptr(), None, gpuprobe.SCALAR, ne, 3), oracle.stream_lanes(ne, 3), "pipeline")
pipe.close()

# multi-GPU protocols, single-GPU group form: 3 ranks as rows of one launch
world = 3
xs = [gpuprobe.Exchange(r, world, 1 << 20) for r in range(world)]
gpuprobe.Exchange.connect_local(xs)
nr = 4096 + 3
shards = [np.random.default_rng(r).integers(0, 1 << 64, size=nr, dtype=np.uint64) for r in range(world)]
dsh = []
for s in shards:
    t = dev(nr)
    t.copy_(torch.from_numpy(s.view(np.float64)))
    dsh.append(t)
outs = [torch.zeros(2, dtype=torch.int64, device="cuda") for _ in range(world)]
wss = [torch.zeros(gpuprobe.workspace_bytes() // 8, dtype=torch.int64, device="cuda") for _ in range(world)]
gpuprobe.Exchange.checksum_u64_group(xs, [t.data_ptr() for t in dsh], [nr] * world, [r * nr for r in range(world)],
                                     [o.data_ptr() for o in outs], [w.data_ptr() for w in wss], st)
torch.cuda.synchronize()
l0 = l1 = 0
for r, s in enumerate(shards):
    x0, x1 = oracle.checksum_u64(s, r * nr)
    l0, l1 = (l0 + x0) % (1 << 64), (l1 + x1) % (1 << 64)
for o in outs:
    expect(lanes_of(o), (l0, l1), "lanes exchange, group form")
cnt = 1027
for algo, name in ((gpuprobe.ALLREDUCE_PULL_PUSH, "pull_push"), (gpuprobe.ALLREDUCE_PUSH, "push")):
    c4 = cnt if algo != gpuprobe.ALLREDUCE_PUSH else 1028
    for r, x in enumerate(xs):
        src = (torch.arange(c4, device="cuda", dtype=torch.float32) % 7 + 1) * float(r + 1)
        gpuprobe.p2p_copy(x.buffer, src.data_ptr(), 4 * c4, st)
    torch.cuda.synchronize()
    gpuprobe.Exchange.allreduce_f32_group(xs, 0, c4, 2, st, algo)
    torch.cuda.synchronize()
    want = (torch.arange(c4, device="cuda", dtype=torch.float32) % 7 + 1) * 6.0
    for x in xs:
        got = torch.empty(c4, dtype=torch.float32, device="cuda")
        gpuprobe.p2p_copy(got.data_ptr(), x.buffer, 4 * c4, st)
        torch.cuda.synchronize()
        expect(bool(torch.equal(got, want)), True, f"K4 {name}, group form")
    expect(all(x.status() == 0 for x in xs), True, "exchange status")
for x in xs:
    x.close()
print(f"sanitize_driver ok: {checks} results checked against the oracle")
