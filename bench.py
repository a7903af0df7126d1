#!/usr/bin/env python
"""bench.py -- the probe's headline measurement (BASELINE.json metric/config).

One "step" = one pass of the pod's hot path over one GPU's buffer:
    K1 triad  a = fma(3, c, b)   (24 B/element)  ->  K2 checksum_u64(a)  (8 B/element)
    [N > 1]  + the 16-byte lanes exchange (the path's only collective)
Workload = BASELINE.json configs[1]: 2^29 fp64 elements per array (4 GiB each) per GPU,
generated on the device by K0.  `value` is whole-job GB/s of algorithmic bytes
(32 B/element/step, summed over ranks) with everything resident in HBM; `e2e` is the
same metric through gpu_probe_pipeline_run_host (pinned HOST b,c in, lanes out).

    python bench.py [--gpus N] [--steps K] [--warmup W]           # this repo's CUDA path
    python bench.py --impl reference ...                          # the CPU port (oracle/)

torch is plumbing here: allocator, stream/events, torch.distributed.  Every compute
call goes through the C ABI in lib/libgpuprobe.so (no fallback: a missing library or a
failed launch raises).  oracle/ is used only as the checker and as the timed CPU arm.
"""
from __future__ import annotations

import argparse
import json
import os
import signal
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "kind-gpu-sim_b200"))

METRIC = "probe HBM GB/s/GPU + NVLink allreduce bus GB/s at 1/2/4/8 B200"
FALLBACK_HBM_GBS = 6650.0  # /opt/skills/guides/B200_PROFILING.md fallback
NVLINK_NOMINAL_GBS = 900.0
BYTES_TRIAD, BYTES_CHECKSUM = 24, 8  # algorithmic bytes per element


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


def ncu_traffic():
    """DRAM bytes per launch of the dominant kernel from the committed ncu capture."""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    try:
        with open(p) as f:
            return json.load(f)
    except Exception:
        return None


def mem_available_bytes():
    try:
        with open("/proc/meminfo") as f:
            for line in f:
                if line.startswith("MemAvailable:"):
                    return int(line.split()[1]) * 1024
    except Exception:
        pass
    return 0


class ClockSampler:
    """Samples SM clock / throttle reasons of one GPU while the timed region runs."""

    BAD = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
           "hw_power_brake_slowdown": 0x80}
    NOTED = {"sw_power_cap": 0x4}

    def __init__(self, index: int, period_s: float = 0.02, enabled: bool = True):
        self.index, self.period, self.enabled = index, period_s, enabled
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        self.err = None

    def prepare(self):
        """NVML start-up (tens of ms, takes driver locks): done BEFORE the timed region, never inside it."""
        if not self.enabled:
            self.err = "disabled (--no-clocks)"
            return self
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = self.index
            if vis:
                ids = [v for v in vis.split(",") if v != ""]
                if self.index < len(ids) and ids[self.index].isdigit():
                    idx = int(ids[self.index])
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
            pynvml.nvmlDeviceGetClockInfo(self._h, pynvml.NVML_CLOCK_SM)  # first query pays the set-up cost
            self._ready = True
        except Exception as e:  # NVML missing: report it, do not invent clocks
            self.err = f"{type(e).__name__}: {e}"
        return self

    def __enter__(self):
        if getattr(self, "_ready", False):
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()
        return self

    def _run(self):
        nv = self._nv
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)))
                try:
                    r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self._h))
                except Exception:
                    r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
                for name, bit in {**self.BAD, **self.NOTED}.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception as e:
                self.err = f"{type(e).__name__}: {e}"
                return
            self._stop.wait(self.period)

    def __exit__(self, *exc):
        self._stop.set()
        if self._thr:
            self._thr.join(timeout=2)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                    "samples": 0, "error": self.err or "no samples (timed region shorter than one period)"}
        return {"sm_mhz": int(statistics.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# --------------------------------------------------------------------------------------
# CPU arm: the oracle port timed on host cores (cpu_baseline leg and --impl reference)
# --------------------------------------------------------------------------------------
def cpu_pass_factory(n_sample: int):
    import numpy as np
    import oracle
    b = oracle.init_fill(n_sample, 0, oracle.SEED_B)
    c = oracle.init_fill(n_sample, 0, oracle.SEED_C)
    a = np.empty_like(b)

    def one_pass():
        oracle.triad(b, c, 3.0, out=a)
        return oracle.checksum_u64(a, 0)

    return one_pass, oracle.max_threads()


def host_cores() -> int:
    return len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)


def cpu_baseline(n_sample: int, repeats: int = 3, threads: int = 0):
    import oracle
    if not threads and oracle.max_threads() < host_cores():
        oracle.set_threads(host_cores())  # a launcher (torchrun) pinned OMP_NUM_THREADS=1: this leg runs on rank 0 alone
    all_threads = oracle.max_threads()
    if threads:
        oracle.set_threads(threads)
    try:
        return _cpu_baseline(n_sample, repeats)
    finally:
        if threads:
            oracle.set_threads(all_threads)


def _cpu_baseline(n_sample: int, repeats: int):
    one_pass, cores = cpu_pass_factory(n_sample)
    one_pass()  # first touch of a[]
    best = float("inf")
    for _ in range(repeats):
        t0 = time.perf_counter()
        one_pass()
        best = min(best, time.perf_counter() - t0)
    gbs = (BYTES_TRIAD + BYTES_CHECKSUM) * n_sample / best / 1e9
    return {"value": round(gbs, 3), "unit": "GB/s", "cores": cores, "kind": "port",
            "sample": f"oracle/probe_ref.c array mode, triad+checksum over 2^{n_sample.bit_length() - 1} fp64 elements "
                      f"per array (of 2^29), best of {repeats}, OpenMP x{cores} of {host_cores()} host cores"}


def run_reference(args, rank: int):
    """--impl reference: the reference has no implementation of this path (its slot is an
    echo), so the CPU arm is the oracle port, all host threads, bounded sample per step."""
    if rank != 0:
        return
    avail = mem_available_bytes()
    while args.ref_log2n > 20 and avail and 3 * (8 << args.ref_log2n) > 0.5 * avail:
        args.ref_log2n -= 1
    n_sample = 1 << args.ref_log2n
    import oracle
    # torchrun exports OMP_NUM_THREADS=1 to every rank; rank 0 is the only rank that computes here,
    # so it takes all the host's cores back (what the N=1 launch gets by default)
    oracle.set_threads(host_cores())
    one_pass, cores = cpu_pass_factory(n_sample)
    for _ in range(max(args.warmup, 1)):
        one_pass()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        one_pass()
    dt = time.perf_counter() - t0
    ms = dt / args.steps * 1e3
    gbs = (BYTES_TRIAD + BYTES_CHECKSUM) * n_sample / (ms / 1e3) / 1e9
    sample = (f"oracle/probe_ref.c array mode (CPU port; the reference's own slot is `echo`), triad+checksum over "
              f"2^{args.ref_log2n} fp64 elements per array per step "
              f"({'the whole 2^29 workload' if args.ref_log2n >= 29 else 'bounded sample of the 2^29 workload'}), OpenMP x{cores}")
    line = {
        "impl": "reference", "metric": METRIC, "value": round(gbs, 3), "unit": "GB/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms, 4), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64+u64", "data": "synthetic",
        "config": workload_config(args, 1),
        "cpu_baseline": {"value": round(gbs, 3), "unit": "GB/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": round(gbs, 3), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(args, world):
    return {
        "workload": "BASELINE.json configs[1]: 1-GPU probe pod, 4 GiB stream-triad + uint64 checksum vs CPU reference"
                    + ("" if world == 1 else f" (replicated per GPU, {world} ranks; lanes exchanged every step)"),
        "elements_per_gpu": 1 << args.log2n, "bytes_per_array_per_gpu": 8 << args.log2n,
        "step": "triad(24 B/elem) + checksum_u64(8 B/elem)" + ("" if world == 1 else " + lanes allreduce(16 B, fused into the checksum kernel)"),
        "variant": args.variant,
        "l2": "inputs are 3 x 4 GiB per GPU >> 126 MB L2, so no flush between iterations",
        "parallelism": f"shard{world}" if world > 1 else "single",
    }


# --------------------------------------------------------------------------------------
# Sharding: pure host logic, shared by the GPU arm and the gloo CPU tests
# --------------------------------------------------------------------------------------
def shard_first_index(rank: int, n_per_rank: int) -> int:
    """Rank r owns global indices [r*n, (r+1)*n) of ONE index space (weak scaling)."""
    return rank * n_per_rank


def to_i64(u: int) -> int:
    """u64 bit pattern as the int64 torch/NCCL/gloo sum over (two's complement: same bits)."""
    return u - (1 << 64) if u >= (1 << 63) else u


def from_i64(v: int) -> int:
    return v & ((1 << 64) - 1)


def allreduce_lanes(dist, torch, lanes, device="cpu"):
    """Sum (lane0, lane1) over all ranks mod 2^64 with an int64 SUM allreduce (wrapping
    add is associative and commutative, so any backend and any order gives the same bits)."""
    t = torch.tensor([to_i64(lanes[0]), to_i64(lanes[1])], dtype=torch.int64, device=device)
    dist.all_reduce(t)
    return from_i64(int(t[0].item())), from_i64(int(t[1].item()))


# --------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------
def run_ours(args, rank, local_rank, world):
    import numpy as np
    import torch

    import gpuprobe

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the probe has no CPU fallback (use --impl reference for the CPU arm)")
    gpuprobe.lib()  # raises if lib/libgpuprobe.so is missing
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # Coordination (barriers, max over ranks, handle exchange) runs over gloo.  The NCCL
        # communicator is created AFTER the timed HBM region: creating it costs every local
        # HBM-streaming kernel 3.4 % on this stack (6904 -> 6693 GB/s triad, persists while the
        # communicator lives, independent of NVLS/P2P/cuMem settings; tools/ablate_dist.py), and
        # the probe's own collectives do not need it.  NCCL is the baseline of the sweep below.
        dist.init_process_group("gloo")

    n = 1 << args.log2n
    first = shard_first_index(rank, n)  # each rank owns the shard [rank*n, (rank+1)*n) of one global index space
    variant = gpuprobe.VARIANTS[args.variant]
    st = torch.cuda.current_stream().cuda_stream
    b = torch.empty(n, dtype=torch.float64, device="cuda")
    c = torch.empty(n, dtype=torch.float64, device="cuda")
    a = torch.empty(n, dtype=torch.float64, device="cuda")
    ws = torch.zeros(gpuprobe.workspace_bytes() // 8, dtype=torch.int64, device="cuda")
    lanes = torch.zeros(2, dtype=torch.int64, device="cuda")
    total = torch.zeros(2, dtype=torch.int64, device="cuda")
    gpuprobe.init_fill(b.data_ptr(), n, first, gpuprobe.SEED_B, st)
    gpuprobe.init_fill(c.data_ptr(), n, first, gpuprobe.SEED_C, st)
    torch.cuda.synchronize()

    def barrier():
        if dist is not None:
            dist.barrier()

    def host_reduce(value, op):
        """MAX / MIN over ranks of one float, over the gloo group (CPU tensors)."""
        if dist is None:
            return value
        t = torch.tensor([value], dtype=torch.float64)
        dist.all_reduce(t, op=op)
        return float(t.item())

    def calibrate():
        """The ONLY blocking part of AUTO's pacing, made explicitly and outside every timed region."""
        if variant == gpuprobe.VARIANT_AUTO:
            gpuprobe.triad_calibrate(a.data_ptr(), b.data_ptr(), c.data_ptr(), n, st)
        return gpuprobe.triad_pace_ns()

    nccl = {"group": None}

    def nccl_group():
        """Created on first use by ALL ranks at the same point (after the HBM region)."""
        if nccl["group"] is None:
            nccl["group"] = dist.new_group(backend="nccl", device_id=torch.device("cuda", local_rank))
            warm = torch.ones(4, device="cuda")
            dist.all_reduce(warm, group=nccl["group"])
            torch.cuda.synchronize()
            calibrate()  # what the memory system sustains has just changed: pace again (blocking, untimed)
        return nccl["group"]

    # N > 1: the path's one collective (16 bytes of lanes).  Product path: fused into the
    # checksum kernel over NVLink peer stores (gpu_probe_checksum_u64_exchange); the mailbox
    # segments reach the other ranks as 64-byte handles over torch.distributed (gloo); underneath,
    # gpu_probe_exchange_connect_ipc passes VMM file descriptors over unix sockets.
    # NCCL (int64 allreduce after K2) is timed beside it as the two-launch baseline.
    xchg, combine = None, "none (single GPU)"
    if dist is not None and args.combine == "exchange":
        try:
            xchg = gpuprobe.Exchange(rank, world, 1 << args.sweep_log2max)  # payload: K4's allreduce operand
            mine = torch.tensor(list(xchg.export_handle()), dtype=torch.uint8)
            allh = [torch.empty_like(mine) for _ in range(world)]
            dist.all_gather(allh, mine)
            xchg.connect_ipc(b"".join(bytes(h.tolist()) for h in allh))
            ok = 1.0
        except gpuprobe.ProbeError as e:
            print(f"bench.py: rank {rank}: cross-process exchange unavailable ({e}); combining with NCCL", file=sys.stderr)
            ok = 0.0
        if host_reduce(ok, dist.ReduceOp.MIN) == 0.0:
            xchg = None
        combine = ("p2p_exchange_fused (gpu_probe_checksum_u64_exchange; segments shared across processes as VMM file descriptors)" if xchg is not None
                   else "nccl_allreduce_int64 (cross-process segments unavailable on this box)")
        dist.barrier()
    elif dist is not None:
        combine = "nccl_allreduce_int64"

    def step(ev=None, use_nccl=False):
        if ev:
            ev[0].record()
        gpuprobe.triad(a.data_ptr(), b.data_ptr(), c.data_ptr(), gpuprobe.SCALAR, n, variant, st)
        if ev:
            ev[1].record()
        if xchg is not None and not use_nccl:  # every rank's `total` = lanes of the whole index space
            xchg.checksum_u64(a.data_ptr(), n, first, total.data_ptr(), ws.data_ptr(), variant, st)
            if ev:
                ev[2].record()
            return
        gpuprobe.checksum_u64(a.data_ptr(), n, first, lanes.data_ptr(), ws.data_ptr(), variant, st)
        if ev:
            ev[2].record()
        if dist is not None:  # wrapping int64 sum == wrapping u64 sum, bit for bit
            total.copy_(lanes)
            dist.all_reduce(total, group=nccl_group())

    pace_first = calibrate()
    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()

    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
    t_start, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler = ClockSampler(local_rank, period_s=args.clock_period_ms / 1e3, enabled=not args.no_clocks).prepare()
    barrier()
    torch.cuda.synchronize()
    with sampler as clk:
        t_start.record()
        for k in range(args.steps):
            step(evs[k])
        t_end.record()
        torch.cuda.synchronize()
    barrier()
    ms_total = t_start.elapsed_time(t_end)
    pace_timed = gpuprobe.triad_pace_ns()  # as used in the timed region (creating NCCL later re-calibrates)
    triad_ms = [e[0].elapsed_time(e[1]) for e in evs]
    csum_ms = [e[1].elapsed_time(e[2]) for e in evs]
    # this rank's own shard lanes (untimed; the exchange path only ever holds the combined ones)
    gpuprobe.checksum_u64(a.data_ptr(), n, first, lanes.data_ptr(), ws.data_ptr(), variant, st)
    torch.cuda.synchronize()
    rank_lanes = tuple(int(x) for x in lanes.cpu().numpy().view(np.uint64))
    my_lanes = rank_lanes if dist is None else tuple(int(x) for x in total.cpu().numpy().view(np.uint64))

    nccl_ms_step = None
    per_rank = None
    if dist is not None:
        # every rank's own view of the timed region (rank 0 prints them: shows which GPU the others wait for)
        mine_stats = torch.tensor([ms_total / args.steps, sum(triad_ms) / len(triad_ms), sum(csum_ms) / len(csum_ms),
                                   float(pace_timed), statistics.median(triad_ms), max(triad_ms), float(pace_first)],
                                  dtype=torch.float64)
        all_stats = [torch.empty_like(mine_stats) for _ in range(world)]
        dist.all_gather(all_stats, mine_stats)
        per_rank = [{"rank": r, "ms_per_step": round(float(v[0]), 4), "triad_ms": round(float(v[1]), 4),
                     "triad_ms_median": round(float(v[4]), 4), "triad_ms_max": round(float(v[5]), 4),
                     "checksum_incl_wait_ms": round(float(v[2]), 4), "triad_pace_ns": int(v[3]),
                     # a GPU whose calibration lands at the slow end of the ladder (or on "ring") sustains ~5 % less HBM
                     # bandwidth than its neighbours: seen on GPUs another multi-GPU process used moments before
                     "slow_state": bool(int(v[6]) < 0 or int(v[6]) >= 1052)}
                    for r, v in enumerate(all_stats)]
        ms_total = host_reduce(ms_total, dist.ReduceOp.MAX)
        if xchg is not None:
            if xchg.status() != 0:
                raise SystemExit(f"bench.py: rank {rank}: lanes exchange timed out")
            # the same step with the two-launch NCCL combine, for comparison only
            k2 = min(args.steps, 20)
            nccl_group()  # created here (and the triad pace forgotten): everything below runs with NCCL alive
            for _ in range(max(args.warmup, 3)):
                step(use_nccl=True)
            barrier()
            torch.cuda.synchronize()
            t_start.record()
            for _ in range(k2):
                step(use_nccl=True)
            t_end.record()
            torch.cuda.synchronize()
            nccl_ms_step = host_reduce(t_start.elapsed_time(t_end) / k2, dist.ReduceOp.MAX)
            nccl_lanes = tuple(int(x) for x in total.cpu().numpy().view(np.uint64))
            if nccl_lanes != my_lanes:
                raise SystemExit(f"bench.py: rank {rank}: exchange lanes {my_lanes} != NCCL lanes {nccl_lanes}")
    ms_step = ms_total / args.steps
    step_bytes = (BYTES_TRIAD + BYTES_CHECKSUM) * n * world
    value = step_bytes / (ms_step / 1e3) / 1e9

    # ---- end to end: pinned host b,c -> device -> lanes back, through the pipeline ABI ----
    e2e = None if args.no_e2e else run_e2e(args, gpuprobe, torch, b, c, n, first, rank_lanes, dist, world)

    # ---- NVLink part of the metric (N > 1): allreduce sweep, bus bandwidth ----
    nvlink = run_allreduce_sweep(args, torch, dist, world, xchg, st, nccl_group()) if dist is not None else None

    if rank == 0:
        peak, peak_src = measured_peak()
        t_avg = sum(triad_ms) / len(triad_ms)
        c_avg = sum(csum_ms) / len(csum_ms)
        ach = BYTES_TRIAD * n / (t_avg / 1e3) / 1e9
        traffic = ncu_traffic()
        line = {
            "metric": METRIC, "value": round(value, 2), "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64+u64", "data": "synthetic", "config": workload_config(args, world),
            "roofline": {
                "bound": "hbm", "kernel": "triad", "achieved": round(ach, 1), "peak": peak, "unit": "GB/s",
                "frac": round(ach / peak, 4), "peak_source": peak_src, "frac_of_nominal_8000": round(ach / 8000.0, 4),
                "traffic": (traffic or {}).get("triad_paced_dram_bytes_per_launch_at_2^29" if pace_timed > 0
                                               else "triad_dram_bytes_per_launch"),
                "kernel_variant": "triad_tma_paced<1024,4>" if pace_timed > 0 else "triad_tma<512,4>",
                "algorithmic_bytes_per_launch": BYTES_TRIAD * n,
                "avg_launch_ms": round(t_avg, 4), "best_launch_ms": round(min(triad_ms), 4),
                "median_launch_ms": round(statistics.median(triad_ms), 4), "max_launch_ms": round(max(triad_ms), 4),
                "triad_pace_ns": pace_timed,  # gpu_probe_triad_calibrate's verdict (> 0: paced TMA kernel; -2: drift guard fired)
                "slow_state": bool(pace_first < 0 or pace_first >= 1052),
            },
            "kernels": {
                "triad": {"bytes": BYTES_TRIAD * n, "avg_ms": round(t_avg, 4), "best_ms": round(min(triad_ms), 4),
                          "median_ms": round(statistics.median(triad_ms), 4), "max_ms": round(max(triad_ms), 4),
                          "GBps_avg": round(ach, 1), "frac_measured_peak": round(ach / peak, 4)},
                "checksum_u64": {"bytes": BYTES_CHECKSUM * n, "avg_ms": round(c_avg, 4),
                                 "best_ms": round(min(csum_ms), 4),
                                 "median_ms": round(statistics.median(csum_ms), 4), "max_ms": round(max(csum_ms), 4),
                                 "GBps_avg": round(BYTES_CHECKSUM * n / (c_avg / 1e3) / 1e9, 1),
                                 "frac_measured_peak": round(BYTES_CHECKSUM * n / (c_avg / 1e3) / 1e9 / peak, 4)},
            },
            "clocks": clk.summary(),
            "e2e": e2e,
            "gpu_launches": 2 * args.steps,
            "lanes": [f"0x{my_lanes[0]:016x}", f"0x{my_lanes[1]:016x}"],
        }
        if nvlink is not None:
            line["nvlink"] = nvlink
        if dist is not None:
            line["per_rank"] = per_rank
            line["combine"] = {"how": combine, "ms_per_step_with_nccl_combine": nccl_ms_step,
                               "note": "ranks coordinate over gloo; the NCCL communicator is created after the timed region "
                                       "(creating it slows every local HBM kernel by ~3.4 % for as long as it lives, "
                                       "tools/ablate_dist.py), so the NCCL-combine figure carries that cost",
                               "launches_per_step": 2 if xchg is not None else 4}
            line["gpu_launches"] = (2 if xchg is not None else 4) * args.steps
            # weak scaling over one global index space: the combined lanes have a known answer
            try:
                with open(os.path.join(ROOT, "tests", "golden", "probe_golden.json")) as f:
                    pow2 = json.load(f)["pow2_lanes"]
                want = tuple(int(v, 16) for v in pow2[(n * world).bit_length() - 1]) if (n * world) & (n * world - 1) == 0 else None
            except Exception:
                want = None
            line["parity"] = {"combined_lanes_equal_golden": (want == my_lanes) if want else None,
                              "checked": f"{world} shards x 2^{args.log2n} elements, golden = tests/golden/probe_golden.json pow2_lanes"}
            if want is not None and want != my_lanes:
                print(json.dumps(line), flush=True)
                raise SystemExit("bench.py: combined GPU lanes differ from the golden lanes")
        if world == 1 and not args.no_cpu:
            import oracle
            want = oracle.stream_lanes(n, first)
            line["parity"] = {"gpu_lanes_equal_cpu_oracle": want == my_lanes, "checked": f"2^{args.log2n} elements"}
            avail = mem_available_bytes()
            log2s = args.cpu_log2n
            while log2s > 20 and avail and 3 * (8 << log2s) > 0.5 * avail:
                log2s -= 1
            line["cpu_baseline"] = cpu_baseline(1 << log2s)
            line["cpu_baseline_single_core"] = cpu_baseline(1 << min(log2s, 25), repeats=2, threads=1)
            if want != my_lanes:
                print(json.dumps(line), flush=True)
                raise SystemExit("bench.py: GPU lanes differ from the CPU oracle")
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        if xchg is not None:
            xchg.close()
        del a, b, c
        dist.barrier()
        dist.destroy_process_group()


def restore_hbm_state(world):
    """Leave the GPUs as this run found them.  Runs in rank 0's PARENT process (see main): by now the
    rank's worker -- the only process of this job that ever held a CUDA context here -- is gone.

    A multi-process run's collectives leave every GPU they touched 3.8 % short of its read+write
    HBM bandwidth FOR THE NEXT PROCESS, for minutes (fresh-process triad 7030-7120 -> 6690 GB/s,
    calibration verdict 1082 ns; read-only and write-only streams are unaffected, nothing NVML
    reports moves).  What puts it back, every time it was tried, is bulk peer traffic made by ONE
    process that holds all the GPUs involved and is the last to touch them: the host binary's P2P
    phase or its allreduce sweep to 1 GiB (profiles/slow_state_r2.md sections 3-4).  The same copies
    issued from inside the job only cured the GPUs whose own rank was already gone.  So the cure is
    the probe's own pod command on small arrays, run once after the job: a few seconds, after
    everything that is timed or reported."""
    exe = os.path.join(ROOT, "kind-gpu-sim_b200", "bin", "gpu-probe")
    if world < 2 or not os.path.exists(exe):
        return
    time.sleep(1.5)  # the other ranks are on their way out
    try:
        r = subprocess.run([exe, f"--max-gpus={world}", "--bytes=64m", "--iters=2", "--warmup=0"], capture_output=True,
                           text=True, timeout=240)
        if r.returncode != 0:
            print(f"bench.py: closing HBM-state pass exited {r.returncode}: {r.stdout[-200:]} {r.stderr[-200:]}", file=sys.stderr)
    except Exception as e:  # hygiene only: never fail a finished measurement over it
        print(f"bench.py: closing HBM-state pass skipped ({type(e).__name__}: {e})", file=sys.stderr)


def run_e2e(args, gpuprobe, torch, b, c, n, first, want_lanes, dist, world):
    """Same step through the host-buffer ABI.  Inputs start in pinned HOST memory every
    step (untimed setup copies them out of the device buffers K0 filled)."""
    log2e = min(args.e2e_log2n, args.log2n)
    avail = mem_available_bytes()
    while log2e > 20 and avail and world * 2 * (8 << log2e) > 0.4 * avail:
        log2e -= 1
    ne = 1 << log2e
    hb = torch.empty(ne, dtype=torch.float64).pin_memory()
    hc = torch.empty(ne, dtype=torch.float64).pin_memory()
    hb.copy_(b[:ne])
    hc.copy_(c[:ne])
    torch.cuda.synchronize()
    pipe = gpuprobe.Pipeline(0, 0)
    try:
        for _ in range(max(1, args.e2e_warmup)):
            got = pipe.run_host(hb.data_ptr(), hc.data_ptr(), None, gpuprobe.SCALAR, ne, first)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            got = pipe.run_host(hb.data_ptr(), hc.data_ptr(), None, gpuprobe.SCALAR, ne, first)  # blocking
        dt = time.perf_counter() - t0
    finally:
        pipe.close()
    if ne == n and got != want_lanes:
        raise SystemExit(f"bench.py: e2e lanes {got} differ from the resident path {want_lanes}")
    ms = dt / args.e2e_steps * 1e3
    if dist is not None:
        t = torch.tensor([ms], dtype=torch.float64)  # gloo group: CPU tensor
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    gbs = (BYTES_TRIAD + BYTES_CHECKSUM) * ne * world / (ms / 1e3) / 1e9
    return {"value": round(gbs, 2), "unit": "GB/s", "h2d_bytes_per_step": 16 * ne, "d2h_bytes_per_step": 16,
            "ms_per_step": round(ms, 3), "steps": args.e2e_steps, "elements_per_gpu": ne,
            "api": "gpu_probe_pipeline_run_host (pinned host b,c -> chunked H2D overlapped with fused "
                   "triad+checksum -> lanes D2H); host wall clock around blocking calls",
            "pcie_h2d_GBps": round(16 * ne / (ms / 1e3) / 1e9, 2)}


def run_allreduce_sweep(args, torch, dist, world, xchg, st, nccl):
    """In-place fp32 sum allreduce, 1 KiB .. 1 GiB per rank; busbw = algbw * 2(G-1)/G.
    Two implementations on the same sizes: NCCL (torch.distributed) and the probe's own
    one-kernel allreduce over peer memory (gpu_probe_allreduce_f32, K4)."""
    rank = dist.get_rank()
    max_bytes = 1 << args.sweep_log2max
    buf = torch.ones(max_bytes // 4, dtype=torch.float32, device="cuda")

    def series(fn, size, iters):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64)  # gloo: CPU tensor
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        algbw = size / (ms / 1e3) / 1e9
        return {"bytes": size, "us": round(ms * 1e3, 2), "algbw_GBps": round(algbw, 2),
                "busbw_GBps": round(algbw * 2 * (world - 1) / world, 2)}

    def summary(rows, what):
        top, best = rows[-1], max(rows, key=lambda r: r["busbw_GBps"])
        return {"collective": what, "busbw_GBps_at_max": top["busbw_GBps"], "max_bytes": top["bytes"],
                "busbw_GBps_peak": best["busbw_GBps"], "peak_bytes": best["bytes"],
                "frac_of_nominal_900": round(top["busbw_GBps"] / NVLINK_NOMINAL_GBS, 4),
                "latency_us_1KiB": rows[0]["us"], "sweep": rows}

    rows_nccl, rows_own = [], []
    size = 1 << 10
    while size <= max_bytes:
        view = buf[: size // 4]
        iters = 20 if size >= (1 << 20) else 50
        reps = 3 if size <= (4 << 20) else 1  # short series are noisy: keep the best of three
        rows_nccl.append(min((series(lambda: dist.all_reduce(view, group=nccl), size, iters) for _ in range(reps)),
                             key=lambda r: r["us"]))
        buf.fill_(1.0)
        if xchg is not None:
            rows_own.append(min((series(lambda: xchg.allreduce_f32(0, size // 4, 0, st), size, iters)
                                 for _ in range(reps)), key=lambda r: r["us"]))
        size <<= 1
    own = None
    if xchg is not None:
        # correctness of K4 at the bench's scale: integer-valued inputs, closed-form sum, every rank
        import gpuprobe
        n_chk = min(max_bytes // 4, (1 << 22) + 3)
        src = (torch.arange(n_chk, device="cuda", dtype=torch.float32) % 7 + 1) * float(rank + 1)
        gpuprobe.p2p_copy(xchg.buffer, src.data_ptr(), 4 * n_chk, st)
        torch.cuda.synchronize()
        dist.barrier()
        xchg.allreduce_f32(0, n_chk, 0, st)
        out = torch.empty_like(src)
        gpuprobe.p2p_copy(out.data_ptr(), xchg.buffer, 4 * n_chk, st)
        torch.cuda.synchronize()
        want = (torch.arange(n_chk, device="cuda", dtype=torch.float32) % 7 + 1) * float(world * (world + 1) // 2)
        ok = torch.tensor([int(torch.equal(out, want) and xchg.status() == 0)], dtype=torch.int32)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok.item()) != 1:
            raise SystemExit(f"bench.py: rank {rank}: gpu_probe_allreduce_f32 result differs from the closed-form sum")
        own = summary(rows_own, "gpu_probe_allreduce_f32 (K4, one kernel over NVLink peer memory; AUTO = one-shot push "
                                "below the threshold, then NVSwitch multimem reduce+broadcast if multicast is "
                                "available and G >= 3, else pull-reduce + push-broadcast), max over ranks")
        own["nvswitch_multicast"] = bool(xchg.has_multicast)
        own["checked"] = f"{n_chk} elements against the closed-form sum on every rank"
    if rank != 0:
        return None
    out = summary(rows_nccl, "NCCL allreduce fp32 sum in place (torch.distributed), max over ranks")
    if own is not None:
        out["p2p_kernel"] = own
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--log2n", type=int, default=29, help="fp64 elements per array per GPU (29 = BASELINE 4 GiB)")
    ap.add_argument("--variant", choices=["auto", "ldg128", "tma", "ldg256", "tma_paced"], default="auto")
    ap.add_argument("--e2e-log2n", type=int, default=29)
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--e2e-warmup", type=int, default=1)
    ap.add_argument("--cpu-log2n", type=int, default=29, help="cpu_baseline sample size (elements per array; halved while 3 arrays exceed half of MemAvailable)")
    ap.add_argument("--ref-log2n", type=int, default=29, help="--impl reference elements per array per step (same fallback)")
    ap.add_argument("--sweep-log2max", type=int, default=30)
    ap.add_argument("--combine", choices=["exchange", "nccl"], default="exchange",
                    help="N>1: how the 16-byte lanes are summed over ranks (fused P2P exchange, or NCCL)")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg (profiling runs only)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline / parity leg")
    ap.add_argument("--no-restore", action="store_true", help="N>1: skip the closing pass that puts the GPUs' HBM state back (see restore_hbm_state)")
    ap.add_argument("--no-clocks", action="store_true", help="ablation only: do not sample NVML clocks while timing")
    ap.add_argument("--clock-period-ms", type=float, default=5.0, help="NVML sampling period during the timed region")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3  # timing rule: at least 3 warm-up steps
    rank, local_rank, world = env_int("RANK", 0), env_int("LOCAL_RANK", 0), env_int("WORLD_SIZE", 1)
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit(f"bench.py: --gpus {args.gpus} needs torchrun (one rank per GPU); see the docstring")
        args.gpus = world
    if world > 1 and rank == 0 and not args.no_restore and os.environ.get("GPU_PROBE_BENCH_WORKER") != "1":
        # Rank 0 runs its share of the job in a child, so that THIS process never holds a CUDA context and can
        # run the closing HBM-state pass when nothing of the job is left on the GPUs (restore_hbm_state).
        child = subprocess.Popen([sys.executable, os.path.abspath(__file__)] + sys.argv[1:],
                                 env=dict(os.environ, GPU_PROBE_BENCH_WORKER="1"))
        for sig in (signal.SIGTERM, signal.SIGINT):
            signal.signal(sig, lambda s, f: child.send_signal(s))
        rc = child.wait()
        if rc == 0:
            restore_hbm_state(world)
        sys.exit(rc)
    run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
