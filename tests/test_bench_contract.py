"""The bench line's contract (keys the driver and the judge read), checked on CPU against the
committed lines in profiles/ and against a fresh `--impl reference` run."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
             "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches"}


def _baseline_metric():
    with open(os.path.join(ROOT, "BASELINE.json")) as f:
        return json.load(f)["metric"]


R2_LINES = {1: "scale_r2_norestore_n1.json", 2: "scale_r2_norestore_n2.json", 4: "scale_r2_norestore_n4.json",
            8: "scale_r2_restore_n8.json"}


@pytest.mark.parametrize("n,name", [(n, f"bench_n{n}_r1.json") for n in (1, 2, 4, 8)] + sorted(R2_LINES.items()))
def test_committed_bench_lines_keep_the_contract(n, name):
    with open(os.path.join(ROOT, "profiles", name)) as f:
        d = json.load(f)
    if "_r2" in name:  # round 2 added: median and max per kernel, the slow-state flag, >= 5 clock samples, a first-hand traffic figure
        assert d["roofline"]["median_launch_ms"] <= d["roofline"]["max_launch_ms"]
        assert isinstance(d["roofline"]["slow_state"], bool) and d["clocks"]["samples"] >= 5
        assert {"median_ms", "max_ms"} <= set(d["kernels"]["triad"]) and {"median_ms", "max_ms"} <= set(d["kernels"]["checksum_u64"])
        if d["roofline"]["triad_pace_ns"] > 0:
            assert d["roofline"]["traffic"] in (12815095500, 12849892096)
        if n > 1:
            assert all({"slow_state", "triad_ms_median", "triad_ms_max"} <= set(r) for r in d["per_rank"])
            assert max(r["triad_ms_max"] for r in d["per_rank"]) < 2.5  # no 55 ms calibration inside the timed loop any more
    assert BASE_KEYS <= set(d), sorted(BASE_KEYS - set(d))
    assert d["metric"] == _baseline_metric() and d["unit"] == "GB/s" and d["n_gpus"] == n
    assert d["higher_is_better"] is True and d["scaling"] == "weak" and d["vs_baseline"] is None
    assert d["data"] == "synthetic" and "workload" in d["config"] and "model" not in d["config"]
    assert d["warmup"] >= 3 and d["gpu_launches"] == 2 * d["steps"]
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-3
    assert r["traffic"] is None or 0.98 < r["traffic"] / r["algorithmic_bytes_per_launch"] < 1.02
    e = d["e2e"]
    assert e["unit"] == "GB/s" and e["h2d_bytes_per_step"] == 16 * e["elements_per_gpu"] and e["d2h_bytes_per_step"] == 16
    assert 0 < e["value"] < d["value"]          # host buffers cross PCIe: never the HBM-resident figure
    assert set(d["clocks"]) >= {"sm_mhz", "sm_max_mhz", "reasons"}
    assert not set(d["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    if n == 1:
        c = d["cpu_baseline"]
        assert c["kind"] == "port" and c["cores"] >= 1 and c["value"] > 0 and "sample" in c
        assert d["parity"]["gpu_lanes_equal_cpu_oracle"] is True
    else:
        assert d["parity"]["combined_lanes_equal_golden"] is True
        assert len(d["per_rank"]) == n
        nv = d["nvlink"]
        assert nv["sweep"][0]["bytes"] == 1024 and nv["sweep"][-1]["bytes"] == 1 << 30
        assert nv["p2p_kernel"]["busbw_GBps_at_max"] > 0


def test_reference_arm_line_keeps_the_contract():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "1",
                        "--ref-log2n", "20"], capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    assert d["impl"] == "reference" and d["metric"] == _baseline_metric() and d["unit"] == "GB/s"
    assert d["cpu_baseline"]["value"] == d["value"] == d["e2e"]["value"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0 and d["gpu_launches"] == 0


def test_rank0_parent_passes_the_workers_failure_through_and_skips_the_closing_pass(tmp_path):
    """N > 1: rank 0 runs its share in a child so that the parent, which never holds a CUDA context, can run the
    closing HBM-state pass once the job is gone.  On a machine without a GPU the child fails loudly (no CPU
    fallback); the parent must hand that exit status on and must not start the pod command."""
    env = dict(os.environ, RANK="0", LOCAL_RANK="0", WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT="29591")
    env.pop("GPU_PROBE_BENCH_WORKER", None)
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("GPU present: the N > 1 path is exercised by the scaling jobs")
    except ImportError:
        pass
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--gpus", "2", "--steps", "1"], capture_output=True,
                       text=True, timeout=300, cwd=ROOT, env=env)
    assert r.returncode != 0
    assert r.stderr.count("no CUDA device; the probe has no CPU fallback") == 1   # said once, by the worker
    assert "closing HBM-state pass" not in r.stderr and "{" not in r.stdout
