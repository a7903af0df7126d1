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


@pytest.mark.parametrize("n", [1, 2, 4, 8])
def test_committed_bench_lines_keep_the_contract(n):
    with open(os.path.join(ROOT, "profiles", f"bench_n{n}_r1.json")) as f:
        d = json.load(f)
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
