"""The pod's C host binary (kind-gpu-sim_b200/bin/gpu-probe) on a real GPU: RESULT wire
format, known-answer lanes, exit codes.  The lanes it prints are also recomputed by the
CPU oracle here, so the constants it embeds are checked, not trusted."""
import os
import re
import subprocess

import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "kind-gpu-sim_b200", "bin", "gpu-probe")


def run(*args, timeout=600):
    r = subprocess.run([BIN, *args], capture_output=True, text=True, timeout=timeout)
    results = [dict(kv.split("=", 1) for kv in line.split()[1:]) for line in r.stdout.splitlines()
               if line.startswith("RESULT ")]
    return r, results


def test_one_gpu_probe_passes_with_known_answers():
    log2n = 24
    r, res = run("--max-gpus=1", f"--bytes={8 << log2n}", "--iters=5", "--warmup=2")
    assert r.returncode == 0, r.stdout + r.stderr
    assert re.search(r"^PROBE PASS rc=0 ", r.stdout, re.M)
    kinds = [x["kind"] for x in res]
    assert kinds == ["init_fill", "triad", "checksum", "triad_checksum"]
    want = oracle.stream_lanes(1 << log2n, 0)
    for x in res:
        assert set(x) >= {"kind", "gpus", "gpu", "bytes", "ms", "ms_med", "gbs", "frac_nominal", "frac_measured",
                          "lanes", "match"}
        assert float(x["ms"]) > 0 and float(x["gbs"]) > 0
    for x in res[2:]:
        assert x["match"] == "1"
        assert x["lanes"] == f"{want[0]:016x},{want[1]:016x}"
    assert int(res[1]["bytes"]) == 24 << log2n and int(res[2]["bytes"]) == 8 << log2n


def test_size_without_known_answer_reports_na_but_still_cross_checks():
    n = (1 << 20) + 8
    r, res = run("--max-gpus=1", f"--bytes={8 * n}", "--iters=2", "--warmup=1", "--variant=ldg128")
    assert r.returncode == 0, r.stdout + r.stderr
    want = oracle.stream_lanes(n, 0)
    assert res[2]["match"] == "na" and res[2]["lanes"] == f"{want[0]:016x},{want[1]:016x}"
    assert res[3]["lanes"] == res[2]["lanes"]


def test_injected_fault_is_caught_and_exits_6():
    """--inject-fault flips bit 17 of a[n/2] between K1 and K2: the checksum must disagree with the
    known answer (match=0) and with the fused kernel's lanes, and the pod must fail with rc 6."""
    log2n = 24
    r, res = run("--max-gpus=1", f"--bytes={8 << log2n}", "--iters=2", "--warmup=1", "--inject-fault")
    assert r.returncode == 6, r.stdout + r.stderr
    assert "PROBE FAIL rc=6" in r.stdout and "PROBE INFO inject_fault gpu=0" in r.stdout
    want = oracle.stream_lanes(1 << log2n, 0)
    by = {x["kind"]: x for x in res}
    assert by["checksum"]["match"] == "0" and by["checksum"]["lanes"] != f"{want[0]:016x},{want[1]:016x}"
    # exactly the flipped bit: lane0 differs by +-2^17, nothing else
    got0 = int(by["checksum"]["lanes"].split(",")[0], 16)
    assert (got0 - want[0]) % (1 << 64) in (1 << 17, (1 << 64) - (1 << 17))
    assert by["triad_checksum"]["match"] == "1"  # the fused pass recomputes a[] and is clean again
    # and on a size without a known answer the fused-vs-separate cross-check still catches it
    r, res = run("--max-gpus=1", f"--bytes={8 * ((1 << 20) + 8)}", "--iters=2", "--warmup=1", "--inject-fault")
    assert r.returncode == 6 and "fused lanes differ" in r.stderr


def test_calibration_is_reported_before_anything_is_timed():
    r, _ = run("--max-gpus=1", "--bytes=1g", "--iters=3", "--warmup=1")
    assert r.returncode == 0, r.stdout + r.stderr
    lines = r.stdout.splitlines()
    cal = [i for i, ln in enumerate(lines) if ln.startswith("PROBE INFO triad_calibrate gpu=0 pace_ns=")]
    first_result = min(i for i, ln in enumerate(lines) if ln.startswith("RESULT "))
    assert len(cal) == 1 and cal[0] < first_result
    pace = int(re.search(r"pace_ns=(-?\d+)", lines[cal[0]]).group(1))
    assert pace == -1 or 1000 <= pace <= 1100
    triad = [ln for ln in lines if ln.startswith("RESULT kind=triad ")][0]
    assert f"pace_ns={pace}" in triad  # the verdict stood through the timed launches


def test_device_count_contract():
    have = torch.cuda.device_count()
    r, _ = run(f"--expect-gpus={have + 1}", "--bytes=1m")
    assert r.returncode == 3 and "PROBE FAIL rc=3" in r.stdout
    assert "promised" in r.stderr


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_multi_gpu_probe_sharded_checksum_sweep_and_p2p():
    have = min(torch.cuda.device_count(), 8)
    g = 1 << (have.bit_length() - 1)  # 2, 4 or 8: n*g stays a power of two -> known answer
    log2n = 24
    r, res = run(f"--max-gpus={g}", f"--bytes={8 << log2n}", "--iters=5", "--warmup=2", "--sweep-max-bytes=64m",
                 "--p2p-bytes=64m")
    assert r.returncode == 0, r.stdout[-4000:] + r.stderr[-2000:]
    want = oracle.stream_lanes(g << log2n, 0)
    hexw = f"{want[0]:016x},{want[1]:016x}"
    by = {}
    for x in res:
        by.setdefault(x["kind"], []).append(x)
    assert by["checksum_sum"][0]["lanes"] == hexw and by["checksum_sum"][0]["match"] == "1"
    combos = {x["combine"]: x for x in by["sharded_checksum"]}
    assert set(combos) == {"nccl_allreduce_u64", "p2p_exchange_fused"}
    for x in combos.values():
        assert x["lanes"] == hexw and x["match"] == "1"
    assert all(x["match"] == "1" for x in by["allreduce_check"])
    impls = {x["impl"] for x in by["allreduce_check"]}
    assert {"nccl", "p2p_kernel"} <= impls <= {"nccl", "nccl_registered", "nccl_symmetric", "p2p_kernel"}
    # the registered columns appear whenever the loaded NCCL can register buffers, and then they are checked too
    for impl in impls:
        assert [int(x["bytes"]) for x in by["allreduce"] if x["gpus"] == "2" and x["impl"] == impl] == [1 << k for k in range(10, 27)]
    for impl in ("nccl", "p2p_kernel"):
        sizes = [int(x["bytes"]) for x in by["allreduce"] if x["gpus"] == "2" and x["impl"] == impl]
        assert sizes == [1 << k for k in range(10, 27)], impl
    assert all(x["match"] == "1" for x in by["p2p"])
    assert {x["mode"] for x in by["p2p"]} == {"sm_kernel", "sm_kernel_pull", "copy_engine"}


def _pod(*args):
    import json
    import sys
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "pod_runner.py"), "--json", *args],
                       capture_output=True, text=True, timeout=900)
    return r, json.loads(r.stdout)


def test_probe_pod_1gpu_runs_the_baseline_config_and_passes():
    # BASELINE.json configs[1] exactly as the pod manifest states it: 4 GiB arrays on one GPU
    r, out = _pod(os.path.join(ROOT, "pods", "probe-1gpu.yaml"))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    run0 = out["runs"][0]
    assert out["placement"] == {"nodeSelector_ok": True, "toleration_ok": True}
    assert run0["visible"] == "0" and run0["stdout"][-1].startswith("PROBE PASS rc=0")
    lanes = [ln for ln in run0["stdout"] if ln.startswith("RESULT kind=checksum ")][0]
    assert "lanes=0277599ce8795db9,d761e4ba92530d51 match=1" in lanes  # golden lanes of 2^29 elements
    assert run0["launch_to_first_result_s"] is not None


def test_concurrent_one_gpu_pods_are_isolated():
    # configs[3]: one 1-GPU pod per visible GPU at the same time, each seeing exactly its own device
    g = min(torch.cuda.device_count(), 8)
    r, out = _pod(os.path.join(ROOT, "pods", "probe-1gpu.yaml"), "--replicas", str(g), "--", "--bytes=1g", "--iters=5")
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    seen = set()
    for k, run_k in enumerate(out["runs"]):
        assert run_k["rc"] == 0 and run_k["visible"] == str(k)
        info = [ln for ln in run_k["stdout"] if ln.startswith("PROBE INFO visible_gpus=")][0]
        assert "visible_gpus=1 expect_gpus=1" in info
        seen.add([ln for ln in run_k["stdout"] if ln.startswith("PROBE GPU index=0")][0].split("pci=")[1])
    assert len(seen) == g  # distinct PCI bus ids: no two pods shared a GPU
