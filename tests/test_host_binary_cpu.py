"""CPU-side checks of the pod's C host binary: it builds, validates its arguments with the
documented exit codes, fails loudly without a GPU (no CPU path), and the known answers it
embeds equal the committed golden lanes and the oracle."""
import os
import re
import subprocess

import pytest

import oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "kind-gpu-sim_b200", "bin", "gpu-probe")
HDR = os.path.join(ROOT, "kind-gpu-sim_b200", "host", "golden_lanes.h")


def _no_gpu():
    try:
        import torch
        return not torch.cuda.is_available()
    except Exception:
        return True


def test_usage_errors_exit_2():
    for bad in (["--bogus"], ["--bytes=12"], ["--bytes=abc"], ["--variant=fast"], ["--iters=0"],
                ["--sweep-max-bytes=3000"]):
        r = subprocess.run([BIN, *bad], capture_output=True, text=True)
        assert r.returncode == 2, (bad, r.stdout, r.stderr)
        assert "PROBE FAIL rc=2" in r.stdout and "usage:" in r.stderr
    assert "--inject-fault" in r.stderr and "GPU_PROBE_MEASURED_HBM_GBS" in r.stderr


@pytest.mark.skipif(not _no_gpu(), reason="only meaningful on a machine without a GPU")
def test_without_gpu_fails_loudly_never_computes():
    r = subprocess.run([BIN, "--expect-gpus=1"], capture_output=True, text=True)
    assert r.returncode == 3 and "PROBE FAIL rc=3" in r.stdout and "RESULT" not in r.stdout
    r = subprocess.run([BIN], capture_output=True, text=True)
    assert r.returncode == 4 and "no CPU path" in r.stderr


def test_embedded_known_answers_match_golden_and_oracle(golden):
    rows = re.findall(r"UINT64_C\(1\) << (\d+), UINT64_C\((0x[0-9a-f]{16})\), UINT64_C\((0x[0-9a-f]{16})\)", open(HDR).read())
    assert [int(r[0]) for r in rows] == list(range(34))
    assert [[r[1], r[2]] for r in rows] == golden["pow2_lanes"]
    for k in (0, 1, 7, 12, 20, 24):  # recomputed here; the large ones are pinned by make_golden.py
        l0, l1 = oracle.stream_lanes(1 << k, 0)
        assert rows[k][1:] == (f"0x{l0:016x}", f"0x{l1:016x}")
    for k, v in golden["stream_lanes"].items():  # the older fixture agrees with the new table
        assert golden["pow2_lanes"][int(k).bit_length() - 1] == v


def test_host_binary_links_the_probe_library_and_no_oracle():
    deps = subprocess.run(["ldd", BIN], capture_output=True, text=True).stdout
    assert "libgpuprobe.so" in deps and "probe_ref" not in deps and "libnccl" not in deps
