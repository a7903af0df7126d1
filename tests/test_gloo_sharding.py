"""N > 1 host logic on CPU: world_size-2 and -4 gloo runs (no GPU needed)."""
import json
import os
import subprocess
import sys

import pytest

import bench

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def torchrun(nproc, port, script, *args, timeout=300):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), script, *args]
    env = dict(os.environ, OMP_NUM_THREADS="2")
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env, cwd=ROOT)


@pytest.mark.parametrize("world,port", [(2, 29611), (4, 29612)])
def test_sharded_lanes_combine_bit_exact_over_gloo(world, port):
    r = torchrun(world, port, os.path.join(ROOT, "tests", "mp_gloo_worker.py"))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert f"GLOO_SHARDING_OK world={world}" in r.stdout


def test_i64_u64_round_trip_and_shard_ownership():
    for u in (0, 1, (1 << 63) - 1, 1 << 63, (1 << 64) - 1, 0xD761E4BA92530D51):
        assert bench.from_i64(bench.to_i64(u)) == u
        assert -(1 << 63) <= bench.to_i64(u) < (1 << 63)
    n = 1 << 29
    assert [bench.shard_first_index(r, n) for r in range(4)] == [0, n, 2 * n, 3 * n]


def test_reference_arm_under_torchrun_prints_exactly_one_line():
    r = torchrun(2, 29613, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "2",
                 "--warmup", "1", "--ref-log2n", "20")
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 2 and d["unit"] == "GB/s" and d["value"] > 0
    assert d["cpu_baseline"]["kind"] == "port" and d["e2e"]["h2d_bytes_per_step"] == 0
