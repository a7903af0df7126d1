"""The paced triad's fallback path (drain the ring, restart slower) must never change a bit.

The pace is forced through GPU_PROBE_TRIAD_PACE_NS, which the library reads once, so each case
runs in its own process: an absurdly fast pace (every CTA falls down the ladder again and
again), a pace just past what HBM sustains, and a very slow one.  Lanes of the result are
compared with the golden lanes, and the fused kernel with the separate pair."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SNIPPET = r'''
import json, os, sys
sys.path.insert(0, os.path.join(%(root)r, "kind-gpu-sim_b200"))
import numpy as np, torch, gpuprobe
n = 1 << 27
st = torch.cuda.current_stream().cuda_stream
b = torch.empty(n, dtype=torch.float64, device="cuda"); c = torch.empty_like(b); a = torch.empty_like(b)
ws = torch.zeros(gpuprobe.workspace_bytes() // 8, dtype=torch.int64, device="cuda")
out = torch.zeros(2, dtype=torch.int64, device="cuda")
gpuprobe.init_fill(b.data_ptr(), n, 0, gpuprobe.SEED_B, st)
gpuprobe.init_fill(c.data_ptr(), n, 0, gpuprobe.SEED_C, st)
res = []
for rep in range(2):
    a.fill_(-1.0)
    gpuprobe.triad(a.data_ptr(), b.data_ptr(), c.data_ptr(), gpuprobe.SCALAR, n, gpuprobe.VARIANT_AUTO, st)
    gpuprobe.checksum_u64(a.data_ptr(), n, 0, out.data_ptr(), ws.data_ptr(), gpuprobe.VARIANT_LDG, st)
    torch.cuda.synchronize()
    res.append([int(v) for v in out.cpu().numpy().view(np.uint64)])
    gpuprobe.triad_checksum(a.data_ptr(), b.data_ptr(), c.data_ptr(), gpuprobe.SCALAR, n, 0, out.data_ptr(), ws.data_ptr(),
                            gpuprobe.VARIANT_AUTO, st)
    torch.cuda.synchronize()
    res.append([int(v) for v in out.cpu().numpy().view(np.uint64)])
print(json.dumps({"lanes": res, "pace": gpuprobe.triad_pace_ns()}))
'''


@pytest.mark.parametrize("pace_ns", [1, 700, 980, 5000])
def test_forced_pace_keeps_results_bit_exact(pace_ns, golden):
    env = dict(os.environ, GPU_PROBE_TRIAD_PACE_NS=str(pace_ns))
    r = subprocess.run([sys.executable, "-c", SNIPPET % {"root": ROOT}], capture_output=True, text=True, env=env,
                       timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    got = json.loads(r.stdout.strip().splitlines()[-1])
    want = [int(v, 16) for v in golden["pow2_lanes"][27]]
    assert got["lanes"] == [want] * 4, (pace_ns, got)
