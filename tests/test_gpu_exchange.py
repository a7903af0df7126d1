"""The checksum fused with its collective (gpu_probe_*_exchange), through the C ABI.

Every rank must end with the lanes of the WHOLE buffer, bit-exact against the CPU
oracle -- the same bar as the single-GPU checksum.  With one GPU per rank the tests launch
one kernel per rank and cross NVLink, in one process (connect_local) and in one process
per GPU (file descriptors over unix sockets, handles carried over gloo).  With fewer GPUs
than ranks, kernels that wait for each other must not be separate launches (nothing makes
them run at the same time), so all ranks live on cuda:0 and run as the rows of ONE launch
(gpu_probe_checksum_u64_exchange_group): same device code, same mailbox protocol."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

import gpuprobe
import oracle
from gpu_util import f64, from_np, lanes_buf, lanes_of, workspace

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ALL_ONES = ((1 << 64) - 1, (1 << 64) - 1)


def _devices(world: int) -> list[int]:
    """Rank -> device: one GPU per rank when the box has them, else everyone on cuda:0."""
    return list(range(world)) if torch.cuda.device_count() >= world else [0] * world


class Ranks:
    """`world` ranks in this process: mailbox, stream, workspace and lanes buffer each."""

    def __init__(self, world: int):
        self.world, self.dev = world, _devices(world)
        self.x, self.st, self.ws, self.out = [], [], [], []
        for r in range(world):
            with torch.cuda.device(self.dev[r]):
                self.x.append(gpuprobe.Exchange(r, world))
                self.st.append(torch.cuda.Stream())
                self.ws.append(workspace())
                self.out.append(lanes_buf())
        gpuprobe.Exchange.connect_local(self.x)
        self.grouped = world > 1 and len(set(self.dev)) == 1

    def checksum(self, ptrs, ns, firsts, outs):
        """One sharded checksum + exchange: a launch per rank, or (shared GPU) all ranks in one launch."""
        if self.grouped:
            with torch.cuda.device(0):
                gpuprobe.Exchange.checksum_u64_group(self.x, ptrs, ns, firsts, [o.data_ptr() for o in outs],
                                                     [w.data_ptr() for w in self.ws], self.st[0].cuda_stream)
            return
        for r in range(self.world):
            with torch.cuda.device(self.dev[r]):
                self.x[r].checksum_u64(ptrs[r], ns[r], firsts[r], outs[r].data_ptr(), self.ws[r].data_ptr(),
                                       gpuprobe.VARIANT_AUTO, self.st[r].cuda_stream)

    def sync(self):
        for d in set(self.dev):
            torch.cuda.synchronize(d)

    def close(self):
        self.sync()
        for x in self.x:
            x.close()


def _shards(n: int, world: int, ragged: bool):
    cuts = [n * r // world for r in range(world + 1)]
    if ragged and world > 1 and n > 4 * world:
        cuts = [0] + [min(n, c + 3 * r + 1) for r, c in enumerate(cuts[1:-1])] + [n]
    return list(zip(cuts[:-1], cuts[1:]))


@pytest.mark.parametrize("world", [1, 2, 4, 8])
@pytest.mark.parametrize("n", [0, 5, 4096, (1 << 20) + 3, 148 * 4096 * 2 + 777])
def test_sharded_checksum_exchange_matches_oracle(world, n):
    rng = np.random.default_rng(1000 * world + n % 997)
    w = rng.integers(0, 1 << 64, size=n, dtype=np.uint64)
    first = (1 << 40) + 12345
    want = oracle.checksum_u64(w, first)
    R = Ranks(world)
    try:
        for rep, ragged in enumerate((False, True, True)):  # three calls in a row: both mailbox buffers get reused
            parts = _shards(n, world, ragged)
            ts = []
            for r, (lo, hi) in enumerate(parts):  # all allocation and H2D first: nothing but launches below
                with torch.cuda.device(R.dev[r]):
                    ts.append(from_np(w[lo:hi]))
                    R.out[r].fill_(-1)
            R.sync()
            R.checksum([t.data_ptr() for t in ts], [hi - lo for lo, hi in parts], [first + lo for lo, _ in parts], R.out)
            R.sync()
            for r in range(world):
                assert lanes_of(R.out[r]) == want, (world, n, rep, r)
                assert R.x[r].status() == 0
    finally:
        R.close()


@pytest.mark.parametrize("variant", [gpuprobe.VARIANT_LDG, gpuprobe.VARIANT_TMA, gpuprobe.VARIANT_LDG256])
@pytest.mark.parametrize("world", [2, 3])
def test_fused_triad_checksum_exchange_matches_oracle(world, variant):
    n, first = (1 << 19) + 11, 1 << 33
    want = oracle.stream_lanes(n, first)
    R = Ranks(world)
    try:
        parts, bufs = _shards(n, world, True), []
        for r, (lo, hi) in enumerate(parts):
            with torch.cuda.device(R.dev[r]):
                bufs.append((f64(hi - lo), f64(hi - lo), f64(hi - lo)))
        R.sync()
        for r, (lo, hi) in enumerate(parts):
            m, (a, b, c) = hi - lo, bufs[r]
            with torch.cuda.device(R.dev[r]):
                s = R.st[r].cuda_stream
                gpuprobe.init_fill(b.data_ptr(), m, first + lo, gpuprobe.SEED_B, s)
                gpuprobe.init_fill(c.data_ptr(), m, first + lo, gpuprobe.SEED_C, s)
                if R.grouped:  # shared GPU: triad per rank now, every rank's checksum + exchange in one launch below
                    gpuprobe.triad(a.data_ptr(), b.data_ptr(), c.data_ptr(), gpuprobe.SCALAR, m, variant, s)
                else:
                    R.x[r].triad_checksum(a.data_ptr(), b.data_ptr(), c.data_ptr(), gpuprobe.SCALAR, m, first + lo,
                                          R.out[r].data_ptr(), R.ws[r].data_ptr(), variant, s)
        if R.grouped:
            R.sync()
            R.checksum([bufs[r][0].data_ptr() for r in range(world)], [hi - lo for lo, hi in parts],
                       [first + lo for lo, _ in parts], R.out)
        R.sync()
        for r in range(world):
            assert lanes_of(R.out[r]) == want, (world, variant, r)
    finally:
        R.close()


def test_exchange_many_calls_in_a_row_stay_in_step():
    # 200 back-to-back collectives, no host sync in between: the two-buffer epoch scheme must hold
    world, n = 4, 70001
    rng = np.random.default_rng(5)
    w = rng.integers(0, 1 << 64, size=n, dtype=np.uint64)
    want = oracle.checksum_u64(w, 0)
    R = Ranks(world)
    try:
        parts = _shards(n, world, False)
        ts = []
        for r, (lo, hi) in enumerate(parts):
            with torch.cuda.device(R.dev[r]):
                ts.append(from_np(w[lo:hi]))
        outs = [[torch.zeros(2, dtype=torch.int64, device=f"cuda:{R.dev[r]}") for _ in range(200)] for r in range(world)]
        R.sync()
        for k in range(200):
            R.checksum([t.data_ptr() for t in ts], [hi - lo for lo, hi in parts], [lo for lo, _ in parts],
                       [outs[r][k] for r in range(world)])
        R.sync()
        for r in range(world):
            for k in range(200):
                assert lanes_of(outs[r][k]) == want, (r, k)
    finally:
        R.close()


def test_missing_peer_times_out_loudly():
    R = Ranks(2)
    try:
        R.x[0].set_timeout_ms(50)
        with torch.cuda.device(R.dev[0]):
            t = from_np(np.arange(1000, dtype=np.uint64))
            R.x[0].checksum_u64(t.data_ptr(), 1000, 0, R.out[0].data_ptr(), R.ws[0].data_ptr(),
                                gpuprobe.VARIANT_AUTO, R.st[0].cuda_stream)  # rank 1 never calls
        R.sync()
        assert lanes_of(R.out[0]) == ALL_ONES
        assert R.x[0].status() == gpuprobe.ETIMEDOUT
        assert "timed out" in gpuprobe.error_string(gpuprobe.ETIMEDOUT)
    finally:
        R.close()


def test_exchange_argument_and_state_errors():
    L = gpuprobe.lib()
    import ctypes
    h = ctypes.c_void_p()
    assert L.gpu_probe_exchange_create(ctypes.byref(h), 0, 0, 0) == gpuprobe.EINVAL
    assert L.gpu_probe_exchange_create(ctypes.byref(h), 2, 2, 0) == gpuprobe.EINVAL
    assert L.gpu_probe_exchange_create(ctypes.byref(h), 0, gpuprobe.EXCHANGE_MAX_RANKS + 1, 0) == gpuprobe.EINVAL
    x = gpuprobe.Exchange(0, 2)  # created but never connected
    try:
        t, ws, out = from_np(np.arange(8, dtype=np.uint64)), workspace(), lanes_buf()
        with pytest.raises(gpuprobe.ProbeError) as ei:
            x.checksum_u64(t.data_ptr(), 8, 0, out.data_ptr(), ws.data_ptr())
        assert ei.value.rc == gpuprobe.ENOTREADY
    finally:
        x.close()
    assert L.gpu_probe_exchange_destroy(None) == gpuprobe.ENOTREADY


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs: one process per GPU, segments passed as file descriptors")
def test_exchange_across_processes_over_ipc():
    world = min(torch.cuda.device_count(), 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29577", os.path.join(ROOT, "tests", "mp_exchange_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert f"EXCHANGE_IPC_OK world={world}" in r.stdout


def test_ipc_rendezvous_ignores_strangers():
    """Two processes, both on cuda:0 (only the rendezvous runs, so one GPU is enough): forged
    datagrams on a rank's socket -- no secret, a wrong secret, the peer's secret -- are dropped."""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", "29578", os.path.join(ROOT, "tests", "mp_connect_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "CONNECT_OK world=2" in r.stdout
