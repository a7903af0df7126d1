"""gpu_probe_checksum_u64_noalloc -- the checksum for hosts that cannot allocate.

Some users of this library have no way to cudaMalloc the 65,792-byte workspace the
ordinary checksum asks for: their process handed every byte of HBM to another
allocator at start-up, or policy forbids allocator calls once the workload runs.
For them the library keeps that scratch inside its own device image.  This test
calls the new entry point through the C ABI and requires the same lanes, bit for
bit (the tolerance of every parity test here is zero), as the existing
gpu_probe_checksum_u64 -- twice in a row, because the built-in scratch is shared by
all calls and still holds the previous call's ticket and partial sums.

The file is named so pytest collects it last: a run with -x has then already
executed every other test.  It passes alone in a fresh process as well (the first
call performs the one-time address look-up for the device).
"""
import numpy as np
import pytest
import torch

import gpuprobe
import oracle
from gpu_util import from_np, lanes_buf, lanes_of, stream, workspace

pytestmark = pytest.mark.gpu


def _existing(d, n, first):
    out, ws = lanes_buf(), workspace()
    gpuprobe.checksum_u64(d.data_ptr(), n, first, out.data_ptr(), ws.data_ptr(), gpuprobe.VARIANT_AUTO, stream())
    torch.cuda.synchronize()
    return lanes_of(out)


def _noalloc(d, n, first):
    out = lanes_buf()
    gpuprobe.checksum_u64_noalloc(d.data_ptr(), n, first, out.data_ptr(), stream())
    torch.cuda.synchronize()
    return lanes_of(out)


def test_noalloc_matches_existing_entry_point_twice_in_a_row():
    rng = np.random.default_rng(2026)
    # a large ragged buffer (every CTA of the grid leaves a partial behind) ...
    n_big, first_big = 148 * 4096 * 5 + 4097, (1 << 41) + 3
    w_big = rng.integers(0, 1 << 64, size=n_big, dtype=np.uint64)
    d_big = from_np(w_big)
    # ... then a small one that uses far fewer CTAs, so stale partials from the first
    # call are still lying in the shared scratch when the second call runs
    n_small, first_small = 4096 + 5, 7
    w_small = rng.integers(0, 1 << 64, size=n_small, dtype=np.uint64)
    d_small = from_np(w_small)

    want_big, want_small = _existing(d_big, n_big, first_big), _existing(d_small, n_small, first_small)
    assert want_big == oracle.checksum_u64(w_big, first_big)        # the existing path is itself right
    assert want_small == oracle.checksum_u64(w_small, first_small)

    assert _noalloc(d_big, n_big, first_big) == want_big            # call 1
    assert _noalloc(d_small, n_small, first_small) == want_small    # call 2, right behind it
    assert _noalloc(d_big, n_big, first_big) == want_big            # and back again
    # back-to-back on one stream without a host sync in between
    outs = [lanes_buf() for _ in range(4)]
    for i, out in enumerate(outs):
        d, n, f = (d_big, n_big, first_big) if i % 2 == 0 else (d_small, n_small, first_small)
        gpuprobe.checksum_u64_noalloc(d.data_ptr(), n, f, out.data_ptr(), stream())
    torch.cuda.synchronize()
    assert [lanes_of(o) for o in outs] == [want_big, want_small, want_big, want_small]


def test_noalloc_edge_sizes_and_limit():
    out = lanes_buf()
    gpuprobe.checksum_u64_noalloc(0, 0, 5, out.data_ptr(), stream())   # empty input: lanes are (0, 0)
    torch.cuda.synchronize()
    assert lanes_of(out) == (0, 0)
    w = np.arange(3, dtype=np.uint64) + np.uint64(1 << 63)
    d = from_np(w, 1)                                                   # 8-byte-aligned only
    assert _noalloc(d, 3, (1 << 64) - 2) == oracle.checksum_u64(w, (1 << 64) - 2)
    # beyond the stated maximum nothing is launched and the error says why
    rc = gpuprobe.lib().gpu_probe_checksum_u64_noalloc(d.data_ptr(), gpuprobe.NOALLOC_MAX_WORDS + 1, 0,
                                                       out.data_ptr(), None)
    assert rc == gpuprobe.ETOOBIG
    assert "maximum" in gpuprobe.error_string(rc)
    assert gpuprobe.lib().gpu_probe_checksum_u64_noalloc(d.data_ptr(), 3, 0, None, None) == gpuprobe.EINVAL


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs: the scratch address is looked up per device")
def test_noalloc_on_a_second_device_and_back():
    """Every GPU holds its own copy of the built-in scratch, at an address the library looks up the
    first time the entry point runs with that GPU current.  Device 0, then device 1, then device 0
    again -- each against the CPU oracle, so an address cached for the wrong device (the kernel
    would read another GPU's ticket word, or fault) cannot pass."""
    rng = np.random.default_rng(7)
    n, first = 148 * 4096 * 2 + 33, (1 << 35) + 1
    w = rng.integers(0, 1 << 64, size=n, dtype=np.uint64)
    want = oracle.checksum_u64(w, first)
    try:
        for dev in (0, 1, 0, 1):
            with torch.cuda.device(dev):
                d = from_np(w)
                out = lanes_buf()
                gpuprobe.checksum_u64_noalloc(d.data_ptr(), n, first, out.data_ptr(), stream())
                gpuprobe.checksum_u64_noalloc(d.data_ptr(), n, first, out.data_ptr(), stream())  # twice in a row here too
                torch.cuda.synchronize()
                assert lanes_of(out) == want, dev
    finally:
        torch.cuda.set_device(0)
