"""No kernel writes a byte outside the buffer it was given.

compute-sanitizer is closed on this GPU pool (profiles/sanitizer_r2.md), so out-of-bounds WRITES
are hunted the old way: every output buffer sits between two guard zones filled with a pattern,
at ragged sizes and 8-byte-only alignments that exercise each kernel's tail path and the TMA
rings' last partial tiles, and the zones must come back untouched while the payload equals the
CPU oracle's.  (Out-of-bounds reads of mapped memory cannot change a result that is compared bit
for bit with the oracle; reads of unmapped memory fault, which fails the test.)"""
import numpy as np
import pytest
import torch

import gpuprobe
import oracle
from gpu_util import lanes_buf, lanes_of, stream, workspace

pytestmark = pytest.mark.gpu

GUARD = 4096          # doubles on each side (32 KiB: more than any tile this library moves at once)
PATTERN = -7.25


class Guarded:
    """n doubles at an 8- or 16-byte-aligned address, between two pattern-filled zones."""

    def __init__(self, n, off):
        self.n, self.off = n, off
        self.base = torch.full((n + 2 * GUARD + off + 2,), PATTERN, dtype=torch.float64, device="cuda")
        self.lo = GUARD + off
        self.view = self.base[self.lo:self.lo + n]

    def ptr(self):
        return self.view.data_ptr()

    def intact(self):
        torch.cuda.synchronize()
        return bool((self.base[:self.lo] == PATTERN).all()) and bool((self.base[self.lo + self.n:] == PATTERN).all())


SIZES = [1, 2, 3, 255, 511, 513, 1023, 1025, 4095, 4097, 512 * 296 + 1, 1024 * 296 * 4 + 1023, 4096 * 148 * 3 + 4095,
         (1 << 20) + 77]


@pytest.mark.parametrize("off", [0, 1], ids=["aligned16", "aligned8"])
def test_writers_stay_inside_their_buffers(off):
    for n in SIZES:
        first = (1 << 40) + n
        hb, hc = oracle.init_fill(n, first, oracle.SEED_B), oracle.init_fill(n, first, oracle.SEED_C)
        ha = oracle.triad(hb, hc, 3.0)
        want = oracle.checksum_u64(ha.view(np.uint64), first)
        b, c = Guarded(n, off), Guarded(n, off)
        gpuprobe.init_fill(b.ptr(), n, first, gpuprobe.SEED_B, stream())
        gpuprobe.init_fill(c.ptr(), n, first, gpuprobe.SEED_C, stream())
        assert b.intact() and c.intact(), ("init_fill", n)
        assert np.array_equal(b.view.cpu().numpy(), hb) and np.array_equal(c.view.cpu().numpy(), hc)
        ws, out = workspace(), lanes_buf()
        for name, variant in gpuprobe.VARIANTS.items():
            a = Guarded(n, off)
            gpuprobe.triad(a.ptr(), b.ptr(), c.ptr(), gpuprobe.SCALAR, n, variant, stream())
            assert a.intact(), ("triad", name, n)
            assert np.array_equal(a.view.cpu().numpy(), ha), ("triad", name, n)
            a = Guarded(n, off)
            gpuprobe.triad_checksum(a.ptr(), b.ptr(), c.ptr(), gpuprobe.SCALAR, n, first, out.data_ptr(), ws.data_ptr(),
                                    variant, stream())
            assert a.intact(), ("fused", name, n)
            assert lanes_of(out) == want and np.array_equal(a.view.cpu().numpy(), ha), ("fused", name, n)
        assert b.intact() and c.intact()  # inputs are read-only for every variant


@pytest.mark.parametrize("nbytes", [1, 15, 16, 17, 4096 + 5, (8 << 20) + 16, (8 << 20) + 32768 * 3 + 21, (64 << 20) - 16])
def test_copy_stays_inside_its_destination(nbytes):
    n = (nbytes + 7) // 8
    src = torch.from_numpy(np.random.default_rng(nbytes).integers(0, 1 << 63, size=n, dtype=np.int64)).cuda()
    for off in (0, 1):
        dst = Guarded(n + 1, off)
        gpuprobe.p2p_copy(dst.ptr(), src.data_ptr(), nbytes, stream())
        assert dst.intact(), (nbytes, off)
        got = dst.view.cpu().numpy().view(np.uint8)
        assert np.array_equal(got[:nbytes], src.cpu().numpy().view(np.uint8)[:nbytes])
        # bytes past nbytes inside the last word and the spare word keep the pattern
        tail = dst.view.cpu().numpy().view(np.uint8)[nbytes:]
        want_tail = np.full(n + 1, PATTERN).view(np.uint8)[nbytes:]
        assert np.array_equal(tail, want_tail), (nbytes, off)


def test_repeated_launches_of_the_rings_are_deterministic():
    """A race in a shared-memory ring (a slot refilled while a thread still reads it, a phase bit
    waited on one turn early) shows up as a result that changes from launch to launch.  200
    launches per ring kernel at a size with ~7 tiles per CTA plus a ragged tail: one answer."""
    n, first = 512 * 296 * 7 + 333, 12345
    hb, hc = oracle.init_fill(n, first, oracle.SEED_B), oracle.init_fill(n, first, oracle.SEED_C)
    want = oracle.checksum_u64(oracle.triad(hb, hc, 3.0).view(np.uint64), first)
    b, c, a = Guarded(n, 0), Guarded(n, 0), Guarded(n, 0)
    gpuprobe.init_fill(b.ptr(), n, first, gpuprobe.SEED_B, stream())
    gpuprobe.init_fill(c.ptr(), n, first, gpuprobe.SEED_C, stream())
    ws = workspace()
    outs = torch.zeros((200, 2), dtype=torch.int64, device="cuda")
    for variant in (gpuprobe.VARIANT_TMA, gpuprobe.VARIANT_TMA_PACED):
        for k in range(200):
            if k % 2:
                gpuprobe.triad_checksum(a.ptr(), b.ptr(), c.ptr(), gpuprobe.SCALAR, n, first, outs[k].data_ptr(), ws.data_ptr(),
                                        variant, stream())
            else:
                gpuprobe.triad(a.ptr(), b.ptr(), c.ptr(), gpuprobe.SCALAR, n, variant, stream())
                gpuprobe.checksum_u64(a.ptr(), n, first, outs[k].data_ptr(), ws.data_ptr(), gpuprobe.VARIANT_TMA, stream())
        torch.cuda.synchronize()
        got = {tuple(int(x) for x in row) for row in outs.cpu().numpy().view(np.uint64)}
        assert got == {want}, variant
    assert a.intact() and b.intact() and c.intact()
