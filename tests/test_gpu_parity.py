"""Parity of the sm_100a kernels against the CPU oracle, through the C ABI.

Bit-exact everywhere: the triad is one correctly rounded FMA on both sides and the
checksum is wrapping u64 arithmetic, so there is no tolerance in this file.
Small/ragged sizes compare element by element with the oracle on the same seeds;
BASELINE-size buffers (2^29 elements = 4 GiB per array) are checked against the
committed golden lanes and through size-independent properties (shard additivity,
fused == separate, all variants agree, a flipped bit is caught)."""
import numpy as np
import pytest
import torch

import gpuprobe
import oracle
from gpu_util import (M64, f64, from_np, gpu_checksum, gpu_init, hx, lanes_buf, lanes_of, stream, to_u64, workspace)

pytestmark = pytest.mark.gpu

VARIANTS = [gpuprobe.VARIANT_LDG, gpuprobe.VARIANT_TMA, gpuprobe.VARIANT_LDG256, gpuprobe.VARIANT_TMA_PACED,
            gpuprobe.VARIANT_AUTO]
VNAMES = ["ldg128", "tma", "ldg256", "tma_paced", "auto"]
# ragged sizes around every tiling boundary: vector width, CTA tile (256*U), TMA tiles (1024 / 4096)
SIZES = [0, 1, 2, 3, 4, 5, 255, 256, 257, 1023, 1024, 1025, 2047, 2048, 2049, 4095, 4096, 4096 + 3, 8192 + 1, 65536,
         (1 << 20) + 3, 148 * 4096 * 3 + 4095,
         3 * 1184 * 1024 + 517]
FIRSTS = [0, 7, 1 << 40, (1 << 63) - 3, M64 - 1]


def test_kernel_info_says_no_fallback():
    assert "cpu_fallback=none" in gpuprobe.kernel_info()


@pytest.mark.parametrize("n", SIZES)
def test_init_fill_matches_oracle(n):
    for first, off in ((0, 0), (12345, 1), (M64 - 5, 0)):
        got = gpu_init(n, first, oracle.SEED_B, off)
        torch.cuda.synchronize()
        want = oracle.init_fill(n, first, oracle.SEED_B)
        assert np.array_equal(to_u64(got), want.view(np.uint64)), (n, first, off)


@pytest.mark.parametrize("variant", VARIANTS, ids=VNAMES)
@pytest.mark.parametrize("n", SIZES)
def test_triad_matches_oracle(n, variant):
    b = oracle.init_fill(n, 3, oracle.SEED_B)
    c = oracle.init_fill(n, 3, oracle.SEED_C)
    want = oracle.triad(b, c, 3.0)
    for off in (0, 1):  # 16-byte-aligned (vector/TMA kernels) and 8-byte-only (scalar kernel)
        db, dc, da = from_np(b, off), from_np(c, off), f64(n, off)
        da.fill_(-1.0)
        gpuprobe.triad(da.data_ptr(), db.data_ptr(), dc.data_ptr(), 3.0, n, variant, stream())
        torch.cuda.synchronize()
        assert np.array_equal(to_u64(da), want.view(np.uint64)), (n, variant, off)


def test_triad_rounds_once():
    # FMA-discriminating KAT: the unfused result differs on these inputs
    n = 4096
    b = oracle.init_fill(n, 0, oracle.SEED_B)
    c = oracle.init_fill(n, 0, oracle.SEED_C)
    unfused = oracle.triad_unfused(b, c, 3.0)
    for variant in VARIANTS:
        da, db, dc = f64(n), from_np(b), from_np(c)  # keep db/dc alive until the kernel has run
        gpuprobe.triad(da.data_ptr(), db.data_ptr(), dc.data_ptr(), 3.0, n, variant, stream())
        torch.cuda.synchronize()
        got = to_u64(da)
        assert np.array_equal(got, oracle.triad(b, c, 3.0).view(np.uint64))
        assert (got != unfused.view(np.uint64)).sum() > 100


@pytest.mark.parametrize("variant", VARIANTS, ids=VNAMES)
@pytest.mark.parametrize("n", SIZES)
def test_checksum_matches_oracle(n, variant):
    rng = np.random.default_rng(n + 17)
    w = rng.integers(0, 1 << 64, size=n, dtype=np.uint64)
    ws = workspace()
    for first in FIRSTS:
        for off in (0, 1):
            got = gpu_checksum(from_np(w, off), first, variant, ws)  # same workspace reused: ticket must reset
            assert got == oracle.checksum_u64(w, first), (n, variant, first, off)
    assert int(ws[0].item()) == 0


def test_checksum_extreme_words():
    for pattern in (0, M64, 1 << 63):
        w = np.full(100003, pattern, dtype=np.uint64)
        for variant in VARIANTS:
            assert gpu_checksum(from_np(w), 5, variant) == oracle.checksum_u64(w, 5)


@pytest.mark.parametrize("variant", VARIANTS, ids=VNAMES)
@pytest.mark.parametrize("n", SIZES)
def test_fused_triad_checksum_equals_separate(n, variant):
    first = (1 << 35) + 11
    b = oracle.init_fill(n, first, oracle.SEED_B)
    c = oracle.init_fill(n, first, oracle.SEED_C)
    want_a = oracle.triad(b, c, 3.0)
    want_l = oracle.checksum_u64(want_a, first)
    for off in (0, 1):
        db, dc, da = from_np(b, off), from_np(c, off), f64(n, off)
        out, ws = lanes_buf(), workspace()
        gpuprobe.triad_checksum(da.data_ptr(), db.data_ptr(), dc.data_ptr(), 3.0, n, first, out.data_ptr(),
                                ws.data_ptr(), variant, stream())
        torch.cuda.synchronize()
        assert np.array_equal(to_u64(da), want_a.view(np.uint64))
        assert lanes_of(out) == want_l
        assert lanes_of(out) == oracle.stream_lanes(n, first)


def _gpu_stream_lanes(n, first, variant, fused=False):
    db = gpu_init(n, first, oracle.SEED_B)
    dc = gpu_init(n, first, oracle.SEED_C)
    da = f64(n)
    out, ws = lanes_buf(), workspace()
    if fused:
        gpuprobe.triad_checksum(da.data_ptr(), db.data_ptr(), dc.data_ptr(), 3.0, n, first, out.data_ptr(),
                                ws.data_ptr(), variant, stream())
    else:
        gpuprobe.triad(da.data_ptr(), db.data_ptr(), dc.data_ptr(), 3.0, n, variant, stream())
        gpuprobe.checksum_u64(da.data_ptr(), n, first, out.data_ptr(), ws.data_ptr(), variant, stream())
    torch.cuda.synchronize()
    return lanes_of(out), da


def test_golden_lanes_small_and_ragged(golden):
    for n_s, want in golden["stream_lanes"].items():
        n = int(n_s)
        if n > (1 << 26):
            continue
        for variant in VARIANTS:
            got, _ = _gpu_stream_lanes(n, 0, variant)
            assert hx(got) == want, (n, variant)
    for case in golden["ragged"]:
        for variant in VARIANTS:
            got, _ = _gpu_stream_lanes(case["n"], int(case["first_index"], 16), variant, fused=True)
            assert hx(got) == case["lanes"], (case, variant)


def test_full_size_golden_and_properties(golden):
    """BASELINE config 2: 2^29 fp64 elements per array (4 GiB each)."""
    n = 1 << 29
    want = golden["stream_lanes"][str(n)]
    db = gpu_init(n, 0, oracle.SEED_B)
    dc = gpu_init(n, 0, oracle.SEED_C)
    da = f64(n)
    ws = workspace()
    results = {}
    for variant, name in zip(VARIANTS, VNAMES):
        out = lanes_buf()
        da.zero_()
        gpuprobe.triad(da.data_ptr(), db.data_ptr(), dc.data_ptr(), 3.0, n, variant, stream())
        gpuprobe.checksum_u64(da.data_ptr(), n, 0, out.data_ptr(), ws.data_ptr(), variant, stream())
        torch.cuda.synchronize()
        results[name] = lanes_of(out)
        assert hx(results[name]) == want, name
        # fused pass writes the same a[] and the same lanes
        out2 = lanes_buf()
        da2 = f64(n)
        gpuprobe.triad_checksum(da2.data_ptr(), db.data_ptr(), dc.data_ptr(), 3.0, n, 0, out2.data_ptr(),
                                ws.data_ptr(), variant, stream())
        torch.cuda.synchronize()
        assert hx(lanes_of(out2)) == want, name
        assert torch.equal(da.view(torch.int64), da2.view(torch.int64)), name
        del da2
    # shard additivity at 2/4/8 shards and one ragged split: sum of shard lanes == whole
    for cuts in ([0, n // 2, n], [0, n // 4, n // 2, 3 * n // 4, n], [i * (n // 8) for i in range(9)],
                 [0, 1, 12345677, n // 3, n - 1, n]):
        l0 = l1 = 0
        for lo, hi in zip(cuts[:-1], cuts[1:]):
            s = gpu_checksum(da[lo:hi], lo, gpuprobe.VARIANT_AUTO, ws)
            l0, l1 = (l0 + s[0]) & M64, (l1 + s[1]) & M64
        assert hx((l0, l1)) == want, cuts
    # the 8 committed shard lanes
    for r in range(8):
        s = gpu_checksum(da[r * (n // 8):(r + 1) * (n // 8)], r * (n // 8), gpuprobe.VARIANT_TMA, ws)
        assert hx(s) == golden["shards8"][r]
    # fault injection: one flipped bit anywhere must change the lanes
    idx = 400_000_001
    v = da.view(torch.int64)
    v[idx] ^= 1
    bad = gpu_checksum(da, 0, gpuprobe.VARIANT_AUTO, ws)
    assert hx(bad) != want
    v[idx] ^= 1
    assert hx(gpu_checksum(da, 0, gpuprobe.VARIANT_AUTO, ws)) == want


def test_checksum_is_idempotent_and_stream_ordered():
    n = (1 << 22) + 5
    w = np.random.default_rng(3).integers(0, 1 << 64, size=n, dtype=np.uint64)
    d = from_np(w)
    ws = workspace()
    want = oracle.checksum_u64(w, 9)
    outs = [lanes_buf() for _ in range(8)]
    for i, out in enumerate(outs):  # back-to-back launches sharing one workspace on one stream
        gpuprobe.checksum_u64(d.data_ptr(), n, 9, out.data_ptr(), ws.data_ptr(), VARIANTS[i % 3], stream())
    torch.cuda.synchronize()
    for out in outs:
        assert lanes_of(out) == want


# below 8 MiB: LDG/STG kernel; from 8 MiB with 16-byte-aligned pointers: TMA ring (32 KiB tiles) + plain tail
@pytest.mark.parametrize("nbytes", [0, 1, 15, 16, 17, 4095, 4096, (1 << 20) + 7, (8 << 20) - 16, 8 << 20, (8 << 20) + 16,
                                    (64 << 20) + 24, 148 * 3 * 32768 * 2 + 32768 + 16 + 5])
def test_copy_kernel_local(nbytes):
    src = torch.randint(0, 256, (nbytes + 64,), dtype=torch.uint8, device="cuda")
    for so, do in ((0, 0), (16, 32), (1, 0), (0, 3)):
        dst = torch.zeros(nbytes + 64, dtype=torch.uint8, device="cuda")
        gpuprobe.p2p_copy(dst.data_ptr() + do, src.data_ptr() + so, nbytes, stream())
        torch.cuda.synchronize()
        assert torch.equal(dst[do:do + nbytes], src[so:so + nbytes])
        assert int(dst[:do].sum()) == 0 and int(dst[do + nbytes:].sum()) == 0


@pytest.mark.parametrize("n,chunk,depth,want_a", [(0, 1 << 16, 2, True), (5, 1 << 16, 2, True),
                                                   (3 * (1 << 16) + 77, 1 << 16, 3, True),
                                                   ((1 << 22) + 1, 1 << 18, 4, False), (1 << 24, 0, 0, True)])
def test_host_pipeline_matches_oracle(n, chunk, depth, want_a):
    first = 1 << 33
    b = torch.from_numpy(oracle.init_fill(n, first, oracle.SEED_B)).pin_memory()
    c = torch.from_numpy(oracle.init_fill(n, first, oracle.SEED_C)).pin_memory()
    a = torch.empty(n, dtype=torch.float64).pin_memory() if want_a else None
    with gpuprobe.Pipeline(chunk, depth) as p:
        for _ in range(2):  # second run reuses every staging slot
            lanes = p.run_host(b.data_ptr(), c.data_ptr(), a.data_ptr() if want_a else None, 3.0, n, first)
            assert lanes == oracle.stream_lanes(n, first)
            if want_a and n:
                want = oracle.triad(b.numpy(), c.numpy(), 3.0)
                assert np.array_equal(a.numpy().view(np.uint64), want.view(np.uint64))


def test_beyond_2_32_elements_matches_golden(golden):
    """Maximum size: 2^32 elements per array (3 x 32 GiB), so every tile, vector and word index in
    the kernels crosses 32 bits.  Known answer: the golden lanes of 2^32 elements."""
    n = 1 << 32
    torch.cuda.empty_cache()
    free, _ = torch.cuda.mem_get_info()
    if free < 3 * 8 * n + (4 << 30):
        pytest.skip(f"needs {3 * 8 * n >> 30} GiB of free HBM, have {free >> 30}")
    want = tuple(int(v, 16) for v in golden["pow2_lanes"][32])
    b = torch.empty(n, dtype=torch.float64, device="cuda")
    c = torch.empty(n, dtype=torch.float64, device="cuda")
    a = torch.empty(n, dtype=torch.float64, device="cuda")
    ws, out = workspace(), lanes_buf()
    gpuprobe.init_fill(b.data_ptr(), n, 0, gpuprobe.SEED_B, stream())
    gpuprobe.init_fill(c.data_ptr(), n, 0, gpuprobe.SEED_C, stream())
    gpuprobe.triad_checksum(a.data_ptr(), b.data_ptr(), c.data_ptr(), gpuprobe.SCALAR, n, 0, out.data_ptr(),
                            ws.data_ptr(), gpuprobe.VARIANT_TMA, stream())
    torch.cuda.synchronize()
    assert lanes_of(out) == want, hx(lanes_of(out))
    for variant in (gpuprobe.VARIANT_LDG, gpuprobe.VARIANT_LDG256, gpuprobe.VARIANT_TMA, gpuprobe.VARIANT_TMA_PACED,
                    gpuprobe.VARIANT_AUTO):
        a.zero_()
        gpuprobe.triad(a.data_ptr(), b.data_ptr(), c.data_ptr(), gpuprobe.SCALAR, n, variant, stream())
        assert gpu_checksum(a, 0, variant, ws) == want, variant
    # the upper half alone, addressed by its global index: lanes(whole) - lanes(lower half) mod 2^64
    lower = tuple(int(v, 16) for v in golden["pow2_lanes"][31])
    half = n // 2
    got = gpu_checksum(a[half:], half, gpuprobe.VARIANT_AUTO, ws)
    assert got == ((want[0] - lower[0]) & M64, (want[1] - lower[1]) & M64)
    del a, b, c
    torch.cuda.empty_cache()


def test_auto_triad_is_bit_exact_before_and_after_an_explicit_calibration(golden):
    """AUTO never calibrates by itself: an uncalibrated device runs the ring (pace 0).  After
    gpu_probe_triad_calibrate on the caller's buffers AUTO at >= 2^26 elements may run the paced
    kernel: whatever the verdict, the bits are the same."""
    n = 1 << 27
    want = tuple(int(v, 16) for v in golden["pow2_lanes"][27])
    b, c, a = gpu_init(n, 0, oracle.SEED_B), gpu_init(n, 0, oracle.SEED_C), f64(n)
    ws = workspace()
    gpuprobe.triad_recalibrate()
    assert gpuprobe.triad_pace_ns() == 0
    a.fill_(-1.0)
    gpuprobe.triad(a.data_ptr(), b.data_ptr(), c.data_ptr(), gpuprobe.SCALAR, n, gpuprobe.VARIANT_AUTO, stream())
    assert gpu_checksum(a, 0, gpuprobe.VARIANT_AUTO, ws) == want
    assert gpuprobe.triad_pace_ns() == 0  # launching did not calibrate
    verdict = gpuprobe.triad_calibrate(a.data_ptr(), b.data_ptr(), c.data_ptr(), n, stream())
    assert verdict == -1 or 1000 <= verdict <= 1100, verdict
    for rep in range(4):  # the probe's own step: triad, then the checksum of what it wrote
        gpuprobe.triad(a.data_ptr(), b.data_ptr(), c.data_ptr(), gpuprobe.SCALAR, n, gpuprobe.VARIANT_AUTO, stream())
        assert gpu_checksum(a, 0, gpuprobe.VARIANT_AUTO, ws) == want, rep
    assert gpuprobe.triad_pace_ns() == verdict  # behaved launches do not trip the drift guard
    if verdict > 0:
        ns, over = gpuprobe.triad_pace_report()
        ideal = -(-(n // 1024) // (2 * torch.cuda.get_device_properties(0).multi_processor_count)) * verdict
        assert over == 0 and 0.98 * ideal <= ns <= 1.04 * ideal + 10000, (ns, ideal)
    for rep in range(4):
        # a[] refilled right before: the triad starts into an L2 full of dirty lines, runs 15-22 % over what
        # its pace allows (tools/pace_guard_probe.py) and the guard may send AUTO back to the ring -- same bits
        a.fill_(-1.0)
        gpuprobe.triad(a.data_ptr(), b.data_ptr(), c.data_ptr(), gpuprobe.SCALAR, n, gpuprobe.VARIANT_AUTO, stream())
        assert gpu_checksum(a, 0, gpuprobe.VARIANT_AUTO, ws) == want, rep
    assert gpuprobe.triad_pace_ns() in (verdict, -2)
    out = lanes_buf()
    gpuprobe.triad_checksum(a.data_ptr(), b.data_ptr(), c.data_ptr(), gpuprobe.SCALAR, n, 0, out.data_ptr(), ws.data_ptr(),
                            gpuprobe.VARIANT_TMA_PACED, stream())
    torch.cuda.synchronize()
    assert lanes_of(out) == want
    gpuprobe.triad_recalibrate()


def test_calibrate_rejects_overlapping_buffers_and_ignores_small_problems():
    n = 1 << 26
    buf = f64(2 * n)
    L = gpuprobe.lib()
    gpuprobe.triad_recalibrate()
    # a[] aliases b[] (in place) or overlaps c[] by one element: the launches could not be replayed
    assert L.gpu_probe_triad_calibrate(buf.data_ptr(), buf.data_ptr(), buf.data_ptr() + 8 * n, n, stream()) == gpuprobe.EINVAL
    assert L.gpu_probe_triad_calibrate(buf.data_ptr() + 8 * (n - 1), buf.data_ptr() + 8 * n, buf.data_ptr(), n - 2, stream()) == gpuprobe.EINVAL
    # below 2^26 elements, or 8-byte-aligned only: nothing to calibrate, nothing changes
    small = f64(3 * 1024)
    assert L.gpu_probe_triad_calibrate(small.data_ptr(), small.data_ptr() + 8192, small.data_ptr() + 16384, 1024, stream()) == 0
    assert gpuprobe.triad_pace_ns() == 0
    del buf
    torch.cuda.empty_cache()
