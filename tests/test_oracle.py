"""The oracle against its pins: the published SplitMix64 vector, the independent
pure-Python restatement (tests/pyspec.py), algebraic identities of the lanes, and
the committed golden fixtures.  CPU only."""
import numpy as np
import pytest

import oracle
import pyspec

M64 = (1 << 64) - 1


def test_splitmix64_published_vector():
    # gen(i, seed) must be the (i+1)-th output of the sequential generator
    assert pyspec.splitmix64_sequential(pyspec.SPLITMIX64_KAT_SEED, 5) == pyspec.SPLITMIX64_KAT
    for i, want in enumerate(pyspec.SPLITMIX64_KAT):
        assert oracle.gen(i, pyspec.SPLITMIX64_KAT_SEED) == want
        assert pyspec.gen(i, pyspec.SPLITMIX64_KAT_SEED) == want
    for i, want in enumerate(pyspec.SPLITMIX64_SEED0):
        assert oracle.gen(i, 0) == want


def test_gen_counter_wraps_mod_2_64():
    for i in (0, 1, (1 << 32) - 1, 1 << 32, (1 << 63) + 12345, M64):
        for seed in (0, oracle.SEED_B, oracle.SEED_C, M64):
            assert oracle.gen(i, seed) == pyspec.gen(i, seed)


def test_init_fill_matches_pyspec_and_range():
    for first in (0, 5, 1 << 40):
        got = oracle.init_fill(257, first, oracle.SEED_B)
        want = [pyspec.unit_double(pyspec.gen(first + k, pyspec.SEED_B)) for k in range(257)]
        assert got.tolist() == want
        assert (got >= 0).all() and (got < 1).all()
    assert oracle.init_fill(0, 0, 1).size == 0


def test_triad_is_one_rounding():
    b = oracle.init_fill(4096, 0, oracle.SEED_B)
    c = oracle.init_fill(4096, 0, oracle.SEED_C)
    fused = oracle.triad(b, c)
    want = np.array([pyspec.fma_exact(3.0, ci, bi) for bi, ci in zip(b.tolist(), c.tolist())])
    assert fused.view(np.uint64).tolist() == want.view(np.uint64).tolist()
    # FMA-discriminating KAT: on these inputs a two-rounding triad differs, so a
    # silently unfused (or contracted elsewhere) build cannot pass the lanes check.
    unfused = oracle.triad_unfused(b, c)
    ndiff = int((fused.view(np.uint64) != unfused.view(np.uint64)).sum())
    assert ndiff > 100, ndiff
    assert oracle.checksum_u64(fused) != oracle.checksum_u64(unfused)


def test_checksum_identities():
    assert oracle.checksum_u64(np.empty(0, dtype=np.uint64)) == (0, 0)
    for w, i in ((1, 0), (M64, 3), (0xDEADBEEFCAFEF00D, (1 << 63) + 9), (5, M64)):
        assert oracle.checksum_u64(np.array([w], dtype=np.uint64), i) == (w, (w * (2 * i + 1)) & M64)
    rng = np.random.default_rng(7)
    w = rng.integers(0, 1 << 64, size=10007, dtype=np.uint64)
    assert oracle.checksum_u64(w, 11) == pyspec.lanes_of_words([int(x) for x in w], 11)


@pytest.mark.parametrize("cuts", [[0, 5000, 10007], [0, 2501, 5003, 7504, 10007], [0, 1, 2, 10006, 10007],
                                  list(range(0, 10007, 1251)) + [10007]])
def test_checksum_shard_additivity(cuts):
    rng = np.random.default_rng(11)
    w = rng.integers(0, 1 << 64, size=10007, dtype=np.uint64)
    first = 1 << 33
    whole = oracle.checksum_u64(w, first)
    l0 = l1 = 0
    for lo, hi in zip(cuts[:-1], cuts[1:]):
        s0, s1 = oracle.checksum_u64(np.ascontiguousarray(w[lo:hi]), first + lo)
        l0, l1 = (l0 + s0) & M64, (l1 + s1) & M64
    assert (l0, l1) == whole


def test_stream_mode_equals_array_mode_and_pyspec():
    for n, first in ((0, 0), (1, 0), (33, 9), (5000, 1 << 35)):
        b = oracle.init_fill(n, first, oracle.SEED_B)
        c = oracle.init_fill(n, first, oracle.SEED_C)
        arr = oracle.checksum_u64(oracle.triad(b, c), first)
        assert oracle.stream_lanes(n, first) == arr
        if n <= 33:
            assert pyspec.stream_lanes(n, first) == arr


def test_thread_count_does_not_change_lanes():
    n = (1 << 20) + 3
    oracle.set_threads(1)
    one = oracle.stream_lanes(n, 17)
    oracle.set_threads(oracle.max_threads() if oracle.max_threads() > 1 else 4)
    many = oracle.stream_lanes(n, 17)
    assert one == many


def test_golden_head_and_small_lanes(golden):
    b = oracle.init_fill(8, 0, oracle.SEED_B)
    c = oracle.init_fill(8, 0, oracle.SEED_C)
    a = oracle.triad(b, c)
    for name, arr in (("b", b), ("c", c), ("a", a)):
        assert [f"0x{int(x):016x}" for x in arr.view(np.uint64)] == golden["head"][name]
    # pure-Python restatement agrees with the committed fixtures where it is cheap
    for n in ("1", "2", "32", "1024"):
        l = pyspec.stream_lanes(int(n))
        assert [f"0x{l[0]:016x}", f"0x{l[1]:016x}"] == golden["stream_lanes"][n]
    for case in golden["ragged"]:
        got = oracle.stream_lanes(case["n"], int(case["first_index"], 16))
        assert [f"0x{got[0]:016x}", f"0x{got[1]:016x}"] == case["lanes"]
        if case["n"] <= 4097:
            py = pyspec.stream_lanes(case["n"], int(case["first_index"], 16))
            assert py == got


def test_golden_full_size_shards_sum_to_whole(golden):
    # the 8 committed shard lanes of the 2^29 buffer add up to the committed whole
    l0 = sum(int(s[0], 16) for s in golden["shards8"]) & M64
    l1 = sum(int(s[1], 16) for s in golden["shards8"]) & M64
    assert [f"0x{l0:016x}", f"0x{l1:016x}"] == golden["stream_lanes"][str(1 << 29)]
    # and the oracle reproduces a mid-size committed entry live
    got = oracle.stream_lanes(1 << 24)
    assert [f"0x{got[0]:016x}", f"0x{got[1]:016x}"] == golden["stream_lanes"][str(1 << 24)]
