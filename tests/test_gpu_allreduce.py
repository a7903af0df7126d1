"""K4, the probe's own fp32 allreduce over peer memory (gpu_probe_allreduce_f32), through
the C ABI.  The result is defined exactly: every element is the fp32 sum of the ranks'
values taken in rank order 0..G-1, and all ranks hold identical bits -- so the comparison
with the numpy restatement below is bit-exact, no tolerance.

With one GPU per rank the tests launch one kernel per rank and cross NVLink.  With fewer
GPUs than ranks all ranks live on cuda:0 and run as the rows of ONE cooperative launch
(gpu_probe_allreduce_f32_group): kernels that wait for each other are never separate
launches on one GPU."""
import numpy as np
import pytest
import torch

import gpuprobe

pytestmark = pytest.mark.gpu


def _devices(world):
    return list(range(world)) if torch.cuda.device_count() >= world else [0] * world


class Group:
    def __init__(self, world, buffer_bytes):
        self.world, self.dev = world, _devices(world)
        self.shared = world > 1 and len(set(self.dev)) == 1
        self.x, self.st = [], []
        for r in range(world):
            with torch.cuda.device(self.dev[r]):
                self.x.append(gpuprobe.Exchange(r, world, buffer_bytes))
                self.st.append(torch.cuda.Stream())
        gpuprobe.Exchange.connect_local(self.x)

    @property
    def ctas(self):
        return 8 if self.shared else 0  # shared GPU: world x ctas CTAs must be resident together

    def put(self, r, arr, offset=0):
        with torch.cuda.device(self.dev[r]):
            t = torch.from_numpy(arr).cuda()
            gpuprobe.p2p_copy(self.x[r].buffer + offset, t.data_ptr(), arr.nbytes, 0)
            torch.cuda.synchronize()

    def get(self, r, count, offset=0):
        with torch.cuda.device(self.dev[r]):
            t = torch.empty(count, dtype=torch.float32, device="cuda")
            gpuprobe.p2p_copy(t.data_ptr(), self.x[r].buffer + offset, 4 * count, 0)
            torch.cuda.synchronize()
            return t.cpu().numpy()

    def allreduce(self, count, offset=0, ctas=None, algo=0):
        if self.shared:  # all ranks in one cooperative launch; few CTAs per rank so that all are resident
            with torch.cuda.device(0):
                gpuprobe.Exchange.allreduce_f32_group(self.x, offset, count, self.ctas if ctas is None else ctas,
                                                      self.st[0].cuda_stream, algo)
            return
        for r in range(self.world):
            with torch.cuda.device(self.dev[r]):
                self.x[r].allreduce_f32(offset, count, self.ctas if ctas is None else ctas, self.st[r].cuda_stream, algo)

    def sync(self):
        for d in set(self.dev):
            torch.cuda.synchronize(d)

    def close(self):
        self.sync()
        for x in self.x:
            x.close()


def rank_order_sum(parts):
    s = parts[0].copy()
    for p in parts[1:]:
        s = (s + p).astype(np.float32)
    return s


ALGOS = {"auto": gpuprobe.ALLREDUCE_AUTO, "pull_push": gpuprobe.ALLREDUCE_PULL_PUSH, "push": gpuprobe.ALLREDUCE_PUSH,
         "multimem": gpuprobe.ALLREDUCE_MULTIMEM}


def assert_sum(got, parts, want, exact, ctx):
    """exact: bit-identical to the rank-order fp32 sum.  Otherwise (NVLS: the switch chooses the
    order) within the worst-case fp32 reordering error of a G-term sum."""
    if exact:
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), ctx
        return
    ref = np.sum(np.stack(parts).astype(np.float64), axis=0)
    bound = len(parts) * np.finfo(np.float32).eps * np.sum(np.abs(np.stack(parts)).astype(np.float64), axis=0)
    assert np.all(np.abs(got.astype(np.float64) - ref) <= bound + 1e-30), ctx


@pytest.mark.parametrize("algo", list(ALGOS))
@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("count", [1, 3, 4, 5, 1023, 4096, 65537, 1 << 19, (1 << 20) + 2])
def test_allreduce_is_the_rank_order_fp32_sum_on_every_rank(world, count, algo):
    if algo == "push" and (count % 4 or 4 * count > (2 << 20)):
        pytest.skip("the one-shot algorithm takes whole 16-byte vectors up to 2 MiB")
    if algo == "multimem" and torch.cuda.device_count() < max(world, 2):
        pytest.skip("NVSwitch multicast needs one GPU per rank")
    if algo not in ("pull_push", "multimem") and count >= (1 << 17) and world > 2 and torch.cuda.device_count() < world:
        pytest.skip("the one-shot grid must cover the operand; that many CTAs per rank only fit with one GPU per rank")
    rng = np.random.default_rng(100 * world + count % 89)
    parts = [rng.standard_normal(count).astype(np.float32) * np.float32(10.0 ** rng.integers(-3, 4)) for _ in range(world)]
    want = rank_order_sum(parts)
    G = Group(world, 4 * count + 64)
    try:
        if algo == "multimem" and not G.x[0].has_multicast:
            pytest.skip("this box does not offer NVSwitch multicast")
        # the switch's summation order is its own: NVLS results are checked within the reordering bound
        exact = not (G.x[0].has_multicast and (algo == "multimem" or (algo == "auto" and world >= 3)))
        for rep in range(3):  # repeated calls: the one-shot staging is double-buffered by call parity
            for r in range(world):
                G.put(r, parts[r])
            G.allreduce(count, algo=ALGOS[algo])
            G.sync()
            first = G.get(0, count)
            for r in range(world):
                got = G.get(r, count)
                assert_sum(got, parts, want, exact, (world, count, algo, rep, r))
                assert np.array_equal(got.view(np.uint32), first.view(np.uint32)), "ranks must agree bit for bit"
        for r in range(world):
            assert G.x[r].status() == 0
    finally:
        G.close()


def test_allreduce_at_an_offset_leaves_the_rest_untouched_and_repeats():
    world, count, off = 4, 50000, 4096
    rng = np.random.default_rng(7)
    G = Group(world, off + 4 * count + 4096)
    try:
        total = (off + 4 * count + 4096) // 4
        base = [rng.standard_normal(total).astype(np.float32) for _ in range(world)]
        for r in range(world):
            G.put(r, base[r])
        cur = [b[off // 4: off // 4 + count].copy() for b in base]
        for rep in range(5):  # back-to-back collectives on the same data: values grow, epochs advance
            G.allreduce(count, off)
            s = rank_order_sum(cur)
            cur = [s.copy() for _ in range(world)]
        G.sync()
        for r in range(world):
            got = G.get(r, total)
            assert np.array_equal(got[: off // 4].view(np.uint32), base[r][: off // 4].view(np.uint32))
            assert np.array_equal(got[off // 4 + count:].view(np.uint32), base[r][off // 4 + count:].view(np.uint32))
            assert np.array_equal(got[off // 4: off // 4 + count].view(np.uint32), cur[r].view(np.uint32)), r
    finally:
        G.close()


def test_allreduce_grid_choices_agree():
    world, count = 2, (1 << 18) + 7
    rng = np.random.default_rng(11)
    parts = [rng.standard_normal(count).astype(np.float32) for _ in range(world)]
    want = rank_order_sum(parts)
    G = Group(world, 4 * count)
    try:
        for ctas in (1, 2, 7, 16):
            for r in range(world):
                G.put(r, parts[r])
            G.allreduce(count, 0, ctas)
            G.sync()
            for r in range(world):
                assert np.array_equal(G.get(r, count).view(np.uint32), want.view(np.uint32)), (ctas, r)
    finally:
        G.close()


def test_one_shot_push_with_a_grid_that_changes_between_calls():
    """Consecutive one-shot calls on one exchange with DIFFERENT explicit grids: vector i then moves
    from one CTA to another between calls.  The staging buffer of a call is chosen by the parity of a
    per-exchange call counter, the same in every CTA -- with a per-CTA count (round 1) CTA 1 of the
    second call below would have written into the buffer the peer's CTA 0 was still reading."""
    world, count = 2, 4096  # 1024 vectors: ctas=1 covers them with 2 per thread, ctas=2 with one each
    rng = np.random.default_rng(5)
    G = Group(world, 4 * count)
    try:
        for call, ctas in enumerate([1, 2, 1, 2, 2, 1, 4, 1, 2, 8, 1]):
            parts = [rng.standard_normal(count).astype(np.float32) for _ in range(world)]
            want = rank_order_sum(parts)
            for r in range(world):
                G.put(r, parts[r])
            G.allreduce(count, 0, ctas, gpuprobe.ALLREDUCE_PUSH)
            G.sync()
            for r in range(world):
                assert np.array_equal(G.get(r, count).view(np.uint32), want.view(np.uint32)), (call, ctas, r)
        # and back to back without a host sync in between: each call's input is the previous call's output
        parts = [np.full(count, float(r + 1), dtype=np.float32) for r in range(world)]
        for r in range(world):
            G.put(r, parts[r])
        for ctas in (1, 2, 1, 4, 2, 1):
            G.allreduce(count, 0, ctas, gpuprobe.ALLREDUCE_PUSH)
        G.sync()
        want = np.full(count, 3.0 * 2 ** 5, dtype=np.float32)  # (1 + 2) doubled by each of the five further calls
        for r in range(world):
            assert np.array_equal(G.get(r, count), want), r
            assert G.x[r].status() == 0
    finally:
        G.close()


def test_allreduce_argument_errors_and_missing_peer():
    G = Group(2, 1 << 16)
    try:
        x = G.x[0]
        with torch.cuda.device(G.dev[0]):
            for args, rc in (((8, 4), gpuprobe.EINVAL),               # offset not 16-byte aligned
                             ((0, (1 << 14) + 1), gpuprobe.ETOOBIG),  # beyond the buffer
                             ((1 << 17, 1), gpuprobe.ETOOBIG)):
                with pytest.raises(gpuprobe.ProbeError) as ei:
                    x.allreduce_f32(*args)
                assert ei.value.rc == rc
            nobuf = gpuprobe.Exchange(0, 1)
            with pytest.raises(gpuprobe.ProbeError) as ei:
                nobuf.allreduce_f32(0, 4)
            assert ei.value.rc == gpuprobe.EINVAL
            nobuf.close()
            with pytest.raises(gpuprobe.ProbeError) as ei:   # one-shot limits
                x.allreduce_f32(0, 1022, 0, 0, gpuprobe.ALLREDUCE_PUSH)
            assert ei.value.rc == gpuprobe.ETOOBIG
            with pytest.raises(gpuprobe.ProbeError) as ei:
                x.allreduce_f32(0, 4, 0, 0, 4)
            assert ei.value.rc == gpuprobe.EINVAL
            if not x.has_multicast:
                with pytest.raises(gpuprobe.ProbeError) as ei:
                    x.allreduce_f32(0, 4, 0, 0, gpuprobe.ALLREDUCE_MULTIMEM)
                assert ei.value.rc == gpuprobe.ENOTREADY
            x.set_timeout_ms(50)
            x.allreduce_f32(0, 1024, 4, G.st[0].cuda_stream)  # rank 1 never calls
        G.sync()
        assert x.status() == gpuprobe.ETIMEDOUT
    finally:
        G.close()
