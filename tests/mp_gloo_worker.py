"""Worker for tests/test_gloo_sharding.py: world_size-2 (or more) gloo ranks on CPU.

Exercises bench.py's N > 1 host logic without a GPU: the shard each rank owns, and the
wrapping-u64 combine of the lanes through an int64 SUM allreduce.  The per-shard lanes
come from the CPU oracle (this is a test), the combined lanes must equal the oracle's
lanes of the whole index space and the committed golden value."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench  # noqa: E402
import oracle  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo")
    log2n = 20
    n = 1 << log2n
    first = bench.shard_first_index(rank, n)
    mine = oracle.stream_lanes(n, first)
    combined = bench.allreduce_lanes(dist, torch, mine)
    whole = oracle.stream_lanes(n * world, 0)
    with open(os.path.join(ROOT, "tests", "golden", "probe_golden.json")) as f:
        pow2 = json.load(f)["pow2_lanes"]
    ok = combined == whole
    if world & (world - 1) == 0:
        ok = ok and combined == tuple(int(v, 16) for v in pow2[(n * world).bit_length() - 1])
    # every rank must hold the same combined lanes
    flags = [None] * world
    dist.all_gather_object(flags, (ok, combined))
    if rank == 0:
        if all(f[0] for f in flags) and len({f[1] for f in flags}) == 1:
            print(f"GLOO_SHARDING_OK world={world} lanes=0x{combined[0]:016x},0x{combined[1]:016x}", flush=True)
        else:
            print(f"GLOO_SHARDING_BAD {flags} whole={whole}", flush=True)
            sys.exit(1)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
