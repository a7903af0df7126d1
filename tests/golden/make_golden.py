#!/usr/bin/env python
"""Regenerates tests/golden/probe_golden.json from the C oracle.

Run from the repo root:  python tests/golden/make_golden.py
The reference cannot generate these (it has no arithmetic), so the vectors come
from oracle/probe_ref.c after it passed the published SplitMix64 KAT and the
pure-Python cross-check (tests/test_oracle.py).  Full-size lanes (N = 2^29) are
the figures the shipped probe embeds as its known answers.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "probe_golden.json")


def hx(v):
    return f"0x{v:016x}"


def main():
    g = {"spec": "SURVEY.md 8(d)", "seed_b": hx(oracle.SEED_B), "seed_c": hx(oracle.SEED_C), "s": 3.0}
    # first 8 elements of b, c, a as bit patterns
    b = oracle.init_fill(8, 0, oracle.SEED_B)
    c = oracle.init_fill(8, 0, oracle.SEED_C)
    a = oracle.triad(b, c)
    g["head"] = {k: [hx(int(x)) for x in v.view(np.uint64)] for k, v in (("b", b), ("c", c), ("a", a))}
    # lanes of the triad output for whole buffers of 2^k elements, first_index 0
    g["stream_lanes"] = {}
    for logn in (0, 1, 5, 10, 16, 20, 24, 26, 29):
        l0, l1 = oracle.stream_lanes(1 << logn)
        g["stream_lanes"][str(1 << logn)] = [hx(l0), hx(l1)]
    # ragged sizes and non-zero / huge first_index (index wrap is part of the spec: mod 2^64)
    g["ragged"] = []
    for n, first in ((1, 7), (3, 0), (31, 1000), (33, 1 << 40), (1000003, 12345), (4097, (1 << 63) - 5)):
        l0, l1 = oracle.stream_lanes(n, first)
        g["ragged"].append({"n": n, "first_index": hx(first), "lanes": [hx(l0), hx(l1)]})
    # the 8 equal shards of the full-size buffer (their sum must equal the whole)
    n = 1 << 29
    g["shards8"] = []
    for r in range(8):
        l0, l1 = oracle.stream_lanes(n // 8, r * (n // 8))
        g["shards8"].append([hx(l0), hx(l1)])
    with open(OUT, "w") as f:
        json.dump(g, f, indent=1)
        f.write("\n")
    print("wrote", OUT)


if __name__ == "__main__":
    main()
