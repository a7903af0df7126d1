"""Independent pure-Python (big-integer / exact-rational) restatement of the
probe spec in SURVEY.md section 8(d).  Used only to cross-check the C oracle on
small cases: it shares no code with oracle/probe_ref.c.

PARITY UNPINNED against the reference (it has no arithmetic for this slot,
/root/reference/pods/nvidia-gpu-test-pod.yaml:9).
"""
from __future__ import annotations

import struct
from fractions import Fraction

M64 = (1 << 64) - 1
SEED_B = 0x0123456789ABCDEF
SEED_C = 0xFEDCBA9876543210

# The public SplitMix64 known-answer vector: first five outputs for state 1234567.
SPLITMIX64_KAT_SEED = 1234567
SPLITMIX64_KAT = [
    6457827717110365317,
    3203168211198807973,
    9817491932198370423,
    4593380528125082431,
    16408922859458223821,
]
SPLITMIX64_SEED0 = [0xE220A8397B1DCDAF, 0x6E789E6AA1B965F4, 0x06C45D188009454F]


def splitmix64_sequential(state: int, count: int) -> list[int]:
    """The generator as published: advance state by the golden gamma, then mix."""
    out = []
    for _ in range(count):
        state = (state + 0x9E3779B97F4A7C15) & M64
        z = state
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M64
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M64
        out.append(z ^ (z >> 31))
    return out


def gen(i: int, seed: int) -> int:
    """Counter form: equals splitmix64_sequential(seed, i+1)[-1]."""
    z = (seed + (i + 1) * 0x9E3779B97F4A7C15) & M64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M64
    return z ^ (z >> 31)


def unit_double(x: int) -> float:
    return float(Fraction(x >> 11, 1 << 53))  # exact: 53-bit integer / 2^53


def fma_exact(s: float, c: float, b: float) -> float:
    """Correctly rounded s*c+b: exact rational arithmetic, one rounding at the end."""
    return float(Fraction(s) * Fraction(c) + Fraction(b))


def bits(x: float) -> int:
    return struct.unpack("<Q", struct.pack("<d", x))[0]


def lanes_of_words(words, first_index: int = 0) -> tuple[int, int]:
    l0 = l1 = 0
    for k, w in enumerate(words):
        l0 = (l0 + w) & M64
        l1 = (l1 + w * (2 * (first_index + k) + 1)) & M64
    return l0, l1


def stream_lanes(n: int, first_index: int = 0, seed_b: int = SEED_B, seed_c: int = SEED_C,
                 s: float = 3.0) -> tuple[int, int]:
    words = []
    for k in range(n):
        i = first_index + k
        words.append(bits(fma_exact(s, unit_double(gen(i, seed_c)), unit_double(gen(i, seed_b)))))
    return lanes_of_words(words, first_index)
