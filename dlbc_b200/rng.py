"""Host-side restatement of the reference's random numbers.

Reference: src/dlbc/random.d:19-53 (``alias RNG = Mt19937``; ``seed``; ``shiftSeedByRank``)
and the call sites in src/dlbc/fields/init.d:43,64,72,80 (``uniform(a, b, lrng)``).

The algorithm itself lives in Phobos (std.random, DMD 2.066-2.070, not vendored under the
reference tree): ``Mt19937`` is the standard 32-bit Mersenne Twister seeded with Knuth's
linear recurrence, and ``uniform!"[)"(a, b, rng)`` for doubles is

    _b = nextafter(b, -inf);  result = a + (_b - a) * double(raw32) / 4294967295.0

with one 32-bit draw per call.  The nine known-answer values of random.d:56-87 pin this
(tests/test_host_init.py).
"""
from __future__ import annotations

import numpy as np


class Mt19937:
    """std.random.Mt19937 (32-bit MT, default parameters)."""

    def __init__(self, seed=5489):
        self.seed(seed)

    def seed(self, value):
        key = np.empty(624, dtype=np.uint32)
        x = int(value) & 0xFFFFFFFF
        key[0] = x
        for i in range(1, 624):
            x = (1812433253 * (x ^ (x >> 30)) + i) & 0xFFFFFFFF
            key[i] = x
        self._bg = np.random.MT19937()
        self._bg.state = {"bit_generator": "MT19937", "state": {"key": key, "pos": 624}}

    def raw(self, n):
        """Next n raw 32-bit outputs as float64 (exactly representable)."""
        return self._bg.random_raw(int(n)).astype(np.float64)

    def uniform(self, a, b, n=None):
        """``uniform(a, b, rng)`` of Phobos for doubles, n draws (scalar if n is None)."""
        a = np.float64(a)
        _b = np.nextafter(np.float64(b), -np.inf)
        m = 1 if n is None else n
        r = a + (_b - a) * self.raw(m) / np.float64(4294967295.0)
        return r[0] if n is None else r
