"""Lattice connectivity tables (host side).

Mirrors the compile-time tables of the reference, src/dlbc/lb/connectivity.d:
velocities (:372-418), bounce (:128-142 -> index of -c_q), weights (:250-257 for D3Q19,
D2Q9 analogue), css = c_s^2 = 1/3 (:145-183).  The CUDA library has its own copy of the
D3Q19 tables in constant memory; tests check that the two agree.
"""
from __future__ import annotations

import numpy as np


class Connectivity:
    def __init__(self, name, velocities, weights, css):
        self.name = name
        self.velocities = np.asarray(velocities, dtype=np.int64)
        self.q, self.d = self.velocities.shape
        self.weights = np.asarray(weights, dtype=np.float64)
        self.css = float(css)
        self.bounce = np.empty(self.q, dtype=np.int64)
        for i, ci in enumerate(self.velocities):
            for j, cj in enumerate(self.velocities):
                if not np.any(ci + cj):
                    self.bounce[i] = j
                    break

    def __repr__(self):
        return f"Connectivity({self.name})"


_d3q19_v = [
    (0, 0, 0),
    (1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1),
    (1, 1, 0), (-1, 1, 0), (0, 1, 1), (0, -1, 1), (1, 0, 1), (1, 0, -1),
    (1, -1, 0), (-1, -1, 0), (0, 1, -1), (0, -1, -1), (-1, 0, 1), (-1, 0, -1),
]
d3q19 = Connectivity("d3q19", _d3q19_v, [1.0 / 3.0] + [1.0 / 18.0] * 6 + [1.0 / 36.0] * 12, 1.0 / 3.0)

_d2q9_v = [(0, 0), (1, 0), (-1, 0), (0, 1), (0, -1), (1, 1), (1, -1), (-1, 1), (-1, -1)]
d2q9 = Connectivity("d2q9", _d2q9_v, [4.0 / 9.0] + [1.0 / 9.0] * 4 + [1.0 / 36.0] * 4, 1.0 / 3.0)

by_name = {"d3q19": d3q19, "d2q9": d2q9}
