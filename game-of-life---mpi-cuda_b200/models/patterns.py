"""Initial patterns.  0-4 are the reference's (gol-with-cuda.cu:56-171), 5-6 are new.

The cell coordinates are defined once, in ``csrc/life_math.h::pattern_cell``; the
device initialiser (``pattern_init_kernel``) and this CPU view both evaluate that
predicate, so they cannot drift apart.
"""
from __future__ import annotations

import numpy as np

from .. import _native

PATTERNS = {
    0: "all zeros",                 # gol-with-cuda.cu:56-69
    1: "all ones",                  # gol-with-cuda.cu:72-92
    2: "ten ones, last local row",  # gol-with-cuda.cu:95-120 (cols 127..136 of every rank's last row)
    3: "ones at corners",           # gol-with-cuda.cu:123-147
    4: "spinner at corner",         # gol-with-cuda.cu:150-171
    5: "glider across the first slab boundary",
    6: "random density (seed, density)",
}


def pattern_bytes(pattern: int, width: int, rows: int, rank: int = 0, nranks: int = 1, seed: int = 1,
                  density: float = 0.5) -> np.ndarray:
    """One rank's ``rows x width`` slab of the pattern as 0/1 bytes."""
    if pattern not in PATTERNS:
        raise _native.GolError(f"Pattern {pattern} has not been implemented ")
    return _native.pattern_fill(pattern, width, rows, rank, nranks, seed, density)


def global_pattern(pattern: int, width: int, rows: int, nranks: int, seed: int = 1,
                   density: float = 0.5) -> np.ndarray:
    """The whole ``rows*nranks x width`` torus (rank slabs stacked in rank order, gol-main.c:22)."""
    return np.concatenate(
        [pattern_bytes(pattern, width, rows, r, nranks, seed, density) for r in range(nranks)], axis=0)
