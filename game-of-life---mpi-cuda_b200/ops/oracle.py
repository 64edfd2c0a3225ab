"""CPU oracles.

``oracle_run`` is true torus Life (what the reference *intends*).  ``reference_quirk_run``
models what the as-built reference computes over several generations — stale halo
rows (Q1: gol-with-cuda.cu:40-47 fills the send rows once) and, for R <= 2, ghost rows
delivered the wrong way round (Q2: gol-main.c:97-107) — so its output files can be
explained; it is never used by the product path.
"""
from __future__ import annotations

import numpy as np

from .. import _native


def oracle_run(cells: np.ndarray, generations: int) -> np.ndarray:
    return _native.oracle_run(cells, generations)


def numpy_life_run(cells: np.ndarray, generations: int) -> np.ndarray:
    """True torus Life with numpy ``roll`` on a 0/1 byte grid.  Shares nothing with the native
    code (no ``life_math.h``, no packing): the independent check for wide grids, where the
    byte-per-cell C oracle is slow and the packed CPU stepper shares its math with the kernels."""
    g = np.ascontiguousarray(cells, dtype=np.uint8)
    for _ in range(int(generations)):
        v = g + np.roll(g, 1, axis=0) + np.roll(g, -1, axis=0)         # vertical 3-sums (<= 3)
        n = v + np.roll(v, 1, axis=1) + np.roll(v, -1, axis=1) - g     # 8 neighbours (<= 8)
        g = ((n == 3) | ((g == 1) & (n == 2))).astype(np.uint8)
    return g


def _step_with_ghosts(slab: np.ndarray, top: np.ndarray, bottom: np.ndarray) -> np.ndarray:
    ext = np.vstack([top[None, :], slab, bottom[None, :]]).astype(np.int32)
    n = np.zeros_like(ext)
    for dy in (-1, 0, 1):
        for dx in (-1, 0, 1):
            if dy == 0 and dx == 0:
                continue
            n += np.roll(np.roll(ext, dy, axis=0), dx, axis=1)
    nxt = ((n == 3) | ((ext == 1) & (n == 2))).astype(np.uint8)
    return nxt[1:-1]


def reference_quirk_run(slabs, generations: int):
    """slabs: list of per-rank (H, W) arrays.  Returns the per-rank arrays the reference binary writes."""
    R = len(slabs)
    cur = [s.copy() for s in slabs]
    first0 = [s[0].copy() for s in slabs]   # never refreshed after init (Q1)
    last0 = [s[-1].copy() for s in slabs]
    for _ in range(generations):
        nxt = []
        for r in range(R):
            prev, nex = (r - 1) % R, (r + 1) % R
            if R <= 2:
                # both receives match the same source in posting order (Q2)
                top, bottom = first0[prev], last0[nex]
            else:
                top, bottom = last0[prev], first0[nex]
            nxt.append(_step_with_ghosts(cur[r], top, bottom))
        cur = nxt
    return cur
