"""Rank-file helpers: parse ``Rank_<r>_of_<R>.txt`` (the layout of gol-main.c:17-28,135-139) back
into cells, so that outputs of this framework and of the reference binary can be compared."""
from __future__ import annotations

import os
import re
from typing import List, Tuple

import numpy as np

_HEADER = re.compile(r"^#{25} FINAL WORLD IN RANK (\d+) IS #{31}$")
_ROW = re.compile(r"^Row\s*(-?\d+): (.*)$")


def read_rank_file(path: str) -> Tuple[int, int, np.ndarray]:
    """Returns (rank, first global row, cells[rows, width] as uint8)."""
    with open(path, "r") as f:
        lines = f.read().split("\n")
    m = _HEADER.match(lines[0])
    if not m:
        raise ValueError(f"{path}: not a rank file (bad header line)")
    rank = int(m.group(1))
    rows: List[List[int]] = []
    first = None
    for line in lines[1:]:
        if not line:
            continue
        r = _ROW.match(line)
        if not r:
            raise ValueError(f"{path}: malformed row line {line[:40]!r}")
        if first is None:
            first = int(r.group(1))
        rows.append([int(t) for t in r.group(2).split()])
    return rank, (first or 0), np.array(rows, dtype=np.uint8)


def read_world(directory: str, nranks: int) -> np.ndarray:
    """Stack Rank_0_of_R .. Rank_{R-1}_of_R into the global grid."""
    parts = []
    for r in range(nranks):
        rank, _, cells = read_rank_file(os.path.join(directory, f"Rank_{r}_of_{nranks}.txt"))
        assert rank == r
        parts.append(cells)
    return np.concatenate(parts, axis=0)
