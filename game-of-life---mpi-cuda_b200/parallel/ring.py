"""Periodic ring of ranks: rank r's previous / next neighbour and its global rows.

The reference computes the same thing inline (gol-main.c:84-87 for the ring,
gol-main.c:22 for the global row label).  R = 1 (self-neighbour) and R = 2 (both
neighbours are the same peer) are ordinary cases here: ghost slots are addressed
by side, never by message order (the reference's Q2 defect).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Tuple


@dataclass(frozen=True)
class Ring:
    nranks: int

    def prev(self, rank: int) -> int:
        return (rank - 1 + self.nranks) % self.nranks

    def next(self, rank: int) -> int:
        return (rank + 1) % self.nranks

    def neighbours(self, rank: int) -> Tuple[int, int]:
        return self.prev(rank), self.next(rank)


def slab_rows(rank: int, rows_per_rank: int) -> Tuple[int, int]:
    """Global row range [first, last) owned by ``rank``."""
    return rank * rows_per_rank, (rank + 1) * rows_per_rank
