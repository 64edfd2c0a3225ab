"""BASELINE, not product: per-generation NCCL halo exchange + unfused stencil.

This is the straightforward "recompile for NCCL" of the reference's loop
(gol-main.c:94-116): each generation posts ncclSend/ncclRecv of the first / last packed
row to the ring neighbours (``dist.batch_isend_irecv``), then launches the stand-alone
stencil kernel ``step_rows``.  It exists so the fused path has a measured
library-collective baseline to be compared against (BASELINE.json north star: "a path
that only calls MPI or NCCL for the halo is the baseline, not the product").
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from .. import _native
from ..ops import stateless
from .ring import Ring


class NcclHaloStepper:
    def __init__(self, width: int, rows: int, threads: int = 256, group=None):
        self.width, self.rows, self.threads, self.group = width, rows, threads, group
        self.rank = dist.get_rank(group)
        self.ring = Ring(dist.get_world_size(group))
        self.stride = _native.row_stride_words(width)
        dev = torch.device("cuda", torch.cuda.current_device())
        self.words = torch.zeros((rows, self.stride), dtype=torch.int64, device=dev)
        self.top = torch.zeros(self.stride, dtype=torch.int64, device=dev)
        self.bottom = torch.zeros(self.stride, dtype=torch.int64, device=dev)

    def load(self, cells_u8: torch.Tensor) -> None:
        self.words = stateless.pack(cells_u8.contiguous())

    def step(self, generations: int) -> None:
        prev, nxt = self.ring.neighbours(self.rank)
        for _ in range(generations):
            first = self.words[0].contiguous()
            last = self.words[-1].contiguous()
            if self.ring.nranks == 1:
                self.top.copy_(last)
                self.bottom.copy_(first)
            else:
                ops = [
                    dist.P2POp(dist.isend, first, prev, self.group),
                    dist.P2POp(dist.isend, last, nxt, self.group),
                    dist.P2POp(dist.irecv, self.bottom, nxt, self.group),
                    dist.P2POp(dist.irecv, self.top, prev, self.group),
                ]
                # R == 2: both neighbours are the same peer, NCCL matches in posting order; the
                # peer sends (first, last), so the first receive posted must be my bottom ghost.
                for req in dist.batch_isend_irecv(ops):
                    req.wait()
            self.words = stateless.step_rows(self.words, self.top, self.bottom, self.width, self.threads)

    def cells(self) -> torch.Tensor:
        return stateless.unpack(self.words, self.width)
