"""CPU ranks over gloo: the plumbing configuration (BASELINE.json config 1).

Each process owns one CPU slab of the native world; per generation the two edge rows
travel by ``dist.isend`` / ``dist.irecv`` with *distinct tags per direction*, so the
R = 2 case (both neighbours are the same peer) routes top and bottom correctly — the
case the reference gets wrong by relying on message order (gol-main.c:97-107).
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from .._native import World
from .ring import Ring

TAG_TO_PREV = 11   # my first row -> prev rank's bottom ghost
TAG_TO_NEXT = 12   # my last row  -> next rank's top ghost


class GlooRing:
    def __init__(self, world: World, group=None):
        if world.backend == "cuda" or world.nlocal != 1:
            raise ValueError("GlooRing drives a CPU world with one local slab")
        self.world, self.group = world, group
        self.rank = world.first_rank
        self.ring = Ring(world.nranks)

    def exchange(self) -> None:
        w = self.world
        first = torch.from_numpy(w.cpu_get_edge(0, bottom=False).view("int64").copy())
        last = torch.from_numpy(w.cpu_get_edge(0, bottom=True).view("int64").copy())
        if self.ring.nranks == 1:
            w.cpu_set_ghost(0, False, last.numpy().view("uint64"))
            w.cpu_set_ghost(0, True, first.numpy().view("uint64"))
            return
        prev, nxt = self.ring.neighbours(self.rank)
        top = torch.empty_like(first)
        bottom = torch.empty_like(first)
        reqs = [
            dist.irecv(top, src=prev, group=self.group, tag=TAG_TO_NEXT),
            dist.irecv(bottom, src=nxt, group=self.group, tag=TAG_TO_PREV),
            dist.isend(first, dst=prev, group=self.group, tag=TAG_TO_PREV),
            dist.isend(last, dst=nxt, group=self.group, tag=TAG_TO_NEXT),
        ]
        for r in reqs:
            r.wait()
        w.cpu_set_ghost(0, False, top.numpy().view("uint64"))
        w.cpu_set_ghost(0, True, bottom.numpy().view("uint64"))

    def step(self, generations: int) -> None:
        for _ in range(generations):
            self.exchange()
            self.world.cpu_step_local()
