"""`Life`: the user-facing stepper.

Two rank models, both producing the reference's ``Rank_<r>_of_<R>.txt`` files:

* **single process** (``ranks=R``): one process drives R slabs, one per GPU
  (or several per GPU when R exceeds the device count), ring neighbours reached
  through NVLink peer pointers;
* **process per GPU** (``torchrun``): parity with the reference's MPI model
  (gol-main.c:58-62).  ``torch.distributed`` is used only to bootstrap: ranks
  all-gather CUDA-IPC handles once, after which every halo row travels as a
  direct store into the neighbour's memory from inside the step kernel.
"""
from __future__ import annotations

import os
from typing import List, Optional

import numpy as np

from .. import _native
from .._native import World


def _dist():
    import torch.distributed as dist

    return dist if dist.is_available() and dist.is_initialized() else None


class Life:
    def __init__(self, size: int, rows: Optional[int] = None, *, ranks: Optional[int] = None,
                 backend: str = "cuda", mode: str = "auto", k: int = 0, depth: int = 0, threads: int = 0,
                 distributed: Optional[bool] = None, flags: int = 0):
        self.width = int(size)
        self.rows = int(rows if rows is not None else size)
        dist = _dist()
        if distributed is None:
            distributed = dist is not None and dist.get_world_size() > 1 and ranks is None
        self.distributed = bool(distributed)
        if self.distributed:
            if dist is None:
                raise _native.GolError("distributed=True needs an initialised torch.distributed group")
            self.nranks = dist.get_world_size()
            self.rank = dist.get_rank()
            local_rank = int(os.environ.get("LOCAL_RANK", self.rank))
            if backend == "cuda":
                ndev = int(_native.lib.gol_cuda_device_count())
                local_world = int(os.environ.get("LOCAL_WORLD_SIZE", self.nranks))
                if ndev < min(local_world, self.nranks):
                    # kernels of two processes that wait on each other's halo flags must not share a GPU:
                    # nothing guarantees that both are resident at the same time
                    raise _native.GolError(f"{min(local_world, self.nranks)} ranks on this node but only {ndev} "
                                           "GPU(s): the process-per-GPU model needs one GPU per rank "
                                           "(use Life(..., ranks=R) to host several ranks per GPU in one process)")
            self.world = World(self.width, self.rows, nranks=self.nranks, first_rank=self.rank, nlocal=1,
                               backend=backend, mode=mode, k=k, depth=depth, threads=threads,
                               device_base=local_rank, flags=flags)
            if backend == "cuda":
                from ..parallel.peer import connect_over_dist

                connect_over_dist(self.world)
            self._cpu_ring = None
            if backend != "cuda":
                from ..parallel.gloo_cpu import GlooRing

                self._cpu_ring = GlooRing(self.world)
        else:
            self.nranks = int(ranks or 1)
            self.rank = 0
            self.world = World(self.width, self.rows, nranks=self.nranks, backend=backend, mode=mode, k=k,
                               depth=depth, threads=threads, flags=flags)
            self.world.connect()
            self._cpu_ring = None
        self.backend = backend
        self.generation = 0

    # ------------------------------------------------------------------ state
    @property
    def plan(self) -> dict:
        return self.world.plan

    @property
    def local_ranks(self) -> List[int]:
        return list(range(self.world.first_rank, self.world.first_rank + self.world.nlocal))

    def init(self, pattern: int, seed: int = 1, density: float = 0.5) -> "Life":
        self.world.init_pattern(pattern, seed, density)
        self.generation = 0
        return self

    def load(self, cells: np.ndarray, local: int = 0) -> "Life":
        self.world.load_bytes(local, cells)
        return self

    def load_packed(self, words, local: int = 0, asynchronous: bool = False) -> "Life":
        self.world.load_packed(local, words, asynchronous=asynchronous)
        return self

    def cells(self, local: int = 0) -> np.ndarray:
        return self.world.read_bytes(local)

    def all_local_cells(self) -> np.ndarray:
        return np.concatenate([self.world.read_bytes(i) for i in range(self.world.nlocal)], axis=0)

    # --------------------------------------------------------------- stepping
    def step(self, generations: int = 1) -> "Life":
        if self._cpu_ring is not None:
            self._cpu_ring.step(generations)
        else:
            self.world.step(generations)
        self.generation += generations
        return self

    def sync(self) -> "Life":
        self.world.sync()
        return self

    def step_timed(self, generations: int) -> float:
        """Device-timed step: milliseconds, max over this process's slabs and over all ranks."""
        if self._cpu_ring is not None:
            import time

            t0 = time.perf_counter()
            self._cpu_ring.step(generations)
            ms = (time.perf_counter() - t0) * 1e3
        else:
            ms = max(self.world.step_timed(generations))
        self.generation += generations
        dist = _dist()
        if self.distributed and dist is not None:
            import torch

            dev = "cuda" if (self.backend == "cuda" and dist.get_backend() == "nccl") else "cpu"
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    def run_from_host(self, pinned_words, generations: int, pinned_result) -> int:
        """End-to-end step as a user issues it: upload a packed world from pinned host memory,
        advance it, bring the packed result grid back into pinned host memory.  ``pinned_words`` /
        ``pinned_result`` are one buffer (``rows x stride`` 64-bit words) or a list with one buffer
        per local slab.  Asynchronous and pipelined: uploads, compute and downloads of consecutive
        calls overlap (two slots).  Returns the slot for :meth:`wait_result`."""
        ins = list(pinned_words) if isinstance(pinned_words, (list, tuple)) else [pinned_words]
        outs = list(pinned_result) if isinstance(pinned_result, (list, tuple)) else [pinned_result]
        slot = self.world.run_from_host_async(ins, generations, outs)
        self.generation += generations
        return slot

    def wait_result(self, slot: int) -> "Life":
        """Block until the result of the :meth:`run_from_host` call that returned ``slot`` is in
        host memory."""
        self.world.wait_result(slot)
        return self

    # ---------------------------------------------------------------- metrics
    def population(self, global_sum: bool = True) -> int:
        n = sum(self.world.population(i) for i in range(self.world.nlocal))
        dist = _dist()
        if global_sum and self.distributed and dist is not None:
            import torch

            dev = "cuda" if (self.backend == "cuda" and dist.get_backend() == "nccl") else "cpu"
            t = torch.tensor([n], dtype=torch.int64, device=dev)
            dist.all_reduce(t)
            n = int(t.item())
        return n

    def cell_updates(self, iterations: int) -> int:
        return self.nranks * self.rows * self.width * iterations  # gol-main.c:124-125

    def metrics(self, ms: float, iterations: int) -> dict:
        """GCUPS of a step of ``iterations`` generations that took ``ms`` (whole ring), with the
        per-GPU fraction of the bit-packed HBM roofline at this plan's generations per pass."""
        from ..utils.metrics import gcups, roofline_fraction

        total = gcups(self.cell_updates(iterations), ms * 1e-3)
        k = max(1, int(self.plan.get("temporal_k", 1)))
        roof = roofline_fraction(total * 1e9 / max(1, self.nranks), generations_per_pass=k)
        return {"gcups": total, "ms": ms, "generations": iterations, "mode": self.plan["mode_name"], "k": k,
                "hbm_roofline_fraction": roof["fraction"], "hbm_roofline_of": roof["of"]}

    # --------------------------------------------------------------------- io
    def dump(self, directory: Optional[str] = None) -> List[str]:
        """Write ``Rank_<r>_of_<R>.txt`` for every local slab (gol-main.c:135-139)."""
        names = []
        for i, r in enumerate(self.local_ranks):
            self.world.dump_text(i, directory)
            names.append(os.path.join(directory or ".", _native.rank_file_name(r, self.nranks)))
        return names

    # ------------------------------------------------------------- checkpoint
    def save_checkpoint(self, directory: str) -> List[str]:
        """Write one ``ckpt_rank<r>_of_<R>.npz`` per local slab: the packed words plus geometry and
        the generation counter.  (The reference can only write its final world as text and cannot
        read anything back, gol-main.c:135-139.)"""
        os.makedirs(directory, exist_ok=True)
        self.sync()
        names = []
        for i, r in enumerate(self.local_ranks):
            path = os.path.join(directory, f"ckpt_rank{r}_of_{self.nranks}.npz")
            np.savez_compressed(path, words=self.world.read_packed(i), width=self.width, rows=self.rows,
                                rank=r, nranks=self.nranks, generation=self.generation)
            names.append(path)
        return names

    def load_checkpoint(self, directory: str) -> "Life":
        """Resume from :meth:`save_checkpoint` files (same geometry and rank count)."""
        for i, r in enumerate(self.local_ranks):
            path = os.path.join(directory, f"ckpt_rank{r}_of_{self.nranks}.npz")
            with np.load(path) as z:
                if (int(z["width"]), int(z["rows"]), int(z["nranks"]), int(z["rank"])) != \
                        (self.width, self.rows, self.nranks, r):
                    raise _native.GolError(f"{path} was written for a different geometry")
                self.world.load_packed(i, np.ascontiguousarray(z["words"], dtype=np.uint64))
                self.generation = int(z["generation"])
        return self

    def report(self, seconds: float, iterations: int) -> str:
        """The two stdout lines of rank 0 (gol-main.c:124-125,132)."""
        return _native.format_report(seconds, self.nranks, self.width, self.rows, iterations)

    def close(self) -> None:
        self.world.close()
