"""Product-path wiring for the process-per-GPU model.

Every rank exports one 256-byte blob (CUDA-IPC handle of its slab allocation plus the
offsets of its recv buffers and flags), the blobs are all-gathered once over
``torch.distributed`` (NCCL or gloo — bootstrap only), and each rank maps its two ring
neighbours.  From then on the step kernels store halo rows straight into the
neighbour's memory over NVLink and signal with device-side flags; no MPI / NCCL call
is made per generation (contrast gol-main.c:97-111).
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from .._native import EXPORT_BYTES, World
from .ring import Ring


def gather_blobs(blob: bytes, group=None) -> list:
    world = dist.get_world_size(group)
    use_cuda = dist.get_backend(group) == "nccl"
    dev = torch.device("cuda", torch.cuda.current_device()) if use_cuda else torch.device("cpu")
    mine = torch.frombuffer(bytearray(blob), dtype=torch.uint8).to(dev)
    out = [torch.empty(EXPORT_BYTES, dtype=torch.uint8, device=dev) for _ in range(world)]
    dist.all_gather(out, mine, group=group)
    return [bytes(t.cpu().numpy().tobytes()) for t in out]


def connect_over_dist(world: World, group=None) -> None:
    """All-gather export blobs, import the ring neighbours, connect, barrier."""
    if world.nlocal != 1:
        raise ValueError("process-per-GPU wiring expects exactly one local slab per process")
    torch.cuda.set_device(world.device_of(0))
    blobs = gather_blobs(world.export_blob(0), group)
    ring = Ring(world.nranks)
    me = world.first_rank
    for r in set(ring.neighbours(me)):
        if r != me:
            world.import_blob(r, blobs[r])
    world.connect()
    dist.barrier(group)
