"""Rank topology and the transports that wire ranks together.

* ``ring``       — periodic 1-D ring (gol-main.c:84-87) and slab geometry
* ``peer``       — product path: CUDA-IPC handle exchange over torch.distributed, once
* ``nccl_halo``  — baseline only: ncclSend/ncclRecv halo + unfused stencil per generation
* ``gloo_cpu``   — CPU ranks over gloo (plumbing tests without a GPU)
"""
from .ring import Ring, slab_rows

__all__ = ["Ring", "slab_rows"]
