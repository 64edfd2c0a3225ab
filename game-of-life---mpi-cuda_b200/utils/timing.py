"""L2 flush between timed steps (device timing itself is done with CUDA events in the core:
``World.step_timed``)."""
from __future__ import annotations

import torch


_flush_buf = {}


def flush_l2(device=None, nbytes: int = 256 << 20) -> None:
    """Write a buffer larger than the 126 MB L2 so the next timed step starts cold."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    buf = _flush_buf.get(dev)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        _flush_buf[dev] = buf
    buf.fill_(1)
