"""Device timing with CUDA events on the launching stream (never wall clock)."""
from __future__ import annotations

import torch


class CudaTimer:
    """``with CudaTimer() as t: ...`` then ``t.ms`` (synchronises on both sides)."""

    def __init__(self, stream=None):
        self.stream = stream
        self.ms = 0.0

    def __enter__(self):
        self._a = torch.cuda.Event(enable_timing=True)
        self._b = torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        self._a.record(self.stream)
        return self

    def __exit__(self, *exc):
        self._b.record(self.stream)
        torch.cuda.synchronize()
        self.ms = self._a.elapsed_time(self._b)


_flush_buf = {}


def flush_l2(device=None, nbytes: int = 256 << 20) -> None:
    """Write a buffer larger than the 126 MB L2 so the next timed step starts cold."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    buf = _flush_buf.get(dev)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        _flush_buf[dev] = buf
    buf.fill_(1)
