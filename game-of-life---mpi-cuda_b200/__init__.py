"""B200-native distributed Game of Life (capabilities of shoron-dutta/Game-of-Life---MPI-CUDA).

Layers
------
``_native``   ctypes binding of the C ABI in ``include/gol_b200.h`` (C++/CUDA core, sm_100a)
``ops``       stateless tensor ops (pack / unpack / population / torus step) and the CPU oracle
``models``    :class:`Life` — the user-facing stepper (patterns, step, dump) on one or many GPUs
``parallel``  ring topology, CUDA-IPC peer wiring over ``torch.distributed``, NCCL/gloo baselines
``utils``     device timing, clock sampling, metrics, rank-file helpers
"""
from . import _native
from ._native import GolError, World, launch_count, version
from .models.life import Life
from .models.patterns import PATTERNS, pattern_bytes

__all__ = ["Life", "World", "GolError", "PATTERNS", "pattern_bytes", "launch_count", "version"]
