"""Stateless ops on caller-owned tensors, plus the CPU oracle."""
from .oracle import oracle_run, reference_quirk_run
from .stateless import pack, population, step_rows, step_torus, unpack

__all__ = ["oracle_run", "reference_quirk_run", "pack", "unpack", "population", "step_torus", "step_rows"]
