"""Stateless device ops on torch CUDA tensors (thin wrappers over the C ABI).

Each function launches hand-written sm_100a kernels from ``csrc/kernels.cu`` on the
current torch stream; none falls back to PyTorch math.
"""
from __future__ import annotations

import torch

from .. import _native
from .._native import check, lib


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _need_cuda(t: torch.Tensor, name: str) -> None:
    if not t.is_cuda or not t.is_contiguous():
        raise _native.GolError(f"{name} must be a contiguous CUDA tensor")


def pack(cells: torch.Tensor) -> torch.Tensor:
    """(H, W) uint8 0/1 cells -> (H, stride) int64 packed words (N-K7)."""
    _need_cuda(cells, "cells")
    H, W = cells.shape
    out = torch.zeros((H, _native.row_stride_words(W)), dtype=torch.int64, device=cells.device)
    with torch.cuda.device(cells.device):
        check(lib.gol_cuda_pack(cells.data_ptr(), out.data_ptr(), W, H, _stream()))
    return out


def unpack(words: torch.Tensor, width: int) -> torch.Tensor:
    """(H, stride) packed words -> (H, W) uint8 cells (N-K7)."""
    _need_cuda(words, "words")
    H = words.shape[0]
    out = torch.empty((H, width), dtype=torch.uint8, device=words.device)
    with torch.cuda.device(words.device):
        check(lib.gol_cuda_unpack(words.data_ptr(), out.data_ptr(), width, H, _stream()))
    return out


def population(words: torch.Tensor, width: int) -> torch.Tensor:
    """Number of live cells; returns a 1-element int64 CUDA tensor (no host sync)."""
    _need_cuda(words, "words")
    H = words.shape[0]
    scratch = torch.empty(lib.gol_cuda_population_scratch_bytes() // 8, dtype=torch.int64, device=words.device)
    out = torch.empty(1, dtype=torch.int64, device=words.device)
    with torch.cuda.device(words.device):
        check(lib.gol_cuda_population(words.data_ptr(), width, H, scratch.data_ptr(), out.data_ptr(), _stream()))
    return out


def step_torus(words: torch.Tensor, width: int, threads: int = 256) -> torch.Tensor:
    """One generation of a stand-alone packed torus (no ring, no halo exchange)."""
    _need_cuda(words, "words")
    out = torch.empty_like(words)
    with torch.cuda.device(words.device):
        check(lib.gol_cuda_step_torus(words.data_ptr(), out.data_ptr(), width, words.shape[0], threads, _stream()))
    return out


def step_rows(words: torch.Tensor, top: torch.Tensor, bottom: torch.Tensor, width: int,
              threads: int = 256) -> torch.Tensor:
    """One generation of a slab given explicit ghost rows — the *unfused* stencil used by the
    NCCL-halo baseline (parallel/nccl_halo.py); the product path never calls it."""
    _need_cuda(words, "words")
    _need_cuda(top, "top")
    _need_cuda(bottom, "bottom")
    out = torch.empty_like(words)
    with torch.cuda.device(words.device):
        check(lib.gol_cuda_step_rows(words.data_ptr(), out.data_ptr(), top.data_ptr(), bottom.data_ptr(), width,
                                     words.shape[0], threads, _stream()))
    return out
