"""gol_cuda_run_torus_noalloc — the entry point for callers that cannot allocate device memory
(host applications whose own pool already owns the GPU, latency-critical loops that must stay
clear of the driver's allocator).  Its scratch is a fixed array inside the library, so this
test talks to the C ABI directly: the result of each call must equal what the ordinary
stateless entry point (gol_cuda_step_torus, applied once per generation) produces, bit for bit,
on consecutive calls that reuse the shared scratch with different shapes and kernels.

Collected last on purpose (file name): a failure here cannot hide other tests behind ``-x``.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _lib():
    from gol_b200 import _native

    lib = _native.lib
    fn = lib.gol_cuda_run_torus_noalloc
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_void_p,
                   C.POINTER(C.c_int)]
    return _native, lib, fn


def _reference_result(native, lib, words, W, gens, stream):
    """gens applications of the existing one-generation entry point."""
    import torch

    cur, nxt = words.clone(), torch.empty_like(words)
    for _ in range(gens):
        native.check(lib.gol_cuda_step_torus(cur.data_ptr(), nxt.data_ptr(), W, cur.shape[0], 256, stream))
        cur, nxt = nxt, cur
    return cur


def _run_case(W, H, gens, seed):
    import torch

    native, lib, fn = _lib()
    rng = np.random.default_rng(seed)
    cells = (rng.random((H, W)) < 0.45).astype(np.uint8)
    words = torch.from_numpy(native.pack(cells).view(np.int64)).cuda()
    stream = torch.cuda.current_stream().cuda_stream
    want = _reference_result(native, lib, words, W, gens, stream)

    a, b = words.clone(), torch.full_like(words, -1)
    where = C.c_int(-1)
    rc = fn(a.data_ptr(), b.data_ptr(), W, H, gens, 128, stream, C.byref(where))
    assert rc == 0, native.last_error()
    torch.cuda.synchronize()
    got = b if where.value == 1 else a
    assert where.value in (0, 1)
    assert torch.equal(got, want), f"noalloc result differs for {W}x{H}, {gens} generations"


def test_two_calls_in_a_row_match_the_ordinary_entry_point():
    # first call: fused kernels (width a multiple of 64); second call right after it on the
    # same stream reuses the same scratch with another shape and the generic kernel
    _run_case(2048, 96, 11, seed=1)
    _run_case(200, 37, 5, seed=2)
    # and once more with the first shape, to prove nothing stale from call two leaks in
    _run_case(2048, 96, 4, seed=3)


def test_zero_generations_and_single_row():
    _run_case(128, 1, 3, seed=4)
    _run_case(64, 8, 0, seed=5)


def test_rejects_grids_wider_than_the_scratch():
    import torch

    native, lib, fn = _lib()
    dummy = torch.zeros(8, dtype=torch.int64, device="cuda")
    where = C.c_int(0)
    rc = fn(dummy.data_ptr(), dummy.data_ptr() + 8, 131072 + 64, 1, 1, 128, 0, C.byref(where))
    assert rc != 0 and "exceeds the built-in scratch" in native.last_error()
