"""ctypes binding of the native core (``include/gol_b200.h``).

The shared object is built in-tree by ``make`` (or ``__graft_entry__.build()``)
into ``_lib/libgol_b200.so``.  There is no Python fallback for the CUDA path: if
the library is missing the import fails loudly, and a CUDA world cannot be
created without a GPU.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import List, Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
# GOL_B200_LIB=debug loads the bounds-checking build (libgol_b200_debug.so, -DGOL_DEBUG_BOUNDS=1)
_LIB_PATH = os.path.join(_HERE, "_lib", "libgol_b200_debug.so" if os.environ.get("GOL_B200_LIB") == "debug"
                         else "libgol_b200.so")

BACKEND_CPU, BACKEND_CUDA = 0, 1
MODES = {"auto": 0, "generic": 1, "fused": 2, "persistent": 3, "tma": 4, "resident": 5}
MODE_NAMES = {v: k for k, v in MODES.items()}
FLAG_NO_GRAPH, FLAG_NO_PEER_ACCESS, FLAG_CPU_SCHEDULE = 1, 2, 4
EXPORT_BYTES = 256
MAX_GHOST = 32


class GolError(RuntimeError):
    """A call into the native core failed; the message comes from gol_last_error()."""


def _build_if_missing() -> None:
    if os.path.exists(_LIB_PATH):
        return
    if os.environ.get("GOL_B200_NO_AUTOBUILD"):
        raise ImportError(f"{_LIB_PATH} is missing (run `make` in {_ROOT})")
    proc = subprocess.run(["make", "-C", _ROOT, "-j8"], capture_output=True, text=True)
    if proc.returncode != 0 or not os.path.exists(_LIB_PATH):
        raise ImportError(f"building {_LIB_PATH} failed:\n{proc.stdout[-2000:]}\n{proc.stderr[-2000:]}")


class _Config(C.Structure):
    _fields_ = [
        ("width", C.c_int64), ("rows", C.c_int64),
        ("nranks", C.c_int32), ("first_rank", C.c_int32), ("nlocal", C.c_int32),
        ("backend", C.c_int32), ("mode", C.c_int32), ("temporal_k", C.c_int32),
        ("exchange_depth", C.c_int32), ("threads_per_block", C.c_int32),
        ("ndevices", C.c_int32), ("device_base", C.c_int32), ("flags", C.c_int32),
        ("reserved", C.c_int32),
    ]


class _Plan(C.Structure):
    _fields_ = [
        ("mode", C.c_int32), ("temporal_k", C.c_int32), ("exchange_depth", C.c_int32),
        ("threads_per_block", C.c_int32), ("strip_rows", C.c_int32), ("col_groups", C.c_int32),
        ("ctas_per_slab", C.c_int32), ("regs_hint", C.c_int32),
        ("words_per_row", C.c_int64), ("row_stride", C.c_int64), ("slab_rows", C.c_int64),
    ]


class _PhaseCounters(C.Structure):
    _fields_ = [("halo_wait_ns", C.c_uint64 * 2), ("halo_waits", C.c_uint64 * 2), ("dep_wait_ns", C.c_uint64),
                ("task_ns", C.c_uint64 * 2), ("tasks", C.c_uint64 * 2)]


def _load() -> C.CDLL:
    _build_if_missing()
    lib = C.CDLL(_LIB_PATH, mode=C.RTLD_GLOBAL)
    P, I64, I32, U64, VP = C.POINTER, C.c_int64, C.c_int, C.c_uint64, C.c_void_p
    sigs = {
        "gol_last_error": (C.c_char_p, []),
        "gol_version": (C.c_char_p, []),
        "gol_cuda_device_count": (I32, []),
        "gol_debug_bounds_enabled": (I32, []),
        "gol_debug_bounds_report": (I32, [P(U64), P(U64), P(I32), I32]),
        "gol_debug_bounds_poke": (I32, [VP]),
        "gol_oracle_run": (I32, [VP, VP, I64, I64, I64]),
        "gol_pattern_fill_bytes": (I32, [VP, I32, I64, I64, I32, I32, U64, C.c_double]),
        "gol_words_per_row": (I64, [I64]),
        "gol_row_stride_words": (I64, [I64]),
        "gol_pack_bytes": (I32, [VP, VP, I64, I64]),
        "gol_unpack_words": (I32, [VP, VP, I64, I64]),
        "gol_cpu_run_packed": (I32, [VP, I64, I64, I64]),
        "gol_write_rank_file": (I32, [C.c_char_p, VP, I64, I64, I32]),
        "gol_rank_file_name": (I32, [C.c_char_p, C.c_size_t, I32, I32]),
        "gol_format_report": (I32, [C.c_char_p, C.c_size_t, C.c_double, I32, I64, I64, I64]),
        "gol_world_create": (VP, [P(_Config)]),
        "gol_world_destroy": (None, [VP]),
        "gol_world_plan": (I32, [VP, P(_Plan)]),
        "gol_schedule_describe": (I32, [P(_Config), I64, VP, I32, P(_Plan)]),
        "gol_world_export": (I32, [VP, I32, VP]),
        "gol_world_import": (I32, [VP, I32, VP]),
        "gol_world_connect": (I32, [VP]),
        "gol_world_set_stream": (I32, [VP, I32, VP]),
        "gol_world_init_pattern": (I32, [VP, I32, U64, C.c_double]),
        "gol_world_load_bytes": (I32, [VP, I32, VP]),
        "gol_world_load_packed": (I32, [VP, I32, VP]),
        "gol_world_read_bytes": (I32, [VP, I32, VP]),
        "gol_world_read_packed": (I32, [VP, I32, VP]),
        "gol_world_population": (I32, [VP, I32, P(U64)]),
        "gol_world_load_packed_async": (I32, [VP, I32, VP]),
        "gol_world_population_async": (I32, [VP, I32, VP]),
        "gol_world_read_packed_async": (I32, [VP, I32, VP]),
        "gol_world_run_from_host_async": (I32, [VP, P(VP), I64, P(VP)]),
        "gol_world_wait_result": (I32, [VP, I32]),
        "gol_world_step": (I32, [VP, I64]),
        "gol_world_sync": (I32, [VP]),
        "gol_world_step_timed": (I32, [VP, I64, P(C.c_float)]),
        "gol_world_phase_counters": (I32, [VP, I32, P(_PhaseCounters), I32]),
        "gol_world_device_ptr": (VP, [VP, I32]),
        "gol_world_device_of": (I32, [VP, I32]),
        "gol_world_cpu_get_edge": (I32, [VP, I32, I32, VP]),
        "gol_world_cpu_set_ghost": (I32, [VP, I32, I32, VP]),
        "gol_world_cpu_step_local": (I32, [VP]),
        "gol_world_dump_text": (I32, [VP, I32, C.c_char_p]),
        "gol_launch_count": (U64, []),
        "gol_cuda_pack": (I32, [VP, VP, I64, I64, VP]),
        "gol_cuda_unpack": (I32, [VP, VP, I64, I64, VP]),
        "gol_cuda_population_scratch_bytes": (C.c_size_t, []),
        "gol_cuda_population": (I32, [VP, I64, I64, VP, VP, VP]),
        "gol_cuda_step_torus": (I32, [VP, VP, I64, I64, I32, VP]),
        "gol_cuda_step_rows": (I32, [VP, VP, VP, VP, I64, I64, I32, VP]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()


def last_error() -> str:
    return (lib.gol_last_error() or b"").decode("utf-8", "replace")


def check(rc: int) -> None:
    if rc != 0:
        raise GolError(f"[{rc}] {last_error()}")


def version() -> str:
    return lib.gol_version().decode()


def debug_bounds_enabled() -> bool:
    """True when the loaded library is the bounds-checking build (``GOL_B200_LIB=debug``)."""
    return bool(lib.gol_debug_bounds_enabled())


def debug_bounds_report(reset: bool = False) -> dict:
    n, bad, line = C.c_uint64(0), C.c_uint64(0), C.c_int(0)
    check(lib.gol_debug_bounds_report(C.byref(n), C.byref(bad), C.byref(line), int(reset)))
    return {"violations": int(n.value), "first_bad_address": int(bad.value), "source_line": int(line.value)}


def launch_count() -> int:
    """Kernels launched by the native core in this process (graph replays count per node)."""
    return int(lib.gol_launch_count())


def words_per_row(width: int) -> int:
    return int(lib.gol_words_per_row(width))


def row_stride_words(width: int) -> int:
    return int(lib.gol_row_stride_words(width))


def _ptr(buf) -> int:
    """Raw address of a numpy array, torch tensor or integer address."""
    if isinstance(buf, int):
        return buf
    if isinstance(buf, np.ndarray):
        return buf.ctypes.data
    if hasattr(buf, "data_ptr"):
        return buf.data_ptr()
    raise TypeError(f"cannot take the address of {type(buf)!r}")


class _CudaView:
    """Minimal __cuda_array_interface__ carrier for zero-copy torch views."""

    def __init__(self, ptr: int, shape, typestr: str, owner):
        self.__cuda_array_interface__ = {
            "shape": tuple(shape), "typestr": typestr, "data": (ptr, False), "version": 3, "strides": None,
        }
        self._owner = owner


class World:
    """One process's share of the ring: ``nlocal`` row slabs of a ``width`` x ``rows*nranks`` torus.

    Mirrors what ``gol_initMaster`` + ``init_Ghost_rows`` + the generation loop do
    in the reference (gol-with-cuda.cu:286-328, 32-53; gol-main.c:94-116), as an
    object instead of globals.
    """

    def __init__(self, width: int, rows: Optional[int] = None, nranks: int = 1, first_rank: int = 0,
                 nlocal: Optional[int] = None, backend: str = "cuda", mode: str = "auto", k: int = 0,
                 depth: int = 0, threads: int = 0, ndevices: int = 0, device_base: int = 0,
                 flags: int = 0):
        cfg = _Config()
        cfg.width = int(width)
        cfg.rows = int(rows if rows is not None else width)
        cfg.nranks = int(nranks)
        cfg.first_rank = int(first_rank)
        cfg.nlocal = int(nlocal if nlocal is not None else nranks - first_rank)
        cfg.backend = BACKEND_CUDA if backend == "cuda" else BACKEND_CPU
        cfg.mode = MODES[mode] if isinstance(mode, str) else int(mode)
        cfg.temporal_k = int(k)
        cfg.exchange_depth = int(depth)
        cfg.threads_per_block = int(threads)
        cfg.ndevices = int(ndevices)
        cfg.device_base = int(device_base)
        cfg.flags = int(flags)
        self._cfg = cfg
        self._h = lib.gol_world_create(C.byref(cfg))
        if not self._h:
            raise GolError(last_error())
        self.width, self.rows = cfg.width, cfg.rows
        self.nranks, self.first_rank, self.nlocal = cfg.nranks, cfg.first_rank, cfg.nlocal
        self.backend = backend
        self.stride = row_stride_words(self.width)
        self.words = words_per_row(self.width)

    # -- lifetime
    def close(self) -> None:
        if getattr(self, "_h", None):
            lib.gol_world_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- plan / wiring
    @property
    def plan(self) -> dict:
        p = _Plan()
        check(lib.gol_world_plan(self._h, C.byref(p)))
        d = {name: getattr(p, name) for name, _ in _Plan._fields_}
        d["mode_name"] = MODE_NAMES.get(p.mode, "?")
        return d

    def export_blob(self, local: int = 0) -> bytes:
        buf = C.create_string_buffer(EXPORT_BYTES)
        check(lib.gol_world_export(self._h, local, buf))
        return buf.raw

    def import_blob(self, rank: int, blob: bytes) -> None:
        buf = C.create_string_buffer(bytes(blob), EXPORT_BYTES)
        check(lib.gol_world_import(self._h, rank, buf))

    def connect(self) -> "World":
        check(lib.gol_world_connect(self._h))
        return self

    def set_stream(self, local: int, cuda_stream: int) -> None:
        check(lib.gol_world_set_stream(self._h, local, C.c_void_p(cuda_stream)))

    # -- state
    def init_pattern(self, pattern: int, seed: int = 1, density: float = 0.5) -> None:
        check(lib.gol_world_init_pattern(self._h, int(pattern), int(seed), float(density)))

    def load_bytes(self, local: int, cells) -> None:
        a = np.ascontiguousarray(np.asarray(cells, dtype=np.uint8))
        if a.size != self.width * self.rows:
            raise ValueError(f"expected {self.rows}x{self.width} cells, got {a.shape}")
        check(lib.gol_world_load_bytes(self._h, local, a.ctypes.data))

    def read_bytes(self, local: int = 0) -> np.ndarray:
        out = np.empty((self.rows, self.width), dtype=np.uint8)
        check(lib.gol_world_read_bytes(self._h, local, out.ctypes.data))
        return out

    def _check_packed(self, words, what: str) -> None:
        """A packed slab is rows x stride 64-bit words, contiguous, in HOST memory."""
        if isinstance(words, int):
            return                      # raw address: the caller vouches for it
        n = self.rows * self.stride
        if isinstance(words, np.ndarray):
            if words.dtype.itemsize != 8 or words.size != n or not words.flags["C_CONTIGUOUS"]:
                raise ValueError(f"{what}: expected a contiguous array of {self.rows}x{self.stride} 64-bit words, "
                                 f"got shape {words.shape} dtype {words.dtype}")
            return
        if hasattr(words, "data_ptr"):  # torch tensor
            if words.is_cuda:
                raise ValueError(f"{what}: expected host memory, got a CUDA tensor")
            if words.element_size() != 8 or words.numel() != n or not words.is_contiguous():
                raise ValueError(f"{what}: expected a contiguous tensor of {self.rows}x{self.stride} 64-bit words, "
                                 f"got shape {tuple(words.shape)} dtype {words.dtype}")
            return
        raise TypeError(f"{what}: cannot take the address of {type(words)!r}")

    def load_packed(self, local: int, words, asynchronous: bool = False) -> None:
        """Upload a packed slab (rows x stride uint64, interleaved layout).  Bits beyond column
        ``width`` in the last word of a row must be zero (the kernels rely on it); numpy inputs are
        checked, other inputs are the caller's responsibility."""
        self._check_packed(words, "load_packed")
        if isinstance(words, np.ndarray) and self.width % 64:
            last = words.reshape(self.rows, self.stride)[:, self.words - 1].view(np.uint64)
            t = self.width % 64
            mask = np.uint64(((1 << ((t + 1) >> 1)) - 1) | (((1 << (t >> 1)) - 1) << 32))
            if np.any(last & ~mask):
                raise ValueError("load_packed: bits set beyond the last column of a row")
        fn = lib.gol_world_load_packed_async if asynchronous else lib.gol_world_load_packed
        check(fn(self._h, local, _ptr(words)))

    def read_packed_async(self, local: int, pinned_out) -> None:
        """Queue a device-to-host copy of the current packed slab into pinned host memory."""
        self._check_packed(pinned_out, "read_packed_async")
        check(lib.gol_world_read_packed_async(self._h, local, _ptr(pinned_out)))

    def run_from_host_async(self, pinned_in: Sequence, gens: int, pinned_out: Sequence) -> int:
        """One pipelined end-to-end step (see ``gol_world_run_from_host_async``): returns the slot to
        pass to :meth:`wait_result`."""
        if len(pinned_in) != self.nlocal or len(pinned_out) != self.nlocal:
            raise ValueError(f"expected {self.nlocal} input and output buffers")
        for b in list(pinned_in) + list(pinned_out):
            self._check_packed(b, "run_from_host_async")
        ins = (C.c_void_p * self.nlocal)(*[_ptr(b) for b in pinned_in])
        outs = (C.c_void_p * self.nlocal)(*[_ptr(b) for b in pinned_out])
        slot = lib.gol_world_run_from_host_async(self._h, ins, int(gens), outs)
        if slot < 0:
            raise GolError(f"[{slot}] {last_error()}")
        return int(slot)

    def wait_result(self, slot: int) -> None:
        check(lib.gol_world_wait_result(self._h, int(slot)))

    def read_packed(self, local: int = 0) -> np.ndarray:
        out = np.empty((self.rows, self.stride), dtype=np.uint64)
        check(lib.gol_world_read_packed(self._h, local, out.ctypes.data))
        return out

    def population(self, local: int = 0) -> int:
        v = C.c_uint64(0)
        check(lib.gol_world_population(self._h, local, C.byref(v)))
        return int(v.value)

    def population_async(self, local: int, pinned_out) -> None:
        check(lib.gol_world_population_async(self._h, local, _ptr(pinned_out)))

    # -- stepping
    def step(self, gens: int) -> None:
        check(lib.gol_world_step(self._h, int(gens)))

    def sync(self) -> None:
        check(lib.gol_world_sync(self._h))

    def step_timed(self, gens: int) -> List[float]:
        ms = (C.c_float * self.nlocal)()
        check(lib.gol_world_step_timed(self._h, int(gens), ms))
        return list(ms)

    def phase_counters(self, local: int = 0, reset: bool = False) -> dict:
        """Device-side per-phase counters (see ``gol_phase_counters_t``): exposed halo wait per side,
        strip-dependency wait and edge / interior task time of the persistent kernel."""
        pc = _PhaseCounters()
        check(lib.gol_world_phase_counters(self._h, local, C.byref(pc), int(reset)))
        return {
            "halo_wait_ns": [int(pc.halo_wait_ns[0]), int(pc.halo_wait_ns[1])],
            "halo_waits": [int(pc.halo_waits[0]), int(pc.halo_waits[1])],
            "dep_wait_ns": int(pc.dep_wait_ns),
            "task_ns": {"edge": int(pc.task_ns[0]), "interior": int(pc.task_ns[1])},
            "tasks": {"edge": int(pc.tasks[0]), "interior": int(pc.tasks[1])},
        }

    # -- views / io
    def device_ptr(self, local: int = 0) -> int:
        return int(lib.gol_world_device_ptr(self._h, local) or 0)

    def device_of(self, local: int = 0) -> int:
        return int(lib.gol_world_device_of(self._h, local))

    def device_tensor(self, local: int = 0):
        """Zero-copy ``torch.int64`` view (rows x stride) of the current packed slab."""
        import torch

        if self.backend != "cuda":
            raise GolError("device_tensor needs the CUDA backend")
        view = _CudaView(self.device_ptr(local), (self.rows, self.stride), "<i8", self)
        with torch.cuda.device(self.device_of(local)):
            return torch.as_tensor(view, device=f"cuda:{self.device_of(local)}")

    def dump_text(self, local: int = 0, directory: Optional[str] = None) -> None:
        check(lib.gol_world_dump_text(self._h, local, directory.encode() if directory else None))

    # -- CPU transport hooks
    def cpu_get_edge(self, local: int, bottom: bool) -> np.ndarray:
        row = np.empty(self.stride, dtype=np.uint64)
        check(lib.gol_world_cpu_get_edge(self._h, local, int(bottom), row.ctypes.data))
        return row

    def cpu_set_ghost(self, local: int, bottom: bool, row) -> None:
        a = np.ascontiguousarray(np.asarray(row, dtype=np.uint64))
        check(lib.gol_world_cpu_set_ghost(self._h, local, int(bottom), a.ctypes.data))

    def cpu_step_local(self) -> None:
        check(lib.gol_world_cpu_step_local(self._h))


# ---- thin functional wrappers over the CPU helpers -------------------------------------

class _PassDesc(C.Structure):
    _fields_ = [(n_, C.c_int32) for n_ in ("prime", "generic", "prime_depth", "k", "ext_in", "consume", "push",
                                           "reserved")]


def describe_schedule(width: int, rows: int, gens: int, mode: str = "auto", k: int = 0, depth: int = 0) -> dict:
    """The launch sequence ``step(gens)`` runs for this geometry (no GPU needed): a list of dicts
    (``prime`` / pass entries, see ``gol_pass_desc_t``) and the resolved plan."""
    cfg = _Config()
    cfg.width, cfg.rows, cfg.nranks, cfg.nlocal = int(width), int(rows), 1, 1
    cfg.backend, cfg.mode, cfg.temporal_k, cfg.exchange_depth = BACKEND_CUDA, MODES[mode], int(k), int(depth)
    plan = _Plan()
    n_items = lib.gol_schedule_describe(C.byref(cfg), int(gens), None, 0, C.byref(plan))
    if n_items < 0:
        raise GolError(f"[{n_items}] {last_error()}")
    buf = (_PassDesc * max(1, n_items))()
    check(min(0, lib.gol_schedule_describe(C.byref(cfg), int(gens), buf, n_items, C.byref(plan))))
    items = [{f: getattr(buf[i], f) for f, _ in _PassDesc._fields_ if f != "reserved"} for i in range(n_items)]
    return {"items": items, "plan": {name: getattr(plan, name) for name, _ in _Plan._fields_}}


def oracle_run(cells: np.ndarray, gens: int) -> np.ndarray:
    """True torus Life on a 2-D uint8 array (the test oracle; CPU, byte per cell)."""
    a = np.ascontiguousarray(cells, dtype=np.uint8)
    out = np.empty_like(a)
    check(lib.gol_oracle_run(a.ctypes.data, out.ctypes.data, a.shape[1], a.shape[0], int(gens)))
    return out


def pattern_fill(pattern: int, width: int, rows: int, rank: int = 0, nranks: int = 1, seed: int = 1,
                 density: float = 0.5) -> np.ndarray:
    out = np.empty((rows, width), dtype=np.uint8)
    check(lib.gol_pattern_fill_bytes(out.ctypes.data, pattern, width, rows, rank, nranks, seed, density))
    return out


def pack(cells: np.ndarray) -> np.ndarray:
    a = np.ascontiguousarray(cells, dtype=np.uint8)
    out = np.zeros((a.shape[0], row_stride_words(a.shape[1])), dtype=np.uint64)
    check(lib.gol_pack_bytes(a.ctypes.data, out.ctypes.data, a.shape[1], a.shape[0]))
    return out


def unpack(words: np.ndarray, width: int) -> np.ndarray:
    a = np.ascontiguousarray(words, dtype=np.uint64)
    out = np.empty((a.shape[0], width), dtype=np.uint8)
    check(lib.gol_unpack_words(a.ctypes.data, out.ctypes.data, width, a.shape[0]))
    return out


def cpu_run_packed(words: np.ndarray, width: int, gens: int) -> np.ndarray:
    a = np.array(words, dtype=np.uint64, order="C", copy=True)
    check(lib.gol_cpu_run_packed(a.ctypes.data, width, a.shape[0], int(gens)))
    return a


def write_rank_file(path: str, cells: np.ndarray, rank: int) -> None:
    a = np.ascontiguousarray(cells, dtype=np.uint8)
    check(lib.gol_write_rank_file(path.encode(), a.ctypes.data, a.shape[1], a.shape[0], rank))


def rank_file_name(rank: int, nranks: int) -> str:
    buf = C.create_string_buffer(64)
    check(lib.gol_rank_file_name(buf, 64, rank, nranks))
    return buf.value.decode()


def format_report(seconds: float, nranks: int, width: int, rows: int, iterations: int) -> str:
    buf = C.create_string_buffer(512)
    check(lib.gol_format_report(buf, 512, seconds, nranks, width, rows, iterations))
    return buf.value.decode()
