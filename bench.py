#!/usr/bin/env python
"""Headline benchmark: cell updates per second (GCUPS) on N B200s, device-timed, max over ranks.

Configs (BASELINE.json):
  --scaling weak   (default, config 2) 16384 x 16384 cells per GPU, random 50 % density
  --scaling strong (config 3)          65536 x 65536 global grid, 65536 / N rows per GPU
One "step" = 1024 generations of the whole torus (halo exchange fused into the step kernel).

`--impl reference` runs the UNMODIFIED reference sources built by baseline/build_ref.sh through
the reference's own CLI and MPI loop (`mpirun -n N gol <pattern> <size> <gens> 512 0`) at the same
grid size and generation count.  Its default variant is the build with the one debugging
`printf` line (gol-with-cuda.cu:223) deleted at build time: as shipped that line makes one
generation of 16384^2 take hours, so the as-shipped binary is timed at 512^2 x 1 generation and
reported in the side field `as_shipped` (our arm reports the same small config in
`as_shipped_config`).  The reference arm imports nothing of this repository's package.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference|nccl] [--scaling weak|strong]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
"""
from __future__ import annotations

import argparse
import importlib.util
import json
import os
import re
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))

SIZE = 16384             # weak scaling: cells per GPU = SIZE x SIZE
STRONG_SIZE = 65536      # strong scaling: global grid = STRONG_SIZE x STRONG_SIZE
GENS_PER_STEP = 1024
METRIC = "cell updates/sec (GCUPS, device-timed, max over ranks)"
ALU_OPS_PER_64_CELLS = 22   # 20 LOP3 + 2 SHF per generation (csrc/kernels.cu, push_row)


def env_rank():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))


def load_clock_sampler():
    """utils/clocks.py loaded by path: importing the gol_b200 package would map libgol_b200.so,
    which must not happen in the reference arm's process."""
    path = os.path.join(ROOT, "game-of-life---mpi-cuda_b200", "utils", "clocks.py")
    spec = importlib.util.spec_from_file_location("_gol_bench_clocks", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.ClockSampler


def geometry(args, n: int):
    """(width, rows per GPU, scaling label)."""
    if args.scaling == "strong":
        width = args.size or STRONG_SIZE
        if width % n:
            raise SystemExit(f"strong scaling: {width} rows do not divide over {n} GPUs")
        return width, width // n, "strong"
    width = args.size or SIZE
    return width, (args.rows or width), "weak"


# ---------------------------------------------------------------------------- reference arm
def run_reference(args) -> int:
    rank, world, _ = env_rank()
    if world > 1:
        # launched under torchrun like our arm: rank 0 drives the reference's own launcher
        import datetime

        import torch.distributed as dist

        # the other ranks only wait for rank 0: the reference run may take a quarter of an hour
        dist.init_process_group("gloo", timeout=datetime.timedelta(hours=3))
    out = {"impl": "reference"}
    if rank == 0:
        out = reference_measure(args)
    if world > 1:
        import torch.distributed as dist

        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(out), flush=True)
    return 0


def _ref_run(n: int, exe: str, pattern: int, size: int, iters: int, timeout_s: float, rows: int = 0):
    """One run of the reference's own CLI under its launcher.
    Returns (its TOTAL DURATION in seconds, wall-clock seconds of the whole process)."""
    import tempfile

    exe = os.path.join(ROOT, "baseline", "_ref", "bin", exe)
    mpirun = os.path.join(ROOT, "baseline", "_ref", "bin", "mpirun")
    cmd = [mpirun, "-n", str(n), exe, str(pattern), str(size), str(iters), "512", "0"]
    env = dict(os.environ)
    if rows:
        env["GOL_REF_ROWS"] = str(rows)     # read by the rectangular-slab variant only
    # stdout carries the device-printf flood (gol-with-cuda.cu:223): spool it, keep only the tail
    t0 = time.perf_counter()
    with tempfile.TemporaryFile() as fo:
        p = subprocess.run(cmd, stdout=fo, stderr=subprocess.PIPE, timeout=timeout_s, env=env)
        size_b = os.fstat(fo.fileno()).st_size
        fo.seek(max(0, size_b - (1 << 16)))
        tail = fo.read()
    wall = time.perf_counter() - t0
    if p.returncode != 0:
        raise RuntimeError(f"reference exited {p.returncode}: {p.stderr[-500:]!r}")
    m = re.search(rb"TOTAL DURATION : ([0-9.]+), number of cell updates = (\d+)", tail)
    if not m:
        raise RuntimeError("reference printed no TOTAL DURATION line")
    return float(m.group(1)), wall


def _ensure_reference_built() -> bool:
    bindir = os.path.join(ROOT, "baseline", "_ref", "bin")
    need = ("gol_ref", "gol_ref_noprintf", "gol_ref_rect", "gol_ref_rect_noprintf", "mpirun")
    have = lambda: all(os.path.exists(os.path.join(bindir, b)) for b in need)
    if not have():
        # binaries missing (fresh checkout): rebuild from the verbatim copies or from /root/reference
        for src in (os.path.join(ROOT, "baseline", "_ref", "src"), "/root/reference"):
            if os.path.exists(os.path.join(src, "gol-with-cuda.cu")):
                subprocess.run(["bash", os.path.join(ROOT, "baseline", "build_ref.sh"), src],
                               stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
                break
    return have()


def reference_measure(args) -> dict:
    """The reference arm: same grid, same generation count, the reference's own code path."""
    if not _ensure_reference_built():
        return {"impl": "reference", "unavailable": "baseline/_ref/bin is missing: run baseline/build_ref.sh "
                "(pip cannot install the reference: it has no setup.py/pyproject.toml)"}
    ClockSampler = load_clock_sampler()
    n = args.gpus
    width, rows, scaling = geometry(args, n)
    gens = args.gens
    budget = float(os.environ.get("GOL_REF_BUDGET_S", "1000"))   # wall-clock seconds for the timed reference steps
    noprintf = args.ref_variant == "noprintf"
    if scaling == "strong":
        # the reference CLI only knows square slabs (gol-main.c:76); the rectangular variant passes
        # (W, H) = (width, GOL_REF_ROWS) to the unchanged gol_initMaster (gol-with-cuda.cu:286).
        # Pattern 4: the host fill loops of patterns 1 and 2 index with int and break above 2^31.
        exe, pattern = ("gol_ref_rect_noprintf" if noprintf else "gol_ref_rect"), 4
        if width * rows > (1 << 31):
            return {"impl": "reference", "n_gpus": n, "scaling": scaling,
                    "unavailable": f"the reference cannot run a {width}x{rows} slab: its kernel bounds the cell "
                                   "index with a 32-bit product (gol-with-cuda.cu:198), which wraps beyond 2^31 cells; "
                                   "see BASELINE.md section 3 for its 32768^2 single-GPU rate"}
    else:
        exe, pattern = ("gol_ref_noprintf" if noprintf else "gol_ref"), 1
    out = {"impl": "reference", "metric": METRIC, "unit": "GCUPS", "n_gpus": n, "warmup": args.warmup,
           "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "uint8 cells",
           "data": f"synthetic (reference pattern {pattern}: it has no random pattern; its per-cell work is "
                   "data-independent)"}
    sampler = ClockSampler(period_ms=200).start()
    try:
        as_shipped = None
        if noprintf:
            # calibrate: a short run gives the per-generation cost, warm-up runs are short as well
            # (every step is a fresh process, there is no state to warm)
            cal_gens = 8
            t_cal, wall_cal = _ref_run(n, exe, pattern, width, cal_gens, 900, rows if scaling == "strong" else 0)
            for _ in range(max(0, min(args.warmup, 2) - 1)):
                _ref_run(n, exe, pattern, width, cal_gens, 900, rows if scaling == "strong" else 0)
            # a step costs the process start-up (managed allocation, host-side pattern fill: 10-17 s with
            # eight ranks) plus the generation loop
            est = (wall_cal - t_cal) + t_cal / cal_gens * gens * 1.1 + 1.0
            steps_run = int(max(1, min(args.steps, budget // est)))
            timer_s = wall_s = 0.0
            t_lo = time.time()
            for _ in range(steps_run):
                d, wl = _ref_run(n, exe, pattern, width, gens, max(300.0, 6 * est),
                                 rows if scaling == "strong" else 0)
                timer_s += d
                wall_s += wl
            t_hi = time.time()
            size_run, gens_run = width, gens
            if scaling == "weak":
                # side figure: the binary exactly as shipped, at the largest size that finishes
                try:
                    s_small = int(os.environ.get("GOL_REF_SIZE", "512"))
                    d, wl = _ref_run(n, "gol_ref", 1, s_small, 1, 300)
                    as_shipped = {"value": float(n) * s_small * s_small / d * 1e-9, "unit": "GCUPS",
                                  "cells_per_gpu": f"{s_small}x{s_small}", "generations_per_step": 1, "steps": 1,
                                  "seconds": d, "wall_seconds": wl,
                                  "note": "unmodified binary incl. printf(\"BEFORE AND\") from every cell "
                                          "(gol-with-cuda.cu:223): ~35-65 us per cell update, so one generation of "
                                          "16384x16384 would take hours"}
                except Exception as e:  # noqa: BLE001
                    as_shipped = {"unavailable": str(e)}
        else:
            size_run = int(os.environ.get("GOL_REF_SIZE", "512"))
            gens_run = args.ref_gens
            est = size_run * size_run * gens_run * 36e-6 + 1.5        # measured ~35 us per cell update
            steps_run = int(max(1, min(args.steps, budget // est)))
            timer_s = wall_s = 0.0
            t_lo = time.time()
            for _ in range(steps_run):
                d, wl = _ref_run(n, "gol_ref", 1, size_run, gens_run, max(120.0, 4 * est))
                timer_s += d
                wall_s += wl
            t_hi = time.time()
            rows = size_run
    except Exception as e:  # noqa: BLE001
        sampler.stop()
        return {"impl": "reference", "n_gpus": n, "scaling": scaling, "unavailable": f"reference run failed: {e}"}
    clocks = sampler.stop(t_lo, t_hi)
    rows_run = rows if scaling == "strong" else size_run
    updates = float(n) * size_run * rows_run * gens_run * steps_run
    out.update({
        "value": updates / timer_s * 1e-9, "steps": steps_run, "ms_per_step": timer_s / steps_run * 1e3,
        "config": {"model": "conway-life-torus", "cells_per_gpu": f"{size_run}x{rows_run}",
                   "generations_per_step": gens_run, "global_batch": n * size_run * rows_run,
                   "parallelism": f"row-slab ring x{n}, the reference's MPI loop (minimpi shared-memory MPI, "
                                  "baseline tooling: the image has no MPI)",
                   "variant": "printf line (gol-with-cuda.cu:223) deleted at build time, nothing else changed"
                              if noprintf else "as shipped",
                   "timer": "the reference's own TOTAL DURATION (rank-0 MPI_Wtime around its generation loop: "
                            "host wall clock incl. MPI, launches and one cudaDeviceSynchronize per generation; "
                            "our arm is timed with CUDA events, max over ranks)",
                   "steps_requested": args.steps,
                   "l2": "managed memory, 256 MiB per byte grid per rank: larger than L2"},
        "clocks": {k: clocks.get(k) for k in ("sm_mhz", "sm_max_mhz", "reasons", "samples", "scope")},
        "e2e": {"value": updates / wall_s * 1e-9, "unit": "GCUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                "ms_per_step": wall_s / steps_run * 1e3,
                "note": "host wall clock around each whole reference process (start-up, managed allocation, "
                        "pattern fill, generation loop, teardown); managed memory, no explicit copies"},
        "gpu_launches": gens_run * steps_run * n,
    })
    if as_shipped is not None:
        out["as_shipped"] = as_shipped
    return out


# ---------------------------------------------------------------------------- our arm
def _pin_to_gpu_numa(local_rank: int) -> str:
    """Run this rank's host thread (and first-touch its pinned staging) on the CPUs next to its GPU."""
    try:
        import pynvml

        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = {64 * i + b for i, wd in enumerate(words) for b in range(64) if (int(wd) >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if cpus:
            os.sched_setaffinity(0, cpus)
            return f"{len(cpus)} CPUs next to GPU {local_rank}"
    except Exception as e:  # noqa: BLE001
        return f"not pinned ({type(e).__name__})"
    return "not pinned"


def run_ours(args) -> int:
    rank, world, local_rank = env_rank()
    if args.gpus > 1 and world != args.gpus:
        # single-process launch for N > 1: re-exec under torchrun so that it is one rank per GPU
        import socket

        with socket.socket() as s:
            s.bind(("127.0.0.1", 0))
            port = s.getsockname()[1]
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd)
    # nvidia-smi takes a second or two to deliver its first sample: launch it before anything else
    ClockSampler = load_clock_sampler()
    sampler = ClockSampler(gpu_index=local_rank, period_ms=50).start() if rank == 0 else None
    numa = _pin_to_gpu_numa(local_rank)

    import numpy as np
    import torch
    import torch.distributed as dist

    sys.path.insert(0, ROOT)
    torch.cuda.set_device(local_rank)
    distributed = world > 1
    if distributed:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    import gol_b200
    from gol_b200 import Life
    from gol_b200 import _native as nat
    from gol_b200.models.patterns import pattern_bytes
    from gol_b200.ops.oracle import numpy_life_run
    from gol_b200.utils.timing import flush_l2

    def barrier():
        if distributed:
            dist.barrier()
        torch.cuda.synchronize()

    def all_ok(flag: bool) -> bool:
        t = torch.tensor([1 if flag else 0], dtype=torch.int32, device="cuda")
        if distributed:
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return bool(int(t.item()))

    # ---- correctness gate on the very path that is timed below (same mode, k, depth, ring wiring):
    # a small torus through Life -> compared with an independent numpy stepper
    vW, vH, vG = (10240 if args.mode == "tma" else 2048), 256, 37   # mode tma needs >= 144 words per row
    vlife = Life(vW, vH, mode=args.mode, k=args.k, depth=args.depth, threads=args.threads, distributed=distributed)
    vlife.init(6, seed=99, density=0.5)
    vlife.step(20).step(vG - 20).sync()
    got = vlife.cells(0)
    vlife.close()
    init = np.concatenate([pattern_bytes(6, vW, vH, r, world, seed=99, density=0.5) for r in range(world)])
    want = numpy_life_run(init, vG)[rank * vH:(rank + 1) * vH]
    verified = all_ok(bool((got == want).all()))
    if not verified:
        if rank == 0:
            print(json.dumps({"impl": "ours", "error": "verification against the numpy oracle failed", "n_gpus": world}),
                  flush=True)
        return 1

    width, rows, scaling = geometry(args, world)
    gens = args.gens
    life = Life(width, rows, mode=args.mode, k=args.k, depth=args.depth, threads=args.threads,
                distributed=distributed)
    stream = torch.cuda.Stream()          # a capturable (non-legacy) stream shared by torch and the core
    torch.cuda.set_stream(stream)
    life.world.set_stream(0, stream.cuda_stream)
    life.init(6, seed=1234, density=0.5).sync()
    plan = life.plan

    # pinned host staging for the end-to-end arm: this step's input and two result buffers
    stride = life.world.stride
    host_in = torch.empty((rows, stride), dtype=torch.int64).pin_memory()
    nat.check(nat.lib.gol_world_read_packed(life.world._h, 0, host_in.data_ptr()))
    host_out = [torch.empty((rows, stride), dtype=torch.int64).pin_memory() for _ in range(2)]
    slab_bytes = int(host_in.numel() * 8)

    for _ in range(args.warmup):
        life.step(gens)
    barrier()
    if sampler is not None:
        sampler.wait_first_sample(10.0)
    life.world.phase_counters(0, reset=True)

    # ---- kernel-only: K steps, CUDA events on the launching stream, L2 flushed between steps
    ev0 = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    ev1 = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    launches0 = gol_b200.launch_count()
    t_region0 = time.time()
    for i in range(args.steps):
        flush_l2()
        barrier()
        ev0[i].record()
        life.step(gens)
        ev1[i].record()
    barrier()
    life.sync()
    launches = gol_b200.launch_count() - launches0
    total_ms = sum(a.elapsed_time(b) for a, b in zip(ev0, ev1))
    phase = life.world.phase_counters(0)

    # ---- end to end through the public API.  Every step uploads its input from pinned host memory
    # and brings its packed result grid back to pinned host memory; Life.run_from_host pipelines
    # consecutive steps (copy streams + two staging slots), the host consumes result i while step
    # i+1 is in flight.
    for _ in range(min(args.warmup, 2)):
        life.wait_result(life.run_from_host(host_in, gens, host_out[0]))
    barrier()
    t0 = time.perf_counter()
    checks = []
    prev = None
    for i in range(args.steps):
        slot = life.run_from_host(host_in, gens, host_out[i & 1])
        if prev is not None:
            life.wait_result(prev[0])
            checks.append(int(prev[1][0, 0].item()) ^ int(prev[1][-1, -2].item()))   # the host reads the result
        prev = (slot, host_out[i & 1])
    life.wait_result(prev[0])
    checks.append(int(prev[1][0, 0].item()) ^ int(prev[1][-1, -2].item()))
    barrier()
    e2e_s = time.perf_counter() - t0
    # one step at a time (no overlap between steps): the latency a single call sees
    barrier()
    t0 = time.perf_counter()
    nser = min(args.steps, 5)
    for i in range(nser):
        life.wait_result(life.run_from_host(host_in, gens, host_out[0]))
    barrier()
    serial_ms = (time.perf_counter() - t0) * 1e3 / nser
    t_region1 = time.time()
    life.sync()
    pop = life.population()
    clocks = sampler.stop(t_region0, t_region1) if sampler else {}

    t = torch.tensor([total_ms, e2e_s * 1e3, serial_ms], dtype=torch.float64, device="cuda")
    if distributed:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms, serial_ms = float(t[0]), float(t[1]), float(t[2])
    updates = float(world) * width * rows * gens * args.steps
    value = updates / (total_ms * 1e-3) * 1e-9
    e2e_value = updates / (e2e_ms * 1e-3) * 1e-9
    life.close()

    # ---- the config the as-shipped reference can finish (512 x 512 per rank, one generation per
    # step), through the same API: the like-for-like partner of the reference arm's `as_shipped`
    small = None
    if scaling == "weak" and not args.no_small:
        sl = Life(512, 512, mode=args.mode, distributed=distributed).init(6, seed=5)
        for _ in range(5):
            sl.step(1)
        barrier()
        nsm = 200
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(nsm):
            sl.step(1)
        b.record()
        barrier()
        ts = torch.tensor([a.elapsed_time(b)], dtype=torch.float64, device="cuda")
        if distributed:
            dist.all_reduce(ts, op=dist.ReduceOp.MAX)
        sl.close()
        small = {"value": float(world) * 512 * 512 * nsm / (float(ts[0]) * 1e-3) * 1e-9, "unit": "GCUPS",
                 "cells_per_gpu": "512x512", "generations_per_step": 1, "steps": nsm,
                 "ms_per_step": float(ts[0]) / nsm}

    rc = 0
    if rank == 0:
        from gol_b200.utils.metrics import measured_peaks

        peaks = measured_peaks(ROOT)
        per_gpu = value * 1e9 / world                       # cells / s / GPU
        sm_mhz = clocks.get("sm_mhz") or peaks.get("sm_max_mhz", 1965.0)
        sms = torch.cuda.get_device_properties(local_rank).multi_processor_count
        alu_ceiling = sms * 64 * sm_mhz * 1e6 * 64 / ALU_OPS_PER_64_CELLS      # 64 ALU lanes / clk / SM
        k = max(1, plan["temporal_k"])
        ncg = max(1, plan["col_groups"])
        gens_total = gens * args.steps
        out = {
            "metric": METRIC, "value": value, "unit": "GCUPS", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": scaling, "vs_baseline": None, "dtype": "1-bit cells (uint64 x64 packed)",
            "data": "synthetic (random 50% density, counter-based hash, seed 1234)",
            "verified": verified,
            "config": {"model": "conway-life-torus", "cells_per_gpu": f"{width}x{rows}",
                       "generations_per_step": gens, "global_batch": world * width * rows,
                       "parallelism": f"row-slab ring x{world}, one process per GPU, NVLink P2P halo fused in kernel",
                       "mode": plan["mode_name"], "k": plan["temporal_k"], "exchange_depth": plan["exchange_depth"],
                       "threads_per_block": plan["threads_per_block"], "strip_rows": plan["strip_rows"],
                       "l2": "L2 flushed (256 MiB write) before every timed step; within a step the double "
                             "buffer is L2-resident by design when it fits",
                       "verification": f"{vW}x{vH} per rank, {vG} generations through the same Life path vs a numpy "
                                       "roll stepper, before timing",
                       "host_numa": numa},
            "clocks": {k_: clocks.get(k_) for k_ in ("sm_mhz", "sm_max_mhz", "reasons", "power_w_max", "samples", "scope")},
            "e2e": {"value": e2e_value, "unit": "GCUPS", "h2d_bytes_per_step": slab_bytes,
                    "d2h_bytes_per_step": slab_bytes, "ms_per_step": e2e_ms / args.steps,
                    "serial_ms_per_step": serial_ms,
                    "api": "slot = Life.run_from_host(pinned_in, generations, pinned_out); Life.wait_result(slot) "
                           "(packed input grid up, packed result grid back, every step; consecutive steps pipelined "
                           "over copy streams; serial_ms_per_step = one call at a time)"},
            "gpu_launches": int(launches),
            "impl": "ours",
            "roofline": {"per_gpu_tcells_per_s": per_gpu * 1e-12,
                         "alu_ceiling_tcells_per_s": alu_ceiling * 1e-12,
                         "alu_fraction": per_gpu / alu_ceiling,
                         "alu_model": f"{ALU_OPS_PER_64_CELLS} ALU-pipe instructions per 64 cells per generation, "
                                      f"64 lanes/clk/SM x {sms} SMs at the sampled SM clock",
                         "hbm_single_pass_bound_tcells_per_s": peaks["hbm_gbs"] * 1e9 / 0.25 * 1e-12,
                         "hbm_k_pass_bound_tcells_per_s": peaks["hbm_gbs"] * 1e9 / 0.25 * k * 1e-12,
                         "hbm_fraction_single_pass": per_gpu / (peaks["hbm_gbs"] * 1e9 / 0.25),
                         "hbm_fraction_k_pass": per_gpu / (peaks["hbm_gbs"] * 1e9 / 0.25 * k),
                         "hbm_of": peaks.get("source")},
            "phase": {"halo_wait_us_per_gen": (phase["halo_wait_ns"][0] * 1e-3 / max(1, gens_total))
                      if plan["mode_name"] == "resident" else
                      (phase["halo_wait_ns"][0] + phase["halo_wait_ns"][1]) * 1e-3 / (2 * ncg) / max(1, gens_total),
                      "halo_waits": phase["halo_waits"],
                      "note": "rank 0: time edge warps spent spinning on a neighbour's halo flag, mean over the "
                              "edge warps of both sides, per generation (resident mode: time the CTA at the ring edge "
                              "spent between pushing its rows and having both halos, per generation)"},
            "population_after_run": pop,
            "host_checks": len(checks),
        }
        if small is not None:
            out["as_shipped_config"] = small
        if not clocks.get("samples"):
            out["error"] = "no clock samples were taken during the timed region"
            rc = 3
        print(json.dumps(out), flush=True)
    if distributed:
        dist.barrier()
        dist.destroy_process_group()
    return rc


# ---------------------------------------------------------------------------- NCCL baseline arm
def run_nccl_baseline(args) -> int:
    """Library-collective baseline: ncclSend/ncclRecv halo + unfused stencil kernel per generation
    (gol_b200.parallel.nccl_halo).  Not the product path; measured for comparison only."""
    import torch
    import torch.distributed as dist

    sys.path.insert(0, ROOT)
    rank, world, local_rank = env_rank()
    torch.cuda.set_device(local_rank)
    if "RANK" not in os.environ:      # plain `python bench.py --impl nccl`: a one-rank group
        os.environ.update(RANK="0", WORLD_SIZE="1", LOCAL_RANK="0", MASTER_ADDR="127.0.0.1",
                          MASTER_PORT=os.environ.get("MASTER_PORT", "29533"))
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    from gol_b200.models.patterns import pattern_bytes
    from gol_b200.parallel.nccl_halo import NcclHaloStepper

    width, rows, scaling = geometry(args, world)
    gens = min(args.gens, 64)
    st = NcclHaloStepper(width, rows)
    small = torch.from_numpy(pattern_bytes(6, width, 64, rank, world, seed=1234)).cuda()
    st.load(small.repeat(rows // 64, 1))
    for _ in range(args.warmup):
        st.step(8)
    dist.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(args.steps):
        st.step(gens)
    b.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b)], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t[0])
    if rank == 0:
        updates = float(world) * width * rows * gens * args.steps
        print(json.dumps({"impl": "nccl-baseline", "metric": METRIC, "value": updates / (ms * 1e-3) * 1e-9,
                          "unit": "GCUPS", "n_gpus": world, "steps": args.steps, "ms_per_step": ms / args.steps,
                          "scaling": scaling,
                          "config": {"cells_per_gpu": f"{width}x{rows}", "generations_per_step": gens,
                                     "halo": "dist.batch_isend_irecv (ncclSend/ncclRecv) every generation",
                                     "stencil": "step_rows_kernel (unfused, bit-packed, 1 generation)"}}), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    return 0


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "nccl"])
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: --size x --size cells per GPU (16384); strong: --size x --size global (65536)")
    ap.add_argument("--size", type=int, default=0, help="grid width (default 16384 weak / 65536 strong)")
    ap.add_argument("--rows", type=int, default=0, help="weak scaling: rows per GPU (default: square)")
    ap.add_argument("--gens", type=int, default=GENS_PER_STEP)
    ap.add_argument("--mode", default="auto")
    ap.add_argument("--k", type=int, default=0)
    ap.add_argument("--depth", type=int, default=0)
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--no-small", action="store_true", help="skip the 512x512 one-generation side figure")
    ap.add_argument("--ref-gens", type=int, default=int(os.environ.get("GOL_REF_GENS", "1")),
                    help="--ref-variant as-shipped only: generations per step (it prints from every cell: ~9 s per "
                         "generation already at 512x512)")
    ap.add_argument("--ref-variant", default=os.environ.get("GOL_REF_VARIANT", "noprintf"),
                    choices=["as-shipped", "noprintf"])
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    if args.impl == "reference":
        return run_reference(args)
    if args.impl == "nccl":
        return run_nccl_baseline(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
