"""``python -m gol_b200.cli`` — the reference's five-positional command line (gol-main.c:43-53)
for the process-per-GPU model: launch with ``torchrun --nproc-per-node R`` where the
reference would use ``mpirun -n R``.  Without torchrun it runs ``--ranks`` slabs from one
process, like the native ``bin/gol``.
"""
from __future__ import annotations

import argparse
import os
import sys

USAGE = ("GOL requires 5 arguments: pattern number, sq size of the world and the number of itterations, "
         "threads per block and output-on-off e.g. ./gol 0 32 2 512 0 \n")


def main(argv=None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    ap = argparse.ArgumentParser(add_help=False)
    ap.add_argument("pos", nargs="*")
    ap.add_argument("--ranks", type=int, default=None)
    ap.add_argument("--rows", type=int, default=None)
    ap.add_argument("--mode", default="auto")
    ap.add_argument("--k", type=int, default=0)
    ap.add_argument("--depth", type=int, default=0)
    ap.add_argument("--density", type=float, default=0.5)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--outdir", default=None)
    ap.add_argument("--cpu", action="store_true")
    ns = ap.parse_args(argv)
    if len(ns.pos) != 5:          # gol-main.c:43-47: usage before any runtime is initialised
        sys.stdout.write(USAGE)
        return 255
    pattern, size, iters, threads, on_off = (int(v) for v in ns.pos)
    if pattern < 0 or pattern >= 7:
        sys.stdout.write(f"Pattern {pattern} has not been implemented \n")   # gol-with-cuda.cu:324-326
        return 255

    import torch
    import torch.distributed as dist

    from . import GolError, Life

    under_launcher = "RANK" in os.environ and "WORLD_SIZE" in os.environ and ns.ranks is None
    if under_launcher:
        backend = "gloo" if ns.cpu else "nccl"
        if not ns.cpu:
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
        dist.init_process_group(backend)
    rank = dist.get_rank() if under_launcher else 0
    try:
        life = Life(size, ns.rows, ranks=None if under_launcher else (ns.ranks or 1),
                    backend="cpu" if ns.cpu else "cuda", mode=ns.mode, k=ns.k, depth=ns.depth,
                    threads=threads, distributed=under_launcher)
        life.init(pattern, ns.seed, ns.density).sync()
        ms = life.step_timed(iters)
        if rank == 0:
            sys.stdout.write(life.report(ms * 1e-3, iters))
        if on_off == 1:
            life.dump(ns.outdir)
        life.close()
    except GolError as e:
        sys.stderr.write(f"gol: {e}\n")
        return 255
    finally:
        if under_launcher and dist.is_initialized():
            dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
