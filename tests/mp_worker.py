"""torchrun worker for the multi-GPU tests: one process per GPU, CUDA-IPC peer wiring."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
import torch
import torch.distributed as dist

from gol_b200 import Life
from gol_b200 import _native as n
from gol_b200.models.patterns import global_pattern, pattern_bytes
from gol_b200.parallel.nccl_halo import NcclHaloStepper


def main() -> int:
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    cases = [
        # mode, k, depth, W, H, gens, pattern
        ("fused", 4, 4, 2048, 256, 37, 6),
        ("fused", 4, 16, 1024, 128, 50, 6),
        ("fused", 8, 8, 4096, 512, 24, 6),
        ("fused", 1, 1, 256, 64, 19, 4),
        ("tma", 4, 4, 10240, 256, 29, 6),
        ("persistent", 4, 4, 2048, 256, 43, 6),
        ("resident", 4, 4, 4096, 512, 43, 6),
        ("resident", 3, 3, 2048, 200, 20, 5),
        ("persistent", 2, 8, 1024, 128, 40, 5),
        ("generic", 0, 0, 200, 33, 9, 6),
        ("auto", 0, 0, 256, 256, 24, 5),
        ("auto", 0, 0, 256, 256, 24, 3),
    ]
    for (mode, k, depth, W, H, gens, pattern) in cases:
        life = Life(W, H, mode=mode, k=k, depth=depth, distributed=True)
        life.init(pattern, seed=5, density=0.5).sync()
        dist.barrier()
        # two step() calls so that the second starts from a state left by the first
        life.step(gens // 2)
        life.step(gens - gens // 2)
        life.sync()
        got = life.cells(0)
        g = global_pattern(pattern, W, H, world, seed=5, density=0.5)
        want = n.unpack(n.cpu_run_packed(n.pack(g), W, gens), W)[rank * H:(rank + 1) * H]
        bad = int((got != want).sum())
        assert bad == 0, f"rank {rank}: {mode} k={k} depth={depth} {W}x{H}: {bad} cells differ"
        pop = life.population()
        assert pop == int(n.unpack(n.cpu_run_packed(n.pack(g), W, gens), W).sum())
        dist.barrier()
        life.close()
        dist.barrier()

    # NCCL-halo baseline (library collectives + unfused stencil) must agree too
    W, H, gens = 512, 64, 12
    st = NcclHaloStepper(W, H)
    st.load(torch.from_numpy(pattern_bytes(6, W, H, rank, world, seed=9)).cuda())
    st.step(gens)
    g = global_pattern(6, W, H, world, seed=9)
    want = n.oracle_run(g, gens)[rank * H:(rank + 1) * H]
    assert (st.cells().cpu().numpy() == want).all(), "NCCL baseline differs from the oracle"

    # device-timed step, max over ranks
    life = Life(4096, 4096, distributed=True).init(6)
    ms = life.step_timed(64)
    assert ms > 0
    dist.barrier()
    life.close()
    if rank == 0:
        print(f"MP_WORKER_OK world={world}", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
