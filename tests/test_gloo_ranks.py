"""Process-per-rank plumbing over gloo (world_size 2 and 3) — no GPU needed."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
sys.path.insert(0, os.environ["GOL_ROOT"])
import numpy as np, torch.distributed as dist
from gol_b200 import Life
dist.init_process_group("gloo")
life = Life(64, 32, backend="cpu", distributed=True).init(5)
life.step(16)
np.save(os.path.join(os.environ["GOL_OUT"], f"r{dist.get_rank()}.npy"), life.cells(0))
pop = life.population()
life.dump(os.environ["GOL_OUT"])
assert pop == 5, pop
dist.destroy_process_group()
'''


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("nproc", [2, 3])
def test_gloo_ranks_match_oracle(tmp_path, nproc):
    from gol_b200 import _native as n
    from gol_b200.models.patterns import global_pattern

    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, GOL_ROOT=ROOT, GOL_OUT=str(tmp_path), OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(free_port()), str(script)]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-3000:]
    got = np.concatenate([np.load(tmp_path / f"r{i}.npy") for i in range(nproc)])
    want = n.oracle_run(global_pattern(5, 64, 32, nproc), 16)
    assert (got == want).all()
    assert (tmp_path / f"Rank_0_of_{nproc}.txt").exists()
