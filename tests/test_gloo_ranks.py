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


def test_python_cli_under_torchrun_writes_reference_layout(tmp_path):
    """`torchrun -m gol_b200.cli` is the process-per-rank form of the reference's `mpirun ./gol`."""
    from test_writer import expected_text

    from gol_b200 import _native as n
    from gol_b200.models.patterns import global_pattern

    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""), OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", str(free_port()), "-m", "gol_b200.cli",
           "5", "64", "16", "256", "1", "--cpu", "--rows", "32", "--outdir", str(tmp_path)]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-3000:]
    assert "TOTAL DURATION : " in r.stdout and "number of cell updates = 65536" in r.stdout
    assert "This is the Game of Life running in parallel on a GPU on multiple ranks." in r.stdout
    want = n.oracle_run(global_pattern(5, 64, 32, 2), 16)
    for rank in range(2):
        got = (tmp_path / f"Rank_{rank}_of_2.txt").read_text()
        assert got == expected_text(want[rank * 32:(rank + 1) * 32], rank)
