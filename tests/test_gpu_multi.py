"""Multi-GPU tests (need >= 2 devices; skipped otherwise): single-process NVLink P2P and
one-process-per-GPU CUDA-IPC wiring, every step mode, against the CPU stepper."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def ngpu():
    import torch

    return torch.cuda.device_count()


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("mode,k,depth", [("fused", 4, 4), ("fused", 2, 8), ("persistent", 4, 4), ("generic", 0, 0)])
def test_single_process_peer_to_peer(mode, k, depth):
    if ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    from gol_b200 import World
    from gol_b200 import _native as n
    from gol_b200.models.patterns import global_pattern

    R = min(ngpu(), 4)
    W, H, gens = (1024, 128, 29) if mode != "generic" else (200, 31, 7)
    w = World(W, H, nranks=R, backend="cuda", mode=mode, k=k, depth=depth)   # one slab per device
    w.connect()
    assert len({w.device_of(i) for i in range(R)}) == R
    w.init_pattern(6, 3, 0.5)
    w.step(gens)
    w.sync()
    got = np.concatenate([w.read_bytes(i) for i in range(R)])
    want = n.unpack(n.cpu_run_packed(n.pack(global_pattern(6, W, H, R, seed=3)), W, gens), W)
    assert (got == want).all()
    w.close()


def test_process_per_gpu_over_cuda_ipc():
    if ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    nproc = ngpu()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(free_port()),
           os.path.join(ROOT, "tests", "mp_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "MP_WORKER_OK" in r.stdout, (r.stdout[-2000:] + r.stderr[-4000:])


def test_native_cli_process_per_rank_and_thread_per_device(tmp_path):
    """bin/gol --spawn R: one native process per rank (fork, export blobs through shared memory,
    CUDA-IPC mapping), the reference's process model (gol-main.c:58-62) without Python or MPI.
    Its rank files must be byte-identical to the single-process run's (--ranks R: one launch
    thread per device), and both must pass the oracle check."""
    if ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    R = min(ngpu(), 4)
    exe = os.path.join(ROOT, "bin", "gol")
    outs = []
    for flag in ("--spawn", "--ranks"):
        d = tmp_path / flag.strip("-")
        d.mkdir()
        r = subprocess.run([exe, "6", "1024", "53", "256", "1", "--rows", "192", flag, str(R), "--check", "--outdir", str(d)],
                           capture_output=True, text=True, timeout=300)
        assert r.returncode == 0 and "CHECK: OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
        assert r.stdout.count("TOTAL DURATION") == 1
        outs.append([(d / f"Rank_{i}_of_{R}.txt").read_bytes() for i in range(R)])
    assert outs[0] == outs[1]


def test_native_spawn_refuses_to_share_a_gpu():
    exe = os.path.join(ROOT, "bin", "gol")
    r = subprocess.run([exe, "6", "256", "8", "256", "0", "--spawn", str(ngpu() + 1)], capture_output=True, text=True,
                       timeout=120)
    assert r.returncode == 255 and "one GPU per rank" in r.stderr
