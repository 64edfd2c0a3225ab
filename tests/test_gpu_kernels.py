"""GPU tests: every CUDA kernel against the CPU oracle (bit-exact).

Several logical ranks are hosted on ONE device (``ndevices=1``): one launch covers all
slabs and halo waits only ever refer to the previous launch, so the multi-rank protocol
(recv buffers, flags, R = 1 / 2 / >= 3 routing) is exercised without co-scheduling
kernels that wait on one another.
"""
import os

import numpy as np
import pytest

from gol_b200 import GolError, Life, World, launch_count
from gol_b200 import _native as n
from gol_b200.models.patterns import global_pattern

pytestmark = pytest.mark.gpu


def expected(pattern, W, H, R, gens, seed=1, density=0.5):
    g = global_pattern(pattern, W, H, R, seed=seed, density=density)
    return n.unpack(n.cpu_run_packed(n.pack(g), W, gens), W)


def run_gpu(pattern, W, H, R, gens, mode="auto", k=0, depth=0, threads=0, flags=0, seed=1, density=0.5,
            chunks=None):
    w = World(W, H, nranks=R, backend="cuda", mode=mode, k=k, depth=depth, threads=threads, ndevices=1,
              flags=flags)
    w.connect()
    w.init_pattern(pattern, seed, density)
    for g in (chunks or [gens]):
        w.step(g)
    w.sync()
    out = np.concatenate([w.read_bytes(i) for i in range(R)])
    plan = w.plan
    w.close()
    return out, plan


# ------------------------------------------------------------------ stateless ops
def test_pack_unpack_population_step_torus():
    import torch

    from gol_b200 import ops

    rng = np.random.default_rng(3)
    for (H, W) in [(5, 1), (9, 70), (33, 128), (17, 1000), (64, 4096)]:
        g = (rng.random((H, W)) < 0.4).astype(np.uint8)
        cells = torch.from_numpy(g).cuda()
        words = ops.pack(cells)
        assert (words.cpu().numpy().view(np.uint64) == n.pack(g)).all()
        assert (ops.unpack(words, W).cpu().numpy() == g).all()
        assert int(ops.population(words, W).item()) == int(g.sum())
        nxt = ops.step_torus(words, W)
        assert (ops.unpack(nxt, W).cpu().numpy() == n.oracle_run(g, 1)).all()


# ------------------------------------------------------------------ generic kernel (N-K2)
@pytest.mark.parametrize("W,H", [(1, 1), (1, 3), (33, 5), (70, 2), (130, 1), (200, 7), (64, 64), (32, 32)])
@pytest.mark.parametrize("R", [1, 2, 3])
def test_generic_any_shape(W, H, R):
    got, plan = run_gpu(6, W, H, R, 6, mode="generic", seed=4, density=0.45)
    assert plan["mode_name"] == "generic"
    assert (got == expected(6, W, H, R, 6, seed=4, density=0.45)).all()


@pytest.mark.parametrize("threads", [1, 7, 32, 512, 1024])
def test_generic_honours_threads_per_block_literally(threads):
    got, plan = run_gpu(4, 96, 20, 2, 9, mode="generic", threads=threads)
    assert plan["threads_per_block"] == threads
    assert (got == expected(4, 96, 20, 2, 9)).all()


def test_auto_picks_generic_for_unaligned_width_and_fused_otherwise():
    _, plan = run_gpu(6, 100, 16, 1, 1)
    assert plan["mode_name"] == "generic"
    _, plan = run_gpu(6, 128, 16, 1, 1)
    assert plan["mode_name"] == "fused"
    with pytest.raises(GolError, match="multiple of 64"):
        World(100, 16, backend="cuda", mode="fused")


# ------------------------------------------------------------------ fused kernel (N-K1)
@pytest.mark.parametrize("K", [1, 2, 3, 4, 6, 8])
@pytest.mark.parametrize("R", [1, 2, 3])
def test_fused_k_generations_per_pass(K, R):
    W, H, gens = 2048, 40, 2 * K + 3          # 32 words: two column groups, second nearly empty
    got, plan = run_gpu(6, W, H, R, gens, mode="fused", k=K, seed=9)
    assert plan["temporal_k"] == K and plan["mode_name"] == "fused"
    assert (got == expected(6, W, H, R, gens, seed=9)).all()


@pytest.mark.parametrize("W", [64, 128, 1920, 1984, 4096])
@pytest.mark.parametrize("H", [8, 37, 100])
def test_fused_shapes(W, H):
    got, _ = run_gpu(6, W, H, 2, 13, mode="fused", k=4, seed=W + H)
    assert (got == expected(6, W, H, 2, 13, seed=W + H)).all()


@pytest.mark.parametrize("k,depth", [(1, 4), (2, 8), (4, 8), (4, 16), (8, 32), (3, 6)])
def test_fused_deep_ghost_zones(k, depth):
    # halo exchanged every `depth` generations, k generations per pass (temporal blocking)
    W, H, R, gens = 256, 48, 3, 2 * depth + 5
    got, plan = run_gpu(6, W, H, R, gens, mode="fused", k=k, depth=depth, seed=21)
    assert plan["exchange_depth"] == depth
    assert (got == expected(6, W, H, R, gens, seed=21)).all()


@pytest.mark.parametrize("pattern", [1, 2, 3, 4, 5])
@pytest.mark.parametrize("R", [1, 2, 4])
def test_fused_reference_patterns_cross_every_boundary(pattern, R):
    W, H, gens = 256, 256 // R if pattern != 2 else 256, 24
    got, _ = run_gpu(pattern, W, H, R, gens, mode="fused", k=4)
    assert (got == expected(pattern, W, H, R, gens)).all()


def test_fused_without_graph_and_in_chunks_is_identical():
    a, _ = run_gpu(6, 512, 64, 2, 40, mode="fused", k=4, seed=5)
    b, _ = run_gpu(6, 512, 64, 2, 40, mode="fused", k=4, seed=5, flags=n.FLAG_NO_GRAPH)
    c, _ = run_gpu(6, 512, 64, 2, 40, mode="fused", k=4, seed=5, chunks=[1, 3, 16, 20])
    want = expected(6, 512, 64, 2, 40, seed=5)
    assert (a == want).all() and (b == want).all() and (c == want).all()


@pytest.mark.parametrize("threads", [32, 64, 128, 256, 1024])
def test_fused_block_size(threads):
    got, plan = run_gpu(6, 1024, 64, 2, 9, mode="fused", k=4, threads=threads, seed=8)
    assert plan["threads_per_block"] == min(threads, 256)   # clamped to the kernel's register budget
    assert (got == expected(6, 1024, 64, 2, 9, seed=8)).all()


def test_larger_world_against_packed_cpu_stepper():
    got, plan = run_gpu(6, 8192, 1024, 2, 32, seed=77)
    assert plan["mode_name"] == "resident"        # 1 MiB slab, 1024 rows: AUTO keeps it in shared memory
    assert (got == expected(6, 8192, 1024, 2, 32, seed=77)).all()
    got, plan = run_gpu(6, 8192, 1024, 2, 32, mode="fused", seed=77)
    assert (got == expected(6, 8192, 1024, 2, 32, seed=77)).all()


# ------------------------------------------------------------------ TMA-fed fused kernel
@pytest.mark.parametrize("K", [1, 2, 3, 4])
@pytest.mark.parametrize("W,H,R", [(10240, 100, 2), (9216, 64, 1), (16384, 300, 3), (11264, 41, 2)])
def test_tma_fed_kernel(K, W, H, R):
    # rows arrive through cp.async.bulk + mbarrier ring; 144-word CTA segments wrap the torus seam
    gens = 3 * K + 5
    got, plan = run_gpu(6, W, H, R, gens, mode="tma", k=K, seed=W + K)
    assert plan["mode_name"] == "tma" and plan["threads_per_block"] in (96, 128)
    assert (got == expected(6, W, H, R, gens, seed=W + K)).all()


def test_tma_mode_deep_halo_and_rejections():
    got, _ = run_gpu(6, 10240, 96, 2, 37, mode="tma", k=4, depth=16, seed=3)
    assert (got == expected(6, 10240, 96, 2, 37, seed=3)).all()
    with pytest.raises(GolError, match="mode tma"):
        World(1024, 64, backend="cuda", mode="tma")          # too narrow for a 124-word segment
    with pytest.raises(GolError, match="mode tma"):
        World(16384, 64, backend="cuda", mode="tma", k=8)


# ------------------------------------------------------------------ persistent kernel (N-K3)
@pytest.mark.parametrize("k,depth,R", [(4, 4, 1), (4, 4, 2), (2, 8, 3), (8, 8, 2), (1, 1, 2)])
def test_persistent_kernel(k, depth, R):
    W, H, gens = 1024, 96, 5 * depth + 3
    got, plan = run_gpu(6, W, H, R, gens, mode="persistent", k=k, depth=depth, seed=31)
    assert plan["mode_name"] == "persistent"
    assert (got == expected(6, W, H, R, gens, seed=31)).all()


# ------------------------------------------------------------------ world plumbing
def test_load_step_read_population_and_views():
    import torch

    rng = np.random.default_rng(5)
    W, H, R = 256, 32, 2
    g = (rng.random((H * R, W)) < 0.3).astype(np.uint8)
    w = World(W, H, nranks=R, backend="cuda", ndevices=1)
    w.connect()
    for i in range(R):
        w.load_bytes(i, g[i * H:(i + 1) * H])
    assert sum(w.population(i) for i in range(R)) == int(g.sum())
    before = launch_count()
    w.step(10)
    w.sync()
    assert launch_count() > before
    want = n.oracle_run(g, 10)
    got = np.concatenate([w.read_bytes(i) for i in range(R)])
    assert (got == want).all()
    view = w.device_tensor(0)
    assert view.is_cuda and tuple(view.shape) == (H, w.stride)
    assert (view.cpu().numpy().view(np.uint64) == w.read_packed(0)).all()
    # packed upload path
    w.load_packed(1, n.pack(g[:H]))
    assert (w.read_bytes(1) == g[:H]).all()
    w.close()


def test_device_formatted_dump_is_byte_exact(tmp_path):
    from test_writer import expected_text

    for (W, H, R) in [(64, 12, 2), (70, 9, 3), (200, 101, 1)]:
        w = World(W, H, nranks=R, backend="cuda", ndevices=1)
        w.connect()
        w.init_pattern(6, 3, 0.5)
        w.step(4)
        w.sync()
        for i in range(R):
            w.dump_text(i, str(tmp_path))
            got = (tmp_path / f"Rank_{i}_of_{R}.txt").read_text()
            assert got == expected_text(w.read_bytes(i), i)
        w.close()


def test_life_facade_on_gpu():
    life = Life(256, 64, ranks=2).init(5)
    ms = life.step_timed(64)
    assert ms > 0 and life.population() == 5
    life.close()


def test_native_cli_on_gpu(tmp_path):
    import subprocess

    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "bin", "gol")
    r = subprocess.run([exe, "4", "64", "16", "256", "1", "--check", "--outdir", str(tmp_path)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "CHECK: OK" in r.stdout
    r = subprocess.run([exe, "4", "64", "16", "256", "1", "--check", "--ranks", "2", "--devices", "1",
                        "--outdir", str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 0 and "CHECK: OK" in r.stdout, r.stdout + r.stderr


def test_checkpoint_resume_on_gpu(tmp_path):
    a = Life(512, 64, ranks=2).init(6, seed=2)
    a.step(9)
    a.save_checkpoint(str(tmp_path))
    a.step(11).sync()
    b = Life(512, 64, ranks=2).load_checkpoint(str(tmp_path))
    b.step(11).sync()
    assert (a.all_local_cells() == b.all_local_cells()).all()
    a.close()
    b.close()


def test_watchdog_reports_a_dead_neighbour(monkeypatch):
    """Failure detection: a rank whose neighbour never sends must not hang the GPU.  World `a`
    owns rank 0 of 2; its neighbour is wired to an idle world `b` in the same process, so the
    halo words `a` waits for never arrive.  The flag wait gives up after GOL_WATCHDOG_MS and
    sync() reports which side timed out."""
    monkeypatch.setenv("GOL_WATCHDOG_MS", "300")
    a = World(256, 32, nranks=2, first_rank=0, nlocal=1, backend="cuda", ndevices=1)
    b = World(256, 32, nranks=2, first_rank=1, nlocal=1, backend="cuda", ndevices=1)
    a.import_blob(1, b.export_blob(0))
    a.connect()
    a.init_pattern(6)
    a.step(4)
    with pytest.raises(GolError, match="did not arrive within the watchdog"):
        a.sync()
    a.close()
    b.close()


def test_long_run_is_cut_into_graph_chunks():
    # step() replays a cached graph per chunk of 4096 exchange periods; 4096 + 37 crosses one cut
    got, _ = run_gpu(6, 64, 16, 2, 4096 + 37, mode="fused", k=1, seed=12)
    assert (got == expected(6, 64, 16, 2, 4096 + 37, seed=12)).all()


def test_slab_larger_than_2_pow_31_cells():
    """The reference indexes cells with 32-bit ints and silently does nothing beyond 2^31 cells per
    rank (gol-with-cuda.cu:198).  Here sizes are 64-bit: a 131072 x 32768 slab (2^32 cells, 512 MiB
    per buffer) keeps the wrap-around blinker (pattern 4) and the glider (pattern 5) alive."""
    W, H = 131072, 32768
    w = World(W, H, nranks=1, backend="cuda")
    w.connect()
    w.init_pattern(4)
    w.step(6)
    w.sync()
    assert w.population(0) == 3
    row0 = w.read_packed(0)[0]
    first, last = int(row0[0]), int(row0[W // 64 - 1])
    assert first == (1 | (1 << 32)) and last == (1 << 63)      # columns 0, 1 and W-1 (interleaved layout)
    w.init_pattern(5)
    w.step(16)
    w.sync()
    assert w.population(0) == 5
    w.init_pattern(6, 3, 0.5)
    pop = w.population(0)
    assert abs(pop / float(W * H) - 0.5) < 1e-3
    w.close()


# ------------------------------------------------------------------ flag-in-data halo rows
@pytest.mark.parametrize("mode,k,depth,R", [("fused", 4, 4, 1), ("fused", 4, 4, 2), ("fused", 4, 4, 3),
                                            ("fused", 2, 8, 2), ("fused", 8, 32, 3), ("fused", 1, 1, 2),
                                            ("persistent", 4, 4, 2), ("persistent", 2, 4, 3), ("tma", 4, 4, 2)])
def test_flag_in_data_halo_protocol(monkeypatch, mode, k, depth, R):
    """Halo rows between devices travel as {32 data bits | exchange number} words with no flag and no
    fence.  GOL_FORCE_LL=1 routes same-device neighbours through that path too, so one GPU covers
    it; the rings of two and three ranks exercise both parities of the receive slots."""
    monkeypatch.setenv("GOL_FORCE_LL", "1")
    W, H = (10240, 96) if mode == "tma" else (2048, 80)
    gens = 5 * depth + 3
    got, plan = run_gpu(6, W, H, R, gens, mode=mode, k=k, depth=depth, seed=41, chunks=[depth + 1, gens - depth - 1])
    assert plan["mode_name"] == mode
    assert (got == expected(6, W, H, R, gens, seed=41)).all()


# ------------------------------------------------------------------ resident kernel (N-K3b)
def test_auto_picks_the_resident_kernel_for_small_tall_slabs():
    got, plan = run_gpu(6, 4096, 1024, 2, 19, seed=3, chunks=[7, 12])
    assert plan["mode_name"] == "resident" and plan["temporal_k"] == 4
    assert (got == expected(6, 4096, 1024, 2, 19, seed=3)).all()
    _, plan = run_gpu(6, 4096, 512, 1, 1)          # too short to give every SM a strip: stays fused
    assert plan["mode_name"] == "fused"


@pytest.mark.parametrize("pattern", [1, 2, 3, 4, 5])
def test_resident_reference_patterns(pattern):
    got, plan = run_gpu(pattern, 256, 256, 2, 24, mode="resident", k=4)
    assert plan["mode_name"] == "resident"
    assert (got == expected(pattern, 256, 256, 2, 24)).all()
