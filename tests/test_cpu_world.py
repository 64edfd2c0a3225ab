"""Slab / ring plumbing on the CPU backend: R = 1, 2 (the cases the reference routes
wrongly, gol-main.c:97-107) and R >= 3, rectangular slabs, odd widths."""
import numpy as np
import pytest

from gol_b200 import Life, World
from gol_b200 import _native as n
from gol_b200.models.patterns import global_pattern
from gol_b200.parallel import Ring, slab_rows


def run_world(pattern, W, H, R, gens, **kw):
    w = World(W, H, nranks=R, backend="cpu")
    w.init_pattern(pattern, **kw)
    w.step(gens)
    return np.concatenate([w.read_bytes(i) for i in range(R)])


@pytest.mark.parametrize("R", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("pattern", [1, 3, 4, 5, 6])
def test_cpu_world_matches_oracle(pattern, R):
    W, H, gens = 64, 16, 12
    got = run_world(pattern, W, H, R, gens, seed=11, density=0.4)
    want = n.oracle_run(global_pattern(pattern, W, H, R, seed=11, density=0.4), gens)
    assert (got == want).all()


def test_baseline_config_1_glider_64x64_two_ranks_16_generations():
    # BASELINE.json config 1: 64x64 glider, 2 ranks, 16 generations, CPU stepper
    got = run_world(5, 64, 32, 2, 16)
    want = n.oracle_run(global_pattern(5, 64, 32, 2), 16)
    assert (got == want).all() and got.sum() == 5


@pytest.mark.parametrize("W,H", [(1, 3), (33, 5), (70, 2), (130, 1), (200, 7)])
def test_odd_shapes(W, H):
    got = run_world(6, W, H, 3, 5, seed=2, density=0.5)
    want = n.oracle_run(global_pattern(6, W, H, 3, seed=2, density=0.5), 5)
    assert (got == want).all()


def test_ring_topology():
    r = Ring(4)
    assert r.neighbours(0) == (3, 1) and r.neighbours(3) == (2, 0)
    assert Ring(1).neighbours(0) == (0, 0)
    assert Ring(2).neighbours(1) == (0, 0)
    assert slab_rows(2, 16) == (32, 48)


def test_life_facade_cpu_population_and_report():
    life = Life(64, 32, ranks=2, backend="cpu").init(5)
    life.step(8).sync()
    assert life.population() == 5
    assert life.cell_updates(8) == 2 * 32 * 64 * 8
    assert "cell updates = 32768" in life.report(0.1, 8)


def test_partial_cpu_world_refuses_plain_step():
    from gol_b200 import GolError

    w = World(64, 8, nranks=2, first_rank=0, nlocal=1, backend="cpu")
    with pytest.raises(GolError, match="cpu_step_local"):
        w.step(1)


def test_checkpoint_resume_round_trip(tmp_path):
    a = Life(96, 20, ranks=3, backend="cpu").init(6, seed=8, density=0.4)
    a.step(5)
    a.save_checkpoint(str(tmp_path))
    a.step(7)
    b = Life(96, 20, ranks=3, backend="cpu").load_checkpoint(str(tmp_path))
    assert b.generation == 5
    b.step(7)
    assert (a.all_local_cells() == b.all_local_cells()).all()
    from gol_b200 import GolError

    with pytest.raises(GolError, match="different geometry"):
        Life(96, 21, ranks=3, backend="cpu").load_checkpoint(str(tmp_path))
