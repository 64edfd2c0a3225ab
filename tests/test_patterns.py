"""Pattern cell coordinates vs the reference initialisers (gol-with-cuda.cu:56-171)."""
import numpy as np
import pytest

from gol_b200 import GolError
from gol_b200.models.patterns import PATTERNS, global_pattern, pattern_bytes


def coords(a):
    return sorted(zip(*np.nonzero(a)))


def test_pattern_0_and_1():
    assert pattern_bytes(0, 40, 30, 1, 3).sum() == 0
    assert pattern_bytes(1, 40, 30, 2, 3).sum() == 40 * 30


@pytest.mark.parametrize("rank", [0, 1, 2])
def test_pattern_2_is_ten_cells_on_the_last_local_row_of_every_rank(rank):
    a = pattern_bytes(2, 256, 256, rank, 3)
    assert coords(a) == [(255, c) for c in range(127, 137)]


def test_pattern_2_rejects_narrow_worlds():
    with pytest.raises(GolError, match="137"):
        pattern_bytes(2, 64, 64)


def test_pattern_3_corners_including_the_else_if():
    N = 32
    assert coords(pattern_bytes(3, N, N, 0, 3)) == [(0, 0), (0, N - 1)]
    assert coords(pattern_bytes(3, N, N, 1, 3)) == []
    assert coords(pattern_bytes(3, N, N, 2, 3)) == [(N - 1, 0), (N - 1, N - 1)]
    # with one rank only the rank-0 branch runs (gol-with-cuda.cu:134-142)
    assert coords(pattern_bytes(3, N, N, 0, 1)) == [(0, 0), (0, N - 1)]


def test_pattern_4_spinner_on_rank_0_only():
    N = 32
    assert coords(pattern_bytes(4, N, N, 0, 2)) == [(0, 0), (0, 1), (0, N - 1)]
    assert pattern_bytes(4, N, N, 1, 2).sum() == 0


def test_pattern_5_glider_straddles_the_first_boundary():
    g = global_pattern(5, 64, 32, 2)
    assert g.sum() == 5
    assert g[:32].sum() > 0 and g[32:].sum() > 0


def test_pattern_6_density_and_determinism():
    a = pattern_bytes(6, 256, 256, 0, 2, seed=3, density=0.5)
    b = pattern_bytes(6, 256, 256, 0, 2, seed=3, density=0.5)
    c = pattern_bytes(6, 256, 256, 1, 2, seed=3, density=0.5)
    assert (a == b).all() and not (a == c).all()
    assert 0.47 < a.mean() < 0.53
    assert 0.08 < pattern_bytes(6, 256, 256, density=0.1).mean() < 0.12


def test_unknown_pattern_message():
    with pytest.raises(GolError, match="has not been implemented"):
        pattern_bytes(9, 32, 32)
    assert set(PATTERNS) == set(range(7))
