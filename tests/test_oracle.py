"""Oracle and bit-sliced CPU stepper (oracle + unit levels)."""
import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from gol_b200 import _native as n
from gol_b200.ops.oracle import oracle_run, reference_quirk_run


def numpy_life(a, gens):
    a = a.astype(np.int32)
    for _ in range(gens):
        s = sum(np.roll(np.roll(a, dy, 0), dx, 1) for dy in (-1, 0, 1) for dx in (-1, 0, 1)) - a
        a = ((s == 3) | ((a == 1) & (s == 2))).astype(np.int32)
    return a.astype(np.uint8)


def test_rule_truth_table_all_512_neighbourhoods():
    # every 3x3 neighbourhood, embedded in a 5x5 torus of zeros; check the centre cell
    for code in range(512):
        g = np.zeros((5, 5), dtype=np.uint8)
        bits = [(code >> i) & 1 for i in range(9)]
        g[1:4, 1:4] = np.array(bits, dtype=np.uint8).reshape(3, 3)
        cnt = sum(bits) - bits[4]
        want = 1 if (cnt == 3 or (bits[4] and cnt == 2)) else 0
        assert oracle_run(g, 1)[2, 2] == want
        packed = n.cpu_run_packed(n.pack(g), 5, 1)
        assert n.unpack(packed, 5)[2, 2] == want


def test_still_life_and_oscillator():
    g = np.zeros((8, 8), dtype=np.uint8)
    g[2:4, 2:4] = 1                       # block
    assert (oracle_run(g, 5) == g).all()
    b = np.zeros((8, 8), dtype=np.uint8)
    b[3, 2:5] = 1                         # blinker
    assert (oracle_run(b, 2) == b).all()
    assert oracle_run(b, 1)[2:5, 3].sum() == 3


def test_glider_translates_by_one_every_four_generations():
    g = np.zeros((16, 16), dtype=np.uint8)
    g[1, 2] = g[2, 3] = g[3, 1] = g[3, 2] = g[3, 3] = 1
    assert (oracle_run(g, 4) == np.roll(np.roll(g, 1, 0), 1, 1)).all()


@settings(max_examples=60, deadline=None)
@given(w=st.integers(1, 200), h=st.integers(1, 40), gens=st.integers(0, 6), seed=st.integers(0, 2**31))
def test_oracle_and_packed_stepper_agree_with_numpy(w, h, gens, seed):
    rng = np.random.default_rng(seed)
    g = (rng.random((h, w)) < 0.45).astype(np.uint8)
    want = numpy_life(g, gens)
    assert (oracle_run(g, gens) == want).all()
    got = n.unpack(n.cpu_run_packed(n.pack(g), w, gens), w)
    assert (got == want).all()


@pytest.mark.parametrize("w", [1, 2, 31, 32, 63, 64, 65, 127, 128, 129, 200])
def test_x_wrap_across_word_boundaries(w):
    # a vertical line at column 0 and one at column w-1 interact only through the x wrap
    g = np.zeros((9, w), dtype=np.uint8)
    g[:, 0] = 1
    g[4, w - 1] = 1
    for gens in (1, 2, 3):
        got = n.unpack(n.cpu_run_packed(n.pack(g), w, gens), w)
        assert (got == numpy_life(g, gens)).all()


def test_pack_unpack_round_trip_and_padding_is_zero():
    rng = np.random.default_rng(7)
    for w in (1, 5, 64, 70, 128, 190):
        g = (rng.random((6, w)) < 0.5).astype(np.uint8)
        p = n.pack(g)
        assert p.shape == (6, n.row_stride_words(w))
        assert (n.unpack(p, w) == g).all()
        # bits beyond column w-1 must be zero: the packed population equals the cell count
        assert int(np.unpackbits(p.view(np.uint8)).sum()) == int(g.sum())


def test_word_layout_is_even_odd_interleaved():
    # column 64w + j lives at bit (j >> 1) + 32 * (j & 1): even columns low half, odd columns high half
    for j in (0, 1, 2, 3, 62, 63):
        g = np.zeros((1, 128), dtype=np.uint8)
        g[0, 64 + j] = 1
        p = n.pack(g)
        assert int(p[0, 0]) == 0 and int(p[0, 1]) == 1 << ((j >> 1) + 32 * (j & 1))


def test_reference_quirk_model_matches_survey_table():
    """The as-built reference diverges from true Life exactly where SURVEY.md section 2f says."""
    from gol_b200.models.patterns import pattern_bytes

    def first_divergence(pattern, R, N, gens=16):
        slabs = [pattern_bytes(pattern, N, N, r, R) for r in range(R)]
        world = np.concatenate(slabs)
        for g in range(1, gens + 1):
            ref = np.concatenate(reference_quirk_run(slabs, g))
            if not (ref == oracle_run(world, g)).all():
                return g
        return None

    assert first_divergence(0, 4, 64) is None
    assert first_divergence(3, 3, 64) is None      # 2x2 block through the torus corner: still life
    assert first_divergence(3, 1, 64) == 1
    assert first_divergence(4, 2, 64) == 1         # swapped ghosts at R <= 2
    assert first_divergence(4, 3, 64) == 2         # stale halos from generation 2 on
    assert first_divergence(1, 4, 64) == 2


@pytest.mark.parametrize("w,h,gens", [(1000, 20, 5), (2048, 24, 7), (4096, 17, 6), (4100, 9, 4), (1920, 33, 9)])
def test_packed_stepper_at_multi_kiloword_widths_vs_numpy(w, h, gens):
    """The packed CPU stepper is what most GPU tests compare with, and it shares life_math.h with the
    kernels: pin it to the independent numpy stepper at the widths where column groups, tail
    masks and the x seam matter (it was only pinned for widths <= 200)."""
    from gol_b200.ops.oracle import numpy_life_run

    rng = np.random.default_rng(w + h)
    g = (rng.random((h, w)) < 0.5).astype(np.uint8)
    want = numpy_life_run(g, gens)
    assert (numpy_life(g, gens) == want).all()
    assert (n.unpack(n.cpu_run_packed(n.pack(g), w, gens), w) == want).all()
    if w <= 2048:
        assert (oracle_run(g, gens) == want).all()
