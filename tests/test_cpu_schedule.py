"""The GPU pass schedule executed on the CPU (GOL_FLAG_CPU_SCHEDULE): temporal depth k, halos
deeper than one pass (margins + trapezoid recompute), double-buffered recv rows, prime/push/consume
pairing and ring routing for R = 1, 2, 3... are checked against the oracle without a GPU."""
import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from gol_b200 import World
from gol_b200 import _native as n


def run(W, H, R, gens, k, depth, seed, chunks=None, mode="fused"):
    rng = np.random.default_rng(seed)
    g = (rng.random((H * R, W)) < 0.45).astype(np.uint8)
    w = World(W, H, nranks=R, backend="cpu", mode=mode, k=k, depth=depth, flags=n.FLAG_CPU_SCHEDULE)
    for i in range(R):
        w.load_bytes(i, g[i * H:(i + 1) * H])
    for c in (chunks or [gens]):
        w.step(c)
    got = np.concatenate([w.read_bytes(i) for i in range(R)])
    plan = w.plan
    w.close()
    return got, n.oracle_run(g, gens), plan


@pytest.mark.parametrize("k,depth", [(1, 1), (2, 2), (4, 4), (8, 8), (1, 4), (2, 8), (4, 16), (8, 32), (3, 6), (6, 12)])
@pytest.mark.parametrize("R", [1, 2, 3])
def test_schedule_semantics_on_cpu(k, depth, R):
    got, want, plan = run(128, 40, R, 2 * depth + 3, k, depth, seed=k * 100 + depth + R)
    assert plan["temporal_k"] == k and plan["exchange_depth"] == depth and plan["mode_name"] == "fused"
    assert (got == want).all()


def test_chunked_steps_and_generic_schedule():
    got, want, _ = run(192, 33, 2, 29, 4, 8, seed=5, chunks=[1, 7, 8, 13])
    assert (got == want).all()
    got, want, plan = run(70, 9, 3, 6, 0, 0, seed=6, mode="auto")      # width not a multiple of 64
    assert plan["mode_name"] == "generic" and (got == want).all()


@settings(max_examples=40, deadline=None)
@given(k=st.sampled_from([1, 2, 3, 4, 6, 8]), mult=st.sampled_from([1, 2, 4]), R=st.integers(1, 4),
       extra=st.integers(0, 20), gens=st.integers(0, 50), seed=st.integers(0, 10**6))
def test_schedule_property(k, mult, R, extra, gens, seed):
    depth = k * mult if k * mult <= 32 else k
    got, want, _ = run(64, depth + extra, R, gens, k, depth, seed)
    assert (got == want).all()
