"""Property test on the GPU: random shapes, rank counts, modes, temporal depths and chunkings
must all reproduce the CPU stepper bit for bit ("kernel" level)."""
import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings, strategies as st

from gol_b200 import World
from gol_b200 import _native as n

pytestmark = pytest.mark.gpu

KS = [1, 2, 3, 4, 6, 8]


@st.composite
def cases(draw):
    mode = draw(st.sampled_from(["fused", "fused", "persistent", "generic", "auto"]))
    R = draw(st.integers(1, 4))
    if mode == "generic" or (mode == "auto" and draw(st.booleans())):
        W = draw(st.integers(1, 300))
        H = draw(st.integers(1, 40))
        k = depth = 0
    else:
        W = 64 * draw(st.integers(1, 70))
        k = draw(st.sampled_from(KS))
        depth = k * draw(st.sampled_from([1, 1, 2, 4]))
        if depth > 32:
            depth = k
        H = draw(st.integers(depth, depth + 90))
    gens = draw(st.integers(0, 40))
    cuts = sorted(draw(st.lists(st.integers(0, gens), max_size=2)))
    chunks = [b - a for a, b in zip([0] + cuts, cuts + [gens])]
    threads = draw(st.sampled_from([0, 32, 96, 128, 256]))
    seed = draw(st.integers(0, 2**20))
    density = draw(st.sampled_from([0.1, 0.35, 0.5, 0.9]))
    return mode, R, W, H, k, depth, gens, chunks, threads, seed, density


@settings(max_examples=60, deadline=None, suppress_health_check=list(HealthCheck))
@given(cases())
def test_any_configuration_matches_the_cpu_stepper(case):
    mode, R, W, H, k, depth, gens, chunks, threads, seed, density = case
    rng = np.random.default_rng(seed)
    g = (rng.random((H * R, W)) < density).astype(np.uint8)
    w = World(W, H, nranks=R, backend="cuda", mode=mode, k=k, depth=depth, threads=threads, ndevices=1)
    w.connect()
    for i in range(R):
        w.load_bytes(i, g[i * H:(i + 1) * H])
    for c in chunks:
        w.step(c)
    w.sync()
    got = np.concatenate([w.read_bytes(i) for i in range(R)])
    w.close()
    want = n.unpack(n.cpu_run_packed(n.pack(g), W, gens), W)
    assert (got == want).all(), f"{case}: {(got != want).sum()} cells differ"
