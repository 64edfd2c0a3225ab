"""GPU kernels against an INDEPENDENT oracle at real widths.

Most GPU tests compare with ``cpu_run_packed``, which shares ``life_math.h`` (x-neighbour and rule
helpers) with the generic and stand-alone kernels.  Here every step mode runs at the widths the
benchmarks use (8192 / 16384 / 65536 columns: 5, 9 and 35 warp column groups, torus seam inside a
group) and is compared with ``numpy_life_run`` - numpy ``roll`` on a byte grid, no packing, no
shared code (task 9)."""
import numpy as np
import pytest

from gol_b200 import World
from gol_b200.ops.oracle import numpy_life_run

pytestmark = pytest.mark.gpu


def run(mode, W, H, R, gens, k=4, depth=0, seed=1, chunks=None):
    rng = np.random.default_rng(seed)
    g = (rng.random((H * R, W)) < 0.5).astype(np.uint8)
    w = World(W, H, nranks=R, backend="cuda", mode=mode, k=k, depth=depth, ndevices=1)
    w.connect()
    for i in range(R):
        w.load_bytes(i, g[i * H:(i + 1) * H])
    for c in (chunks or [gens]):
        w.step(c)
    w.sync()
    got = np.concatenate([w.read_bytes(i) for i in range(R)])
    w.close()
    want = numpy_life_run(g, gens)
    bad = int((got != want).sum())
    assert bad == 0, f"{mode} {W}x{H} R={R} gens={gens}: {bad} cells differ from the numpy stepper"


@pytest.mark.parametrize("W", [8192, 16384, 65536])
@pytest.mark.parametrize("R", [1, 2, 3])
@pytest.mark.parametrize("mode", ["fused", "persistent"])
def test_fused_and_persistent_at_benchmark_widths(mode, W, R):
    run(mode, W, 96 if W < 65536 else 64, R, gens=2 * 4 + 3, seed=W + R)


@pytest.mark.parametrize("W,R", [(16384, 1), (16384, 3), (65536, 2), (10240, 2)])
def test_tma_at_benchmark_widths(W, R):
    run("tma", W, 80, R, gens=13, seed=W * 3 + R)


@pytest.mark.parametrize("W,R", [(8192, 2), (16384, 1), (65536, 3), (8200, 2), (16383, 1)])
def test_generic_at_benchmark_widths(W, R):
    run("generic", W, 24, R, gens=5, seed=W + 7 * R)


@pytest.mark.parametrize("k,depth", [(1, 1), (2, 8), (3, 6), (8, 8), (4, 16)])
def test_temporal_depths_at_16384_columns(k, depth):
    run("fused", 16384, 128, 2, gens=2 * depth + 5, k=k, depth=depth, seed=k * 100 + depth)
    run("persistent", 16384, 128, 2, gens=2 * depth + 5, k=k, depth=depth, seed=k * 100 + depth + 1,
        chunks=[depth + 2, depth + 3])


def test_persistent_many_strips_and_passes():
    # tall enough for hundreds of interior strips per pass and tens of passes in flight
    run("persistent", 8192, 2048, 1, gens=83, seed=5)
    run("persistent", 2048, 1500, 3, gens=61, k=2, depth=2, seed=6)


# ------------------------------------------------------------------ resident kernel (clusters / DSMEM)
@pytest.mark.parametrize("W,H,R", [(4096, 512, 1), (4096, 300, 2), (2048, 257, 3), (8192, 96, 1), (64, 40, 2),
                                   (1984, 130, 2)])
@pytest.mark.parametrize("K", [1, 4, 5])
def test_resident_kernel_vs_numpy(W, H, R, K):
    """Strips stay in shared memory for the whole step; halo rows travel as flag-in-data words over
    DSMEM inside a cluster, L2 mailboxes between clusters and ll_recv buffers between ranks."""
    gens = 3 * K + 2                     # the last pass is a partial one
    run("resident", W, H, R, gens=gens, k=K, seed=W + H + K, chunks=[K + 1, gens - K - 1])


@pytest.mark.parametrize("cluster", [1, 2, 4, 8])
@pytest.mark.parametrize("bulk", [True, False])
def test_resident_cluster_sizes_and_copy_paths(monkeypatch, cluster, bulk):
    monkeypatch.setenv("GOL_RESIDENT_CLUSTER", str(cluster))
    if not bulk:
        monkeypatch.setenv("GOL_RESIDENT_NO_BULK", "1")
    run("resident", 4096, 1024, 1, gens=37, k=4, seed=cluster)
    run("resident", 2048, 640, 2, gens=21, k=2, seed=cluster + 10)


def test_resident_rejects_slabs_that_do_not_fit():
    from gol_b200 import GolError

    with pytest.raises(GolError, match="resident"):
        World(65536, 16384, backend="cuda", mode="resident")
