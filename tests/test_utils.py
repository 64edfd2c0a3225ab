"""Utility layer: metrics arithmetic, clock sampler without a GPU, package surface."""
import gol_b200
from gol_b200.utils import ClockSampler, gcups, roofline_fraction
from gol_b200.utils.metrics import measured_peaks


def test_gcups_matches_the_reference_definition():
    # numRank * H * W * iterations / duration  (gol-main.c:124-125)
    assert abs(gcups(8 * 16384 * 16384 * 1024, 0.01) - 8 * 16384 * 16384 * 1024 / 0.01 * 1e-9) < 1e-6
    assert gcups(1, 0) == 0.0


def test_roofline_fraction_uses_measured_peaks():
    peaks = measured_peaks()
    r = roofline_fraction(26.2e12, generations_per_pass=1, hbm_gbs=6553.3)
    assert 0.99 < r["fraction"] < 1.01          # 0.25 B per cell per pass at 6553 GB/s = 26.2 Tcell/s
    assert roofline_fraction(26.2e12, generations_per_pass=4, hbm_gbs=6553.3)["fraction"] < 0.26
    assert peaks["source"] in ("measured", "fallback")


def test_clock_sampler_degrades_gracefully_without_nvidia_smi(monkeypatch):
    monkeypatch.setenv("PATH", "/nonexistent")
    s = ClockSampler(period_ms=50).start()
    out = s.stop()
    assert out["samples"] == 0 and out["reasons"] == []


def test_package_surface():
    assert gol_b200.version().startswith("gol_b200")
    for name in ("Life", "World", "GolError", "PATTERNS", "pattern_bytes", "launch_count"):
        assert hasattr(gol_b200, name)
    assert gol_b200._native.MODES == {"auto": 0, "generic": 1, "fused": 2, "persistent": 3, "tma": 4, "resident": 5}
    assert gol_b200.launch_count() >= 0


def test_life_metrics_on_the_cpu_backend():
    from gol_b200 import Life

    life = Life(64, 32, ranks=2, backend="cpu").init(5)
    ms = life.step_timed(8)
    m = life.metrics(ms, 8)
    assert m["gcups"] > 0 and m["generations"] == 8 and 0 < m["hbm_roofline_fraction"] < 1
    life.close()
